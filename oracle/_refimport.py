"""Import the real reference (only possible in the build container, where
/root/reference exists) with the shims SURVEY.md section 8c lists.  Used by
oracle/refcheck.py and tests/golden/make_golden.py; never at run time on the
GPU box.  TEST INFRASTRUCTURE."""
import os
import sys
import types

REF = os.environ.get('NES_REFERENCE_DIR', '/root/reference')


def available():
    return os.path.isdir(os.path.join(REF, 'src', 'algorithm'))


def load():
    import numpy as np
    if not hasattr(np, 'float'):
        np.float = float            # optimizers.py:41,75-76 / nic_nes_master.py:191 use the removed alias
    for p in (os.path.join(REF, 'src'), os.path.join(REF, 'cider')):
        if p not in sys.path:
            sys.path.insert(0, p)
    # absent third-party modules that the math never touches
    for name in ('redis', 'h5py', 'matplotlib', 'matplotlib.pyplot', 'memory_profiler', 'skimage', 'skimage.io'):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                m = types.ModuleType(name)
                if name == 'memory_profiler':
                    m.profile = lambda *a, **k: (lambda f: f)
                sys.modules[name] = m
    if 'matplotlib' in sys.modules and not hasattr(sys.modules['matplotlib'], 'use'):
        sys.modules['matplotlib'].use = lambda *a, **k: None
        sys.modules['matplotlib'].pyplot = sys.modules['matplotlib.pyplot']
    from algorithm.policies import ModelOptions
    from captioning.nets import FCModel
    from algorithm.nic_nes import optimizers
    from algorithm.nic_nes.nic_nes_master import NESMaster
    return types.SimpleNamespace(ModelOptions=ModelOptions, FCModel=FCModel,
                                 optimizers=optimizers, NESMaster=NESMaster)
