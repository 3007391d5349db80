"""Generate tests/golden/golden_schedule_v1.json: the REAL reference's iteration bookkeeping (imported from /root/reference:
algorithm/tools/iteration.py, podium.py, snapshot.py, statistics.py, nic_es/iteration.py, nic_nes/iteration.py) driven through
tests/schedule_scenario.py, every to_dict() / snapshot / directory listing on the way recorded with paths relative to the root.

Run in the build container only:  python tests/golden/make_golden_schedule.py
"""
import json
import os
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import _refimport  # noqa: E402
import schedule_scenario as sc  # noqa: E402


def reference_namespace():
    _refimport.load()
    from types import SimpleNamespace
    from algorithm.nets import PolicyNet
    from algorithm.nic_es.iteration import ESIteration
    from algorithm.nic_nes.iteration import NESIteration
    from algorithm.tools.snapshot import save_infos
    from algorithm.tools.statistics import Statistics
    from algorithm.tools.utils import Config, mkdir_p

    class RefModel(PolicyNet):                       # NESIteration.set_model asserts isinstance(model, PolicyNet)
        def __init__(self):
            pass
        serialize = sc.FakeModel.serialize

    class StubExperiment:                            # what save_infos reads of an Experiment (tools/snapshot.py:23-37)
        def __init__(self, directory, lth, algorithm, orig_bs):
            self._dir, self._lth, self._alg, self._bs = directory, lth, algorithm, orig_bs
        def snapshot_dir(self): return self._dir
        def orig_trainloader_lth(self): return self._lth
        def to_dict(self): return {'trainloader_lth': self._lth, 'algorithm': self._alg, 'orig_bs': self._bs}

    def save(directory, stats, it, lth, algorithm, orig_bs):
        mkdir_p(directory)
        return save_infos(StubExperiment(directory, lth, algorithm, orig_bs), it, stats)

    def load(path):
        with open(path) as f:
            return json.load(f)

    return SimpleNamespace(Config=Config, ESIteration=ESIteration, NESIteration=NESIteration, Statistics=Statistics, save=save, load=load), RefModel


def main():
    ns, RefModel = reference_namespace()
    out = {}
    with tempfile.TemporaryDirectory() as root:
        sc.FakeModel.count = 0
        out['nes'] = sc.run_nes(ns, root, make_model=RefModel)
    with tempfile.TemporaryDirectory() as root:
        sc.FakeModel.count = 0
        out['es'] = sc.run_es(ns, root)
    path = os.path.join(HERE, 'golden_schedule_v1.json')
    with open(path, 'w') as f:
        json.dump(out, f, indent=0, sort_keys=True)
    print(path, os.path.getsize(path), 'bytes;', len(out['nes']), 'NES records,', len(out['es']), 'ES records')


if __name__ == '__main__':
    main()
