"""CPU-only checks: the C-ABI library loads and exports every declared symbol, fails loudly without a GPU,
the host-side layout equals the oracle's, and the population sharding / fitness all-gather logic works
across 2 processes (gloo)."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope='module')
def lib_path():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g
    return g.build()


def test_library_exports_every_declared_symbol(lib_path):
    hdr = open(os.path.join(ROOT, 'include', 'nes_b200.h')).read()
    names = sorted(set(re.findall(r'\b(nes_[a-z0-9_]+)\s*\(', hdr)))
    assert len(names) >= 35
    lib = ctypes.CDLL(lib_path)
    assert [n for n in names if not hasattr(lib, n)] == []
    import nes_b200
    assert set(nes_b200._lib.SIGNATURES) == set(names)          # the ctypes binding covers the whole header
    lib.nes_version.restype = ctypes.c_int
    assert lib.nes_version() >= 100


def test_no_cpu_fallback(lib_path):
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    import nes_b200
    with pytest.raises(nes_b200.NesError):
        nes_b200.NesContext(counts=[1000], table_len=4096)
    with pytest.raises(nes_b200.NesError):
        nes_b200.gemm_selftest(torch.zeros(8, 64), torch.zeros(8, 64))


def test_product_code_never_imports_the_oracle():
    pkg = os.path.join(ROOT, 'nes-img-captioning_b200')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', src, re.M), f


def test_host_layout_equals_oracle_layout():
    import nes_b200
    from oracle import noise
    for V, E, R, Fd, ln, aff in [(9487, 512, 512, 2048, False, False), (9487, 128, 128, 2048, True, True), (199, 64, 64, 128, False, False),
                                 (999, 64, 128, 256, True, False)]:
        G = nes_b200.layout.granule(E, R, Fd)
        assert G == noise.granularity(E, R, Fd)
        specs, D, D_pad = nes_b200.layout.build_layout(nes_b200.layout.fc_tensors(V, E, R, Fd, ln, aff), G)
        play, D2, D_pad2 = noise.padded_layout(noise.flat_layout(V, E, R, Fd, ln, aff), G)
        assert (D, D_pad) == (D2, D_pad2)
        assert [(t.name, t.shape, t.start, t.count, t.pstart) for t in specs] == [tuple(x) for x in play]
    assert nes_b200.layout.build_layout(nes_b200.layout.fc_tensors(9487, 128, 128, 2048), 2048)[1] == 2865808   # tools/utils.py:180
    # vbn_e: the Sequential wrappers rename and reorder the embedding tensors (captioning/nets.py:156-165)
    for ln, aff, vbn, vaff in [(0, 0, False, True), (2, 1, True, True), (0, 0, False, False), (1, 1, False, True)]:
        specs, D, D_pad = nes_b200.layout.build_layout(nes_b200.layout.fc_tensors(300, 64, 64, 128, ln, aff, True, vaff), 128)
        play, D2, D_pad2 = noise.padded_layout(noise.flat_layout(300, 64, 64, 128, ln == 1, bool(aff) and ln == 1, vbn, vaff, True), 128)
        assert (D, D_pad) == (D2, D_pad2) and [(t.name, t.start, t.pstart) for t in specs] == [(x[0], x[2], x[4]) for x in play]


def test_shard_bounds_cover_the_population():
    import nes_b200
    for F in (0, 1, 7, 5000, 10001):
        for W in (1, 2, 3, 8):
            b = [nes_b200.shard_bounds(F, W, r) for r in range(W)]
            assert b[0][0] == 0 and b[-1][1] == F
            assert all(b[i][1] == b[i + 1][0] for i in range(W - 1))
            assert max(h - l for l, h in b) - min(h - l for l, h in b) <= 1


WORKER = r'''
import os, sys
sys.path.insert(0, sys.argv[1])
import numpy as np, torch, torch.distributed as dist
import nes_b200
rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
dist.init_process_group('gloo')
F = 11
full = torch.from_numpy(np.random.RandomState(0).randn(F, 2))
lo, hi = nes_b200.shard_bounds(F, world, rank)
got = nes_b200.all_gather_fitness(full[lo:hi].clone(), F, world, rank)
assert torch.equal(got, full), (rank, got, full)
# every rank then ranks the identical vector: identical centred ranks on every rank
from oracle import nes_math
w = torch.from_numpy(nes_math.rank_weights(got.numpy()))
ws = [torch.empty_like(w) for _ in range(world)]
dist.all_gather(ws, w)
assert all(torch.equal(ws[0], x) for x in ws)
# the second exchange: each rank holds its slice of the gradient and zeros elsewhere; the all-reduce gives every rank the whole, exactly
D = 1001
g_full = torch.from_numpy(np.random.RandomState(1).randn(D).astype(np.float32))
per = (D + world - 1) // world
mine = torch.zeros(D, dtype=torch.float32)
mine[rank * per:(rank + 1) * per] = g_full[rank * per:(rank + 1) * per]
got_g = nes_b200.all_reduce_gradient(mine, world)
assert torch.equal(got_g, g_full)
dist.destroy_process_group()
print('rank', rank, 'ok')
'''


def test_fitness_all_gather_world_size_2_gloo(tmp_path):
    script = tmp_path / 'worker.py'
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR='127.0.0.1', MASTER_PORT='29617', WORLD_SIZE='2', PYTHONPATH=ROOT)
    procs = [subprocess.Popen([sys.executable, str(script), ROOT], env=dict(env, RANK=str(r)), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs


def test_es_selection_is_the_references_stable_descending_sort():
    """ESMaster.selection (nic_es_master.py:155-167): scored_models.sort(key=fitness, reverse=True) keeps equal fitnesses in
    submission order; the top pop_size - num_elites become parents"""
    import nes_b200
    fit = [3.0, 7.0, 7.0, -1.0, 7.0, 5.0, 3.0]
    want = sorted(range(len(fit)), key=lambda i: fit[i], reverse=True)        # Python's own stable reverse sort
    keep, scores = nes_b200.ESMaster.selection(torch.tensor(fit, dtype=torch.float64), 6, 2)
    assert keep.tolist() == want[:4] == [1, 2, 4, 5]
    assert scores.tolist() == sorted(fit, reverse=True)
    keep, _ = nes_b200.ESMaster.selection(np.array(fit), 3, 3)
    assert keep.numel() == 0


def test_mnist_net_init_rule_on_a_stub_context():
    """MnistNet.initialize_params (src/algorithm/nets.py:55-69): parameters whose name contains '1' (conv1, fc1) keep torch's default
    uniform init, conv2 is Xavier-normal with a zero bias; layout = MnistNet.named_parameters(), D = 8490.  Host logic only: the
    context is a stub that records what set_theta receives."""
    import importlib
    import numpy as np
    nets = importlib.import_module('nes-img-captioning_b200.nets')
    lay = importlib.import_module('nes-img-captioning_b200.layout')

    class StubCtx:
        dims = None
        dev = 'cpu'

        def __init__(self):
            self.tensors, self.D, self.D_pad = lay.build_layout(lay.mnist_tensors(), 64)
            self.theta = None

        def set_mutation(self, kind):
            self.mutation = kind

        def set_theta(self, vector, parent=0):
            self.theta = vector.clone()

        def get_theta(self, parent=0, dtype=None):
            return self.theta.clone()

    ctx = StubCtx()
    net = nets.MnistNet(ctx, rng_state=5, options={'safe_mutations': '', 'safe_mutation_underflow': 0.2})
    assert ctx.D == 8490 and net.get_nb_learnable_params() == 8490 and ctx.mutation == ''
    th = ctx.theta.numpy()
    seg = {t.name: th[t.start:t.start + t.count] for t in ctx.tensors}
    assert np.abs(seg['conv1.weight']).max() <= 1 / 5 and np.abs(seg['conv1.bias']).max() <= 1 / 5 and np.abs(seg['conv1.bias']).max() > 0
    assert np.abs(seg['fc1.weight']).max() <= 1 / np.sqrt(320) + 1e-7 and np.abs(seg['fc1.bias']).max() > 0
    assert np.all(seg['conv2.bias'] == 0)
    assert abs(seg['conv2.weight'].std() - np.sqrt(2.0 / (250 + 500))) < 0.003           # Xavier-normal, fan_in 250, fan_out 500
    sd = net.state_dict()
    assert list(sd) == ['conv1.weight', 'conv1.bias', 'conv2.weight', 'conv2.bias', 'fc1.weight', 'fc1.bias']
    assert tuple(sd['conv1.weight'].shape) == (10, 1, 5, 5) and tuple(sd['fc1.weight'].shape) == (10, 320)
    same = nets.MnistNet(StubCtx(), rng_state=5, options={})
    assert np.array_equal(same.parameter_vector().numpy(), th)                               # seeded init is reproducible


def test_clf_policy_host_logic_on_a_stub_context():
    """ClfPolicy.rollout / accuracy_on / set_model (src/classification/policies.py:16-86, src/algorithm/policies.py:76-111) with the device
    call replaced by the oracle: the float the reference returns, the mean of per-batch accuracies, state_dict hand-over"""
    import importlib
    import numpy as np
    import torch
    from oracle import mnist
    pkg = importlib.import_module('nes-img-captioning_b200')
    lay = pkg.layout

    class StubCtx:
        dims, dev = None, 'cpu'

        def __init__(self):
            self.tensors, self.D, self.D_pad = lay.build_layout(lay.mnist_tensors(), 64)
            self.theta = {}

        def set_mutation(self, kind):
            pass

        def set_theta(self, vector, parent=0):
            self.theta[parent] = vector.clone()

        def get_theta(self, parent=0, dtype=None):
            return self.theta[parent].clone()

        def mnist_rollout(self, offsets, signs, images, labels, sigma, parents=None, shared_batch=None, out_device='cpu', want_logits=False):
            assert sigma == 0.0 and len(offsets) == 1
            f, c = mnist.rollout(self.theta[int(parents[0])].numpy(), np.asarray(images), np.asarray(labels))
            return torch.tensor([f], dtype=torch.float64), torch.tensor([c], dtype=torch.int32)

    exp = {'policy_options': {'net': 'mnist', 'vbn': False, 'model_options': {'safe_mutations': ''}}}
    policy = pkg.PolicyFactory.create(pkg.SuppDataset.MNIST, exp, ctx=StubCtx())
    assert isinstance(policy, pkg.ClfPolicy) and policy.net == pkg.Net.MNIST and policy.nb_learnable_params() == 8490
    theta = mnist.seeded_theta(3, scale=2.0)
    policy.set_from_parameter_vector(theta)
    rs = np.random.RandomState(0)
    batches = [(torch.from_numpy(rs.randn(n, 1, 28, 28).astype(np.float32)), torch.from_numpy(rs.randint(0, 10, n))) for n in (16, 16, 5)]
    fit = policy.rollout(None, batches[0], None)
    want, _ = mnist.rollout(theta, batches[0][0].numpy(), batches[0][1].numpy())
    assert isinstance(fit, float) and fit == float(np.float32(want))
    cfg = type('Cfg', (), {'num_val_batches': 2})()
    accs = [mnist.rollout(theta, b[0].numpy(), b[1].numpy())[1] / len(b[1]) for b in batches]
    assert abs(policy.accuracy_on(batches, cfg, None) - np.mean(accs[:2])) < 1e-12           # num_val_batches caps the loop (:57-60)
    cfg.num_val_batches = 0
    assert abs(policy.accuracy_on(batches, cfg, None) - np.mean(accs)) < 1e-12              # unweighted mean, the reference's own caveat
    # set_model with a state_dict (src/algorithm/policies.py:80-88)
    sd = policy.get_model().state_dict()
    sd['fc1.bias'] = sd['fc1.bias'] + 1.0
    policy.set_model(sd)
    assert np.array_equal(policy.parameter_vector().numpy()[-10:], theta[-10:] + 1.0)
    with pytest.raises(AssertionError):
        policy.set_from_parameter_vector(np.zeros(10, dtype=np.float32))
