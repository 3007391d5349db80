"""Pin the oracle against the REAL reference, imported from /root/reference.

Run in the build container:  python -m oracle.refcheck
Every comparison is bit-exact (np.array_equal) -- the oracle calls the same
torch/numpy CPU kernels the reference does.  TEST INFRASTRUCTURE.
"""
import sys
import numpy as np
import torch

from . import _refimport, fc_caption, nes_math, noise


def check_decode(ref, V, E, R, Fd, B, ln=False, affine=False, seed=0, vbn=False, vbn_e=False):
    opt = ref.ModelOptions(vocab_size=V, input_encoding_size=E, rnn_size=R, fc_feat_size=Fd,
                           layer_n=ln, layer_n_affine=affine and ln, vbn=vbn, vbn_e=vbn_e, vbn_affine=affine and (vbn or vbn_e))
    torch.manual_seed(seed)
    model = ref.FCModel(options=opt, vbn=vbn)
    if (vbn or vbn_e) and affine:            # the reference leaves BN at 1 / 0: give the affine parameters something to do
        for n, p_ in model.named_parameters():
            if 'bn' in n or '.1.' in n:
                p_.data.add_(torch.randn(p_.shape) * 0.1)
    dims = fc_caption.Dims(V, E, R, Fd, layer_n=ln, layer_n_affine=affine and ln, vbn=vbn, vbn_affine=affine and (vbn or vbn_e), vbn_e=vbn_e)
    # layout: names, shapes and order must match named_parameters()
    got = [(n, tuple(p.shape)) for n, p in model.named_parameters()]
    want = [(n, tuple(s)) for n, s, _, _ in dims.layout()]
    assert got == want, (got, want)
    assert model.nb_learnable_params == dims.num_params()
    theta = model.parameter_vector().numpy().copy()
    fc = torch.randn(B, Fd)
    seq_r, lp_r = model(fc, greedy=True)
    seq_o, lp_o = fc_caption.decode_greedy(dims, theta, fc.numpy())
    assert np.array_equal(seq_r.numpy(), seq_o), 'seq differs'
    assert np.array_equal(lp_r.numpy(), lp_o), 'logprobs differ'
    # perturbed member through the reference's own set_from_vector
    G = noise.granularity(E, R, Fd)
    play, D, D_pad = noise.padded_layout(dims.layout(), G)
    eps = noise.noise_row(77, 3 * G, play, D)
    for sign in (+1, -1):
        th, nz = noise.member_theta(theta, eps, 0.01, sign)
        model.set_from_vector(torch.from_numpy(th.copy()))
        seq_r, lp_r = model(fc, greedy=True)
        seq_o, lp_o = fc_caption.decode_greedy(dims, th, fc.numpy())
        assert np.array_equal(seq_r.numpy(), seq_o) and np.array_equal(lp_r.numpy(), lp_o)
    return True


def check_sample_decode(ref, V=199, E=64, R=64, Fd=128, B=7, seed=11):
    """FCModel._sample(greedy=False) with the global numpy generator seeded == decode_sample fed the same generator's
    uniforms (B draws per step starting with the image step, whose samples the reference discards)"""
    opt = ref.ModelOptions(vocab_size=V, input_encoding_size=E, rnn_size=R, fc_feat_size=Fd, layer_n=False, layer_n_affine=False)
    torch.manual_seed(seed)
    model = ref.FCModel(options=opt)
    dims = fc_caption.Dims(V, E, R, Fd)
    theta = model.parameter_vector().numpy().copy()
    fc = torch.randn(B, Fd)
    T = dims.seq_length
    for s in range(3):
        np.random.seed(100 + s)
        seq_r, lp_r = model(fc, greedy=False)
        draws = np.random.RandomState(100 + s).random_sample((T + 1, B))       # [t, b]; row 0 = the image step
        seq_o, lp_o, gap = fc_caption.decode_sample(dims, theta, fc.numpy(), draws[1:].T.copy())
        assert np.array_equal(seq_r.numpy(), seq_o), 'sampled seq differs'
        assert np.array_equal(lp_r.numpy(), lp_o), 'sampled logprobs differ'
    return True


def check_master_math(ref, F=37, D=1000, seed=0):
    rs = np.random.RandomState(seed)
    m = object.__new__(ref.NESMaster)      # the math methods use no instance state
    fit = rs.randn(F, 2) * 10
    nz = rs.randn(F, D).astype(np.float32) * 0.01
    assert np.array_equal(m.compute_centered_ranks(fit), nes_math.compute_centered_ranks(fit))
    assert np.array_equal(m.compute_ranks(fit.ravel()), nes_math.compute_ranks(fit.ravel()))
    g_r = m.gradient_estimate(fit, nz)
    g_o = nes_math.gradient_estimate(fit, nz)
    assert g_r.dtype == g_o.dtype and np.array_equal(g_r, g_o)
    # > 500 rows exercises the chunking of batched_weighted_sum
    F2 = 1203
    fit2 = rs.randn(F2, 2); nz2 = rs.randn(F2, 64).astype(np.float32)
    assert np.array_equal(m.gradient_estimate(fit2, nz2), nes_math.gradient_estimate(fit2, nz2))
    kat = m.compute_centered_ranks(np.array([[101, 200], [2, 100]]))
    assert np.allclose(kat, [[0.16666667, 0.5], [-0.5, -0.16666667]])
    return True


def check_optimizers(ref, D=513, steps=5, seed=0):
    rs = np.random.RandomState(seed)
    theta0 = rs.randn(D).astype(np.float32)
    for name, kw in (('Adam', dict(stepsize=0.001)), ('SGD', dict(stepsize=0.02, momentum=0.9))):
        o_r = getattr(ref.optimizers, name)(theta0.copy(), **kw)
        o_o = getattr(nes_math, name)(theta0.copy(), **kw)
        th_r = theta0.copy(); th_o = theta0.copy()
        for _ in range(steps):
            g = rs.randn(D).astype(np.float32) * 0.1
            r1, th_r = o_r.update(-g + 1e-7 * th_r)
            r2, th_o = nes_math.master_update(o_o, th_o, g, 1e-7)
            assert r1 == r2 and np.array_equal(th_r, th_o) and th_r.dtype == th_o.dtype
    return True


def check_cider(seed=0):
    """oracle/cider.py against the real pyciderevalcap scorer fed like CaptPolicy.compute_ciders (captioning/policies.py:145-193)"""
    import os
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden'))
    import make_golden_cider as mg
    from . import cider
    for c in mg.CASES.values():
        seq, gts = mg.case_inputs(c)
        score_r, scores_r = mg.reference_scores(seq, gts, c['spi'])
        score_o, scores_o = cider.cider_scores(seq, gts, c['spi'])
        assert score_r == score_o and np.array_equal(scores_r, scores_o)
    return True


def check_shaped_fitness(seed=0):
    """oracle/fitness.py against the imported captioning.fitness criteria"""
    from captioning import fitness as rf
    from . import fitness as of
    cls = {'greedy_logprob': rf.AltLogFitnessCriterion, 'greedy_expprob': rf.ExpFitnessCriterion, 'greedy_linprob': rf.LinFitnessCriterion,
           'greedy_avgprob': rf.AvgLogFitnessCriterion, 'sc_loss': rf.LogFitnessCriterion}
    rs = np.random.RandomState(seed)
    rows, T = 40, 16
    seq = rs.randint(0, 9, (rows, T))
    for r in range(rows):
        z = np.nonzero(seq[r] == 0)[0]
        if len(z):
            seq[r, z[0]:] = 0
    lp = -np.abs(rs.randn(rows, T)).astype(np.float32) * 3
    sc = np.abs(rs.randn(rows)) * 2
    for k, c in cls.items():
        ciders = np.repeat(sc[:, None], T, 1)
        want = float(c()(torch.from_numpy(lp), torch.from_numpy(seq), torch.from_numpy(ciders.copy()).float()).item())
        assert of.shaped_fitness(k, lp, seq, sc) == want, k
    return True


def check_proportional(ref, seed=3):
    """SM-PROPORTIONAL scale of oracle/noise.py against the reference's PolicyNet.evolve (algorithm/nets.py:101-119): the same
    torch normal draw goes through both.  Non-zero parameters: bit-exact; exactly-zero parameters are scaled by mean|theta|,
    where torch's float32 mean and this build's float64-accumulated mean may differ in the last bit (checked to 1e-6)."""
    from algorithm.nets import Mutation
    opt = ref.ModelOptions(vocab_size=199, input_encoding_size=64, rnn_size=64, fc_feat_size=128, safe_mutations='SM-PROPORTIONAL')
    torch.manual_seed(seed)
    model = ref.FCModel(options=opt)
    assert model.mutations == Mutation.SAFE_PROPORTIONAL
    theta = model.parameter_vector().numpy().copy()
    torch.manual_seed(seed + 1)
    raw = torch.empty(theta.size).normal_(mean=0.0, std=0.25).numpy().copy()       # what evolve draws first
    torch.manual_seed(seed + 1)
    got = model.evolve(0.25)
    scale = noise.proportional_scale(theta)
    want = raw * scale
    nzm = theta != 0
    assert nzm.sum() > 0 and (~nzm).sum() > 0
    assert np.array_equal(got[nzm], want[nzm])
    assert np.allclose(got[~nzm], want[~nzm], rtol=1e-6, atol=0)
    assert np.array_equal(model.parameter_vector().numpy()[nzm], (theta + want)[nzm])
    return True


def check_sensitivity(ref, seed=5):
    """SM-VECTOR: oracle sensitivity_vector + division against the reference's PolicyNet.evolve with a sensitivity file
    (algorithm/nets.py:49-50,104-105, safe_mutations.py:27-32): bit-exact"""
    import tempfile, os
    from algorithm.nets import Mutation
    rs = np.random.RandomState(seed)
    D = fc_caption.Dims(199, 64, 64, 128).num_params()
    raw_sens = np.abs(rs.randn(D)).astype(np.float32) * 0.05          # many entries below the 0.01 underflow
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, 'sens.pt')
        torch.save(torch.from_numpy(raw_sens.copy()), path)
        opt = ref.ModelOptions(vocab_size=199, input_encoding_size=64, rnn_size=64, fc_feat_size=128, safe_mutations='SM-VECTOR',
                               safe_mutation_vector=path, safe_mutation_underflow=0.01)
        torch.manual_seed(seed)
        model = ref.FCModel(options=opt)
    assert model.mutations == Mutation.SAFE_VECTOR
    theta = model.parameter_vector().numpy().copy()
    sens = noise.sensitivity_vector(raw_sens, 0.01)
    assert np.array_equal(model.get_sensitivity_vector().numpy(), sens)
    torch.manual_seed(seed + 1)
    raw = torch.empty(theta.size).normal_(mean=0.0, std=0.25).numpy().copy()
    torch.manual_seed(seed + 1)
    got = model.evolve(0.25)
    assert np.array_equal(got, raw / sens)
    assert np.array_equal(model.parameter_vector().numpy(), theta + raw / sens)
    return True


def check_grad_sensitivity(ref):
    """SM-G-SUM: the package's autograd restatement (nes-img-captioning_b200/sensitivity.py, torch on the CPU here) against the
    reference's Sensitivity.calc_sensitivity over FCModel.forward_for_sensitivity (safe_mutations.py:34-115, captioning/nets.py:22-69),
    plain / LayerNorm / batch-norm cells: bit-exact on this torch build.  This is the one check that exercises product code: the
    function is device-agnostic torch and has no separate oracle restatement."""
    import importlib, tempfile
    pkg_sm = importlib.import_module('nes-img-captioning_b200.sensitivity')
    lay = importlib.import_module('nes-img-captioning_b200.layout')
    for cfg in [dict(V=199, E=64, R=64, Fd=128, ln=0, aff=0, vbn=0, uf=1e-4), dict(V=299, E=32, R=64, Fd=128, ln=1, aff=1, vbn=0, uf=0.01),
                dict(V=230, E=32, R=32, Fd=64, ln=0, aff=1, vbn=1, uf=0.01), dict(V=230, E=32, R=32, Fd=64, ln=1, aff=0, vbn=0, uf=0.01),
                dict(V=230, E=32, R=32, Fd=64, ln=0, aff=0, vbn=0, uf=0.01, vbn_e=1, vaff=1), dict(V=230, E=32, R=32, Fd=64, ln=0, aff=1, vbn=1, uf=0.01, vbn_e=1, vaff=1)]:
        vbn_e, vaff = bool(cfg.get('vbn_e')), bool(cfg.get('vaff'))
        opt = ref.ModelOptions(vocab_size=cfg['V'], input_encoding_size=cfg['E'], rnn_size=cfg['R'], fc_feat_size=cfg['Fd'],
                               safe_mutations='SM-G-SUM', safe_mutation_underflow=cfg['uf'], layer_n=bool(cfg['ln']),
                               layer_n_affine=bool(cfg['aff'] and cfg['ln']), vbn=bool(cfg['vbn']), vbn_e=vbn_e,
                               vbn_affine=bool(cfg['aff'] and cfg['vbn']) or vaff)
        torch.manual_seed(3)
        model = ref.FCModel(options=opt, vbn=bool(cfg['vbn']))
        theta = model.parameter_vector().detach().clone()
        fc = np.repeat(np.random.RandomState(1).randn(6, cfg['Fd']).astype(np.float32), 5, axis=0)
        with tempfile.TemporaryDirectory() as d:
            model.calc_sensitivity(0, 0, {'fc_feats': fc}, 4, d)
            want = model.get_sensitivity_vector().numpy().copy()
        torch.set_grad_enabled(False)                    # the reference leaves it off (safe_mutations.py:67)
        layer_n = 2 if cfg['vbn'] else cfg['ln']
        dims = dict(vocab_size=cfg['V'], input_encoding_size=cfg['E'], rnn_size=cfg['R'], fc_feat_size=cfg['Fd'], layer_n=layer_n, layer_n_affine=cfg['aff'], vbn_e=vbn_e)
        tensors, D, _ = lay.build_layout(lay.fc_tensors(cfg['V'], cfg['E'], cfg['R'], cfg['Fd'], layer_n, cfg['aff'], vbn_e, vbn_e and vaff), 64)
        assert D == theta.numel()
        got = pkg_sm.sum_sensitivity(theta, tensors, dims, fc, 4)
        got[got < cfg['uf']] = cfg['uf']
        got /= cfg['uf']
        assert np.array_equal(got.numpy(), want), np.max(np.abs(got.numpy() - want) / want)
        assert 0.1 < (want == 1).mean() < 0.9
    return True


def check_mnist(ref):
    """oracle/mnist.py against the imported MnistNet / ClfPolicy.rollout (classification/nets.py:25-37, policies.py:16-48): logits bit-exact,
    fitness to float32 round-off, the parameter order and count (8490)"""
    from classification.nets import MnistNet
    from classification.policies import ClfPolicy
    from oracle import mnist
    net = MnistNet(options=None)
    assert [n for n, _ in net.named_parameters()] == [n for n, _, _, _ in mnist.layout()] and net.nb_learnable_params == mnist.D
    policy = object.__new__(ClfPolicy)
    policy.policy_net, policy.vbn = net, False
    cfg = type('Cfg', (), {'cuda': False})()
    rs = np.random.RandomState(2)
    for k in range(3):
        theta = mnist.seeded_theta(k, scale=1.0 + k)
        images = rs.randn(33, 1, 28, 28).astype(np.float32)
        labels = rs.randint(0, 10, 33).astype(np.int64)
        net.set_from_vector(torch.from_numpy(theta.copy()))
        want = net(torch.from_numpy(images)).numpy()
        assert np.array_equal(mnist.forward(theta, images), want)
        fit = policy.rollout(torch.empty(images.shape), (torch.from_numpy(images), torch.from_numpy(labels)), cfg)
        got, cor = mnist.rollout(theta, images, labels)
        assert got == fit and cor == int((want.argmax(1) == labels).sum())
    return True


def main():
    if not _refimport.available():
        print('reference not present; nothing checked'); return 1
    ref = _refimport.load()
    torch.set_grad_enabled(False)
    n = 0
    for cfg in [dict(V=199, E=64, R=64, Fd=128, B=8), dict(V=999, E=64, R=128, Fd=256, B=5),
                dict(V=300, E=128, R=128, Fd=512, B=12, ln=True),
                dict(V=300, E=128, R=128, Fd=512, B=12, ln=True, affine=True),
                dict(V=300, E=128, R=128, Fd=512, B=12, vbn=True), dict(V=300, E=128, R=128, Fd=512, B=12, vbn=True, affine=True),
                dict(V=300, E=128, R=128, Fd=512, B=12, vbn_e=True, affine=True), dict(V=300, E=64, R=64, Fd=128, B=9, vbn_e=True),
                dict(V=300, E=128, R=128, Fd=512, B=12, vbn=True, vbn_e=True, affine=True), dict(V=300, E=64, R=64, Fd=128, B=9, ln=True, vbn_e=True, affine=True),
                dict(V=9487, E=128, R=128, Fd=2048, B=4)]:
        check_decode(ref, **cfg); n += 1
    # the parameter-count checksum the reference states (src/algorithm/tools/utils.py:180)
    assert fc_caption.Dims(9487, 128, 128, 2048).num_params() == 2865808
    assert fc_caption.Dims(9487, 512, 512, 2048).num_params() == 13400848
    check_master_math(ref); check_optimizers(ref); check_cider(); check_shaped_fitness(); check_proportional(ref); check_sample_decode(ref); check_sensitivity(ref); check_grad_sensitivity(ref); check_mnist(ref)
    print('oracle pinned against the imported reference: %d decode configs, master math, optimizers, CIDEr-D, shaped fitnesses, SM-PROPORTIONAL, SM-VECTOR, SM-G-SUM sensitivity, sampled decode, MNIST net OK' % n)
    return 0


if __name__ == '__main__':
    sys.exit(main())
