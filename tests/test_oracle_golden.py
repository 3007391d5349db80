"""Oracle vs the golden vectors the REAL reference produced (CPU, no GPU)."""
import os
import numpy as np
import pytest

from oracle import fc_caption, nes_math, noise


@pytest.mark.parametrize('name', ['g1', 'g2', 'g3', 'g5', 'g6'])
def test_decode_matches_reference_golden(name, golden, golden_mod):
    c = golden_mod.CASES[name]
    dims, G, play, D, D_pad, L, theta, fc, offs = golden_mod.case_inputs(c)
    assert np.float64(theta.astype(np.float64).sum()) == golden[name + '_theta_sum']
    s0, lp0 = fc_caption.decode_greedy(dims, theta, fc)
    assert np.array_equal(s0, golden[name + '_seq_theta'])
    assert np.array_equal(lp0, golden[name + '_logprobs_theta'])
    for i in range(c['F']):
        eps = noise.noise_row(golden_mod.TABLE_SEED, offs[i], play, D)
        for j, sign in enumerate((+1, -1)):
            th, _ = noise.member_theta(theta, eps, c['sigma'], sign)
            s, lp = fc_caption.decode_greedy(dims, th, fc)
            assert np.array_equal(s, golden[name + '_seq'][i, j])
            assert np.array_equal(lp, golden[name + '_logprobs'][i, j])


def test_centered_ranks_known_answer(golden):
    # the reference's only published known answer: nic_nes_master.py:189
    got = nes_math.compute_centered_ranks(np.array([[101, 200], [2, 100]]))
    assert np.allclose(got, [[0.16666667, 0.5], [-0.5, -0.16666667]], atol=1e-8)
    assert np.array_equal(got, golden['mm_kat_out'])


def test_gradient_and_ranks_match_reference_golden(golden, golden_mod):
    c = golden_mod.CASES['g1']
    dims, G, play, D, D_pad, L, theta, fc, _ = golden_mod.case_inputs(c)
    fit, offs = golden['mm_fitness'], golden['mm_offsets']
    assert np.array_equal(nes_math.compute_centered_ranks(fit), golden['mm_centered_ranks'])
    nz = np.stack([noise.member_theta(theta, noise.noise_row(golden_mod.TABLE_SEED, o, play, D), 0.01, +1)[1]
                   for o in offs])
    assert np.array_equal(nes_math.gradient_estimate(fit, nz), golden['mm_grad'])


@pytest.mark.parametrize('oname', ['adam', 'sgd'])
def test_optimizer_trajectory_matches_reference_golden(oname, golden, golden_mod):
    c = golden_mod.CASES['g1']
    dims, G, play, D, D_pad, L, theta, fc, _ = golden_mod.case_inputs(c)
    g = golden['mm_grad']
    opt = {'adam': lambda: nes_math.Adam(theta.copy(), stepsize=0.001),
           'sgd': lambda: nes_math.SGD(theta.copy(), stepsize=0.02, momentum=0.9)}[oname]()
    th = theta.copy()
    ratios = []
    for step in range(3):
        gi = (g * np.float32(1.0 + 0.5 * step)).astype(np.float32)
        r, th = nes_math.master_update(opt, th, gi, 1e-7)
        ratios.append(r)
        if step == 0:
            assert np.array_equal(th[::16], golden['opt_%s_theta1_s16' % oname])
    assert np.array_equal(np.array(ratios), golden['opt_%s_ratios' % oname])
    assert np.array_equal(th[::16], golden['opt_%s_theta3_s16' % oname])
    assert th.sum() == golden['opt_%s_theta3_sum' % oname]


def test_rank_ties_use_index_order():
    # D7: the reference's default argsort leaves tie order unspecified; this build's rule is stable
    x = np.array([3., 1., 3., 1., 2.])
    assert list(nes_math.compute_ranks(x)) == [3, 0, 4, 1, 2]


def test_early_break_leaves_zeros():
    # a model whose argmax is token 0 at the first step stops immediately (nets.py:242-243)
    dims = fc_caption.Dims(9, 8, 8, 8)
    theta = np.zeros(dims.num_params(), dtype=np.float32)
    lay = {n: (s, c) for n, _, s, c in dims.layout()}
    s, c = lay['logit.bias']
    theta[s] = 5.0
    seq, lp = fc_caption.decode_greedy(dims, theta, np.ones((3, 8), dtype=np.float32))
    assert (seq == 0).all() and (lp[:, 1:] == 0).all() and (lp[:, 0] < 0).all()


@pytest.mark.parametrize('name', ['g1', 'g2', 'g5'])
def test_sampled_decode_matches_reference_golden(name, golden_mod):
    """oracle decode_sample fed the uniforms of the seeded numpy generator == the reference's FCModel._sample(greedy=False)"""
    import make_golden_sample as mgs
    gs = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'golden_sample_v1.npz'))
    c = golden_mod.CASES[name]
    dims, G, play, D, D_pad, L, theta, fc, offs = golden_mod.case_inputs(c)
    eps = noise.noise_row(golden_mod.TABLE_SEED, offs[0], play, D)
    for k, seed in enumerate(mgs.SEEDS):
        u = mgs.uniforms(seed, dims.seq_length, c['B'])
        for j, sign in enumerate((+1, -1)):
            th, _ = noise.member_theta(theta, eps, c['sigma'], sign)
            s, lp, _ = fc_caption.decode_sample(dims, th, fc, u)
            assert np.array_equal(s, gs[name + '_seq'][k, j])
            assert np.array_equal(lp, gs[name + '_logprobs'][k, j])


@pytest.mark.parametrize('name', ['v1', 'v2', 'v3'])
def test_vbn_decode_matches_reference_golden(name):
    """the batch-norm ("vbn") cell of the oracle against the real reference's outputs"""
    import make_golden_vbn as mv
    gv = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'golden_vbn_v1.npz'))
    c = mv.CASES[name]
    dims, G, play, D, D_pad, L, theta, fc, offs = mv.case_inputs(c)
    eps = noise.noise_row(mv.TABLE_SEED, offs[0], play, D)
    for j, sign in enumerate((+1, -1)):
        th, _ = noise.member_theta(theta, eps, c['sigma'], sign)
        s, lp = fc_caption.decode_greedy(dims, th, fc)
        assert np.array_equal(s, gv[name + '_seq'][j]) and np.array_equal(lp, gv[name + '_logprobs'][j])


@pytest.mark.parametrize('name,members', [('t5a', [0, 1, 2, 3]), ('t5b', [0, 1]), ('t4', [1]), ('t2', [6])])
def test_oracle_matches_reference_at_the_timed_shapes(name, members):
    """the restatement against the goldens at the shapes the benchmarks time (tests/golden/make_golden_timed.py): LayerNorm with 150 / 200 /
    512 rows, and one E = R = 512 / V = 9487 member of a parent other than parent 0 with a negative sign"""
    import os
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, os.path.join(here, 'golden'))
    import make_golden_timed as mt
    from oracle import fc_caption, noise
    gold = np.load(os.path.join(here, 'golden', 'golden_timed_v1.npz'))
    c = mt.CASES[name]
    dims, G, play, D, D_pad, L, thetas, fc, offs = mt.case_inputs(c)
    for i in members:
        th, _ = noise.member_theta(thetas[c['parents'][i]], noise.noise_row(mt.TABLE_SEED, offs[i], play, D), c['sigma'], c['signs'][i])
        s, lp = fc_caption.decode_greedy(dims, th, fc)
        assert np.array_equal(s, gold[name + '_seq'][i].astype(np.int64))
        assert np.allclose(lp, gold[name + '_logprobs'][i], rtol=2e-6, atol=2e-6)     # MKL's summation order follows the thread count


@pytest.mark.parametrize('name', ['e1', 'e2', 'e3', 'e4'])
def test_oracle_vbn_e_matches_reference_golden(name):
    """batch norm on the embeddings (src/captioning/nets.py:156-165), with every cell variant, against the real reference's goldens"""
    import os
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, os.path.join(here, 'golden'))
    import make_golden_vbne as me
    from oracle import fc_caption, noise
    gold = np.load(os.path.join(here, 'golden', 'golden_vbne_v1.npz'))
    c = me.CASES[name]
    dims, G, play, D, D_pad, L, theta, fc, offs = me.case_inputs(c)
    eps = noise.noise_row(me.TABLE_SEED, offs[0], play, D)
    for j, sign in enumerate((+1, -1)):
        th, _ = noise.member_theta(theta, eps, c['sigma'], sign)
        s, lp = fc_caption.decode_greedy(dims, th, fc)
        assert np.array_equal(s, gold[name + '_seq'][j].astype(np.int64))
        assert np.allclose(lp, gold[name + '_logprobs'][j], rtol=2e-6, atol=2e-6)
