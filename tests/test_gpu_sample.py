"""Sampled decode (FCModel._sample(greedy=False), src/captioning/nets.py:210-231) and the fitnesses built on it
(Fitness.SAMPLE / SELF_CRITICAL / SC_LOSS, src/captioning/policies.py:27-30,113-125,164-186) through the C ABI against the
oracle, whose sampler is pinned against the imported reference with the global numpy generator seeded (oracle/refcheck.py).

The reference draws from the unseeded generator; here the draws are an input.  Given the same draws the tokens are the
oracle's unless a draw lies within float32 round-off of a cdf edge (the device logits are within 1e-3 of the oracle's, not
bit-identical): the tests demand identical tokens up to the first such edge case of a row and bound how close to the
edge it was."""
import numpy as np
import pytest
import torch

from oracle import cider, fc_caption, noise
from oracle import fitness as ofit

pytestmark = pytest.mark.gpu

EDGE_TOL = 2e-5        # a token may differ only where |u - cdf edge| is below this


def _compare_rows(seq_d, lp_d, seq_o, lp_o, gap):
    """-> number of rows that diverged at an edge case; asserts everything else is identical"""
    diverged = 0
    for b in range(seq_o.shape[0]):
        neq = np.nonzero(seq_d[b] != seq_o[b])[0]
        upto = seq_o.shape[1] if len(neq) == 0 else int(neq[0])
        assert np.allclose(lp_d[b, :upto], lp_o[b, :upto], atol=1e-3), (b, np.abs(lp_d[b, :upto] - lp_o[b, :upto]).max())
        if len(neq):
            assert gap[b, upto] < EDGE_TOL, 'row %d step %d: tokens %d / %d with the draw %.3g away from the cdf edge' % (
                b, upto, seq_d[b, upto], seq_o[b, upto], gap[b, upto])
            diverged += 1
    return diverged


@pytest.mark.parametrize('V,E,R,Fd,B,ln', [(199, 64, 64, 128, 6, False), (999, 64, 128, 256, 5, False), (300, 128, 128, 512, 7, True)])
def test_sampled_decode_against_oracle(V, E, R, Fd, B, ln):
    import nes_b200
    dims = fc_caption.Dims(V, E, R, Fd, layer_n=ln, layer_n_affine=ln)
    G = noise.granularity(E, R, Fd)
    play, D, D_pad = noise.padded_layout(dims.layout(), G)
    L = D_pad + 63 * G
    ctx = nes_b200.NesContext(dict(vocab_size=V, input_encoding_size=E, rnn_size=R, fc_feat_size=Fd, layer_n=ln, layer_n_affine=ln),
                              seed=31, table_len=L, wave_members=4, rows=B)          # 6 members over waves of 4
    theta = fc_caption.xavier_theta(dims, 5)
    ctx.set_theta(theta)
    fc = noise.table_slice(8, 0, B * Fd).astype(np.float32).reshape(B, Fd)
    offs = np.repeat(ctx.sample_offsets(2, 3).numpy(), 2)
    signs = np.tile(np.array([1.0, -1.0], dtype=np.float32), 3)
    T = dims.seq_length
    u = np.random.RandomState(4).random_sample((6, B, T))
    sigma = 0.02
    seq, lp = ctx.rollout(offs, signs, fc, sigma, uniforms=u)
    seq2, lp2 = ctx.rollout(torch.from_numpy(offs).cuda(), torch.from_numpy(signs).cuda(), torch.from_numpy(fc).cuda(), sigma,
                            uniforms=torch.from_numpy(u).cuda(), out_device='cuda')
    assert torch.equal(seq, seq2.cpu()) and torch.equal(lp, lp2.cpu())                 # deterministic; device and host pointers alike
    seq, lp = seq.numpy(), lp.numpy()
    diverged = 0
    for m in range(6):
        th, _ = noise.member_theta(theta, noise.noise_row(31, offs[m], play, D), sigma, signs[m])
        seq_o, lp_o, gap = fc_caption.decode_sample(dims, th, fc, u[m])
        diverged += _compare_rows(seq[m], lp[m], seq_o, lp_o, gap)
    assert diverged <= 1
    # not the greedy captions
    g, _ = ctx.rollout(offs, signs, fc, sigma, precision=nes_b200.PREC_HYBRID)
    assert not np.array_equal(g.numpy(), seq)
    ctx.close()


def test_sampled_tokens_follow_the_distribution():
    """draws through k_sample reproduce the step-1 distribution of one row: chi-square over 4096 independent draws"""
    import nes_b200
    V, E, R, Fd = 49, 64, 64, 128
    dims = fc_caption.Dims(V, E, R, Fd)
    ctx = nes_b200.NesContext(dict(vocab_size=V, input_encoding_size=E, rnn_size=R, fc_feat_size=Fd), seed=3, table_len=1 << 20, wave_members=8, rows=128)
    theta = fc_caption.xavier_theta(dims, 1) * 3.0
    ctx.set_theta(theta)
    fc = np.tile(noise.table_slice(8, 0, Fd).astype(np.float32)[None], (128, 1))        # 128 identical rows
    n = 32
    u = np.random.RandomState(0).random_sample((n, 128, dims.seq_length))
    seq, _ = ctx.rollout(np.zeros(n, dtype=np.int64), np.ones(n, dtype=np.float32), fc, 0.0, uniforms=u)
    first = seq.numpy()[:, :, 0].reshape(-1)
    _, _, tr = fc_caption.decode_greedy(dims, theta, fc[:1], trace=True)
    lg = tr['logits'][0][0].astype(np.float64)
    p = np.exp(lg - lg.max()); p /= p.sum()
    cnt = np.bincount(first, minlength=V + 1).astype(np.float64)
    big = p * len(first) >= 5
    chi2 = (((cnt - p * len(first)) ** 2) / (p * len(first)))[big].sum()
    assert chi2 < 2.0 * big.sum() + 20, (chi2, big.sum())
    ctx.close()


@pytest.mark.parametrize('kind', ['sample', 'self_critical', 'sc_loss'])
def test_sampling_fitnesses_through_capt_policy(kind):
    """CaptPolicy.rollout_population_fitness for the sampling fitnesses == the oracle's recipe on the device's own tokens:
    CIDEr-D of the sampled captions (minus the greedy baseline; corpus statistics over 2 x batch documents as compute_ciders
    builds them), mean * 100 or LogFitnessCriterion"""
    import nes_b200
    V, E, R, Fd, n_img, spi = 199, 64, 64, 128, 4, 3
    exp = dict(policy_options=dict(net='fc_caption', fitness=kind,
                                   model_options=dict(vocab_size=V, input_encoding_size=E, rnn_size=R, fc_feat_size=Fd)))
    pol = nes_b200.PolicyFactory.create(nes_b200.SuppDataset.MSCOCO, exp, seq_per_img=spi, seed=9, table_len=1 << 22, wave_members=4, rows=n_img * spi)
    pol.scorer = nes_b200.CiderD(pol.ctx, df='corpus')
    pol.precision = nes_b200.PREC_HYBRID
    ctx = pol.ctx
    rs = np.random.RandomState(2)
    fc_u = rs.randn(n_img, Fd).astype(np.float32)
    rows, T = n_img * spi, 16
    offs = np.repeat(ctx.sample_offsets(1, 2).numpy(), 2)
    signs = np.tile(np.array([1.0, -1.0], dtype=np.float32), 2)
    u = rs.random_sample((4, rows, T))
    # references that share words with what the members say, so the scores are not all zero
    seq_s, lp_s = ctx.rollout(offs, signs, np.repeat(fc_u, spi, axis=0), 0.05, uniforms=u)
    seq_g, _ = ctx.rollout(offs, signs, fc_u, 0.05, precision=nes_b200.PREC_HYBRID)
    seq_s, lp_s, seq_g = seq_s.numpy(), lp_s.numpy(), seq_g.numpy()
    gts = [np.stack([seq_s[0, i * spi], seq_g[1, i], rs.randint(0, V, T)]) for i in range(n_img)]
    data = {'fc_feats': np.repeat(fc_u, spi, axis=0), 'gts': gts}
    got = pol.rollout_population_fitness(offs, signs, data, 0.05, uniforms=u)
    assert got.shape == (4,)
    for m in range(4):
        if kind == 'sample':
            mean, scores = cider.cider_scores(seq_s[m], gts, seq_per_img=spi)
        else:
            both = np.concatenate([seq_s[m], np.repeat(seq_g[m], spi, axis=0)])
            _, s2 = cider.cider_scores(both, gts + gts, seq_per_img=spi)
            scores = s2[:rows] - s2[rows:]
            mean = np.mean(scores)
        if kind == 'sc_loss':
            want = ofit.shaped_fitness('sc_loss', lp_s[m], seq_s[m], scores)
            assert abs(got[m] - want) <= 1e-5 * max(abs(want), 1e-3), (m, got[m], want)
        else:
            assert got[m] == float(mean * 100), (m, got[m], float(mean * 100))
    assert np.abs(got).max() > 0
    # the F = 1 form the reference calls
    one = pol.rollout(None, data, None)
    assert np.isfinite(one)
    ctx.close()


@pytest.mark.parametrize('name', ['g1', 'g2', 'g5'])
def test_sampled_decode_against_reference_golden(name, golden_mod):
    """the device sampler against what the REAL reference sampled with its numpy generator seeded (tests/golden/golden_sample_v1.npz)"""
    import os
    import nes_b200
    import make_golden_sample as mgs
    gs = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'golden_sample_v1.npz'))
    c = golden_mod.CASES[name]
    dims, G, play, D, D_pad, L, theta, fc, offs = golden_mod.case_inputs(c)
    ln, aff = c.get('ln', False), c.get('affine', False)
    ctx = nes_b200.NesContext(dict(vocab_size=c['V'], input_encoding_size=c['E'], rnn_size=c['R'], fc_feat_size=c['Fd'], layer_n=ln, layer_n_affine=aff),
                              seed=golden_mod.TABLE_SEED, table_len=L, wave_members=4, rows=c['B'])
    ctx.set_theta(theta)
    eps = noise.noise_row(golden_mod.TABLE_SEED, offs[0], play, D)
    diverged = 0
    for k, seed in enumerate(mgs.SEEDS):
        u = mgs.uniforms(seed, dims.seq_length, c['B'])
        seq, lp = ctx.rollout(np.repeat(offs[:1], 2), np.array([1.0, -1.0], dtype=np.float32), fc, c['sigma'], uniforms=np.stack([u, u]))
        for j, sign in enumerate((+1, -1)):
            th, _ = noise.member_theta(theta, eps, c['sigma'], sign)
            _, _, gap = fc_caption.decode_sample(dims, th, fc, u)
            diverged += _compare_rows(seq[j].numpy(), lp[j].numpy(), gs[name + '_seq'][k, j].astype(np.int64), gs[name + '_logprobs'][k, j], gap)
    assert diverged <= 1
    ctx.close()
