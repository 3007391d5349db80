"""GPU: the device CIDEr-D scorer (k_cider through nes_cider_set_tables / nes_cider_scores) against the reference goldens and
the pinned oracle -- bit-exact, it is float64 arithmetic in the reference's order."""
import os
import sys

import numpy as np
import pytest
import torch

from oracle import cider, fc_caption, noise

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'tests', 'golden'))
import make_golden_cider as mg  # noqa: E402


@pytest.fixture(scope='module')
def ctx():
    import nes_b200
    c = nes_b200.NesContext(dict(vocab_size=199, input_encoding_size=64, rnn_size=64, fc_feat_size=128), seed=3, table_len=1 << 22, wave_members=8, rows=8)
    yield c
    c.close()


@pytest.mark.parametrize('name', sorted(mg.CASES))
def test_device_scores_equal_reference_golden(name, ctx):
    import nes_b200
    gold = np.load(os.path.join(ROOT, 'tests', 'golden', 'golden_cider_v1.npz'))
    c = mg.CASES[name]
    seq, gts = mg.case_inputs(c)
    sc = nes_b200.CiderD(ctx, df='corpus').set_refs(gts, seq_per_img=c['spi'], rows_per_img=c['spi'])
    s = sc.scores(seq[None]).numpy()[0]
    assert np.array_equal(s, gold[name + '_scores']), np.abs(s - gold[name + '_scores']).max()
    assert float(np.mean(s)) == float(gold[name + '_score'])
    # unique rows decoded once + replication on the host gives the same fitness scalar
    sc1 = nes_b200.CiderD(ctx, df='corpus').set_refs(gts, seq_per_img=c['spi'], rows_per_img=1)
    f = sc1.fitness(1, len(gts), seq=seq[::c['spi']][None])
    same_rows = all(np.array_equal(seq[i], seq[i - i % c['spi']]) for i in range(len(seq)))
    if same_rows:
        assert f[0] == float(gold[name + '_score'] * 100)


def test_many_members_edge_captions_and_reference_style_call(ctx):
    import nes_b200
    rs = np.random.RandomState(0)
    T, n_img, P = 16, 9, 40
    gts = []
    for i in range(n_img):
        k = rs.randint(1, 6)
        refs = np.zeros((k, T), dtype=np.int64)
        for j in range(k):
            ln = rs.randint(0, T + 1)
            refs[j, :ln] = rs.randint(1, 9, ln)
        gts.append(refs)
    seq = rs.randint(0, 9, (P, n_img, T)).astype(np.int64)
    seq[0, 0] = 0                      # empty caption
    seq[1, 1] = 3                      # full length, one repeated word, no terminator
    seq[2, 2] = gts[2][0]              # exact copy of a reference
    for m in range(P):                 # decoder invariant: zeros after the first zero
        for r in range(n_img):
            z = np.nonzero(seq[m, r] == 0)[0]
            if len(z):
                seq[m, r, z[0]:] = 0
    sc = nes_b200.CiderD(ctx, df='corpus').set_refs(gts)
    got = sc.scores(torch.from_numpy(seq).cuda(), device='cuda').cpu().numpy()
    for m in range(P):
        _, want = cider.cider_scores(seq[m], gts)
        assert np.array_equal(got[m], want), (m, np.abs(got[m] - want).max())
    # the reference's own calling convention: dicts of space-separated token strings
    def to_str(a):
        return ' '.join(str(t) for t in cider.caption_tokens(a))
    res = [{'image_id': i, 'caption': [to_str(seq[3, i])]} for i in range(n_img)]
    g = {i: [to_str(r) for r in gts[i]] for i in range(n_img)}
    mean, scores = nes_b200.CiderD(ctx, df='corpus').compute_score(g, res)
    want_mean, want = cider.cider_scores(seq[3], gts)
    assert np.array_equal(scores, want) and mean == want_mean


def test_precomputed_document_frequencies_and_resident_tokens(ctx):
    """df from a pickle-like dict (the reference's coco-train-idxs mode) and scoring of the tokens the last rollout left on the device"""
    import nes_b200
    dims = fc_caption.Dims(199, 64, 64, 128)
    G = noise.granularity(64, 64, 128)
    play, D, D_pad = noise.padded_layout(dims.layout(), G)
    ctx.set_theta(fc_caption.xavier_theta(dims, 1))
    fc = noise.table_slice(8, 0, 8 * 128).astype(np.float32).reshape(8, 128)
    offs = ctx.sample_offsets(0, 3).numpy()
    seq, _ = ctx.rollout(offs, np.ones(3, dtype=np.float32), fc, 0.05, precision=nes_b200.PREC_HYBRID)
    rs = np.random.RandomState(4)
    gts = [np.concatenate([seq[rs.randint(3), i].numpy()[None], rs.randint(0, 199, (2, 16))]) for i in range(8)]
    for refs in gts:
        for r in refs:
            z = np.nonzero(r == 0)[0]
            if len(z):
                r[z[0]:] = 0
    df = {}
    for refs in gts:
        for g_ in set(g_ for r in refs for g_ in cider.precook(cider.caption_tokens(r))):
            df[tuple(str(t) for t in g_)] = df.get(tuple(str(t) for t in g_), 0) + 3
    sc = nes_b200.CiderD(ctx, df='coco-train-idxs', document_frequency=df, ref_len=1000).set_refs(gts)
    got = sc.scores(None, 3, 8).numpy()
    df_int = {tuple(int(t) for t in k): float(v) for k, v in df.items()}
    for m in range(3):
        _, want = cider.cider_scores(seq[m].numpy(), gts, df=df_int, ref_len=np.log(1000.0))
        assert np.array_equal(got[m], want)
    assert got.max() > 0


def test_capt_policy_rollout_with_device_scorer_equals_host_scorer():
    """CaptPolicy.rollout (captioning/policies.py:86-128, greedy fitness = CIDEr-D * 100) with the device scorer against the
    same decode scored by the oracle on the host; 5x replicated loader rows enter through seq_per_img"""
    import nes_b200
    exp = dict(policy_options=dict(net='fc_caption', fitness='greedy',
                                   model_options=dict(vocab_size=199, input_encoding_size=64, rnn_size=64, fc_feat_size=128)))
    pol = nes_b200.PolicyFactory.create(nes_b200.SuppDataset.MSCOCO, exp, seq_per_img=5, seed=9, table_len=1 << 22, wave_members=6, rows=6)
    pol.scorer = nes_b200.CiderD(pol.ctx, df='corpus')
    pol.precision = nes_b200.PREC_HYBRID
    rs = np.random.RandomState(1)
    fc_u = rs.randn(6, 128).astype(np.float32)
    data = {'fc_feats': np.repeat(fc_u, 5, axis=0), 'gts': [rs.randint(0, 199, (rs.randint(1, 6), 16)) for _ in range(6)]}
    fit = pol.rollout(None, data, None)
    seq, _ = pol.rollout_tokens(data)                         # [30, 16], rows replicated like the loader's
    want, _ = cider.cider_scores(seq.numpy(), data['gts'], seq_per_img=5)
    assert fit == float(want * 100)
    offs = pol.ctx.sample_offsets(1, 3).numpy()
    f3 = pol.rollout_population_fitness(np.repeat(offs, 2), np.tile(np.array([1.0, -1.0], dtype=np.float32), 3), data, 0.05)
    assert f3.shape == (6,) and np.isfinite(f3).all()


@pytest.mark.parametrize('kind', ['greedy_logprob', 'greedy_expprob', 'greedy_linprob', 'greedy_avgprob', 'sc_loss'])
def test_shaped_fitness_against_the_pinned_oracle(kind, ctx):
    """src/captioning/fitness.py criteria on the resident log-probs / tokens with CIDEr-D rewards: device (float32 elements,
    float64 sum) vs the oracle (the reference's torch float32 ops, bit-identical to the imported reference): 1e-5 relative"""
    import nes_b200
    from oracle import fitness as ofit
    dims = fc_caption.Dims(199, 64, 64, 128)
    ctx.set_theta(fc_caption.xavier_theta(dims, 2))
    fc = noise.table_slice(12, 0, 8 * 128).astype(np.float32).reshape(8, 128)
    offs = ctx.sample_offsets(5, 4).numpy()
    seq, lp = ctx.rollout(offs, np.ones(4, dtype=np.float32), fc, 0.05, precision=nes_b200.PREC_HYBRID)
    rs = np.random.RandomState(7)
    gts = [np.concatenate([seq[rs.randint(4), i].numpy()[None], rs.randint(0, 199, (2, 16))]) for i in range(8)]
    sc = nes_b200.CiderD(ctx, df='corpus').set_refs(gts)
    scores = sc.scores(None, 4, 8).numpy()
    got = ctx.shaped_fitness(kind, 4, 8).numpy()                            # rewards = the scores just computed, still resident
    got2 = ctx.shaped_fitness(kind, 4, 8, scores=scores).numpy()            # rewards passed from the host
    for m in range(4):
        want = ofit.shaped_fitness(kind, lp[m].numpy(), seq[m].numpy(), scores[m])
        assert abs(got[m] - want) <= 1e-5 * max(abs(want), 1e-3), (m, got[m], want)
        assert got2[m] == got[m]
