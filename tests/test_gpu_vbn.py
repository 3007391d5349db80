"""The reference's "virtual batch norm" cell (BatchNorm1d over the rows of the batch without running statistics,
src/captioning/nets.py:87-91,99-101,124-125) through the C ABI: tokens against the real reference's goldens,
per-step hidden states / logits teacher-forced against the pinned oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import fc_caption, noise

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _mo(c):
    return dict(vocab_size=c['V'], input_encoding_size=c['E'], rnn_size=c['R'], fc_feat_size=c['Fd'], vbn=True, vbn_affine=c['affine'])


@pytest.mark.parametrize('name', ['v1', 'v2', 'v3'])
@pytest.mark.parametrize('prec', [1, 2], ids=['exact', 'hybrid'])
def test_vbn_tokens_match_reference_golden(name, prec):
    import sys
    sys.path.insert(0, os.path.join(HERE, 'golden'))
    import make_golden_vbn as mv
    import nes_b200
    gv = np.load(os.path.join(HERE, 'golden', 'golden_vbn_v1.npz'))
    c = mv.CASES[name]
    dims, G, play, D, D_pad, L, theta, fc, offs = mv.case_inputs(c)
    ctx = nes_b200.NesContext(_mo(c), seed=mv.TABLE_SEED, table_len=L, wave_members=2, rows=c['B'])
    assert ctx.D == D and [t.name for t in ctx.tensors] == [n for n, _, _, _ in dims.layout()]
    ctx.set_theta(theta)
    seq, lp = ctx.rollout(np.repeat(offs[:1], 2), np.array([1.0, -1.0], dtype=np.float32), fc, c['sigma'], precision=prec)
    for j in range(2):
        want = gv[name + '_seq'][j].astype(np.int64)
        assert np.array_equal(seq[j].numpy(), want), '%d tokens differ' % (seq[j].numpy() != want).sum()
        assert np.allclose(lp[j].numpy(), gv[name + '_logprobs'][j], rtol=1e-4, atol=1e-5)
    ctx.close()


@pytest.mark.parametrize('name', ['v1', 'v3'])
def test_vbn_hidden_states_teacher_forced(name):
    """h and logits of every step within 2e-5 (norm-wise) of the oracle on the oracle's token path; the statistics couple the rows
    of the batch, so padding rows (12 of 128, 150 of 256) must not enter them"""
    import sys
    sys.path.insert(0, os.path.join(HERE, 'golden'))
    import make_golden_vbn as mv
    import nes_b200
    c = mv.CASES[name]
    dims, G, play, D, D_pad, L, theta, fc, offs = mv.case_inputs(c)
    ctx = nes_b200.NesContext(_mo(c), seed=mv.TABLE_SEED, table_len=L, wave_members=2, rows=c['B'])
    rs = np.random.RandomState(1)
    theta = theta.copy()
    for n, shape, start, cnt in dims.layout():                       # affine parameters away from 1 / 0
        if 'bn' in n:
            theta[start:start + cnt] += (rs.randn(cnt) * 0.2).astype(np.float32)
    ctx.set_theta(theta)
    eps = noise.noise_row(mv.TABLE_SEED, offs[0], play, D)
    forced, traces = [], []
    for sgn in (+1, -1):
        th, _ = noise.member_theta(theta, eps, c['sigma'], sgn)
        s, lp, tr = fc_caption.decode_greedy(dims, th, fc, trace=True)
        forced.append(s); traces.append(tr)
    forced = np.stack(forced).astype(np.int64)
    seq, lp, h_tr, l_tr = ctx.rollout(np.repeat(offs[:1], 2), np.array([1.0, -1.0], dtype=np.float32), fc, c['sigma'], precision=1,
                                      forced_tokens=forced, trace=True)
    h_tr, l_tr = h_tr.cpu().numpy(), l_tr.cpu().numpy()
    tol = 2e-5
    for m in range(2):
        for t in range(len(traces[m]['h'])):
            ref_h = traces[m]['h'][t]
            assert np.linalg.norm(h_tr[t, m] - ref_h) <= tol * np.linalg.norm(ref_h), ('h', m, t)
            assert np.max(np.abs(h_tr[t, m] - ref_h)) <= 3 * tol * np.max(np.abs(ref_h)), ('h max', m, t)
        for t in range(len(traces[m]['logits'])):
            ref_l = traces[m]['logits'][t]
            assert np.linalg.norm(l_tr[t, m] - ref_l) <= tol * np.linalg.norm(ref_l), ('logits', m, t)
    ctx.close()


def test_vbn_through_the_policy_interface():
    """policy_options.vbn (src/algorithm/policies.py:62-64) and model_options.vbn_e reach the context"""
    import nes_b200
    exp = dict(policy_options=dict(net='fc_caption', fitness='greedy', vbn=True,
                                   model_options=dict(vocab_size=99, input_encoding_size=64, rnn_size=64, fc_feat_size=64, vbn_affine=True)))
    pol = nes_b200.PolicyFactory.create(nes_b200.SuppDataset.MSCOCO, exp, seed=2, table_len=1 << 20, wave_members=2, rows=8)
    assert pol.ctx.dims['layer_n'] == 2 and any('core.c_bn.weight' == t.name for t in pol.ctx.tensors)
    fc = np.random.RandomState(0).randn(8, 64).astype(np.float32)
    seq, lp = pol.policy_net(torch.from_numpy(fc), greedy=True)
    dims = fc_caption.Dims(99, 64, 64, 64, vbn=True, vbn_affine=True)
    s, l = fc_caption.decode_greedy(dims, pol.parameter_vector().numpy(), fc)
    assert np.array_equal(seq.numpy(), s) and np.allclose(lp.numpy(), l, atol=1e-4)
    pol.ctx.close()
    exp['policy_options']['model_options'].update(vbn_e=True)
    pol = nes_b200.PolicyFactory.create(nes_b200.SuppDataset.MSCOCO, exp, seed=2, table_len=1 << 20, wave_members=2, rows=8)
    assert pol.ctx.dims['vbn_e'] == 1 and [t.name for t in pol.ctx.tensors][:7] == [
        'img_embed.0.weight', 'img_embed.0.bias', 'img_embed.1.weight', 'img_embed.1.bias', 'embed.0.weight', 'embed.1.weight', 'embed.1.bias']
    seq, lp = pol.policy_net(torch.from_numpy(fc), greedy=True)
    dims = fc_caption.Dims(99, 64, 64, 64, vbn=True, vbn_affine=True, vbn_e=True)
    s, l = fc_caption.decode_greedy(dims, pol.parameter_vector().numpy(), fc)
    assert np.array_equal(seq.numpy(), s) and np.allclose(lp.numpy(), l, atol=1e-4)
    pol.ctx.close()


@pytest.mark.parametrize('name', ['e1', 'e2', 'e3', 'e4'])
@pytest.mark.parametrize('prec', [1, 2], ids=['exact', 'hybrid'])
def test_vbn_e_tokens_match_reference_golden(name, prec):
    """vbn_e: batch norm over the rows on the outputs of img_embed and embed (src/captioning/nets.py:156-165), alone (e1, e3: 150 rows,
    no affine), with the batch-norm cell (e2) and with the LayerNorm cell (e4); tokens / log-probs of the REAL reference"""
    import sys
    sys.path.insert(0, os.path.join(HERE, 'golden'))
    import make_golden_vbne as me
    import nes_b200
    gv = np.load(os.path.join(HERE, 'golden', 'golden_vbne_v1.npz'))
    c = me.CASES[name]
    dims, G, play, D, D_pad, L, theta, fc, offs = me.case_inputs(c)
    ctx = nes_b200.NesContext(me.model_options(c), seed=me.TABLE_SEED, table_len=L, wave_members=2, rows=c['B'])
    assert ctx.D == D and [t.name for t in ctx.tensors] == [n for n, _, _, _ in dims.layout()]
    ctx.set_theta(theta)
    seq, lp = ctx.rollout(np.repeat(offs[:1], 2), np.array([1.0, -1.0], dtype=np.float32), fc, c['sigma'], precision=prec)
    for j in range(2):
        want = gv[name + '_seq'][j].astype(np.int64)
        assert np.array_equal(seq[j].numpy(), want), '%d tokens differ' % (seq[j].numpy() != want).sum()
        assert np.allclose(lp[j].numpy(), gv[name + '_logprobs'][j], rtol=1e-4, atol=1e-5)
    # the state_dict carries the reference's Sequential names and round-trips
    model = nes_b200.FCModel(ctx, options=me.model_options(c), rng_state=1)
    sd = model.state_dict()
    assert list(sd)[:2] == ['img_embed.0.weight', 'img_embed.0.bias'] and ('embed.1.weight' in sd) == bool(c['affine'])
    if c['affine']:
        assert torch.all(sd['img_embed.1.weight'] == 1) and torch.all(sd['embed.1.bias'] == 0)      # algorithm/nets.py:62-64 leaves them at 1 / 0
    ctx.close()


def test_vbn_e_hidden_states_teacher_forced():
    """h and logits of every step within 2e-5 (norm-wise) of the oracle on the oracle's token path, vbn_e + batch-norm cell"""
    import sys
    sys.path.insert(0, os.path.join(HERE, 'golden'))
    import make_golden_vbne as me
    import nes_b200
    c = me.CASES['e2']
    dims, G, play, D, D_pad, L, theta, fc, offs = me.case_inputs(c)
    ctx = nes_b200.NesContext(me.model_options(c), seed=me.TABLE_SEED, table_len=L, wave_members=2, rows=c['B'])
    ctx.set_theta(theta)
    eps = noise.noise_row(me.TABLE_SEED, offs[0], play, D)
    forced, traces = [], []
    for sgn in (+1, -1):
        th, _ = noise.member_theta(theta, eps, c['sigma'], sgn)
        s, lp, tr = fc_caption.decode_greedy(dims, th, fc, trace=True)
        forced.append(s); traces.append(tr)
    forced = np.stack(forced).astype(np.int64)
    seq, lp, h_tr, l_tr = ctx.rollout(np.repeat(offs[:1], 2), np.array([1.0, -1.0], dtype=np.float32), fc, c['sigma'], precision=1,
                                      forced_tokens=forced, trace=True)
    h_tr, l_tr = h_tr.cpu().numpy(), l_tr.cpu().numpy()
    for m in range(2):
        for t in range(len(traces[m]['h'])):
            ref_h = traces[m]['h'][t]
            assert np.linalg.norm(h_tr[t, m] - ref_h) <= 2e-5 * np.linalg.norm(ref_h), ('h', m, t)
        for t in range(len(traces[m]['logits'])):
            ref_l = traces[m]['logits'][t]
            assert np.linalg.norm(l_tr[t, m] - ref_l) <= 2e-5 * np.linalg.norm(ref_l), ('logits', m, t)
    ctx.close()
