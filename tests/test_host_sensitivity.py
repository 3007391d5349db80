"""SM-G-SUM output-gradient sensitivity (src/algorithm/safe_mutations.py:34-115, src/captioning/nets.py:22-69): the
autograd restatement in the package (device-agnostic torch, run on the CPU here) against vectors the REAL reference produced
(tests/golden/make_golden_sens.py), and the sens_t*_p*_*.pt file protocol the workers share."""
import importlib
import os
import sys
import types

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, 'golden'))
import make_golden_sens as ms  # noqa: E402

sm = importlib.import_module('nes-img-captioning_b200.sensitivity')
lay = importlib.import_module('nes-img-captioning_b200.layout')
GOLD = np.load(os.path.join(HERE, 'golden', 'golden_sens_v1.npz'))


def _layout(c):
    mo = ms.model_options(c)
    layer_n = 2 if mo['vbn'] else int(mo['layer_n'])
    affine = int(mo['vbn_affine'] if mo['vbn'] else mo['layer_n_affine'])
    tensors, D, _ = lay.build_layout(lay.fc_tensors(c['V'], c['E'], c['R'], c['Fd'], layer_n, affine), 64)
    dims = dict(vocab_size=c['V'], input_encoding_size=c['E'], rnn_size=c['R'], fc_feat_size=c['Fd'], layer_n=layer_n, layer_n_affine=affine)
    return tensors, D, dims


def _post(raw, underflow):
    s = raw.clone()
    s[s < underflow] = underflow
    s /= underflow
    return s.numpy()


@pytest.mark.parametrize('name', sorted(ms.CASES))
def test_sum_sensitivity_matches_reference_golden(name):
    c = ms.CASES[name]
    _, theta, fc5 = ms.case_inputs(c)
    tensors, D, dims = _layout(c)
    assert D == theta.size
    raw = sm.sum_sensitivity(torch.from_numpy(theta.copy()), tensors, dims, fc5, c['bs'])
    assert not raw.requires_grad and raw.dtype == torch.float32
    got, want = _post(raw, c['underflow']), GOLD[name + '_sens']
    assert 0.2 < (want == 1).mean() < 0.8                     # both the clamped and the free regime are exercised
    assert np.allclose(got, want, rtol=1e-5, atol=0), np.max(np.abs(got - want) / want)


def test_forward_for_sensitivity_shapes_and_row_selection():
    """every 5th row of fc_feats, at most orig_bs of them, or the single row i (captioning/nets.py:41-46); V1 % split == 0 adds a whole
    group of zeros (:62-63)"""
    c = ms.CASES['s4']
    _, theta, fc5 = ms.case_inputs(c)
    tensors, D, dims = _layout(c)
    th = torch.from_numpy(theta)
    with torch.no_grad():
        full = sm.forward_for_sensitivity(th, tensors, dims, fc5, 0)
        first2 = sm.forward_for_sensitivity(th, tensors, dims, fc5, 2)
        row1 = sm.forward_for_sensitivity(th, tensors, dims, fc5, 0, i=1)
    assert tuple(full.shape) == (3, 3) and tuple(first2.shape) == (2, 3) and tuple(row1.shape) == (1, 3)
    assert torch.all(full[:, 2] == 0)
    assert torch.allclose(full[:2], first2, rtol=1e-5) and torch.allclose(full[1:2], row1, rtol=1e-5)


class _FakeNet:
    def __init__(self, c):
        _, self.theta, self.fc5 = ms.case_inputs(c)
        tensors, D, dims = _layout(c)
        self.ctx = types.SimpleNamespace(tensors=tensors, dims=dims, dev=torch.device('cpu'))

    def parameter_vector(self):
        return torch.from_numpy(self.theta.copy())

    def sensitivity_forward(self, experiences, orig_bs):                  # as FCModel.sensitivity_forward
        return lambda th, i: sm.forward_for_sensitivity(th, self.ctx.tensors, self.ctx.dims, experiences['fc_feats'], orig_bs, i=i)


def test_sensitivity_file_protocol(tmp_path):
    """first call computes and writes sens_t{task}_p{parent}_{pid}.pt; the same (task, parent) again is a no-op; another worker loads the
    file instead of recomputing; a new task id recomputes (safe_mutations.py:34-91)"""
    nets = importlib.import_module('nes-img-captioning_b200.nets')
    c = ms.CASES['s1']
    net = _FakeNet(c)
    w = sm.Sensitivity(net, c['underflow'], nets.Mutation.SAFE_GRAD_SUM)
    assert w.get_sensitivity() is None
    assert w.calc_sensitivity(3, 1, {'fc_feats': net.fc5}, c['bs'], str(tmp_path)) is True
    assert os.listdir(tmp_path) == ['sens_t3_p1_%d.pt' % os.getpid()]
    assert np.allclose(w.get_sensitivity().numpy(), GOLD['s1_sens'], rtol=1e-5, atol=0)
    assert np.array_equal(torch.load(tmp_path / os.listdir(tmp_path)[0]).numpy(), w.get_sensitivity().numpy())
    assert w.calc_sensitivity(3, 1, {'fc_feats': net.fc5}, c['bs'], str(tmp_path)) is False
    w2 = sm.Sensitivity(net, c['underflow'], nets.Mutation.SAFE_GRAD_SUM)
    w2._calc_sensitivity = lambda *_: (_ for _ in ()).throw(AssertionError('must come from the file'))
    assert w2.calc_sensitivity(3, 1, None, c['bs'], str(tmp_path)) is True
    assert np.array_equal(w2.get_sensitivity().numpy(), w.get_sensitivity().numpy())
    # orig_batch_size sticks to the first call's value (:57-58): a later, larger batch_size still uses the first 4 images
    assert w.calc_sensitivity(4, 1, {'fc_feats': net.fc5}, 6, str(tmp_path)) is True
    assert len(os.listdir(tmp_path)) == 2 and w._orig_batch_size == c['bs']
    assert np.array_equal(w.get_sensitivity().numpy(), w2.get_sensitivity().numpy())
    # SM-VECTOR's set_sensitivity: clamp, divide by the minimum (:27-32)
    w.set_sensitivity(torch.tensor([0.00001, 0.02, 0.5]))
    assert np.allclose(w.get_sensitivity().numpy(), np.array([1.0, 0.02 / c['underflow'], 0.5 / c['underflow']], dtype=np.float32))


def test_abs_sensitivity_runs_and_bounds_sum_variant():
    """SM-G-ABS is unpinned (the reference's method raises); |J| averaged per sample can only be >= |sum of J| / batch, term by term"""
    c = dict(V=99, E=16, R=16, Fd=32, rows=2, bs=2, underflow=0.01, scale=2.0)
    _, theta, fc5 = ms.case_inputs(c)
    tensors, D, dims = _layout(c)
    a = sm.abs_sensitivity(torch.from_numpy(theta.copy()), tensors, dims, fc5, 2)
    s = sm.sum_sensitivity(torch.from_numpy(theta.copy()), tensors, dims, fc5, 2)
    assert a.shape == s.shape and torch.isfinite(a).all()
    assert torch.all(a >= s * (1 - 1e-3) - 1e-7)                 # float32 round-off of 2 x 1 forward passes
