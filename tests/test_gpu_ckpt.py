"""Checkpoint compatibility: files written by the REAL reference (tests/golden/ckpt_v1, made by
tests/golden/make_golden_ckpt.py with /root/reference imported) load into this package, resume to the reference's next
step bit for bit, and what this package writes back has the keys, order, dtypes and shapes of the reference's own files."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
CK = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'ckpt_v1')
DIMS = dict(vocab_size=49, input_encoding_size=64, rnn_size=64, fc_feat_size=64)


def test_reference_state_dict_file_loads_and_round_trips(tmp_path):
    """PolicyNet.from_serialized / serialize (src/algorithm/nets.py:153-161) on a .pth the reference wrote"""
    import nes_b200
    exp = np.load(os.path.join(CK, 'expect.npz'))
    ref_sd = torch.load(os.path.join(CK, 'model.pth'), map_location='cpu')
    model = nes_b200.FCModel(options=DIMS, from_param_file=os.path.join(CK, 'model.pth'), table_len=1 << 20, wave_members=2, rows=4)
    th = model.parameter_vector().numpy()
    assert th.size == int(exp['model_D']) == model.nb_learnable_params
    assert np.array_equal(th[::16], exp['model_theta_s16'])
    assert np.float64(th.astype(np.float64).sum()) == exp['model_theta_sum'] and np.float64(np.abs(th.astype(np.float64)).sum()) == exp['model_theta_abs']
    out = model.serialize(path=str(tmp_path / 'out.pth'))
    mine = torch.load(out, map_location='cpu')
    assert list(mine.keys()) == list(ref_sd.keys())                       # same names, same (registration) order
    for k in ref_sd:
        assert mine[k].dtype == ref_sd[k].dtype and torch.equal(mine[k], ref_sd[k]), k
    # the loaded weights decode: same tokens as the oracle on that theta
    from oracle import fc_caption
    dims = fc_caption.Dims(49, 64, 64, 64)
    fc = np.random.RandomState(0).randn(4, 64).astype(np.float32)
    seq, lp = model(torch.from_numpy(fc), greedy=True)
    seq_o, lp_o = fc_caption.decode_greedy(dims, th, fc)
    assert np.array_equal(seq.numpy(), seq_o) and np.allclose(lp.numpy(), lp_o, atol=1e-4)
    model.ctx.close()


@pytest.mark.parametrize('oname', ['adam', 'sgd'])
def test_reference_optimizer_file_resumes_to_the_reference_next_step(oname, tmp_path):
    """Optimizer.load_from_file / save_to_file (src/algorithm/nic_nes/optimizers.py:49-70,85-107): resume from the reference's file
    after its 2nd update, take the 3rd update, compare with the theta the reference reached; the file written back has the
    reference's keys and float64 state"""
    import nes_b200
    exp = np.load(os.path.join(CK, 'expect.npz'))
    ctx = nes_b200.NesContext(counts=[1000], seed=1, table_len=1 << 16)
    ctx.set_theta(exp[oname + '_theta2'])                                    # float64 theta, as the master holds it after an update
    opt = (nes_b200.Adam if oname == 'adam' else nes_b200.SGD)(ctx, None, stepsize=123.0)
    path = os.path.join(CK, oname + '.tar')
    opt.load_from_file(path)
    ref_state = torch.load(path, map_location='cpu', weights_only=False)
    assert opt.t == ref_state['t'] == 2 and opt.stepsize == ref_state['stepsize']
    th2 = exp[oname + '_theta2']
    _, th3 = opt.update(-exp['g2'] + 1e-7 * th2)                              # nic_nes_master.py:123-133
    assert np.array_equal(th3, exp[oname + '_theta3'])
    out = str(tmp_path / 'optimizer.tar')
    opt.save_to_file(out)
    mine = torch.load(out, map_location='cpu', weights_only=False)
    assert set(mine) == set(ref_state) and mine['t'] == 3 and mine['dim'] == ref_state['dim']
    for k in ('m', 'v'):
        if k in ref_state:
            assert np.asarray(mine[k]).dtype == np.float64 and np.asarray(mine[k]).shape == np.asarray(ref_state[k]).shape
    ctx.close()
