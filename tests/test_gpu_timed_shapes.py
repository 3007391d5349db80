"""Parity at the shapes the benchmarks time (parity gaps 1-2): tokens bit-exact and log-probs to 1e-4 against what the
REAL reference's FCModel produced (tests/golden/golden_timed_v1.npz, made by tests/golden/make_golden_timed.py):
  t2   BASELINE configs[1]: V 9487, E = R = 512, 128 rows, 16 members over 4 parents, decoded in two waves (wave_members 8)
  t4   BASELINE configs[3]: V 9487, E = R = 128, LayerNorm + affine, 512 rows (4 M-tiles per member)
  t5a / t5b  LayerNorm with 150 / 200 ragged rows."""
import os
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, 'golden'))
import make_golden_timed as mt  # noqa: E402

GOLD = np.load(os.path.join(HERE, 'golden', 'golden_timed_v1.npz'))


def _run(name, prec, wave, one_wave=False):
    import nes_b200
    c = mt.CASES[name]
    dims, G, play, D, D_pad, L, thetas, fc, offs = mt.case_inputs(c)
    assert np.allclose(thetas.astype(np.float64).sum(1), GOLD[name + '_theta_sum'], rtol=0, atol=1e-9)      # same inputs as the generator had
    assert np.array_equal(offs, GOLD[name + '_offsets'])
    mo = dict(vocab_size=c['V'], input_encoding_size=c['E'], rnn_size=c['R'], fc_feat_size=c['Fd'], layer_n=c.get('ln', False), layer_n_affine=c.get('affine', False))
    ctx = nes_b200.NesContext(mo, seed=mt.TABLE_SEED, table_len=L, n_parents=c['n_parents'], wave_members=wave, rows=c['B'])
    for p in range(c['n_parents']):
        ctx.set_theta(thetas[p], p)
    P = len(c['parents'])
    assert P > wave or name != 't2' or one_wave                          # t2 must take more than one wave
    seq, lp = ctx.rollout(offs, np.asarray(c['signs'], dtype=np.float32), fc, c['sigma'], parents=np.asarray(c['parents'], dtype=np.int32), precision=prec)
    want_seq, want_lp = GOLD[name + '_seq'].astype(np.int64), GOLD[name + '_logprobs']
    bad = (seq.numpy() != want_seq)
    assert not bad.any(), '%s: %d of %d tokens differ, first at %s' % (name, bad.sum(), bad.size, np.argwhere(bad)[0])
    assert np.allclose(lp.numpy(), want_lp, rtol=1e-4, atol=1e-5), np.abs(lp.numpy() - want_lp).max()
    assert len(np.unique(want_seq[:, :, 0])) > 1
    ctx.close()


@pytest.mark.parametrize('prec', [2, 1], ids=['hybrid', 'exact'])
def test_config2_shape_16_members_4_parents_two_waves(prec):
    _run('t2', prec, wave=8)


def test_config2_shape_few_units_split_the_vocabulary_sweep():
    """16 units in a workspace of 64: every unit's vocabulary sweep is cut into 4 ranges with their own candidate lists (decode.cu,
    the launch shape of the elite-candidate rollout of a NIC-ES generation)"""
    _run('t2', 2, wave=64, one_wave=True)


@pytest.mark.parametrize('prec', [2, 1], ids=['hybrid', 'exact'])
def test_config4_shape_layernorm_512_rows(prec):
    _run('t4', prec, wave=4)


@pytest.mark.parametrize('name', ['t5a', 't5b'])
def test_layernorm_ragged_rows(name):
    _run(name, 2, wave=3)
    _run(name, 1, wave=8)
