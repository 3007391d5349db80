"""GPU parity: noise table (K1), offsets / parents, theta copies, member materialisation -- through the C ABI."""
import numpy as np
import pytest
import torch

from oracle import fc_caption, noise

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def ctx_small():
    import nes_b200
    mo = dict(vocab_size=199, input_encoding_size=64, rnn_size=64, fc_feat_size=128)
    c = nes_b200.NesContext(mo, seed=1234, table_len=1 << 22, n_parents=3, wave_members=8, rows=8)
    yield c
    c.close()


def test_table_matches_philox_boxmuller_spec(ctx_small):
    c = ctx_small
    for start, count in [(0, 4096), (1001, 777), (c.table_len - 4096, 4096), (123456, 50000)]:
        got = c.table_slice(start, count)
        want = noise.table_slice(1234, start, count)
        assert got.dtype == np.float16
        assert np.array_equal(got.view(np.uint16), want.view(np.uint16)), (start, count)
    assert c.table_slice(5, 0).shape == (0,)


def test_full_table_statistics(ctx_small):
    t = ctx_small.table_slice(0, ctx_small.table_len).astype(np.float64)
    assert abs(t.mean()) < 2e-3 and abs(t.std() - 1.0) < 2e-3 and np.isfinite(t).all()


def test_offsets_and_parents_bit_exact(ctx_small):
    c = ctx_small
    for it, F in [(0, 1), (7, 1000), (1 << 33, 257), (3, 0)]:
        got = c.sample_offsets(it, F).numpy()
        want = noise.sample_offsets(1234, it, F, c.table_len, c.D_pad, c.G)
        assert np.array_equal(got, want)
        if F:
            assert np.array_equal(c.sample_parents(it, F, 10).numpy(), noise.sample_parents(1234, it, F, 10))
    o = c.sample_offsets(7, 1000).numpy()
    assert (o % c.G == 0).all() and (o + c.D_pad <= c.table_len).all()


def test_layout_agrees_with_oracle(ctx_small):
    c = ctx_small
    dims = fc_caption.Dims(199, 64, 64, 128)
    play, D, D_pad = noise.padded_layout(dims.layout(), noise.granularity(64, 64, 128))
    assert (c.D, c.D_pad) == (D, D_pad)
    assert [(t.name, t.start, t.count, t.pstart) for t in c.tensors] == [(n, s, cnt, ps) for n, _, s, cnt, ps in play]


def test_set_get_theta_roundtrip_and_length_assert(ctx_small):
    c = ctx_small
    rs = np.random.RandomState(0)
    th = rs.randn(c.D).astype(np.float32)
    c.set_theta(th, parent=1)
    assert np.array_equal(c.get_theta(1).numpy(), th)
    assert np.array_equal(c.get_theta(1, dtype=torch.float64).numpy(), th.astype(np.float64))
    th64 = rs.randn(c.D)
    c.set_theta(th64, parent=2)
    assert np.array_equal(c.get_theta(2, dtype=torch.float64).numpy(), th64)
    assert np.array_equal(c.get_theta(2).numpy(), th64.astype(np.float32))
    with pytest.raises(AssertionError):
        c.set_theta(th[:-1])
    # device-resident source
    c.set_theta(torch.from_numpy(th).cuda(), parent=0)
    assert np.array_equal(c.get_theta(0).numpy(), th)


def test_evolve_matches_reference_member_formula(ctx_small):
    c = ctx_small
    dims = fc_caption.Dims(199, 64, 64, 128)
    play, D, D_pad = noise.padded_layout(dims.layout(), c.G)
    theta = fc_caption.xavier_theta(dims, 11)
    c.set_theta(theta, parent=0)
    off = int(c.sample_offsets(3, 5)[2])
    eps = noise.noise_row(1234, off, play, D)
    for sign in (+1.0, -1.0):
        member, nz = c.evolve(off, 0.01, sign)
        want_m, want_n = noise.member_theta(theta, eps, 0.01, sign)
        assert np.array_equal(nz.numpy(), want_n)
        assert np.array_equal(member.numpy(), want_m)
    # survivor materialisation into another parent slot
    c.evolve(off, 0.01, 1.0, parent=0, dst_parent=2, want_member=False, want_noise=False)
    assert np.array_equal(c.get_theta(2).numpy(), noise.member_theta(theta, eps, 0.01, +1)[0])
    import nes_b200
    with pytest.raises(nes_b200.NesError):
        c.evolve(off + 1, 0.01)          # not slot aligned
