"""Size-independent properties at BASELINE.json's full sizes (configs[1] / configs[2] shapes: V 9487, E = R = 512, 128 rows, D = 13.4 M,
F = 5000), where the oracle would need minutes per case:
  * a population decoded with sigma = 0 must give, for EVERY member of every wave, the tokens the REAL reference produced for the
    unperturbed theta (tests/golden/golden_v1.npz, case g4) -- whatever its offset, sign and place in the launch;
  * a member's result does not depend on which other members are decoded with it;
  * K7 (nes_gradient): invariant under a permutation of the rows (bit for bit), exact under scaling the weights by 2, additive,
    equal to the sum of its parameter-axis shards (bit for bit), and equal to a float64 sum over table slices at spot-checked elements."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MO = dict(vocab_size=9487, input_encoding_size=512, rnn_size=512, fc_feat_size=2048)


def test_full_population_with_sigma_zero_reproduces_the_reference_decode(golden, golden_mod):
    import nes_b200
    c = golden_mod.CASES['g4']
    dims, G, play, D, D_pad, L, theta, fc, offs = golden_mod.case_inputs(c)
    P, wave = 600, 256                                    # three waves, the last one ragged (88 members)
    ctx = nes_b200.NesContext(MO, seed=golden_mod.TABLE_SEED, table_len=1 << 27, n_parents=3, wave_members=wave, rows=c['B'])
    for p in range(3):
        ctx.set_theta(theta, p)
    rs = np.random.RandomState(3)
    o = ctx.sample_offsets(11, P)
    signs = torch.from_numpy(np.where(rs.rand(P) < 0.5, 1.0, -1.0).astype(np.float32))
    parents = torch.from_numpy(np.sort(rs.randint(0, 3, P)).astype(np.int32))
    want_seq = golden['g4_seq_theta'].astype(np.int64)
    want_lp = golden['g4_logprobs_theta']
    for prec in (nes_b200.PREC_HYBRID, nes_b200.PREC_EXACT):
        seq, lp = ctx.rollout(o, signs, fc, 0.0, parents=parents, precision=prec)
        bad = (seq.numpy() != want_seq[None]).any(axis=(1, 2))
        assert not bad.any(), 'precision %d: members %s differ from the reference decode of theta' % (prec, np.nonzero(bad)[0][:10])
        assert np.allclose(lp.numpy(), np.broadcast_to(want_lp[None], lp.shape), rtol=1e-4, atol=1e-5)
        assert torch.equal(lp, lp[:1].expand_as(lp))      # and the members agree among themselves to the bit
    ctx.close()


def test_full_size_members_do_not_depend_on_their_neighbours():
    import nes_b200
    P, mu, rows = 400, 8, 128
    ctx = nes_b200.NesContext(MO, seed=1234, table_len=1 << 27, n_parents=mu, wave_members=256, rows=rows)
    nes_b200.FCModel(ctx, parent=0, rng_state=0)
    for k in range(1, mu):
        ctx.evolve(int(ctx.sample_offsets(999, mu)[k]), 0.01, 1.0, parent=0, dst_parent=k, want_member=False, want_noise=False)
    fc = torch.randn(rows, 2048, generator=torch.Generator().manual_seed(0))
    offs = ctx.sample_offsets(1, P)
    parents, _ = torch.sort(ctx.sample_parents(1, P, mu))
    signs = torch.ones(P)
    seq, lp = ctx.rollout(offs, signs, fc, 0.005, parents=parents, precision=nes_b200.PREC_HYBRID)
    for sel in (slice(37, 53), slice(0, P, 9), slice(250, 400)):        # inside one wave, strided over both, the tail across the wave cut
        s2, l2 = ctx.rollout(offs[sel].contiguous(), signs[sel].contiguous(), fc, 0.005, parents=parents[sel].contiguous(), precision=nes_b200.PREC_HYBRID)
        assert torch.equal(s2, seq[sel]), 'tokens of a member changed with its neighbours'
        assert torch.allclose(l2, lp[sel], rtol=1e-6, atol=1e-7)
    assert len(torch.unique(seq[:, :, 0])) > 1
    ctx.close()


def test_k7_full_size_properties():
    import nes_b200
    ctx = nes_b200.NesContext(MO, seed=77, table_len=1 << 30, n_parents=1, wave_members=2, rows=8)
    D, F, sigma = ctx.D, 5000, 0.01
    assert D == 13400848
    offs = torch.unique(ctx.sample_offsets(3, F))                       # unique rows: the table order is then a strict order
    F = offs.numel()
    assert F > 4900
    g = torch.Generator().manual_seed(5)
    offs = offs[torch.randperm(F, generator=g)]
    w1 = torch.randn(F, generator=g)
    w2 = torch.randn(F, generator=g)
    dev = ctx.dev
    g1 = ctx.gradient(offs, w1, sigma, out_device=dev)
    assert torch.isfinite(g1).all() and float(g1.abs().max()) > 0
    # rows are summed in table order whatever order they are given in
    perm = torch.randperm(F, generator=g)
    assert torch.equal(ctx.gradient(offs[perm], w1[perm], sigma, out_device=dev), g1)
    # scaling every weight by 2 scales every product and every partial sum exactly
    assert torch.equal(ctx.gradient(offs, 2.0 * w1, sigma, out_device=dev), 2.0 * g1)
    # additivity to float32 round-off of a 5000-term sum
    g2 = ctx.gradient(offs, w2, sigma, out_device=dev)
    g12 = ctx.gradient(offs, w1 + w2, sigma, out_device=dev)
    scale = float(g12.abs().max())
    assert float((g12 - (g1 + g2)).abs().max()) <= 1e-5 * scale
    # parameter-axis shards (multi-GPU K7): disjoint supports that add up to the whole, bit for bit
    total = torch.zeros_like(g1)
    support = torch.zeros(D, dtype=torch.int32, device=dev)
    for r in range(8):
        part = ctx.gradient(offs, w1, sigma, out_device=dev, shard=(r, 8))
        support += (part != 0).to(torch.int32)
        total += part
    assert int(support.max()) == 1 and torch.equal(total, g1)
    # spot checks against a float64 sum over table slices: first elements, a run inside logit.weight, the last elements of theta
    tens = {t.name: t for t in ctx.tensors}
    last = ctx.tensors[-1]
    runs = [(ctx.tensors[0], 0), (tens['logit.weight'], 1234567), (last, last.count - 64)]
    w64 = w1.double().numpy()
    o_np = offs.numpy()
    for t, j in runs:
        acc = np.zeros(64)
        for i in range(F):
            acc += w64[i] * ctx.table_slice(int(o_np[i]) + t.pstart + j, 64).astype(np.float64)
        want = np.float32(sigma) * acc / (2 * F)
        got = g1[t.start + j: t.start + j + 64].cpu().numpy()
        assert np.allclose(got, want, rtol=2e-5, atol=1e-5 * np.abs(want).max()), (t.name, np.abs(got - want).max())
    ctx.close()


@pytest.mark.parametrize('D', [350000, 2865808])
def test_k7_every_launch_shape_gives_the_same_bits(D, monkeypatch):
    """the sum over the rows runs in table order per element whatever the launch shape (vectors per thread x rows in flight, bands):
    every shape the library can pick (NES_K7 forces one) must reproduce the default's bits, and the rows as sampled (NES_K7_SORT=0,
    another sum order) must agree to float32 round-off"""
    import nes_b200
    ctx = nes_b200.NesContext(counts=[D - 1000, 1000], seed=9, table_len=1 << 26)
    F = 700
    g = torch.Generator().manual_seed(D)
    offs = torch.unique(ctx.sample_offsets(2, F))
    offs = offs[torch.randperm(offs.numel(), generator=g)]
    w = torch.randn(offs.numel(), generator=g)
    monkeypatch.delenv('NES_K7', raising=False)
    monkeypatch.delenv('NES_K7_SORT', raising=False)
    want = ctx.gradient(offs, w, 0.02, out_device=ctx.dev)
    assert float(want.abs().max()) > 0
    for shape in ('1x16', '1x8', '2x8', '2x4', '3x4', '4x4', '4x2', '5x2', '6x2', '7x1', '8x2', '8x1'):
        monkeypatch.setenv('NES_K7', shape)
        assert torch.equal(ctx.gradient(offs, w, 0.02, out_device=ctx.dev), want), shape
        parts = sum(ctx.gradient(offs, w, 0.02, out_device=ctx.dev, shard=(r, 3)) for r in range(3))
        assert torch.equal(parts, want), shape
    monkeypatch.delenv('NES_K7')
    monkeypatch.setenv('NES_K7_SORT', '0')
    got = ctx.gradient(offs, w, 0.02, out_device=ctx.dev)
    assert float((got - want).abs().max()) <= 1e-5 * float(want.abs().max())
    ctx.close()
