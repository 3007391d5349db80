"""GPU parity: ranks (K6), gradient (K7), optimizers (K8) against the pinned oracle and the reference's goldens."""
import numpy as np
import pytest
import torch

from oracle import fc_caption, nes_math, noise

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def g1(golden_mod):
    import nes_b200
    c = golden_mod.CASES['g1']
    dims, G, play, D, D_pad, L, theta, fc, offs = golden_mod.case_inputs(c)
    ctx = nes_b200.NesContext(dict(vocab_size=c['V'], input_encoding_size=c['E'], rnn_size=c['R'], fc_feat_size=c['Fd']),
                              seed=golden_mod.TABLE_SEED, table_len=L, wave_members=0)
    yield ctx, dims, play, D, theta
    ctx.close()


def test_centered_ranks_known_answer_and_golden(g1, golden):
    ctx = g1[0]
    import nes_b200
    m = nes_b200.NESMaster(ctx)
    got = m.compute_centered_ranks(np.array([[101., 200.], [2., 100.]]))
    assert np.allclose(got, [[0.16666667, 0.5], [-0.5, -0.16666667]], atol=1e-8)       # nic_nes_master.py:189
    assert np.array_equal(got, golden['mm_kat_out'])
    assert np.array_equal(m.compute_centered_ranks(golden['mm_fitness']), golden['mm_centered_ranks'])


@pytest.mark.parametrize('n', [1, 2, 37, 1024, 1025, 10000])
def test_ranks_bit_exact_tie_free(g1, n):
    ctx = g1[0]
    x = np.random.RandomState(n).randn(n) * 10
    assert np.array_equal(ctx.compute_ranks(x).numpy(), nes_math.compute_ranks(x))


def test_ranks_ties_use_index_order(g1):
    ctx = g1[0]
    x = np.array([3., 1., 3., 1., 2.])
    assert list(ctx.compute_ranks(x).numpy()) == [3, 0, 4, 1, 2]
    x = np.round(np.random.RandomState(1).randn(4000) * 3)          # tie heavy
    assert np.array_equal(ctx.compute_ranks(x).numpy(), nes_math.compute_ranks(x, stable=True))


def test_gradient_matches_reference_golden(g1, golden):
    ctx, dims, play, D, theta = g1
    fit, offs = golden['mm_fitness'], golden['mm_offsets']
    cen, w = ctx.rank_weights(fit)
    assert np.array_equal(cen.numpy(), golden['mm_centered_ranks'])
    assert np.array_equal(w.numpy(), nes_math.rank_weights(fit).astype(np.float32))
    g = ctx.gradient(offs, w, 0.01).numpy()
    ref = golden['mm_grad']
    assert g.dtype == np.float32 and g.shape == ref.shape
    # float32 summation order differs from np.dot: 1e-5 of the gradient's scale
    assert np.max(np.abs(g - ref)) <= 1e-5 * np.max(np.abs(ref))


@pytest.mark.parametrize('F', [1, 3, 500, 501, 1203])
def test_gradient_against_oracle_various_F(g1, F):
    ctx, dims, play, D, theta = g1
    rs = np.random.RandomState(F)
    fit = rs.randn(F, 2)
    offs = ctx.sample_offsets(100 + F, F).numpy()
    nz = np.stack([noise.member_theta(theta, noise.noise_row(1234, o, play, D), 0.02, +1)[1] for o in offs])
    ref = nes_math.gradient_estimate(fit, nz)
    import nes_b200
    g = nes_b200.NESMaster(ctx).gradient_estimate(fit, offs, 0.02)
    assert np.max(np.abs(g - ref)) <= 1e-5 * np.max(np.abs(ref))
    # linearity: doubling sigma doubles g exactly (power of two)
    g2 = nes_b200.NESMaster(ctx).gradient_estimate(fit, offs, 0.04)
    assert np.array_equal(g2, 2 * g)


def test_gradient_empty_population_is_zero(g1):
    ctx = g1[0]
    g = ctx.gradient(np.zeros(0, dtype=np.int64), np.zeros(0, dtype=np.float32), 0.01).numpy()
    assert (g == 0).all()


@pytest.mark.parametrize('oname', ['adam', 'sgd'])
def test_optimizer_trajectory_matches_reference_golden(g1, golden, oname):
    import nes_b200
    ctx, dims, play, D, theta = g1
    g = golden['mm_grad']
    opt_o = {'adam': lambda: nes_math.Adam(theta.copy(), stepsize=0.001),
             'sgd': lambda: nes_math.SGD(theta.copy(), stepsize=0.02, momentum=0.9)}[oname]()
    opt_g = {'adam': lambda: nes_b200.Adam(ctx, theta.copy(), stepsize=0.001),
             'sgd': lambda: nes_b200.SGD(ctx, theta.copy(), stepsize=0.02, momentum=0.9)}[oname]()
    th_o = theta.copy()
    ratios = []
    for step in range(3):
        gi = (g * np.float32(1.0 + 0.5 * step)).astype(np.float32)
        r_o, th_o = nes_math.master_update(opt_o, th_o, gi, 1e-7)
        r_g, th_g = opt_g.step(gi, 1e-7, fetch=True)
        ratios.append(r_g)
        # float64 state, the reference's operation order: theta' agrees to the last bit
        assert th_g.dtype == np.float64 and np.array_equal(th_g, np.asarray(th_o, dtype=np.float64)), step
        assert abs(r_g - r_o) <= 1e-6 * abs(r_o)      # the reference takes |theta| of a float32 array at step 0
        if step == 0:
            assert np.array_equal(th_g[::16], golden['opt_%s_theta1_s16' % oname])
    assert np.allclose(ratios, golden['opt_%s_ratios' % oname], rtol=1e-6, atol=0)
    assert np.array_equal(th_g[::16], golden['opt_%s_theta3_s16' % oname])
    # the float32 evaluation copy is the rounded master copy (what the reference's workers load)
    assert np.array_equal(ctx.get_theta().numpy(), th_g.astype(np.float32))


def test_optimizer_update_signature_and_state_roundtrip(g1, tmp_path):
    import nes_b200
    ctx, dims, play, D, theta = g1
    rs = np.random.RandomState(3)
    G = rs.randn(D).astype(np.float32) * 0.1
    o_ref = nes_math.Adam(theta.copy(), stepsize=0.002)
    o_gpu = nes_b200.Adam(ctx, theta.copy(), stepsize=0.002)
    r1, t1 = o_ref.update(G)
    r2, t2 = o_gpu.update(G)                 # Optimizer.update(globalg) -> (ratio, theta')  optimizers.py:15-22
    assert np.array_equal(np.asarray(t1, dtype=np.float64), t2) and abs(r1 - r2) <= 1e-6 * r1
    p = str(tmp_path / 'optimizer.tar')
    o_gpu.save_to_file(p)
    st = torch.load(p, weights_only=False)
    assert set(st) == {'dim', 't', 'stepsize', 'beta1', 'beta2', 'epsilon', 'm', 'v'} and st['t'] == 1
    assert np.array_equal(st['m'], o_ref.m) and np.array_equal(st['v'], o_ref.v)
    o_gpu2 = nes_b200.Adam(ctx, None, stepsize=0.5)
    o_gpu2.load_from_file(p)
    r1, t1 = o_ref.update(G * 2)
    r2, t2 = o_gpu2.update(G * 2)
    assert np.array_equal(np.asarray(t1), t2)
    o_gpu2.stepsize = 0.001                  # annealing: stepsize /= stepsize_divisor (nic_nes_master.py:139-141)
    o_ref.stepsize = 0.001
    assert np.array_equal(np.asarray(o_ref.update(G)[1]), o_gpu2.update(G)[1])


def test_master_step_equals_the_three_calls(g1):
    import nes_b200
    ctx, dims, play, D, theta = g1
    F = 64
    fit = np.random.RandomState(9).randn(F, 2)
    offs = ctx.sample_offsets(5, F)
    nes_b200.Adam(ctx, theta.copy(), stepsize=0.001)
    ratio_a = ctx.master_step(fit, offs, 0.01, 1e-7)
    th_a = ctx.get_theta(dtype=torch.float64).numpy()
    opt = nes_b200.Adam(ctx, theta.copy(), stepsize=0.001)
    g = nes_b200.NESMaster(ctx).gradient_estimate(fit, offs, 0.01)
    ratio_b, th_b = opt.step(g, 1e-7, fetch=True)
    assert ratio_a == ratio_b and np.array_equal(th_a, th_b)


def test_flat_context_generic_parameter_vector():
    """master math on a generic flat theta (MnistNet has 8 490 parameters)"""
    import nes_b200
    ctx = nes_b200.NesContext(counts=[250, 10, 4500, 20, 3200, 10, 500], flat_granule=64, seed=5, table_len=1 << 20)
    assert ctx.D == 8490
    F = 50
    offs = ctx.sample_offsets(0, F).numpy()
    fit = np.random.RandomState(0).randn(F, 2)
    w = nes_math.rank_weights(fit).astype(np.float32)
    g = ctx.gradient(offs, w, 0.05).numpy()
    rows = []
    for o in offs:
        row = np.empty(ctx.D, dtype=np.float16)
        for t in ctx.tensors:
            row[t.start:t.start + t.count] = noise.table_slice(5, int(o) + t.pstart, t.count)
        rows.append(np.float32(0.05) * row.astype(np.float32))
    ref = nes_math.gradient_estimate(fit, np.stack(rows))
    assert np.max(np.abs(g - ref)) <= 1e-5 * np.max(np.abs(ref))
    ctx.close()


def test_gradient_with_proportional_scale():
    """K7 under SM-PROPORTIONAL: g = sum_i w_i (sigma eps_i (.) s) / (2F) (nic_nes_master.py:178-182 on the scaled noise vectors
    of algorithm/nets.py:104-111) -- the common scale is applied to the sum, within float32 round-off of the reference order"""
    import nes_b200
    from oracle import nes_math, noise
    D = 5000
    ctx = nes_b200.NesContext(counts=[D], seed=11, table_len=1 << 20)
    theta = (np.random.RandomState(0).randn(D) * 0.05).astype(np.float32)
    theta[::7] = 0.0
    ctx.set_theta(theta)
    ctx.set_mutation('SM-PROPORTIONAL')
    F = 37
    offs = ctx.sample_offsets(2, F).numpy()
    fit = np.random.RandomState(1).randn(F, 2)
    scale = noise.proportional_scale(theta)
    play, D_, D_pad = noise.padded_layout([('t0', (D,), 0, D)], 64)
    nz = np.stack([noise.member_theta(theta, noise.noise_row(11, o, play, D), 0.1, +1, scale=scale)[1] for o in offs])
    want = nes_math.gradient_estimate(fit, nz)
    got = nes_b200.NESMaster(ctx).gradient_estimate(fit, offs, 0.1)
    assert np.allclose(got, want, rtol=1e-5, atol=1e-9)
    ctx.close()


@pytest.mark.parametrize('n_shards', [2, 3, 8])
def test_gradient_shards_are_disjoint_and_sum_to_the_whole(g1, n_shards):
    """multi-GPU K7 (nes_gradient_shard): every element is produced by exactly one shard with the single-GPU arithmetic, zeros
    elsewhere, so the all-reduce over ranks equals nes_gradient bit for bit; sharded_master_step at world size 1 is master_step"""
    import nes_b200
    ctx, dims, play, D, theta = g1
    F = 97
    rs = np.random.RandomState(3)
    w = rs.randn(F).astype(np.float32)
    offs = ctx.sample_offsets(9, F)
    whole = ctx.gradient(offs, w, 0.01, out_device='cuda')
    parts = [ctx.gradient(offs, w, 0.01, out_device='cuda', shard=(r, n_shards)) for r in range(n_shards)]
    owners = sum((p != 0).to(torch.int32) for p in parts)
    assert int(owners.max()) <= 1
    assert torch.equal(sum(parts), whole)
    nz = [torch.nonzero(p).flatten() for p in parts]
    for a, b in zip(nz[:-1], nz[1:]):                           # contiguous slices in rank order
        if len(a) and len(b):
            assert int(a.max()) < int(b.min())
    fit = rs.randn(F, 2)
    nes_b200.Adam(ctx, theta.copy(), stepsize=0.001)
    r_a = ctx.master_step(fit, offs, 0.01, 1e-7)
    th_a = ctx.get_theta(dtype=torch.float64).numpy()
    nes_b200.Adam(ctx, theta.copy(), stepsize=0.001)
    r_b = nes_b200.sharded_master_step(ctx, torch.from_numpy(fit).cuda(), offs, 0.01, 1e-7, world_size=1)
    assert r_a == r_b and np.array_equal(th_a, ctx.get_theta(dtype=torch.float64).numpy())
    # the multi-rank code path on one GPU: weights left on the device, slice by slice, summed, then the update
    nes_b200.Adam(ctx, theta.copy(), stepsize=0.001)
    ctx.rank_weights(fit, want_centered=False, want_weights=False)
    g = sum(ctx.gradient(offs, None, 0.01, out_device='cuda', shard=(r, n_shards)) for r in range(n_shards))
    r_c = ctx.opt_update(g, 1e-7)
    assert r_a == r_c and np.array_equal(th_a, ctx.get_theta(dtype=torch.float64).numpy())


def test_gradient_with_sensitivity_divisor():
    """K7 under the sensitivity mutations: g = sum_i w_i (sigma eps_i / sens) / (2F); the common divisor is applied to the sum"""
    import nes_b200
    from oracle import nes_math, noise
    D = 5000
    ctx = nes_b200.NesContext(counts=[D], seed=11, table_len=1 << 20)
    theta = (np.random.RandomState(0).randn(D) * 0.05).astype(np.float32)
    ctx.set_theta(theta)
    ctx.set_mutation('SM-VECTOR')
    sens = noise.sensitivity_vector(np.abs(np.random.RandomState(2).randn(D)).astype(np.float32), 0.01)
    ctx.set_sensitivity(sens)
    F = 37
    offs = ctx.sample_offsets(2, F).numpy()
    fit = np.random.RandomState(1).randn(F, 2)
    play, D_, D_pad = noise.padded_layout([('t0', (D,), 0, D)], 64)
    nz = np.stack([noise.member_theta(theta, noise.noise_row(11, o, play, D), 0.1, +1, sens=sens)[1] for o in offs])
    want = nes_math.gradient_estimate(fit, nz)
    got = nes_b200.NESMaster(ctx).gradient_estimate(fit, offs, 0.1)
    assert np.allclose(got, want, rtol=1e-5, atol=1e-9)
    ctx.close()
