"""NIC-ES on the device: parent selection (uniform / tournament), the `parent is None` first generation, truncation
selection + survivors without a host round trip, the podium-fed elites -- ESPopulation on a real NesContext.  The bookkeeping itself is
pinned against the real reference classes in tests/test_host_es.py; here the PARAMETERS in the slots are checked against the oracle."""
import json
import os
import sys

import numpy as np
import pytest
import torch

from oracle import fc_caption, mnist, noise

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)


def _small_fc(n_parents, wave=8, B=6, seed=21):
    import nes_b200
    V, E, R, Fd = 199, 64, 64, 128
    dims = fc_caption.Dims(V, E, R, Fd)
    G = noise.granularity(E, R, Fd)
    play, D, D_pad = noise.padded_layout(dims.layout(), G)
    L = D_pad + 255 * G
    ctx = nes_b200.NesContext(dict(vocab_size=V, input_encoding_size=E, rnn_size=R, fc_feat_size=Fd), seed=seed, table_len=L,
                              n_parents=n_parents, wave_members=wave, rows=B)
    fc = noise.table_slice(6, 0, B * Fd).astype(np.float32).reshape(B, Fd)
    return ctx, dims, play, D, fc


def test_tournament_sampler_equals_spec():
    ctx, *_ = _small_fc(1)
    for it, F, n, t in [(0, 1000, 50, 5), (7, 333, 10, 3), (2, 64, 3, 8), (5, 200, 4096, 17), (9, 100, 1, 1)]:
        got = ctx.sample_parents_tournament(it, F, n, t).numpy()
        assert np.array_equal(got, noise.sample_parents_tournament(21, it, F, n, t)), (it, F, n, t)
        assert np.array_equal(ctx.sample_parents_tournament(it, F, n, t, device=True).cpu().numpy(), got)
    ctx.close()


def test_select_evolve_is_the_references_stable_descending_sort():
    """nes_es_select_evolve against sorted(key=fitness, reverse=True) (nic_es_master.py:160) with ties and a NaN, host and device inputs;
    the survivors' float32 parameters bit-identical to theta[parent] + sigma * eps (nic_es_worker.py:175,188)"""
    ctx, dims, play, D, fc = _small_fc(12)
    thetas = [fc_caption.xavier_theta(dims, 40 + k) for k in range(4)]
    for k in range(4):
        ctx.set_theta(thetas[k], k)
    n, n_keep, sigma = 37, 6, 0.05
    rs = np.random.RandomState(3)
    fit = np.round(rs.randn(n) * 2)                       # many ties
    fit[5] = np.nan
    offs = ctx.sample_offsets(3, n).numpy()
    pars = rs.randint(0, 4, n).astype(np.int32)
    want = sorted([i for i in range(n) if i != 5], key=lambda i: fit[i], reverse=True)[:n_keep]
    for dev in (False, True):
        f, o, p = (torch.from_numpy(x).cuda() if dev else torch.from_numpy(x) for x in (fit, offs, pars))
        order = ctx.es_select_evolve(f, o, p, n_keep, sigma, 6)
        assert order.cpu().tolist() == want
        for r, i in enumerate(want):
            th, _ = noise.member_theta(thetas[pars[i]], noise.noise_row(21, offs[i], play, D), sigma, +1)
            assert np.array_equal(ctx.get_theta(6 + r).numpy(), th), (dev, r)
    # a rollout from the new slots sees the new parameters (split-fp16 shadow and HYBRID row bound refreshed)
    import nes_b200
    seq, lp = ctx.rollout(np.zeros(2, dtype=np.int64), np.ones(2, dtype=np.float32), fc, 0.0, parents=np.array([6, 8], dtype=np.int32), precision=nes_b200.PREC_HYBRID)
    for j, r in enumerate((0, 2)):
        th, _ = noise.member_theta(thetas[pars[want[r]]], noise.noise_row(21, offs[want[r]], play, D), sigma, +1)
        s_, l_ = fc_caption.decode_greedy(dims, th, fc)
        assert np.array_equal(seq[j].numpy(), s_) and np.allclose(lp[j].numpy(), l_, rtol=1e-4, atol=1e-5)
    with pytest.raises(nes_b200.NesError):
        ctx.es_select_evolve(torch.from_numpy(fit), torch.from_numpy(offs), torch.from_numpy(pars), 6, sigma, 9)     # 9 + 6 > 12 slots
    ctx.close()


def test_copy_parent_copies_every_derived_buffer():
    import nes_b200
    ctx, dims, play, D, fc = _small_fc(3)
    th = fc_caption.xavier_theta(dims, 50)
    ctx.set_theta(th, 0)
    ctx.copy_parent(0, 2)
    assert np.array_equal(ctx.get_theta(2).numpy(), th) and np.array_equal(ctx.get_theta(2, dtype=torch.float64).numpy(), th.astype(np.float64))
    off = ctx.sample_offsets(1, 2).numpy()
    for prec in (nes_b200.PREC_HYBRID, nes_b200.PREC_EXACT):
        a = ctx.rollout(off, np.ones(2, dtype=np.float32), fc, 0.05, parents=np.zeros(2, dtype=np.int32), precision=prec)
        b = ctx.rollout(off, np.ones(2, dtype=np.float32), fc, 0.05, parents=np.full(2, 2, dtype=np.int32), precision=prec)
        assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    ctx.close()


def _check_generation(pop, ctx, gen_inputs, member_of, prev_thetas):
    """after end_generation: every new (non-podium) parent slot holds member_of(parent theta, offset) of the offspring the stable sort ranks
    there; returns {slot: theta} of all live slots"""
    fit, offs, slots = gen_inputs
    fit, offs, slots = fit.cpu().numpy(), offs.cpu().numpy(), slots.cpu().numpy()
    order = sorted(range(len(fit)), key=lambda i: fit[i], reverse=True)[:pop.mu - pop.E]
    n_pod = len(pop.podium())
    new_slots = pop.parent_slots()[n_pod:]
    assert len(new_slots) == len(order)
    for r, i in enumerate(order):
        want = member_of(prev_thetas[int(slots[i])], int(offs[i]))
        got = ctx.get_theta(new_slots[r]).numpy()
        assert np.array_equal(got, want), (r, i, np.abs(got - want).max())
    return order


def test_es_population_generations_on_the_device():
    """three generations from init_from_zero: generation 0's offspring are fresh random models (blank + Xavier std * eps) whose fitness
    equals the oracle decode of exactly that model; survivors, elite candidates and podium hold the parameters the bookkeeping says"""
    import nes_b200
    exp = {'nb_offspring': 14, 'population_size': 5, 'num_elites': 2, 'num_elite_cands': 2, 'selection': 'tournament', 'tournament_size': 3,
           'config': {'noise_stdev': 0.05, 'batch_size': 6, 'patience': 0}}
    ctx, dims, play, D, fc = _small_fc(nes_b200.ESPopulation.slots_needed(exp), wave=16)
    fc_dev = torch.from_numpy(fc).cuda()
    pop = nes_b200.ESPopulation(ctx, exp)
    pop.init_from_zero()
    blank, std = nes_b200.layout.fresh_model_vectors(ctx.tensors)

    def snapshot():
        live = set(pop.parent_slots() or []) | set(pop.elite_candidate_slots()) | {s for s, _ in pop.podium()} | {pop.blank}
        return {s: ctx.get_theta(s).numpy() for s in live}

    # the initial elite candidates are fresh models too
    for s in pop.elite_candidate_slots():
        th = ctx.get_theta(s).numpy()
        bias = [t for t in ctx.tensors if t.name.endswith('bias')][0]
        assert np.all(th[bias.start:bias.start + bias.count] == 0) and abs(th[:1000].std() / std[0] - 1) < 0.2
    scores = iter([(0.3, 0.5), (0.9, 0.1), (0.2, 0.95)])
    for gen in range(3):
        before = snapshot()
        cand_before = [before[s] for s in pop.elite_candidate_slots()]
        fit, offs, slots = nes_b200.es_generation(pop, fc_dev, elite_score_fn=lambda cand: next(scores))
        if gen == 0:
            assert set(slots.cpu().tolist()) == {pop.blank}
            member_of = lambda th, off: noise.member_theta(th, noise.noise_row(21, off, play, D), 1.0, +1, scale=std)[0]
        else:
            assert set(slots.cpu().tolist()) <= set(before) and ctx.mutation == ''
            member_of = lambda th, off: noise.member_theta(th, noise.noise_row(21, off, play, D), 0.05, +1)[0]
        # every offspring's fitness is the oracle's greedy log-prob sum of exactly that member
        f = fit.cpu().numpy()
        for i in (0, 5, 13):
            th = member_of(before[int(slots[i])], int(offs[i]))
            _, lp = fc_caption.decode_greedy(dims, th, fc)
            assert abs(f[i] - lp.astype(np.float64).sum()) <= 1e-4 * abs(f[i]), (gen, i)
        order = _check_generation(pop, ctx, (fit, offs, slots), member_of, before)
        # the next elite candidates are copies of the two best survivors; the podium holds what the candidates' scores put there
        n_pod = len(pop.podium())
        for j, s in enumerate(pop.elite_candidate_slots()):
            assert np.array_equal(ctx.get_theta(s).numpy(), ctx.get_theta(pop.parent_slots()[n_pod + j]).numpy())
        assert pop.tags()['parents'][n_pod:] == ['g%d:o%d' % (gen, i) for i in order]
        if gen == 0:            # podium = the two initial candidates, better one first
            assert [sc for _, sc in pop.podium()] == [0.5, 0.3]
            assert np.array_equal(ctx.get_theta(pop.podium()[0][0]).numpy(), cand_before[1]) and np.array_equal(ctx.get_theta(pop.podium()[1][0]).numpy(), cand_before[0])
        if gen == 1:            # 0.9 enters on top: [cand0 0.9, (old top) 0.5] -- and the reference's sequential copy leaves cand0's
            assert [sc for _, sc in pop.podium()] == [0.9, 0.5]           # parameters in BOTH slots (tools/podium.py:43-62)
            assert np.array_equal(ctx.get_theta(pop.podium()[0][0]).numpy(), cand_before[0]) and np.array_equal(ctx.get_theta(pop.podium()[1][0]).numpy(), cand_before[0])
    ctx.close()


def test_es_generation_from_given_parents_uniform_selection():
    """init_from_singles semantics + uniform parent pick: two generations, survivors bit-identical to the oracle's members"""
    import nes_b200
    exp = {'nb_offspring': 14, 'population_size': 4, 'num_elites': 1, 'num_elite_cands': 1, 'selection': 'uniform',
           'config': {'noise_stdev': 0.05, 'batch_size': 6, 'patience': 0}}
    ctx, dims, play, D, fc = _small_fc(nes_b200.ESPopulation.slots_needed(exp), wave=16)
    pop = nes_b200.ESPopulation(ctx, exp)
    thetas = [fc_caption.xavier_theta(dims, 70 + k) for k in range(4)]
    pop.init_from_thetas(thetas)
    assert [np.array_equal(ctx.get_theta(s).numpy(), thetas[k]) for k, s in enumerate(pop.parent_slots())] == [True] * 4
    member_of = lambda th, off: noise.member_theta(th, noise.noise_row(21, off, play, D), 0.05, +1)[0]
    fc_dev = torch.from_numpy(fc).cuda()
    for gen in range(2):
        before = {s: ctx.get_theta(s).numpy() for s in pop.parent_slots()}
        fit, offs, slots = nes_b200.es_generation(pop, fc_dev)          # elite score = fitness of the unperturbed candidate on the batch
        _check_generation(pop, ctx, (fit, offs, slots), member_of, before)
        assert len(pop.parent_slots()) == len(pop.podium()) + 3
    ctx.close()


def test_mnist_es_generations_and_learning():
    """experiments/mnist_es.json's loop (nb_offspring 50, population_size 10, num_elites 3, num_elite_cands 2; plain mutation at 0.01 here)
    through es_generation_mnist: two generations against the oracle (same survivors, next parents bit-identical), then 25 more on the
    fixed batch must raise the best fitness (the reference's -CrossEntropyLoss) and the podium must have changed hands"""
    import nes_b200
    sys.path.insert(0, os.path.join(HERE, 'golden'))
    import make_golden_mnist as mg
    play, D, D_pad, L0, theta, _, _, _ = mg.case_inputs()
    exp = {'nb_offspring': 50, 'population_size': 10, 'num_elites': 3, 'num_elite_cands': 2, 'selection': 'uniform',
           'config': {'noise_stdev': 0.01, 'batch_size': 64, 'patience': 0}}
    L = D_pad + 4095 * mg.G
    ctx = nes_b200.NesContext(tensors=nes_b200.layout.mnist_tensors(), flat_granule=mg.G, seed=99, table_len=L, n_parents=nes_b200.ESPopulation.slots_needed(exp))
    pop = nes_b200.ESPopulation(ctx, exp)
    pop.init_from_thetas([mnist.seeded_theta(100 + k) for k in range(10)])
    rs = np.random.RandomState(8)
    labels = rs.randint(0, 10, 64).astype(np.int64)
    protos = rs.randn(10, 784).astype(np.float32)
    images = (protos[labels] + 0.3 * rs.randn(64, 784).astype(np.float32)).reshape(64, 1, 28, 28)      # a learnable synthetic task
    member_of = lambda th, off: noise.member_theta(th, noise.noise_row(99, off, play, D), 0.01, +1)[0]
    for gen in range(2):
        before = {s: ctx.get_theta(s).numpy() for s in pop.parent_slots()}
        fit, offs, slots = nes_b200.es_generation_mnist(pop, images, labels)
        f = fit.cpu().numpy()
        for i in (0, 17, 49):
            want = mnist.rollout(member_of(before[int(slots[i])], int(offs[i])), images, labels)[0]
            assert abs(f[i] - want) <= 1e-5 * abs(want)
        _check_generation(pop, ctx, (fit, offs, slots), member_of, before)
    first_best = float(fit.max())
    podium0 = [ctx.get_theta(s).numpy().copy() for s, _ in pop.podium()]
    for gen in range(2, 27):
        fit, offs, slots = nes_b200.es_generation_mnist(pop, images, labels)
    assert float(fit.max()) > first_best + 0.15, (first_best, float(fit.max()))
    # the elites are the podium's best validated candidates, not the first slots carried over for ever
    assert len(pop.podium()) == 3 and not any(np.array_equal(ctx.get_theta(s).numpy(), p) for (s, _), p in zip(pop.podium(), podium0))
    assert pop.podium()[0][1] >= pop.podium()[1][1] >= pop.podium()[2][1]
    ctx.close()


def test_stale_sensitivity_is_refused_after_a_parent_changed():
    """under a sensitivity mutation a survivor written into a slot must not inherit the divisors of the slot's previous
    tenant (the reference recomputes them per (task, parent), nic_es_worker.py:172-173)"""
    import nes_b200
    ctx, dims, play, D, fc = _small_fc(8)
    ctx.set_mutation('SM-VECTOR')
    th = fc_caption.xavier_theta(dims, 60)
    ctx.set_theta(th, 0)
    sens = noise.sensitivity_vector(np.abs(noise.table_slice(5, 0, D).astype(np.float32)) + 0.1, 0.2)
    ctx.set_sensitivity(sens)
    off = ctx.sample_offsets(0, 3).numpy()
    ctx.evolve_batch([0], [int(off[0])], [0.05], [2])
    with pytest.raises(nes_b200.NesError, match='stale sensitivity'):
        ctx.evolve_batch([2], [int(off[1])], [0.05], [3])
    member, nz = ctx.evolve(int(off[1]), 0.05, 1.0, parent=2)
    assert np.all(np.isnan(nz.numpy()))                                   # the poisoned divisors make any use loud
    ctx.set_sensitivity(sens, parent=2)
    ctx.evolve_batch([2], [int(off[1])], [0.05], [3])
    want = noise.member_theta(ctx.get_theta(2).numpy(), noise.noise_row(21, off[1], play, D), 0.05, +1, sens=sens)[0]
    assert np.array_equal(ctx.get_theta(3).numpy(), want)
    # ESPopulation refuses to start a generation under a sensitivity mutation without a way to refresh the new parents' vectors
    exp = {'nb_offspring': 4, 'population_size': 2, 'num_elites': 1, 'num_elite_cands': 1, 'config': {'noise_stdev': 0.05, 'batch_size': 6}}
    pop = nes_b200.ESPopulation(ctx, exp)
    pop.init_from_thetas([th, th])
    with pytest.raises(AssertionError, match='sensitivity_fn'):
        pop.begin_generation()
    seen = []
    pop.sensitivity_fn = lambda slot: (seen.append(slot), ctx.set_sensitivity(sens, parent=slot))
    pop.begin_generation()
    assert sorted(seen) == sorted(set(pop.parent_slots()))
    ctx.close()
