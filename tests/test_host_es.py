"""NIC-ES master bookkeeping on device slots (nes_b200.ESPopulation) against the REAL reference's ESMaster.selection + ESIteration +
Podium (tests/golden/golden_es_v1.json, made by tests/golden/make_golden_es.py from /root/reference): same scripted fitnesses and
elite scores (tests/es_scenario.py), and after every generation the same models -- by where their content came from -- must be the
parents (in order), the elite candidates and the podium elites, incl. the podium's sequential-copy aliasing (tools/podium.py:43-62).

The context here is a stub that mimics what the device calls do to the CONTENT of the parent slots (copy_parent, es_select_evolve,
evolve into a slot); tests/test_gpu_es.py repeats the run on the B200 and checks the slots' actual parameters."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import es_scenario as sc  # noqa: E402


class StubContext:
    """content[slot] = where the slot's parameters came from"""
    MUTATIONS = {'': 0, 'SM-PROPORTIONAL': 1, 'SM-VECTOR': 2, 'SM-G-SUM': 2, 'SM-G-ABS': 2, 'SCALE': 3}

    def __init__(self, n_parents):
        import nes_b200
        self.n_parents, self.dev, self.mutation = n_parents, torch.device('cpu'), ''
        self.tensors, self.D, _ = nes_b200.layout.build_layout(nes_b200.layout.fc_tensors(19, 64, 64, 64), 64)
        self.content = {}
        self.lineage = {}                  # slot -> content of the parent slot at the moment the survivor was derived from it
        self.gen = -1
        self.n_fresh = 0
        self.modes = []

    def set_mutation(self, kind): self.mutation = kind; self.modes.append(kind)
    def set_theta(self, vec, parent=0): self.content[parent] = 'blank' if np.count_nonzero(np.asarray(vec)) == 0 else 'theta'
    def set_sensitivity(self, sens, parent=None, underflow=None): pass
    def sample_offsets(self, iteration, F, device=False): return torch.arange(F, dtype=torch.int64) * 64

    def sample_parents(self, iteration, F, n_parents, device=False):
        return torch.from_numpy(np.random.RandomState(iteration).randint(0, n_parents, F).astype(np.int32))

    def evolve(self, offset, sigma, sign=1.0, parent=0, want_member=True, want_noise=True, dst_parent=-1, device='cpu'):
        assert self.mutation == 'SCALE' and self.content[parent] == 'blank' and sigma == 1.0
        self.content[dst_parent] = 'init_elite:%d' % self.n_fresh
        self.n_fresh += 1
        return None, None

    def copy_parent(self, src, dst): self.content[dst] = self.content[src]

    def es_select_evolve(self, fitness, offsets, parents, n_keep, sigma, dst_base, want_order=True):
        fit = np.asarray(fitness, dtype=np.float64)
        order = np.argsort(-fit, kind='stable')[:n_keep]                 # Python's stable sort with reverse=True
        for r, k in enumerate(order):
            assert not (dst_base <= int(parents[k]) < dst_base + n_keep), 'a survivor would be written over its own parent'
            self.content[dst_base + r] = 'g%d:o%d' % (self.gen, k)
            self.lineage[dst_base + r] = self.content[int(parents[k])]
        return torch.from_numpy(order.astype(np.int64))


def test_es_population_follows_the_reference_master_bookkeeping():
    import nes_b200
    golden = json.load(open(os.path.join(ROOT, 'tests', 'golden', 'golden_es_v1.json')))
    assert golden['exp'] == sc.ES_EXP
    exp = sc.ES_EXP
    ctx = StubContext(nes_b200.ESPopulation.slots_needed(exp))
    pop = nes_b200.ESPopulation(ctx, exp)
    pop.init_from_zero()
    assert pop.parent_slots() is None and pop.sigma == 1.0 and len(pop.elite_candidate_slots()) == exp['num_elite_cands']
    for want in golden['generations']:
        gen = want['gen']
        ctx.gen = gen
        pop.begin_generation()
        offs, slots = pop.sample()
        assert offs.numel() == exp['nb_offspring'] == slots.numel()
        if gen == 0:
            assert set(slots.tolist()) == {pop.blank} and ctx.mutation == 'SCALE'      # parent is None: a fresh model per offspring
        else:
            assert ctx.mutation == '' and set(slots.tolist()) <= set(pop.parent_slots())
        evaluated_from = {k: ctx.content[int(s)] for k, s in enumerate(slots.tolist())}      # what each offspring's parent held when it was evaluated
        info = pop.end_generation(torch.from_numpy(sc.fitnesses(gen)), offs, slots, sc.elite_scores(gen))
        # a survivor is its parent AS EVALUATED plus noise (the worker wrote the file then, nic_es_worker.py:188): the podium update of the
        # same generation, which overwrites elite slots that served as parents, must not leak into it
        for r, k in enumerate(info['order'].tolist()):
            assert ctx.lineage[pop.parent_slots()[len(pop.podium()) + r]] == evaluated_from[k], (gen, r, k)
        got = pop.tags()
        assert got['parents'] == want['parents'], (gen, got['parents'], want['parents'])
        assert got['elites_to_evaluate'] == want['elites_to_evaluate'], gen
        assert [[t, float(s)] for t, s in got['podium']] == want['podium'], (gen, got['podium'], want['podium'])
        # ... and the slots really hold that content (what the device calls did), not just the store's labels
        assert [ctx.content[s] for s in pop.parent_slots()] == want['parents'], gen
        assert [ctx.content[s] for s in pop.elite_candidate_slots()] == want['elites_to_evaluate'], gen
        assert [ctx.content[s] for s, _ in pop.podium()] == [t for t, _ in want['podium']], gen
        assert len(set(pop.parent_slots()[len(pop.podium()):])) == exp['population_size'] - exp['num_elites']
        assert info['best_evaluated_score'] == want['best_evaluated'][1]
        assert pop.it.noise_stdev() == want['noise_stdev'] and pop.it.batch_size() == want['batch_size']
        assert pop.it.to_dict()['bad_generations'] == want['bad_generations'] and info['patience_reached'] == want['patience_reached']
    assert any(g['patience_reached'] for g in golden['generations']), 'the scenario must exercise the patience rule'
    assert ctx.modes[0] == 'SCALE' and ctx.modes[-1] == ''


def test_tournament_rule_matches_numpy_choice_distribution():
    """the lowest of t distinct uniform indices: P(k) = C(n-k-1, t-1) / C(n, t) -- what rs.choice(n, t, replace=False).min() follows
    (nic_es_worker.py:150-158); n_parents < tournament_size degenerates to index 0"""
    import math
    from oracle import noise
    n, t, F = 12, 4, 120000
    idx = noise.sample_parents_tournament(99, 3, F, n, t)
    freq = np.bincount(idx, minlength=n) / F
    want = np.array([math.comb(n - k - 1, t - 1) / math.comb(n, t) if n - k - 1 >= t - 1 else 0.0 for k in range(n)])
    assert np.max(np.abs(freq - want)) < 5e-3
    rs = np.random.RandomState(0)
    ref = np.array([rs.choice(n, t, replace=False).min() for _ in range(20000)])
    assert np.max(np.abs(np.bincount(ref, minlength=n) / 20000 - want)) < 1.5e-2
    assert np.all(noise.sample_parents_tournament(1, 0, 100, 3, 8) == 0)
    assert not np.array_equal(noise.sample_parents_tournament(99, 3, 50, n, t), noise.sample_parents_tournament(99, 4, 50, n, t))
