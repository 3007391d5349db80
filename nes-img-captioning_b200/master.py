"""Master-side math with the reference's interface: NESMaster.gradient_estimate /
compute_centered_ranks / compute_ranks (src/algorithm/nic_nes/nic_nes_master.py:170-205), the
SGD / Adam optimizers (src/algorithm/nic_nes/optimizers.py) and one whole NIC-NES iteration
(run_master :120-137 + NESWorker.fitness nic_nes_worker.py:131-154) as an in-process, population-
sharded step: every rank evaluates its slice of antithetic pairs, fitness scalars are all-gathered
(NCCL on GPUs; gloo in the CPU tests of the sharding logic), and every rank applies the identical update.
"""
from collections import namedtuple

import numpy as np
import torch

from .context import NesContext, OPT_ADAM, OPT_SGD, PREC_EXACT, PREC_HYBRID


class NESMaster:
    """Only the math of the reference's NESMaster; the Redis loop around it is out of scope."""

    def __init__(self, ctx: NesContext):
        self.ctx = ctx

    def compute_ranks(self, x):
        x = np.asarray(x)
        assert x.ndim == 1
        return self.ctx.compute_ranks(x.astype(np.float64)).numpy()

    def compute_centered_ranks(self, x):
        x = np.asarray(x, dtype=np.float64)
        assert x.ndim == 2 and x.shape[1] == 2
        cen, _ = self.ctx.rank_weights(x)
        return cen.numpy()

    def gradient_estimate(self, fitnesses, offsets, sigma):
        """fitnesses float64 [F, 2]; offsets int64 [F] name the noise-table rows that the reference
        would have shipped as noise_vecs [F, D] (nic_nes/iteration.py:69-70)."""
        _, w = self.ctx.rank_weights(np.asarray(fitnesses, dtype=np.float64), want_centered=False)
        return self.ctx.gradient(offsets, w, sigma).numpy()


class Optimizer:
    """theta lives in the context (float64 master copy + float32 / split-fp16 evaluation copies)."""
    KIND = None

    def __init__(self, ctx: NesContext, theta=None):
        self.ctx = ctx
        if theta is not None:
            ctx.set_theta(theta)
        self.dim = ctx.D
        self.t = 0

    def update(self, globalg):
        """-> (ratio, new_theta float64 numpy) like optimizers.py:15-22"""
        self.t += 1
        ratio = self.ctx.opt_update_direct(globalg)
        return ratio, self.ctx.get_theta(dtype=torch.float64).numpy()

    def step(self, grad_estimate=None, l2coeff=0.0, fetch=False):
        """G = -grad_estimate + l2coeff*theta fused with the update (nic_nes_master.py:123-133);
        grad_estimate None = the gradient resident from the last gradient() call."""
        self.t += 1
        ratio = self.ctx.opt_update(grad_estimate, l2coeff)
        return (ratio, self.ctx.get_theta(dtype=torch.float64).numpy()) if fetch else ratio

    @property
    def stepsize(self):
        return self._stepsize

    @stepsize.setter
    def stepsize(self, v):
        self._stepsize = float(v)
        self.ctx.opt_set_stepsize(self._stepsize)


class SGD(Optimizer):
    KIND = OPT_SGD

    def __init__(self, ctx, theta=None, stepsize=0.01, momentum=0.9):
        super().__init__(ctx, theta)
        self.momentum = momentum
        ctx.opt_init(OPT_SGD, stepsize, momentum)
        self._stepsize = float(stepsize)

    def save_to_file(self, path):
        _, v, t = self.ctx.opt_get_state()
        torch.save({'dim': self.dim, 't': t, 'momentum': self.momentum, 'stepsize': self.stepsize, 'v': v.numpy()}, path)

    def load_from_file(self, path):
        s = torch.load(path, map_location='cpu', weights_only=False)
        self.dim, self.t, self.momentum = s['dim'], s['t'], s['momentum']
        self.ctx.opt_init(OPT_SGD, s['stepsize'], self.momentum)
        self._stepsize = float(s['stepsize'])
        self.ctx.opt_set_state(None, np.asarray(s['v'], dtype=np.float64), self.t)


class Adam(Optimizer):
    KIND = OPT_ADAM

    def __init__(self, ctx, theta=None, stepsize=0.001, beta1=0.9, beta2=0.999, epsilon=1e-08):
        super().__init__(ctx, theta)
        self.beta1, self.beta2, self.epsilon = beta1, beta2, epsilon
        ctx.opt_init(OPT_ADAM, stepsize, beta1, beta2, epsilon)
        self._stepsize = float(stepsize)

    def save_to_file(self, path):
        m, v, t = self.ctx.opt_get_state()
        torch.save({'dim': self.dim, 't': t, 'stepsize': self.stepsize, 'beta1': self.beta1, 'beta2': self.beta2,
                    'epsilon': self.epsilon, 'm': m.numpy(), 'v': v.numpy()}, path)

    def load_from_file(self, path):
        s = torch.load(path, map_location='cpu', weights_only=False)
        self.dim, self.t = s['dim'], s['t']
        self.beta1, self.beta2, self.epsilon = s['beta1'], s['beta2'], s['epsilon']
        self.ctx.opt_init(OPT_ADAM, s['stepsize'], self.beta1, self.beta2, self.epsilon)
        self._stepsize = float(s['stepsize'])
        self.ctx.opt_set_state(np.asarray(s['m'], dtype=np.float64), np.asarray(s['v'], dtype=np.float64), self.t)


# ---------------------------------------------------------------------------------------------
# population sharding: rank r evaluates noise rows [lo, hi); both signs of a pair stay together
def shard_bounds(F, world_size, rank):
    base, rem = divmod(int(F), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def all_gather_fitness(local_fit, F, world_size, rank, group=None):
    """local_fit float64 [hi-lo, 2] (torch, on the device the process group communicates on) -> [F, 2].
    The only per-iteration message of the design (80 KB at F = 5000)."""
    import torch.distributed as dist
    if world_size == 1:
        return local_fit
    sizes = [shard_bounds(F, world_size, r) for r in range(world_size)]
    maxn = max(hi - lo for lo, hi in sizes)
    pad = torch.zeros((maxn, 2), dtype=torch.float64, device=local_fit.device)
    pad[:local_fit.shape[0]] = local_fit
    out = [torch.empty_like(pad) for _ in range(world_size)]
    dist.all_gather(out, pad, group=group)
    return torch.cat([o[:hi - lo] for o, (lo, hi) in zip(out, sizes)], 0)


def all_reduce_gradient(g, world_size, group=None):
    """sum of the ranks' parameter-axis slices of K7 (nes_gradient_shard: zeros outside a rank's own slice, so the sum is exact and
    every rank ends with the single-GPU vector).  The second and last exchange step of an iteration: 4 * D bytes."""
    if world_size > 1:
        import torch.distributed as dist
        dist.all_reduce(g, op=dist.ReduceOp.SUM, group=group)
    return g


def sharded_master_step(ctx: NesContext, fit, offsets, sigma, l2coeff, world_size=1, rank=0, group=None, want_ratio=True):
    """NESMaster's half of an iteration (nic_nes_master.py:120-137) on every rank: centred-rank weights from the gathered
    fitnesses, K7 over this rank's slice of the parameter axis, all-reduce, then the identical optimizer update everywhere."""
    if world_size == 1:
        return ctx.master_step(fit, offsets, sigma, l2coeff, want_ratio=want_ratio)
    ctx.rank_weights(fit, want_centered=False, want_weights=False)
    g = ctx.gradient(offsets, None, sigma, out_device=ctx.dev, shard=(rank, world_size))
    all_reduce_gradient(g, world_size, group)
    return ctx.opt_update(g, l2coeff, want_ratio=want_ratio)


def nes_iteration(ctx: NesContext, optimizer: Optimizer, iteration, F, fc, sigma, l2coeff, fitness_fn,
                  world_size=1, rank=0, precision=PREC_EXACT, group=None):
    """One NIC-NES iteration.  fitness_fn(seq [P,rows,T], logprobs [P,rows,T]) -> float64 [P] scores the
    decoded captions (CIDEr-D in the reference, captioning/policies.py:117-125; host-side, off the CUDA path).
    Returns (update_ratio, fitness [F,2])."""
    offsets = ctx.sample_offsets(iteration, F)                       # identical on every rank (same seed)
    lo, hi = shard_bounds(F, world_size, rank)
    n = hi - lo
    offs2 = offsets[lo:hi].repeat_interleave(2)
    signs = torch.tensor([1.0, -1.0], dtype=torch.float32).repeat(n)
    seq, lp = ctx.rollout(offs2, signs, fc, sigma, precision=precision, out_device=ctx.dev)
    fit = torch.as_tensor(fitness_fn(seq, lp), dtype=torch.float64).reshape(n, 2)
    fit = all_gather_fitness(fit.to(ctx.dev), F, world_size, rank, group)
    ratio = sharded_master_step(ctx, fit, offsets, sigma, l2coeff, world_size, rank, group)
    optimizer.t += 1
    return ratio, fit


# ---------------------------------------------------------------------------------------------
# NIC-ES generation (ESWorker.fitness nic_es_worker.py:142-192 + ESMaster.selection nic_es_master.py:155-167), in process
class ESMaster:
    """Only the math of the reference's ESMaster."""

    @staticmethod
    def selection(fitness, pop_size, num_elites):
        """indices of the pop_size - num_elites best offspring, best first, and all scores sorted descending
        (nic_es_master.py:155-167: Python's stable sort with reverse=True, so equal fitnesses keep their submission order).
        fitness: torch float64 [n] (any device) or array-like.  Returns (LongTensor indices, sorted scores) on fitness's device."""
        fit = fitness if torch.is_tensor(fitness) else torch.as_tensor(np.asarray(fitness, dtype=np.float64))
        order = torch.sort(-fit.to(torch.float64), stable=True).indices
        return order[:max(0, int(pop_size) - int(num_elites))], fit[order]


EvalES = namedtuple('EvalES', ['evaluated_cand_id', 'evaluated_cand', 'score'])       # the fields of ESResult that record_eval_result reads


class ESPopulation:
    """One NIC-ES run on a device: the reference master's bookkeeping (ESIteration + Podium + ESMaster.selection in the order of
    ESMaster.run_master, nic_es_master.py:66-125) driving models that are parent slots of a NesContext instead of `.pth` files.

    What the reference does per generation and where it happens here:
      workers pick a parent (uniform or tournament, nic_es_worker.py:150-162), mutate it and evaluate      sample() + the caller's rollout
      `parent is None` (first generation after init_from_zero): a fresh random model per offspring (:166-169)   sample(): blank slot, sigma 1,
                                                                                           per-parameter Xavier std as the noise scale
      workers evaluate the elite candidates on validation data (:114-141)                                    the caller's elite_scores
      process_evaluated_elites -> Podium.record_elites (tools/iteration.py:117-148, tools/podium.py:43-62)  end_generation(), SlotStore copies
      ESMaster.selection, set_next_elites_to_evaluate, record_parents (nic_es_master.py:119-125)            end_generation(): nes_es_select_evolve
                                                                                           + the same ESIteration calls
    The parents of the next generation are the podium's elites followed by the population_size - num_elites best offspring, best first.

    Slots: [0, mu) and [mu, 2 mu) take the survivors of alternate generations; slot 2 mu is the blank model of a fresh initialisation,
    the num_elites + num_elite_cands slots behind it hold the podium and the elite candidates.  `ESPopulation.slots_needed(exp)` is the n_parents the context must have."""

    def __init__(self, ctx, exp, config=None, sensitivity_fn=None, root='/nes_es'):
        from .schedule import Config, ESIteration, SlotStore
        self.ctx, self.exp = ctx, dict(exp)
        self.mu = int(exp.get('population_size', exp['nb_offspring']))
        self.E, self.C = int(exp['num_elites']), int(exp['num_elite_cands'])
        self.selection = exp.get('selection', 'uniform')
        self.tournament_size = int(exp.get('tournament_size') or 0)
        assert ctx.n_parents >= self.slots_needed(exp), 'the context needs n_parents >= %d' % self.slots_needed(exp)
        assert self.mu > self.E >= 0 and self.C <= self.mu - self.E
        config = config if config is not None else Config(**exp.get('config', {}))
        self.store = SlotStore(ctx, free_slots=range(2 * self.mu + 1, 2 * self.mu + 1 + self.E + self.C))
        self.blank = 2 * self.mu
        self.it = ESIteration(config, dict(exp, log_dir=root), store=self.store)
        self.block = 0                      # the half that receives the next survivors
        self.sensitivity_fn = sensitivity_fn
        self.mutation = getattr(ctx, 'mutation', '')
        self._fresh_mode = False
        self.last = {}

    @staticmethod
    def slots_needed(exp):
        mu = int(exp.get('population_size', exp['nb_offspring']))
        return 2 * mu + 1 + int(exp['num_elites']) + int(exp['num_elite_cands'])

    # ---- ESIteration.init_from_zero / init_from_singles (nic_es/iteration.py:50-81)
    def _enter_fresh_mode(self):
        if not self._fresh_mode:
            from .layout import fresh_model_vectors
            blank, std = fresh_model_vectors(self.ctx.tensors)
            self.ctx.set_mutation('SCALE')
            self.ctx.set_theta(blank, self.blank)
            self.ctx.set_sensitivity(std, parent=self.blank)
            self._fresh_mode = True

    def _leave_fresh_mode(self):
        if self._fresh_mode:
            self.ctx.set_mutation(self.mutation)
            self._fresh_mode = False

    def init_from_zero(self, seed_iteration=(1 << 41)):
        """parents = population_size x None, num_elite_cands freshly initialised elite candidates"""
        pop = self
        self._enter_fresh_mode()
        offsets = self.ctx.sample_offsets(int(seed_iteration), self.C)

        class _Fresh:
            def __init__(self, j): self.j = j

            def serialize(self, path=''):
                slot = pop.store.free.pop(0)
                pop.store.adopt(path, slot, tag='init_elite:%d' % self.j)
                pop.ctx.evolve(int(offsets[self.j]), 1.0, 1.0, parent=pop.blank, want_member=False, want_noise=False, dst_parent=slot)
                return path

        class _Policy:
            n = 0

            def generate_model(self, from_param_file=None, start_rng=None):
                _Policy.n += 1
                return _Fresh(_Policy.n - 1)

        self.it.init_from_zero(self.exp, _Policy())

    def init_from_thetas(self, thetas):
        """init_from_singles with the models given as flat vectors: every one becomes a parent, the first num_elite_cands also elite
        candidates"""
        pop = self

        class _Given:
            def __init__(self, k): self.k = k

            def serialize(self, path=''):
                if path.startswith(pop.it.offspring_dir()):      # a parent: the half that the first survivors do NOT go to
                    slot = pop.mu + self.k
                elif self.k >= pop.C:                            # the reference writes an elite file for every single and keeps the first
                    return path                                  # num_elite_cands (nic_es/iteration.py:70-81): no slot for the dropped ones
                else:
                    slot = pop.store.free.pop(0)
                pop.store.adopt(path, slot, tag='single:%d' % self.k)
                pop.ctx.set_theta(thetas[self.k], slot)
                return path

        class _Policy:
            def generate_model(self, from_param_file=None, start_rng=None):
                return _Given(int(from_param_file))

        assert len(thetas) <= self.mu
        self.it.init_from_singles(list(range(len(thetas))), self.exp, _Policy())

    # ---- readers
    def parent_slots(self):
        """slot of every parent in the reference's order (best first), or None while the parents are still `None`"""
        ps = self.it.parents()
        return None if ps and ps[0][1] is None else [self.store.slot(p) for _, p in ps]

    def elite_candidate_slots(self):
        return [self.store.slot(p) for _, p in self.it.elites_to_evaluate()]

    def podium(self):
        return [(self.store.slot(p), sc) for p, sc in self.it.best_elites() if p]

    @property
    def sigma(self):
        """the noise_stdev a rollout of the current generation uses (1 while the offspring are fresh random models)"""
        return 1.0 if self.parent_slots() is None else float(self.it.noise_stdev())

    def tags(self):
        """{parents, elites_to_evaluate, podium: where each model came from}; resolves the device-side selection order on demand"""
        def tag(path):
            t = self.store.tags.get(path)
            if isinstance(t, tuple):
                gen, rank, order = t
                t = 'g%d:o%d' % (gen, int(order[rank]))
            return t
        return {'parents': [tag(p) for _, p in self.it.parents()], 'elites_to_evaluate': [tag(p) for _, p in self.it.elites_to_evaluate()],
                'podium': [[tag(p), sc] for p, sc in self.it.best_elites() if p]}

    # ---- one generation
    def begin_generation(self):
        self.it.incr_iteration()
        slots = self.parent_slots()
        if slots is None:
            self._enter_fresh_mode()
        elif self.ctx.MUTATIONS.get(self.mutation) == 2:
            # the reference recomputes / reloads the sensitivity for every (task, parent) (nic_es_worker.py:172-173)
            assert self.sensitivity_fn is not None, 'a sensitivity mutation needs sensitivity_fn(slot) for the new parents'
            for s in sorted(set(slots)):
                self.sensitivity_fn(s)

    def sample(self, n_offspring=None):
        """(offsets int64 [n], parent slots int32 [n]) on the device, identical on every rank (shared seed); members of one parent are
        adjacent so that their CTAs share the parent's theta through L2 (the order of the offspring carries no meaning)"""
        n = int(n_offspring or self.exp['nb_offspring'])
        gen = self.it.iteration()
        offs = self.ctx.sample_offsets(gen, n, device=True)
        slots = self.parent_slots()
        if slots is None:
            return offs, torch.full((n,), self.blank, dtype=torch.int32, device=self.ctx.dev)
        if self.selection == 'tournament':
            idx = self.ctx.sample_parents_tournament(gen, n, len(slots), self.tournament_size, device=True)
        else:
            idx = self.ctx.sample_parents(gen, n, len(slots), device=True)
        idx, perm = torch.sort(idx, stable=True)
        table = torch.tensor(slots, dtype=torch.int32).to(self.ctx.dev, non_blocking=True)
        return offs[perm].contiguous(), table[idx.long()].contiguous()

    def end_generation(self, fitness_all, offs, slots, elite_scores):
        """fitness_all float64 [n] (all ranks' offspring, any device), offs / slots as sample() returned them, elite_scores: one
        validation score per elite candidate (ESWorker.accuracy).  Order of operations = nic_es_master.py:116-125."""
        it = self.it
        cands = it.elites_to_evaluate()
        assert len(elite_scores) == len(cands)
        # The survivors first: an offspring IS theta[parent] + sigma * eps with the parent as it was when the offspring was evaluated (the
        # worker wrote its file then, nic_es_worker.py:188), so they are re-derived before the podium copies below overwrite elite slots
        # that served as parents.  They go to the idle half of the slots, which nothing else touches.
        n_keep = min(self.mu - self.E, int(fitness_all.numel()))
        base = self.block * self.mu
        order = self.ctx.es_select_evolve(fitness_all, offs, slots, n_keep, self.sigma, base)
        for (cid, cpath), score in zip(cands, elite_scores):
            it.record_eval_result(EvalES(cid, cpath, float(score)))
        best_sc, best_ind = it.process_evaluated_elites()
        self._leave_fresh_mode()
        gen = it.iteration()
        paths = []
        for r in range(n_keep):
            path = '%s/r%d_%d_offspring_params.pth' % (it.offspring_dir(), r, gen)
            self.store.adopt(path, base + r, tag=(gen - 1, r, order))
            paths.append(path)
        it.set_next_elites_to_evaluate(paths[:self.C])
        it.record_parents(paths)
        self.block ^= 1
        self.last = {'best_evaluated_score': best_sc, 'best_evaluated_slot': None if best_ind is None else self.store.slot_of.get(best_ind),
                     'order': order, 'patience_reached': it.patience_reached(), 'schedule_reached': it.schedule_reached()}
        return self.last


def es_generation(pop: ESPopulation, fc, fitness_fn=None, elite_score_fn=None, n_local=None, world_size=1, rank=0, precision=PREC_HYBRID,
                  group=None, evaluate=None):
    """One NIC-ES generation of `pop` (ESWorker.fitness nic_es_worker.py:142-192 for every offspring + the master's half).  Every rank
    evaluates n_local offspring of the n_local * world_size sampled; fitness_fn(seq, logprobs) -> float64 [n_local] (None: the sum of greedy
    log-probs, on the device); evaluate(offsets, signs, slots, sigma) -> float64 [n] replaces the fc_caption rollout (the MNIST net);
    elite_score_fn(slots) -> one score per elite candidate (the reference's validation run, ESWorker.accuracy; None: the fitness of the
    unperturbed candidate on the same batch).  Returns (all fitnesses, offsets, parent slots)."""
    ctx = pop.ctx
    pop.begin_generation()
    n_local = int(n_local or pop.exp['nb_offspring'])
    n_all = n_local * world_size
    offs, slots = pop.sample(n_all)
    lo = rank * n_local
    ones = torch.ones(n_local, dtype=torch.float32, device=ctx.dev)

    def run(o, sg, sl, sigma):
        if evaluate is not None:
            return torch.as_tensor(evaluate(o, sg, sl, sigma), dtype=torch.float64).to(ctx.dev)
        seq, lp = ctx.rollout(o, sg, fc, sigma, parents=sl, precision=precision, out_device=ctx.dev)
        if fitness_fn is None:
            return ctx.logprob_fitness(o.numel(), seq.shape[1], device=ctx.dev)
        return torch.as_tensor(fitness_fn(seq, lp), dtype=torch.float64).to(ctx.dev)

    # the elite candidates first: their scores are only needed by the host after the offspring are done
    cand = pop.elite_candidate_slots()
    if elite_score_fn is not None:
        elite = elite_score_fn(cand)
    elif cand:
        cs = torch.tensor(cand, dtype=torch.int32, device=ctx.dev)
        elite = run(torch.zeros(len(cand), dtype=torch.int64, device=ctx.dev), torch.ones(len(cand), dtype=torch.float32, device=ctx.dev), cs, 0.0)
    else:
        elite = []
    fit = run(offs[lo:lo + n_local], ones, slots[lo:lo + n_local], pop.sigma)
    if world_size > 1:
        import torch.distributed as dist
        out = torch.empty(n_all, dtype=torch.float64, device=ctx.dev)
        dist.all_gather_into_tensor(out, fit.contiguous(), group=group)
        fit = out
    elite = elite.cpu().tolist() if torch.is_tensor(elite) else list(elite)            # the master's one sync point of a generation
    pop.end_generation(fit, offs, slots, elite)
    return fit, offs, slots


def es_generation_mnist(pop: ESPopulation, images, labels, **kw):
    """One NIC-ES generation of the MNIST classifier as experiments/mnist_es.json runs it (nb_offspring 50, population_size 10,
    num_elites 3, batch 64; nic_es_worker.py:142-192 + ClfPolicy.rollout + nic_es_master.py:116-125): every offspring's
    fitness = -CrossEntropyLoss on the batch from one k_mnist_rollout launch"""
    ctx = pop.ctx
    images = torch.as_tensor(images).to(ctx.dev)
    labels = torch.as_tensor(labels).to(ctx.dev)
    return es_generation(pop, None, evaluate=lambda o, s, p, sg: ctx.mnist_rollout(o, s, images, labels, sg, parents=p, out_device=ctx.dev)[0], **kw)
