"""Master-side math with the reference's interface: NESMaster.gradient_estimate /
compute_centered_ranks / compute_ranks (src/algorithm/nic_nes/nic_nes_master.py:170-205), the
SGD / Adam optimizers (src/algorithm/nic_nes/optimizers.py) and one whole NIC-NES iteration
(run_master :120-137 + NESWorker.fitness nic_nes_worker.py:131-154) as an in-process, population-
sharded step: every rank evaluates its slice of antithetic pairs, fitness scalars are all-gathered
(NCCL on GPUs; gloo in the CPU tests of the sharding logic), and every rank applies the identical update.
"""
import numpy as np
import torch

from .context import NesContext, OPT_ADAM, OPT_SGD, PREC_EXACT


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


def es_sample_generation(ctx: NesContext, iteration, n_offspring, mu):
    """(offset, parent) of every offspring of a generation, identical on every rank (shared seed), grouped by parent: members
    of one parent are adjacent so that their CTAs hit the same theta tiles in L2 (offspring order carries no meaning in ES)."""
    offs = ctx.sample_offsets(iteration, n_offspring, device=True)
    pars = ctx.sample_parents(iteration, n_offspring, mu, device=True)
    pars, perm = torch.sort(pars, stable=True)
    return offs[perm].contiguous(), pars.contiguous()


def es_select_and_materialise(ctx: NesContext, fitness_all, offs, pars, mu, num_elites, base, sigma):
    """truncation selection on the gathered fitnesses, then the next generation's parents are written into the other half of
    the context's 2*mu parent slots: the first num_elites parents carried over unchanged, then the selected offspring
    re-derived from (parent, offset) -- no weights ever leave the device.  Returns the new base slot."""
    keep, _ = ESMaster.selection(fitness_all, mu, num_elites)
    sel = torch.stack([keep, offs[keep], pars[keep].to(torch.int64)], 1).cpu().numpy()      # the master's one sync point
    nxt = mu - base
    n = num_elites + len(sel)
    src = np.concatenate([base + np.arange(num_elites), base + sel[:, 2]]).astype(np.int32)
    off = np.concatenate([np.zeros(num_elites, dtype=np.int64), sel[:, 1]]).astype(np.int64)
    sg = np.concatenate([np.zeros(num_elites), np.full(len(sel), sigma)]).astype(np.float32)      # sigma 0: the elite itself
    ctx.evolve_batch(src, off, sg, nxt + np.arange(n))
    return nxt


def es_generation(ctx: NesContext, iteration, n_local, mu, num_elites, fc, sigma, base=0, fitness_fn=None,
                  world_size=1, rank=0, precision=PREC_EXACT, group=None, evaluate=None):
    """One NIC-ES generation on a context created with n_parents = 2*mu (current parents in slots [base, base+mu)).
    Every rank evaluates n_local offspring of the n_local*world_size sampled; fitness_fn(seq, logprobs) -> float64 [n_local]
    (None: sum of greedy log-probs, computed on the device).  evaluate(offsets, signs, parent_slots, sigma) -> float64 [n_local] replaces
    the fc_caption rollout altogether (the MNIST net: es_generation_mnist).  Returns (new base, all fitnesses, offsets, parents)."""
    assert ctx.n_parents >= 2 * mu and base in (0, mu)
    n_all = n_local * world_size
    offs, pars = es_sample_generation(ctx, iteration, n_all, mu)
    lo = rank * n_local
    ones = torch.ones(n_local, dtype=torch.float32, device=ctx.dev)
    if evaluate is not None:
        fit = torch.as_tensor(evaluate(offs[lo:lo + n_local], ones, pars[lo:lo + n_local] + base, sigma), dtype=torch.float64).to(ctx.dev)
    else:
        seq, lp = ctx.rollout(offs[lo:lo + n_local], ones, fc, sigma, parents=pars[lo:lo + n_local] + base, precision=precision, out_device=ctx.dev)
        rows = seq.shape[1]
        fit = ctx.logprob_fitness(n_local, rows, device=ctx.dev) if fitness_fn is None else torch.as_tensor(fitness_fn(seq, lp), dtype=torch.float64).to(ctx.dev)
    if world_size > 1:
        import torch.distributed as dist
        out = torch.empty(n_all, dtype=torch.float64, device=ctx.dev)
        dist.all_gather_into_tensor(out, fit.contiguous(), group=group)
        fit = out
    new_base = es_select_and_materialise(ctx, fit, offs, pars, mu, num_elites, base, sigma)
    return new_base, fit, offs, pars


def es_generation_mnist(ctx: NesContext, iteration, n_local, mu, num_elites, images, labels, sigma, base=0, **kw):
    """One NIC-ES generation of the MNIST classifier as experiments/mnist_es.json runs it (nb_offspring 50, population_size 10,
    num_elites 3, batch 64; nic_es_worker.py:142-192 + ClfPolicy.rollout + nic_es_master.py:155-167): every offspring's
    fitness = -CrossEntropyLoss on the batch from one k_mnist_rollout launch, selection and survivors as es_generation"""
    images = torch.as_tensor(images).to(ctx.dev)
    labels = torch.as_tensor(labels).to(ctx.dev)
    return es_generation(ctx, iteration, n_local, mu, num_elites, None, sigma, base=base,
                         evaluate=lambda o, s, p, sg: ctx.mnist_rollout(o, s, images, labels, sg, parents=p, out_device=ctx.dev)[0], **kw)
