#!/usr/bin/env python
"""Headline benchmark: fc_caption perturbed-member evals/s (bs128 greedy) -- BASELINE.json configs[1]:
NIC-ES, fc_feat 2048, E = R = 512, V = 9487, 128 rows per rollout, population 1000 (per GPU: weak scaling),
mu = 50 parents, synthetic features, random-init weights.

One step = one ES generation as ESMaster.run_master + ESWorker.fitness run it (nic_es_master.py:66-125, nic_es_worker.py:142-192):
sample (parent, noise offset) per offspring (parents = podium elites + best offspring of the last generation) -> evaluate the
num_elite_cands elite candidates (the reference's validation run; here their greedy fitness on a held-out batch) -> decode all
offspring (img_embed, 17 LSTM steps, 16 vocabulary projections with fused argmax/log-prob) -> fitness -> [all-gather of fitness
scalars when N > 1] -> podium update -> stable truncation selection + survivors written as the next parents (on the device).
The first warm-up generation is the reference's `parent is None` one: every offspring a fresh random model.  CIDEr-D scoring is
host-side Python in the reference and outside the path (BASELINE.json); the fitness here is the sum of greedy log-probs.

With N > 1 (and at N = 1) the line also carries `nes_config3`: BASELINE configs[2], NIC-NES / Adam / antithetic, population 5000
(F = 5000 noise rows = 10 000 rollouts per iteration, STRONG scaling: rank r evaluates pairs [r F/N, (r+1) F/N)), fitness all-gather,
K7 sharded over the parameter axis + all-reduce, identical Adam step on every rank, and a 64-bit hash of theta' compared across
ranks and with the committed 1-rank hash (tests/golden/config3_theta_hash.json).

  python bench.py [--gpus N --steps K --warmup W] [--impl reference] [--precision exact|fast]
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

V, E, R, FD, T = 9487, 512, 512, 2048, 16
FLOP_PER_ROW = 2 * FD * E + 17 * 2 * (E + R) * 5 * R + 16 * 2 * R * (V + 1)       # 246 677 504


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--population', type=int, default=1000, help='offspring per GPU (weak scaling)')
    ap.add_argument('--parents', type=int, default=50)
    ap.add_argument('--elites', type=int, default=3)
    ap.add_argument('--elite-cands', type=int, default=2)
    ap.add_argument('--rows', type=int, default=128)
    ap.add_argument('--sigma', type=float, default=0.005)
    ap.add_argument('--precision', default='hybrid', choices=['hybrid', 'exact', 'fast'])
    ap.add_argument('--table-log2', type=int, default=32, help='noise table of 2^k fp16 elements (2^32 = 8 GB: 2.1 M distinct rows, K7 rows do not overlap in L2)')
    ap.add_argument('--nes-pairs', type=int, default=5000, help='configs[2]: noise rows per NIC-NES iteration (0: skip)')
    ap.add_argument('--write-theta-hash', action='store_true', help='N = 1 only: store the config-3 theta hash as the golden')
    ap.add_argument('--cpu-seconds', type=float, default=15.0)
    ap.add_argument('--no-extras', action='store_true')
    ap.add_argument('--sections', action='store_true', help='experiment: CUDA-event time of the phases of a step (sample / rollout / fitness+gather / selection)')
    ap.add_argument('--debug-same-offset', action='store_true', help='experiment only: every member reads the same noise row (eps L2-resident)')
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# clocks during the timed region (B200_PROFILING.md recipe)
class ClockSampler:
    Q = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q, '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(',')]))

    def window(self, t0, t1):
        """keep the samples taken inside the timed region (the sampler is started before the warm-up because nvidia-smi
        needs up to a second to produce its first line on an 8-GPU box)"""
        self.t0, self.t1 = t0, t1

    def stop(self):
        if not self.proc:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        t0, t1 = getattr(self, 't0', None), getattr(self, 't1', None)
        rows = [r for t, r in self.rows if t0 is None or (t0 <= t <= t1 + 0.1)]
        if not rows:                                   # region shorter than the sampling period: the samples nearest to it
            rows = [r for _, r in self.rows[-3:]]
        for r in rows:
            try:
                sm.append(float(r[1])); mx = float(r[2])
                for name, v in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), r[5:9]):
                    if v.lower().startswith('active'):
                        reasons.add(name)
            except Exception:
                pass
        sm.sort()
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': mx, 'reasons': sorted(reasons), 'samples': len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference's worker recipe (nic_es_worker.py:142-175 / nic_nes_worker.py:131-154 minus disk / Redis /
# CIDEr): member = theta + noise with noise = torch.empty_like(theta).normal_(0, sigma) exactly as PolicyNet.evolve forms it
# (algorithm/nets.py:101-113), theta from the reference's Xavier init, then FCModel._sample(greedy).  One single-threaded process per
# host core, exactly how the reference uses a node (src/main.py:8-11,105-106).
def _cpu_worker(args):
    seed, n_members, rows, sigma = args
    import numpy as np
    import torch
    torch.set_num_threads(1)
    from oracle import fc_caption
    dims = fc_caption.Dims(V, E, R, FD)
    theta = torch.from_numpy(fc_caption.xavier_theta(dims, 11))
    torch.manual_seed(seed)
    fc = torch.randn(rows, FD)
    t0 = time.perf_counter()
    for _ in range(n_members):
        noise = torch.empty_like(theta).normal_(mean=0, std=sigma)          # algorithm/nets.py:101-102
        member = theta + noise                                               # :113 (param_vector + noise)
        fc_caption.decode_greedy(dims, member.numpy(), fc.numpy())
    return time.perf_counter() - t0


def cpu_evals_per_s(rows, sigma, budget_s, per_worker=None):
    import multiprocessing as mp
    cores = len(os.sched_getaffinity(0)) if hasattr(os, 'sched_getaffinity') else (os.cpu_count() or 1)
    ctx = mp.get_context('fork')
    with ctx.Pool(cores) as pool:
        t1 = max(pool.map(_cpu_worker, [(100 + i, 1, rows, sigma) for i in range(cores)]))      # calibrate: 1 member each
        n = per_worker or max(1, int(budget_s / max(t1, 1e-3)))
        t0 = time.perf_counter()
        pool.map(_cpu_worker, [(200 + i, n, rows, sigma) for i in range(cores)])
        wall = time.perf_counter() - t0
    return cores * n / wall, cores, n


def cpu_master_leg(F=256, D=2865808):
    """the master's half on the host as the reference computes it (oracle port of NESMaster.gradient_estimate nic_nes_master.py:170-221 --
    centred ranks, np.dot over chunks of 500 float32 rows -- and Adam.update optimizers.py:15-22,78-83 in float64), on an [F, D] float32
    noise stack that fits host memory (D = the shipped E = R = 128 model)"""
    import numpy as np
    import torch
    from oracle import nes_math
    g = torch.Generator().manual_seed(3)
    noise = torch.randn(F, D, generator=g).numpy()
    fit = np.random.RandomState(4).randn(F, 2)
    theta = np.zeros(D, dtype=np.float32) + 0.01
    opt = nes_math.Adam(theta.copy(), stepsize=0.001)
    t0 = time.perf_counter()
    grad = nes_math.gradient_estimate(fit, noise)
    t1 = time.perf_counter()
    nes_math.master_update(opt, theta, grad, 1e-7)
    t2 = time.perf_counter()
    return {'F': F, 'D': D, 'gradient_estimate_ms': (t1 - t0) * 1e3, 'gradient_GBps': F * D * 4 / (t1 - t0) / 1e9,
            'adam_update_ms': (t2 - t1) * 1e3, 'threads': torch.get_num_threads(),
            'what': 'oracle port of NESMaster.gradient_estimate + Adam.update (numpy, float32 stack / float64 state as the reference)'}


METRIC = 'fc_caption perturbed-member evals/sec (bs128 greedy)'


def run_reference(a, rank, world):
    if rank != 0:
        return
    per_step = None
    vals = []
    cores = 1
    budget = max(2.0, min(20.0, 120.0 / max(1, a.steps + a.warmup)))
    for i in range(a.warmup + a.steps):
        v, cores, n = cpu_evals_per_s(a.rows, a.sigma, budget, per_step)
        per_step = n
        if i >= a.warmup:
            vals.append(v)
    value = sum(vals) / len(vals)
    sample = ('%d members per step (1 single-threaded process per core x %d members), %d rows, oracle port of the reference worker recipe: torch normal_ noise, '
              'Xavier theta, FCModel greedy decode (17 LSTM steps + 16 vocabulary projections; the reference also computes 1 + 2 it discards)' % (cores * per_step, per_step, a.rows))
    print(json.dumps({
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': 'member-evals/s',
        'n_gpus': a.gpus, 'steps': a.steps, 'warmup': a.warmup, 'ms_per_step': 1000.0 * cores * per_step / value,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': workload_config(a, 1),
        'cpu_baseline': {'value': value, 'unit': 'member-evals/s', 'cores': cores, 'kind': 'port', 'sample': sample},
        'e2e': {'value': value, 'unit': 'member-evals/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}))


def workload_config(a, world):
    return {'workload': 'configs[1]: fc_caption NIC-ES generation, fc_feat 2048, E=R=512, V=9487, %d rows/rollout, population %d per GPU, mu=%d, elites=%d, sigma=%g'
                        % (a.rows, a.population, a.parents, a.elites, a.sigma),
            'population_per_gpu': a.population, 'global_population': a.population * world, 'rows': a.rows,
            'parallelism': 'population-sharded x%d, fitness all-gather' % world,
            'l2': 'per-step inputs (noise-table slices, 247 MB per member) far exceed the 126 MB L2; no flush needed'}


def es_exp(a, world):
    return {'nb_offspring': a.population * world, 'population_size': a.parents, 'num_elites': a.elites, 'num_elite_cands': a.elite_cands,
            'selection': 'uniform', 'tournament_size': 0,                                  # experiments/mscoco_es.json
            'config': {'noise_stdev': a.sigma, 'batch_size': a.rows, 'patience': 0}}


def theta_hash(torch, theta32):
    """two 64-bit wrap-around checksums of the float32 bits of theta (position-weighted and plain): equal iff the vectors are
    bit-identical, up to a 2^-64-ish collision chance"""
    bits = theta32.view(torch.int32).to(torch.int64)
    w = (torch.arange(bits.numel(), device=bits.device, dtype=torch.int64) % 65521) + 1
    return torch.stack([(bits * w).sum(), bits.sum()])


# ------------------------------------------------------------------------------------------------
def main():
    a = parse()
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if a.impl == 'reference':
        return run_reference(a, rank, world)

    import numpy as np
    import torch
    import nes_b200
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    early_sampler = ClockSampler(local) if rank == 0 else None      # started now: nvidia-smi takes a while to print its first line
    if early_sampler:
        early_sampler.start()
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group('nccl', device_id=dev)
    prec = {'exact': nes_b200.PREC_EXACT, 'fast': nes_b200.PREC_FAST, 'hybrid': nes_b200.PREC_HYBRID}[a.precision]
    P, mu, rows = a.population, a.parents, a.rows
    Pg = P * world
    exp = es_exp(a, world)
    ctx = nes_b200.NesContext(dict(vocab_size=V, input_encoding_size=E, rnn_size=R, fc_feat_size=FD), seed=1234,
                              table_len=1 << a.table_log2, n_parents=nes_b200.ESPopulation.slots_needed(exp), wave_members=min(P, 1024), rows=rows, device=local)
    # NIC-ES from scratch (ESIteration.init_from_zero): random-init elite candidates, generation 0 = a fresh Xavier model per offspring
    # (algorithm/nets.py:62-69); synthetic N(0,1) fc features, a second batch stands in for the validation data of the elite candidates
    pop = nes_b200.ESPopulation(ctx, exp)
    state = {'profiling': False}
    if a.precision == 'hybrid':
        pop.init_from_zero()
    else:                              # the fresh-model generation runs on the scaled-noise kernels, which exist in HYBRID precision only
        pop.init_from_thetas([nes_b200.FCModel(ctx, parent=pop.blank, rng_state=k).parameter_vector() for k in range(mu)])
    g = torch.Generator().manual_seed(0)
    fc_host = torch.randn(rows, FD, generator=g).pin_memory()
    fc_val_host = torch.randn(rows, FD, generator=g).pin_memory()
    fc_dev, fc_val = fc_host.to(dev), fc_val_host.to(dev)
    ones = torch.ones(P, dtype=torch.float32, device=dev)
    lo = rank * P
    sections = []

    def mark():
        if a.sections:
            e = torch.cuda.Event(enable_timing=True); e.record(); sections.append(e)

    def elite_scores_device():
        """the reference's workers score the elite candidates on validation data (nic_es_worker.py:114-141); here: greedy fitness of the
        unperturbed candidate on the held-out batch, identical on every rank"""
        cand = pop.elite_candidate_slots()
        cs = torch.tensor(cand, dtype=torch.int32, device=dev)
        if state['profiling']:
            ctx.profile(False)         # keep these 2-member launches out of the per-kernel averages of the roofline object
        ctx.rollout(torch.zeros(len(cand), dtype=torch.int64, device=dev), ones[:len(cand)], fc_val, 0.0, parents=cs, precision=prec, out_device=dev)
        fit = ctx.logprob_fitness(len(cand), rows, device=dev)
        if state['profiling']:
            ctx.profile(True)
        return fit

    def gather(fit):
        if world > 1:
            out = torch.empty(Pg, dtype=torch.float64, device=dev)
            dist.all_gather_into_tensor(out, fit)
            return out
        return fit

    def step_resident():
        mark()
        pop.begin_generation()
        offs, slots = pop.sample(Pg)
        if a.debug_same_offset:
            offs = offs[:1].expand(Pg).contiguous()
        mark()
        elite = elite_scores_device()
        mark()
        ctx.rollout(offs[lo:lo + P], ones, fc_dev, pop.sigma, parents=slots[lo:lo + P], precision=prec, out_device=dev)
        mark()
        fit = gather(ctx.logprob_fitness(P, rows, device=dev))
        mark()
        pop.end_generation(fit, offs, slots, elite.cpu().tolist())          # the one host sync of a generation (2 scalars)
        mark()

    seq_host = torch.empty((P, rows, T), dtype=torch.int64).pin_memory()
    lp_host = torch.empty((P, rows, T), dtype=torch.float32).pin_memory()
    C = a.elite_cands
    seq_c = torch.empty((C, rows, T), dtype=torch.int64).pin_memory()
    lp_c = torch.empty((C, rows, T), dtype=torch.float32).pin_memory()
    zeros_c = torch.zeros(C, dtype=torch.int64)

    def step_e2e():
        """through the host boundary the reference has: features come from host memory, tokens / log-probs go back to the host where the
        scorer lives, the host returns fitness scalars"""
        pop.begin_generation()
        offs, slots = pop.sample(Pg)
        lib, h = ctx.lib, ctx._h
        ptr, check = nes_b200._lib.ptr, nes_b200._lib.check
        cand = torch.tensor(pop.elite_candidate_slots(), dtype=torch.int32)
        check(lib.nes_rollout(h, C, ptr(zeros_c), ptr(ones[:C]), ptr(cand), ptr(fc_val_host), rows, 1, 0.0, int(prec), ptr(seq_c), ptr(lp_c)))
        elite = lp_c.view(C, -1).sum(1, dtype=torch.float64).tolist()
        sl_local = slots[lo:lo + P].contiguous()
        off_local = offs[lo:lo + P].contiguous()
        check(lib.nes_rollout(h, P, ptr(off_local), ptr(ones), ptr(sl_local), ptr(fc_host), rows, 1, float(pop.sigma), int(prec), ptr(seq_host), ptr(lp_host)))
        fit = lp_host.view(P, -1).sum(1, dtype=torch.float64)          # host-side scorer stand-in
        pop.end_generation(gather(fit.to(dev, non_blocking=True)), offs, slots, elite)

    def timed(fn, steps, warmup, sample_clocks=False, profile=False):
        sampler = early_sampler if sample_clocks else None
        for _ in range(warmup):
            fn()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        if profile:
            ctx.profile(True)          # per-kernel CUDA-event pairs cover the timed steps only
            state['profiling'] = True
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = ctx.gpu_launches
        w0 = time.time()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        if sampler:
            sampler.window(w0, time.time())
        if world > 1:
            dist.barrier()
        ms = e0.elapsed_time(e1)
        clocks = sampler.stop() if sampler else None
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item(), ctx.gpu_launches - l0, clocks, ms

    # pass 1: the timed region (no per-kernel events in the stream) -> value; pass 2: the same K steps again with a CUDA-event pair
    # around every launch (nes_profile_*) -> per-kernel averages of the roofline object
    ms, launches, clocks, ms_own = timed(step_resident, a.steps, a.warmup, sample_clocks=True, profile=False)
    ms_prof, _, _, _ = timed(step_resident, a.steps, 0, profile=True)
    prof = ctx.profile_read()
    per_rank = None
    if a.sections:
        n = 6
        tail = sections[-n * a.steps:]
        names = ['sample', 'elite_candidates', 'rollout', 'fitness+gather', 'podium+select+evolve']
        acc = [0.0] * (n - 1)
        for k in range(a.steps):
            for j in range(n - 1):
                acc[j] += tail[k * n + j].elapsed_time(tail[k * n + j + 1])
        between = sum(tail[k * n + n - 1].elapsed_time(tail[(k + 1) * n]) for k in range(a.steps - 1)) / max(1, a.steps - 1)
        mine = {nm: round(v / a.steps, 3) for nm, v in zip(names, acc)}
        mine.update(between_steps=round(between, 3), own_ms_per_step=round(ms_own / a.steps, 3), rank=rank)
        try:
            import pynvml
            pynvml.nvmlInit()
            mine['sm_mhz_now'] = pynvml.nvmlDeviceGetClockInfo(pynvml.nvmlDeviceGetHandleByIndex(local), pynvml.NVML_CLOCK_SM)
        except Exception:
            pass
        if world > 1:
            allr = [None] * world
            dist.all_gather_object(allr, mine)
            per_rank = allr
        else:
            per_rank = [mine]
    hstats = ctx.hybrid_stats()
    ctx.profile(False)
    state['profiling'] = False
    ms_e2e, _, _, _ = timed(step_e2e, a.steps, max(1, a.warmup - 2))
    value = Pg * a.steps / (ms / 1000.0)
    e2e = Pg * a.steps / (ms_e2e / 1000.0)

    out = {'metric': METRIC, 'value': value, 'unit': 'member-evals/s', 'n_gpus': world,
           'steps': a.steps, 'warmup': a.warmup, 'ms_per_step': ms / a.steps, 'higher_is_better': True, 'scaling': 'weak',
           'vs_baseline': None, 'dtype': {'exact': 'f16 operands split hi+lo (~f32 products), f32 accumulate', 'fast': 'f16 operands, f32 accumulate',
                                        'hybrid': 'LSTM: f16 operands split hi+lo (~f32 products); vocabulary projection: f16 operands + f32 re-evaluation of the arg-max candidates; f32 accumulate'}[a.precision],
           'data': 'synthetic', 'config': workload_config(a, world), 'precision': a.precision,
           'clocks': clocks, 'gpu_launches': launches * world,
           'e2e': {'value': e2e, 'unit': 'member-evals/s', 'h2d_bytes_per_step': int(2 * fc_host.numel() * 4 + (P + C) * 16),
                   'd2h_bytes_per_step': int((seq_host.numel() + seq_c.numel()) * 8 + (lp_host.numel() + lp_c.numel()) * 4), 'ms_per_step': ms_e2e / a.steps}}
    if per_rank is not None:
        out['sections_ms_per_rank'] = per_rank
    import traceback

    def guarded(name, fn, *args):
        """the side measurements must not take the headline line down with them; a failure is reported in place"""
        try:
            return fn(*args)
        except Exception as e:
            traceback.print_exc()
            return {'error': '%s failed: %r' % (name, e)}
    nes3 = nes_config3(a, ctx, nes_b200, torch, dist, dev, rank, world, prec) if a.nes_pairs > 0 else None
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        except Exception:
            pass
        out['roofline'] = rooflines(prof, a, min(P, 1024), rows, ms_prof, peaks)
        out['roofline']['measured_in'] = 'a second pass of the same %d steps with per-launch CUDA-event pairs: %.2f ms/step (timed pass: %.2f)' % (a.steps, ms_prof / a.steps, ms / a.steps)
        out['hybrid_stats'] = hstats
        if nes3 is not None:
            out['nes_config3'] = nes3
        if not a.no_extras and world == 1:
            out['roofline']['other_kernels']['k7_gradient'] = guarded('k7', k7_roofline, ctx, torch, dev, peaks)
            out['extras'] = guarded('extras', extras, ctx, nes_b200, torch, dev, peaks)
            ctx.close()                                        # the host legs below fork worker processes: no live device work beside them
            v, cores, n = cpu_evals_per_s(rows, a.sigma, a.cpu_seconds)
            v640, _, n640 = cpu_evals_per_s(5 * rows, a.sigma, a.cpu_seconds / 2)
            out['cpu_baseline'] = {'value': v, 'unit': 'member-evals/s', 'cores': cores, 'kind': 'port',
                                   'sample': '%d members (%d single-threaded processes x %d), %d rows each; oracle port of PolicyNet.evolve (torch normal_) + FCModel greedy decode, Xavier theta' % (cores * n, cores, n, rows),
                                   'value_at_640_rows': v640, 'sample_640': '%d members at %d rows = batch_size x seq_per_img, what the reference really forwards (dataloader.py:38,135-141)' % (cores * n640, 5 * rows),
                                   'master_leg': guarded('cpu master leg', cpu_master_leg)}
        print(json.dumps(out))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def nes_config3(a, ctx, nes_b200, torch, dist, dev, rank, world, prec):
    """BASELINE configs[2]: one NIC-NES iteration (nic_nes_master.py:120-137 + nic_nes_worker.py:131-154) at population F = 5000 noise rows
    = 10 000 rollouts, Adam, antithetic, sharded over the ranks (strong scaling): sample offsets -> this rank's pairs decoded -> fitness
    all-gather (NCCL) -> centred ranks -> K7 over this rank's slice of the parameter axis -> all-reduce -> identical Adam step everywhere.
    theta' must be bit-identical on every rank and equal to the 1-rank result (hash golden)."""
    F, sigma, l2 = a.nes_pairs, 0.01, 1e-7                       # experiments/mscoco_nes.json: noise_stdev 0.01, l2coeff 1e-7... stepsize 0.001
    ctx.set_mutation('')
    nes_b200.FCModel(ctx, parent=0, rng_state=0)                   # theta_0: the reference's Xavier init from a fixed seed, identical on every rank
    nes_b200.Adam(ctx, None, stepsize=0.001)
    g = torch.Generator().manual_seed(1)
    fc = torch.randn(a.rows, FD, generator=g).to(dev)
    lo, hi = nes_b200.shard_bounds(F, world, rank)
    n = hi - lo
    signs = torch.tensor([1.0, -1.0], dtype=torch.float32, device=dev).repeat(n)

    def iteration(it):
        offs = ctx.sample_offsets(it, F, device=True)                                  # identical on every rank (same seed)
        ctx.rollout(offs[lo:hi].repeat_interleave(2), signs, fc, sigma, precision=prec, out_device=dev)
        fit = ctx.logprob_fitness(2 * n, a.rows, device=dev).reshape(n, 2)
        fit = nes_b200.all_gather_fitness(fit, F, world, rank)
        nes_b200.sharded_master_step(ctx, fit, offs, sigma, l2, world, rank, want_ratio=False)

    iteration(0)                                                                       # warm-up, but part of the trajectory that is hashed
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    iteration(1)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    h = theta_hash(torch, ctx.get_theta(0, dtype=torch.float32, device=dev))
    agree_ranks = True
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        hs = [torch.empty_like(h) for _ in range(world)]
        dist.all_gather(hs, h)
        agree_ranks = all(torch.equal(x, hs[0]) for x in hs)
    mine = [int(x) for x in h.tolist()]
    path = os.path.join(ROOT, 'tests', 'golden', 'config3_theta_hash.json')
    key = 'F%d_rows%d_%s' % (F, a.rows, a.precision)
    golden = None
    try:
        golden = json.load(open(path)).get(key)
    except Exception:
        pass
    if a.write_theta_hash and world == 1 and rank == 0:
        d = {}
        try:
            d = json.load(open(path))
        except Exception:
            pass
        d[key] = mine
        os.makedirs(os.path.dirname(path), exist_ok=True)
        json.dump(d, open(path, 'w'), indent=1)
        golden = mine
    ms = t.item()
    return {'workload': 'configs[2]: fc_caption NIC-NES, Adam, antithetic, population %d noise rows = %d rollouts per iteration, %d rows, sharded over %d GPU(s) (strong scaling), '
                        'fitness all-gather + parameter-axis-sharded K7 + all-reduce' % (F, 2 * F, a.rows, world),
            'n_gpus': world, 'pairs': F, 'rollouts': 2 * F, 'iteration_ms': ms, 'rollouts_per_s': 2 * F / (ms * 1e-3), 'scaling': 'strong',
            'theta_hash': mine, 'theta_identical_on_all_ranks': bool(agree_ranks),
            'theta_equals_1rank_golden': None if golden is None else bool(golden == mine),
            'theta_agree': bool(agree_ranks and (golden is None or golden == mine)),
            'timing': 'second of two iterations, CUDA events, max over ranks'}


def rooflines(prof, a, wave, rows, ms_total, peaks):
    """per-kernel roofline of the decode kernels; the headline object is the member GEMM with the larger share of the step.
    `achieved` / `frac` use the ALGORITHMIC bytes SURVEY 8(d) defines: the eps slice of the layer in the fp16 noise table, 2 B per
    element per noise row -- LSTM step (E+R)*5R elements = 5.243 MB, vocabulary projection R*(V+1) = 9.716 MB per member.  What the
    design really streams per launch (the materialised member weights: split fp16 = 4 B / element for the LSTM, fp16 = 2 B for the
    vocabulary projection, plus activations) is `design_bytes_per_launch` / `frac_design_bytes`; ncu's dram bytes per launch is `traffic`."""
    hbm = peaks.get('hbm_gbs', 6650.0)
    tfl = peaks.get('bf16_tflops_sustained', 1400.0)
    src = 'MEASURED_PEAKS.json' if peaks else 'B200_PROFILING.md fallback'
    traffic = {}
    for name in ('r2_traffic.json', 'r1_traffic.json'):
        try:
            traffic = json.load(open(os.path.join(ROOT, 'profiles', name)))
            break
        except Exception:
            pass
    dense_lstm = a.precision in ('hybrid', 'exact') and os.environ.get('NES_FACTORED', '0') in ('', '0')
    dense_logit = a.precision == 'hybrid' and dense_lstm
    ks = {}
    for name, key, flop, bytes_8d, bytes_design, dense, kern in (
            ('lstm', 'lstm', 2.0 * rows * (E + R) * 5 * R * wave, wave * 5.0 * R * (E + R) * 2, wave * (5.0 * R * (E + R) * 4 + rows * (E + R) * 4 + rows * R * 8), dense_lstm,
             ('k_lstm_pair (split-fp16 LSTM step on materialised member weights, clusters of 2 CTAs sharing the activation block by TMA multicast, fused maxout cell)'
              if os.environ.get('NES_LSTM_PAIR', '1') != '0' else 'k_member_dense<LSTM> (split-fp16 LSTM step on materialised member weights, fused maxout cell)')
             if dense_lstm else 'k_member_gemm<LSTM> (factored theta*x + sigma*eps*x, fused maxout cell)'),
            ('logit', 'logit', 2.0 * rows * R * (V + 1) * wave, wave * (V + 1.0) * R * 2, wave * ((V + 1.0) * R * 2 + rows * R * 2), dense_logit,
             'k_member_dense<LOGITC> (fp16 vocabulary projection, smem-resident x block, fused max / sum-exp / candidate lists)' if dense_logit else 'k_member_gemm<LOGIT> (factored vocabulary projection + fused argmax/log-prob)')):
        k = prof[key]
        if not k['count']:
            continue
        t = k['ms'] / k['count'] * 1e-3
        tr = traffic.get(kern.split(' ')[0])
        ks[name] = {'bound': 'hbm', 'kernel': kern, 'achieved': bytes_8d / t / 1e9, 'peak': hbm, 'unit': 'GB/s', 'frac': bytes_8d / t / 1e9 / hbm,
                    'frac_8d': bytes_8d / t / 1e9 / hbm, 'algorithmic_bytes_8d': bytes_8d,
                    'design_bytes_per_launch': bytes_design, 'frac_design_bytes': bytes_design / t / 1e9 / hbm,
                    'peak_source': src + ' hbm_gbs', 'traffic': tr, 'launches': k['count'], 'avg_launch_ms': t * 1e3,
                    'algorithmic_flop_per_launch': flop, 'algorithmic_tflops': flop / t / 1e12,
                    'frac_of_bf16_sustained_on_algorithmic_flop': flop / t / 1e12 / tfl, 'time_share': k['ms'] / ms_total}
    top = max(ks, key=lambda n: ks[n]['time_share'])
    out = dict(ks[top])
    out['other_kernels'] = {n: v for n, v in ks.items() if n != top}
    km = prof.get('materialise')
    if km and km['count']:
        # k_materialise (3 launches per wave: i2h, h2h split fp16; logit.weight fp16): reads the eps slices (2 B / element) and theta (L2),
        # writes 4 B (LSTM) / 2 B (vocabulary) per element
        n_l, n_v = 5.0 * R * (E + R), (V + 1.0) * R
        nbytes = wave * ((n_l + n_v) * 2 + n_l * 4 + n_v * 2)
        n_waves = km['count']                                  # one event pair around all launches of a wave
        t = km['ms'] / n_waves * 1e-3
        out['other_kernels']['materialise'] = {'bound': 'hbm', 'kernel': 'k_materialise x3 + k_materialise_bias x2 per wave (member weights of the LSTM as split fp16, of the vocabulary projection as fp16)',
                                               'achieved': nbytes / t / 1e9, 'peak': hbm, 'unit': 'GB/s', 'frac': nbytes / t / 1e9 / hbm,
                                               'algorithmic_bytes_per_wave': nbytes, 'ms_per_wave': t * 1e3, 'time_share': km['ms'] / ms_total,
                                               'note': 'not on the 8(d) floor at all: this is the price of forming the operands once per rollout instead of per step'}
    out['kernel_time_share'] = {kk: vv['ms'] / ms_total for kk, vv in prof.items()}
    return out


def k7_roofline(ctx, torch, dev, peaks):
    """K7 (NESMaster.gradient_estimate) at F = 5000 rows of this context's table (2^32 elements by default: rows 27 MB long spread over 8 GB,
    no L2 reuse between them) against the HBM roofline: algorithmic bytes = 2 F D read + 4 D written"""
    F = 5000
    offs = ctx.sample_offsets(12345, F, device=True)
    w = torch.randn(F, device=dev)
    ctx.last_gradient_ms()                      # arm the event pair
    for _ in range(2):
        ctx.gradient(offs, w, 0.01, want=False)
    t = []
    for _ in range(5):
        ctx.gradient(offs, w, 0.01, want=False)
        t.append(ctx.last_gradient_ms())
    ms = sum(t) / len(t)
    nbytes = F * ctx.D * 2 + ctx.D * 4
    hbm = peaks.get('hbm_gbs', 6650.0)
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, 'profiles', 'r2_traffic.json'))).get('k_gradient')
    except Exception:
        pass
    return {'bound': 'hbm', 'kernel': 'k_gradient<8,1> (K7: 128-bit streaming weighted reduction over noise-table rows taken in table order, 8 vectors per thread: rows that overlap in the table hit L2, so `achieved` on the algorithmic bytes may exceed the HBM peak; `traffic` is what DRAM really moved)', 'F': F, 'D': ctx.D,
            'table_elements': ctx.table_len, 'avg_launch_ms': ms, 'algorithmic_bytes_8d': nbytes, 'achieved': nbytes / (ms * 1e-3) / 1e9, 'peak': hbm, 'unit': 'GB/s',
            'frac': nbytes / (ms * 1e-3) / 1e9 / hbm, 'frac_of_8TBps_nominal': nbytes / (ms * 1e-3) / 1e9 / 8000.0, 'traffic': traffic}


def extras(ctx, nes_b200, torch, dev, peaks):
    """K8 update, the device CIDEr-D scorer beside the Python one, sampled decode, BASELINE configs[0]"""
    import numpy as np
    ex = {}
    nes_b200.Adam(ctx, None, stepsize=0.001)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx.opt_update(None, 1e-7, want_ratio=False)
    e0.record()
    for _ in range(5):
        ctx.opt_update(None, 1e-7, want_ratio=False)
    e1.record(); torch.cuda.synchronize()
    ex['k8_adam_ms'] = e0.elapsed_time(e1) / 5
    Fp = 500
    g = torch.Generator().manual_seed(1)
    fc = torch.randn(128, FD, generator=g).to(dev)
    signs = torch.tensor([1.0, -1.0], dtype=torch.float32, device=dev).repeat(Fp)
    offs = ctx.sample_offsets(300, Fp, device=True).repeat_interleave(2)
    ctx.rollout(offs, signs, fc, 0.01, precision=nes_b200.PREC_HYBRID, out_device=dev)
    # CIDEr-D reward of one generation's captions: 1000 members x 128 captions against 5 synthetic references per
    # image, on the device from the resident tokens; the reference scores them one rollout at a time in Python
    rs = np.random.RandomState(0)
    gts = []
    for i in range(128):
        refs = np.zeros((5, T), dtype=np.int64)
        for j in range(5):
            ln = rs.randint(6, T)
            refs[j, :ln] = rs.randint(1, V + 1, ln)
        gts.append(refs)
    sc = nes_b200.CiderD(ctx, df='corpus').set_refs(gts, seq_per_img=5)
    n_m = 2 * Fp
    sc.scores(None, n_m, 128, device=dev)
    torch.cuda.synchronize()
    e0.record()
    for i in range(3):
        sc.scores(None, n_m, 128, device=dev)
    e1.record(); torch.cuda.synchronize()
    ms_c = e0.elapsed_time(e1) / 3
    from oracle import cider as ocider
    seq_host = rs.randint(0, V + 1, (128, T))
    t0 = time.perf_counter()
    for i in range(3):
        ocider.cider_scores(seq_host, gts)
    cpu_s = (time.perf_counter() - t0) / 3
    ex['cider_d'] = {'captions': n_m * 128, 'ms': ms_c, 'captions_per_s': n_m * 128 / (ms_c * 1e-3),
                     'cpu_port_captions_per_s_one_core': 128 / cpu_s, 'refs_per_image': 5}
    # sampled decode (Fitness.SAMPLE / SELF_CRITICAL / SC_LOSS): the same 1000 members drawing their tokens from the full
    # distribution (all accurate logits + k_sample) instead of the arg-max
    u = torch.rand((n_m, 128, T), dtype=torch.float64, device=dev)
    ctx.rollout(offs, signs, fc, 0.01, out_device=dev, uniforms=u)
    torch.cuda.synchronize()
    e0.record()
    for i in range(2):
        ctx.rollout(offs, signs, fc, 0.01, out_device=dev, uniforms=u)
    e1.record(); torch.cuda.synchronize()
    ex['sampled_decode'] = {'rollouts': n_m, 'rows': 128, 'ms': e0.elapsed_time(e1) / 2, 'rollouts_per_s': n_m * 2 / (e0.elapsed_time(e1) * 1e-3)}
    ex['mnist_es'] = mnist_extras(nes_b200, torch, dev)
    return ex


def mnist_extras(nes_b200, torch, dev):
    """BASELINE configs[0] (experiments/mnist_es.json: nb_offspring 50, population_size 10, num_elites 3, num_elite_cands 2, batch 64, sigma 0.05;
    plain mutation here -- the shipped SM-G-SUM needs one autograd pass per new parent and generation, tests/test_gpu_mnist.py runs it): one
    NIC-ES generation on the device (k_mnist_rollout + podium + selection + survivors), the rollout alone at population 5000, and the oracle
    port of evolve + ClfPolicy.rollout on ONE host core -- the reference runs this config with a single CPU worker"""
    import numpy as np
    from oracle import mnist as omnist
    exp = {'nb_offspring': 50, 'population_size': 10, 'num_elites': 3, 'num_elite_cands': 2, 'selection': 'uniform',
           'config': {'noise_stdev': 0.05, 'batch_size': 64, 'patience': 0}}
    mu, P, sigma, rows = 10, 50, 0.05, 64
    mctx = nes_b200.NesContext(tensors=nes_b200.layout.mnist_tensors(), seed=4321, table_len=1 << 24, n_parents=nes_b200.ESPopulation.slots_needed(exp))
    pop = nes_b200.ESPopulation(mctx, exp)
    pop.init_from_thetas([omnist.seeded_theta(k) for k in range(mu)])
    g = torch.Generator().manual_seed(2)
    images = torch.randn(rows, 1, 28, 28, generator=g).to(dev)
    labels = torch.randint(0, 10, (rows,), generator=g).to(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for it in range(3):
        nes_b200.es_generation_mnist(pop, images, labels)
    torch.cuda.synchronize()
    e0.record()
    for it in range(3, 13):
        nes_b200.es_generation_mnist(pop, images, labels)
    e1.record(); torch.cuda.synchronize()
    gen_ms = e0.elapsed_time(e1) / 10
    big = 5000
    offs = mctx.sample_offsets(77, big, device=True)
    ones = torch.ones(big, device=dev)
    slots = torch.tensor(pop.parent_slots(), dtype=torch.int32, device=dev)
    pars = slots[mctx.sample_parents(77, big, len(pop.parent_slots()), device=True).long()].contiguous()
    mctx.mnist_rollout(offs, ones, images, labels, sigma, parents=pars, out_device=dev)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        mctx.mnist_rollout(offs, ones, images, labels, sigma, parents=pars, out_device=dev)
    e1.record(); torch.cuda.synchronize()
    roll_ms = e0.elapsed_time(e1) / 5
    # host: the oracle port, one core (torch threads pinned to 1 like src/main.py:8-11)
    nt = torch.get_num_threads()
    torch.set_num_threads(1)
    th = torch.from_numpy(omnist.seeded_theta(0))
    im, lb = images.cpu().numpy(), labels.cpu().numpy()
    t0 = time.perf_counter()
    n_cpu = 200
    for _ in range(n_cpu):
        member = th + torch.empty_like(th).normal_(mean=0, std=sigma)          # PolicyNet.evolve (algorithm/nets.py:101-113)
        omnist.rollout(member.numpy(), im, lb)
    cpu_s = (time.perf_counter() - t0) / n_cpu
    torch.set_num_threads(nt)
    mctx.close()
    return {'workload': 'configs[0]: MnistNet NIC-ES, nb_offspring 50, population_size 10, num_elites 3, num_elite_cands 2, 64 images / rollout, sigma 0.05, plain mutation',
            'generation_ms': gen_ms, 'generation_evals_per_s': P / (gen_ms * 1e-3),
            'rollout_5000_members_ms': roll_ms, 'rollout_evals_per_s': big / (roll_ms * 1e-3),
            'rollout_gflops': big * rows * 2 * (1440 * 100 + 320 * 1000 + 3200) / (roll_ms * 1e-3) / 1e9,
            'cpu_port_evals_per_s_one_core': 1.0 / cpu_s, 'cpu_sample': '%d evolve + rollout pairs, 1 thread' % n_cpu}


if __name__ == '__main__':
    main()
