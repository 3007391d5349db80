#!/usr/bin/env python
"""Headline benchmark: fc_caption perturbed-member evals/s (bs128 greedy) -- BASELINE.json configs[1]:
NIC-ES, fc_feat 2048, E = R = 512, V = 9487, 128 rows per rollout, population 1000 (per GPU: weak scaling),
mu = 50 parents, synthetic features, random-init weights.

One step = one ES generation: sample (parent, noise offset) per offspring -> decode all offspring
(img_embed, 17 LSTM steps, 16 vocabulary projections with fused argmax/log-prob) -> fitness -> [all-gather
of fitness scalars when N > 1] -> rank -> truncation selection -> materialise the survivors as the next
parents.  CIDEr-D scoring is host-side Python in the reference and outside the path (BASELINE.json); the
fitness here is the sum of greedy log-probs.

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
    ap.add_argument('--rows', type=int, default=128)
    ap.add_argument('--sigma', type=float, default=0.005)
    ap.add_argument('--precision', default='hybrid', choices=['hybrid', 'exact', 'fast'])
    ap.add_argument('--table-log2', type=int, default=28)
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
# CPU arm: the oracle port of the reference's worker recipe (nic_es_worker.py:142-175 minus disk / Redis / CIDEr):
# member = parent + sigma*eps ; FCModel._sample(greedy).  One single-threaded process per host core, exactly how
# the reference uses a node (src/main.py:8-11,105-106).
def _cpu_worker(args):
    seed, n_members, rows, sigma = args
    import numpy as np
    import torch
    torch.set_num_threads(1)
    from oracle import fc_caption, noise
    dims = fc_caption.Dims(V, E, R, FD)
    rs = np.random.RandomState(seed)
    theta = (rs.randn(dims.num_params()) * 0.02).astype(np.float32)
    fc = rs.randn(rows, FD).astype(np.float32)
    t0 = time.perf_counter()
    for _ in range(n_members):
        eps = rs.randn(dims.num_params()).astype(np.float32)       # the reference draws a fresh normal per member (nets.py:102)
        th = theta + np.float32(sigma) * eps
        fc_caption.decode_greedy(dims, th, fc)
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
    sample = '%d members per step (1 single-threaded process per core x %d members), %d rows, oracle port of the reference worker recipe' % (cores * per_step, per_step, a.rows)
    print(json.dumps({
        'impl': 'reference', 'metric': 'fc_caption perturbed-member evals/sec (bs128 greedy)', 'value': value, 'unit': 'member-evals/s',
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
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group('nccl', device_id=dev)
    prec = {'exact': nes_b200.PREC_EXACT, 'fast': nes_b200.PREC_FAST, 'hybrid': nes_b200.PREC_HYBRID}[a.precision]
    P, mu, el, rows = a.population, a.parents, a.elites, a.rows
    Pg = P * world
    ctx = nes_b200.NesContext(dict(vocab_size=V, input_encoding_size=E, rnn_size=R, fc_feat_size=FD), seed=1234,
                              table_len=1 << a.table_log2, n_parents=2 * mu, wave_members=min(P, 1024), rows=rows, device=local)
    # random-init parents (Xavier, algorithm/nets.py:62-69), synthetic N(0,1) fc features
    model = nes_b200.FCModel(ctx, parent=0, rng_state=0)
    for k in range(1, mu):
        ctx.evolve(int(ctx.sample_offsets(999, mu)[k]), 0.01, 1.0, parent=0, dst_parent=k, want_member=False, want_noise=False)
    g = torch.Generator().manual_seed(0)
    fc_host = torch.randn(rows, FD, generator=g).pin_memory()
    fc_dev = fc_host.to(dev)
    ones = torch.ones(P, dtype=torch.float32, device=dev)
    lo = rank * P
    state = {'base': 0, 'it': 0}

    def sample_generation(it):
        offs, pars = nes_b200.es_sample_generation(ctx, it, Pg, mu)
        if a.debug_same_offset:
            offs = offs[:1].expand(Pg).contiguous()
        return offs, pars

    def select_and_materialise(fit_all, offs, pars):
        state['base'] = nes_b200.es_select_and_materialise(ctx, fit_all, offs, pars, mu, el, state['base'], a.sigma)

    sections = []

    def mark():
        if a.sections:
            e = torch.cuda.Event(enable_timing=True); e.record(); sections.append(e)

    def step_resident():
        it = state['it']; state['it'] += 1
        mark()
        offs, pars = sample_generation(it)
        mark()
        ctx.rollout(offs[lo:lo + P], ones, fc_dev, a.sigma, parents=pars[lo:lo + P] + state['base'], precision=prec, out_device=dev)
        mark()
        fit = ctx.logprob_fitness(P, rows, device=dev)
        if world > 1:
            out = torch.empty(Pg, dtype=torch.float64, device=dev)
            dist.all_gather_into_tensor(out, fit)
            fit = out
        mark()
        select_and_materialise(fit, offs, pars)
        mark()

    seq_host = torch.empty((P, rows, T), dtype=torch.int64).pin_memory()
    lp_host = torch.empty((P, rows, T), dtype=torch.float32).pin_memory()

    def step_e2e():
        """through the host boundary the reference has: features come from host memory, tokens / log-probs go back
        to the host where the scorer lives, the host returns fitness scalars"""
        it = state['it']; state['it'] += 1
        offs, pars = sample_generation(it)
        lib, h = ctx.lib, ctx._h
        par_local = (pars[lo:lo + P] + state['base']).contiguous()
        off_local = offs[lo:lo + P].contiguous()
        nes_b200._lib.check(lib.nes_rollout(h, P, nes_b200._lib.ptr(off_local), nes_b200._lib.ptr(ones), nes_b200._lib.ptr(par_local),
                                            nes_b200._lib.ptr(fc_host), rows, 1, float(a.sigma), int(prec),
                                            nes_b200._lib.ptr(seq_host), nes_b200._lib.ptr(lp_host)))
        fit = lp_host.view(P, -1).sum(1, dtype=torch.float64)          # host-side scorer stand-in
        fit = fit.to(dev, non_blocking=True)
        if world > 1:
            out = torch.empty(Pg, dtype=torch.float64, device=dev)
            dist.all_gather_into_tensor(out, fit)
            fit = out
        select_and_materialise(fit, offs, pars)

    def timed(fn, steps, warmup, sample_clocks=False, profile=False):
        sampler = early_sampler if sample_clocks else None
        for _ in range(warmup):
            fn()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        if profile:
            ctx.profile(True)          # per-kernel CUDA-event pairs cover the timed steps only
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
        return t.item(), ctx.gpu_launches - l0, clocks

    # pass 1: the timed region (no per-kernel events in the stream) -> value; pass 2: the same K steps again with a CUDA-event pair
    # around every launch (nes_profile_*) -> per-kernel averages of the roofline object
    ms, launches, clocks = timed(step_resident, a.steps, a.warmup, sample_clocks=True, profile=False)
    ms_prof, _, _ = timed(step_resident, a.steps, 0, profile=True)
    prof = ctx.profile_read()
    if a.sections and rank == 0:
        n = 5
        tail = sections[-n * a.steps:]
        names = ['sample', 'rollout', 'fitness+gather', 'select+materialise']
        acc = [0.0] * 4
        for k in range(a.steps):
            for j in range(4):
                acc[j] += tail[k * n + j].elapsed_time(tail[k * n + j + 1])
        between = sum(tail[k * n + 4].elapsed_time(tail[(k + 1) * n]) for k in range(a.steps - 1)) / max(1, a.steps - 1)
        sys.stderr.write('sections ms/step: %s  between-steps %.2f\n' % ({nm: round(v / a.steps, 2) for nm, v in zip(names, acc)}, between))
    hstats = ctx.hybrid_stats()
    ctx.profile(False)
    ms_e2e, _, _ = timed(step_e2e, a.steps, max(1, a.warmup - 2))
    value = Pg * a.steps / (ms / 1000.0)
    e2e = Pg * a.steps / (ms_e2e / 1000.0)

    out = {'metric': 'fc_caption perturbed-member evals/sec (bs128 greedy)', 'value': value, 'unit': 'member-evals/s', 'n_gpus': world,
           'steps': a.steps, 'warmup': a.warmup, 'ms_per_step': ms / a.steps, 'higher_is_better': True, 'scaling': 'weak',
           'vs_baseline': None, 'dtype': {'exact': 'f16 operands split hi+lo (~f32 products), f32 accumulate', 'fast': 'f16 operands, f32 accumulate',
                                        'hybrid': 'LSTM: f16 operands split hi+lo (~f32 products); vocabulary projection: f16 operands + f32 re-evaluation of the arg-max candidates; f32 accumulate'}[a.precision],
           'data': 'synthetic', 'config': dict(workload_config(a, world), precision=a.precision),
           'clocks': clocks, 'gpu_launches': launches * world,
           'e2e': {'value': e2e, 'unit': 'member-evals/s', 'h2d_bytes_per_step': int(fc_host.numel() * 4 + P * 8),
                   'd2h_bytes_per_step': int(seq_host.numel() * 8 + lp_host.numel() * 4 + (mu - el) * 24), 'ms_per_step': ms_e2e / a.steps}}
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        except Exception:
            pass
        out['roofline'] = rooflines(prof, a, min(P, 1024), rows, ms_prof, peaks)
        out['roofline']['measured_in'] = 'a second pass of the same %d steps with per-launch CUDA-event pairs: %.2f ms/step (timed pass: %.2f)' % (a.steps, ms_prof / a.steps, ms / a.steps)
        out['hybrid_stats'] = hstats
        if not a.no_extras and world == 1:
            out['extras'] = extras(ctx, nes_b200, torch, dev, peaks)
            v, cores, n = cpu_evals_per_s(rows, a.sigma, a.cpu_seconds)
            out['cpu_baseline'] = {'value': v, 'unit': 'member-evals/s', 'cores': cores, 'kind': 'port',
                                   'sample': '%d members (%d single-threaded processes x %d), %d rows each; oracle port of evolve + FCModel greedy decode' % (cores * n, cores, n, rows)}
        print(json.dumps(out))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def rooflines(prof, a, wave, rows, ms_total, peaks):
    """per-kernel roofline of the two member GEMMs; the headline object is the one with the larger share of the step.
    Algorithmic work per launch: LSTM step = 2*rows*(E+R)*5R FLOP per member; its bytes are the
    member's LSTM weights at float32 size (the precision the path computes in: split-fp16 hi+lo is 4 B/element) plus
    the activations in / (h, c) out.  Vocabulary projection = 2*rows*R*(V+1) FLOP per member; HYBRID streams the member's
    fp16 logit weights (2 B/element) plus the fp16 activations."""
    hbm = peaks.get('hbm_gbs', 6650.0)
    tfl = peaks.get('bf16_tflops_sustained', 1400.0)
    src = 'MEASURED_PEAKS.json' if peaks else 'B200_PROFILING.md fallback'
    traffic = {}
    try:
        traffic = json.load(open(os.path.join(ROOT, 'profiles', 'r1_traffic.json')))
    except Exception:
        pass
    dense_lstm = a.precision in ('hybrid', 'exact') and os.environ.get('NES_FACTORED', '0') in ('', '0')
    dense_logit = a.precision == 'hybrid' and dense_lstm
    ks = {}
    for name, key, flop, nbytes, dense, kern in (
            ('lstm', 'lstm', 2.0 * rows * (E + R) * 5 * R * wave, wave * (5.0 * R * (E + R) * 4 + rows * (E + R) * 4 + rows * R * 8), dense_lstm,
             'k_member_dense<LSTM> (split-fp16 LSTM step on materialised member weights, fused maxout cell)' if dense_lstm else 'k_member_gemm<LSTM> (factored theta*x + sigma*eps*x, fused maxout cell)'),
            ('logit', 'logit', 2.0 * rows * R * (V + 1) * wave, wave * ((V + 1.0) * R * 2 + rows * R * 2), dense_logit,
             'k_member_dense<LOGITC> (fp16 vocabulary projection, smem-resident x block, fused max / sum-exp / candidate lists)' if dense_logit else 'k_member_gemm<LOGIT> (factored vocabulary projection + fused argmax/log-prob)')):
        k = prof[key]
        if not k['count']:
            continue
        t = k['ms'] / k['count'] * 1e-3
        if dense:
            ach, peak, unit, bound = nbytes / t / 1e9, hbm, 'GB/s', 'hbm'
        else:
            ach, peak, unit, bound = flop / t / 1e12, tfl, 'TFLOP/s', 'tensor'
        tr = traffic.get(kern.split(' ')[0])
        ks[name] = {'bound': bound, 'kernel': kern, 'achieved': ach, 'peak': peak, 'unit': unit, 'frac': ach / peak,
                    'peak_source': src + (' hbm_gbs' if dense else ' bf16_tflops_sustained'),
                    'traffic': tr, 'launches': k['count'], 'avg_launch_ms': t * 1e3,
                    'algorithmic_bytes_per_launch': nbytes, 'algorithmic_flop_per_launch': flop,
                    'algorithmic_tflops': flop / t / 1e12, 'time_share': k['ms'] / ms_total}
    top = max(ks, key=lambda n: ks[n]['time_share'])
    out = dict(ks[top])
    out['other_kernels'] = {n: v for n, v in ks.items() if n != top}
    out['kernel_time_share'] = {kk: vv['ms'] / ms_total for kk, vv in prof.items()}
    return out


def extras(ctx, nes_b200, torch, dev, peaks):
    """second half of the metric (NES iteration pieces) on the same context: K7 aggregation at F = 5000 rows
    against the HBM roofline, K8 update, and one antithetic NES iteration at population 1000 pairs."""
    import numpy as np
    ex = {}
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
    gb = (F * ctx.D * 2 + ctx.D * 4) / 1e9
    hbm = peaks.get('hbm_gbs', 6650.0)
    ex['k7_gradient'] = {'F': F, 'D': ctx.D, 'ms': ms, 'algorithmic_GB': gb, 'achieved_GBps': gb / (ms * 1e-3), 'peak_GBps': hbm,
                         'frac_of_measured_copy_peak': gb / (ms * 1e-3) / hbm, 'frac_of_8TBps_nominal': gb / (ms * 1e-3) / 8000.0}
    nes_b200.Adam(ctx, None, stepsize=0.001)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx.opt_update(None, 1e-7, want_ratio=False)
    e0.record()
    for _ in range(5):
        ctx.opt_update(None, 1e-7, want_ratio=False)
    e1.record(); torch.cuda.synchronize()
    ex['k8_adam_ms'] = e0.elapsed_time(e1) / 5
    # second half of BASELINE.json's metric: one NIC-NES iteration (nic_nes_master.py:120-137 + nic_nes_worker.py:131-154):
    # sample offsets -> F antithetic pairs = 2F rollouts on the shared batch -> fitness -> centred ranks -> K7 -> L2 + Adam
    Fp = min(500, ctx.n_slots if hasattr(ctx, 'n_slots') else 500)
    g = torch.Generator().manual_seed(1)
    fc = torch.randn(128, FD, generator=g).to(dev)
    signs = torch.tensor([1.0, -1.0], dtype=torch.float32, device=dev).repeat(Fp)
    prec = nes_b200.PREC_HYBRID

    def nes_iter(it):
        offs = ctx.sample_offsets(it, Fp, device=True)
        ctx.rollout(offs.repeat_interleave(2), signs, fc, 0.01, precision=prec, out_device=dev)
        fit = ctx.logprob_fitness(2 * Fp, 128, device=dev).reshape(Fp, 2)
        ctx.master_step(fit, offs, 0.01, 1e-7, want_ratio=False)
    for i in range(2):
        nes_iter(100 + i)
    torch.cuda.synchronize()
    e0.record()
    for i in range(3):
        nes_iter(200 + i)
    e1.record(); torch.cuda.synchronize()
    ex['nes_iteration'] = {'pairs': Fp, 'rollouts': 2 * Fp, 'rows': 128, 'optimizer': 'adam', 'precision': 'hybrid',
                           'ms': e0.elapsed_time(e1) / 3, 'rollouts_per_s': 2 * Fp * 3 / (e0.elapsed_time(e1) * 1e-3)}
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
    offs = ctx.sample_offsets(300, Fp, device=True).repeat_interleave(2)
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
    """BASELINE configs[0] (experiments/mnist_es.json: nb_offspring 50, population_size 10, num_elites 3, batch 64, sigma 0.05): one
    NIC-ES generation on the device (k_mnist_rollout + selection + survivors), the rollout alone at population 5000, and the oracle port
    of evolve + ClfPolicy.rollout on ONE host core -- the reference runs this config with a single CPU worker"""
    import numpy as np
    from oracle import mnist as omnist
    mu, el, P, sigma, rows = 10, 3, 50, 0.05, 64
    mctx = nes_b200.NesContext(tensors=nes_b200.layout.mnist_tensors(), seed=4321, table_len=1 << 24, n_parents=2 * mu)
    for k in range(mu):
        mctx.set_theta(omnist.seeded_theta(k), k)
    g = torch.Generator().manual_seed(2)
    images = torch.randn(rows, 1, 28, 28, generator=g).to(dev)
    labels = torch.randint(0, 10, (rows,), generator=g).to(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    base = 0
    for it in range(3):
        base = nes_b200.es_generation_mnist(mctx, it, P, mu, el, images, labels, sigma, base=base)[0]
    torch.cuda.synchronize()
    e0.record()
    for it in range(3, 13):
        base = nes_b200.es_generation_mnist(mctx, it, P, mu, el, images, labels, sigma, base=base)[0]
    e1.record(); torch.cuda.synchronize()
    gen_ms = e0.elapsed_time(e1) / 10
    big = 5000
    offs = mctx.sample_offsets(77, big, device=True)
    ones = torch.ones(big, device=dev)
    pars = mctx.sample_parents(77, big, mu, device=True)
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
    th = omnist.seeded_theta(0)
    im, lb = images.cpu().numpy(), labels.cpu().numpy()
    rs = np.random.RandomState(0)
    t0 = time.perf_counter()
    n_cpu = 200
    for _ in range(n_cpu):
        member = th + np.float32(sigma) * rs.randn(omnist.D).astype(np.float32)          # PolicyNet.evolve
        omnist.rollout(member, im, lb)
    cpu_s = (time.perf_counter() - t0) / n_cpu
    torch.set_num_threads(nt)
    mctx.close()
    return {'workload': 'configs[0]: MnistNet NIC-ES, nb_offspring 50, population_size 10, num_elites 3, 64 images / rollout, sigma 0.05',
            'generation_ms': gen_ms, 'generation_evals_per_s': P / (gen_ms * 1e-3),
            'rollout_5000_members_ms': roll_ms, 'rollout_evals_per_s': big / (roll_ms * 1e-3),
            'rollout_gflops': big * rows * 2 * (1440 * 100 + 320 * 1000 + 3200) / (roll_ms * 1e-3) / 1e9,
            'cpu_port_evals_per_s_one_core': 1.0 / cpu_s, 'cpu_sample': '%d evolve + rollout pairs, 1 thread' % n_cpu}


if __name__ == '__main__':
    main()
