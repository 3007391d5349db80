#!/usr/bin/env python
"""Measure the BASELINE.json configs that are not bench.py's headline line, on one GPU (or one rank of N under torchrun):
  --config 1 : experiments/mnist_es.json -- MnistNet NIC-ES generation (50 offspring, mu 10, 3 elites, 64 images) + the rollout at population 5000
  --config 3 : NIC-NES, Adam, antithetic, population 5000 pairs (10 000 rollouts) per iteration, E = R = 512
  --config 4 : bu_xent shapes E = R = 128, LayerNorm + affine, 512 rows, population `--pairs`
  --config 5 : K7 aggregation + K8 update sweep, D in 1M..50M x F in 1k..20k, GB/s against the HBM roofline
Prints one JSON line per measurement.  Synthetic features, random-init weights, fitness = sum of greedy log-probs."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import nes_b200  # noqa: E402


def time_ms(fn, n, warm=1):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def nes_config(a, E, R, rows, ln, pairs, sigma, label):
    rank, world = int(os.environ.get('RANK', '0')), int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    group = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group('nccl', device_id=dev)
    mo = dict(vocab_size=9487, input_encoding_size=E, rnn_size=R, fc_feat_size=2048, layer_n=ln, layer_n_affine=ln)
    ctx = nes_b200.NesContext(mo, seed=1234, table_len=1 << a.table_log2, n_parents=1, wave_members=a.wave, rows=rows, device=local)
    nes_b200.FCModel(ctx, parent=0, rng_state=0)
    opt = nes_b200.Adam(ctx, None, stepsize=0.001)
    fc = torch.randn(rows, 2048, generator=torch.Generator().manual_seed(0)).to(dev)
    it = [0]

    def fitness_fn(seq, lp):
        return lp.reshape(lp.shape[0], -1).sum(1, dtype=torch.float64)

    def step():
        it[0] += 1
        nes_b200.nes_iteration(ctx, opt, it[0], pairs, fc, sigma, 1e-7, fitness_fn, world_size=world, rank=rank,
                               precision=nes_b200.PREC_HYBRID, group=group)
    ms = time_ms(step, a.steps)
    ctx.profile(True)                      # one more iteration with an event pair around every launch: where the time goes
    step()
    prof = {k: round(v['ms'], 2) for k, v in ctx.profile_read().items() if v['count']}
    ctx.profile(False)
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = t.item()
    if rank == 0:
        print(json.dumps({'config': label, 'n_gpus': world, 'pairs': pairs, 'rollouts': 2 * pairs, 'rows': rows, 'E': E, 'R': R, 'layer_n': ln,
                          'D': ctx.D, 'iteration_ms': ms, 'rollouts_per_s': 2 * pairs / (ms * 1e-3), 'precision': 'hybrid', 'kernel_ms_rank0': prof}))
    ctx.close()
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def sweep(a):
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:
        pass
    hbm = peaks.get('hbm_gbs', 6650.0)
    rank, world = int(os.environ.get('RANK', '0')), int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group('nccl', device_id=dev)
    for D in (1_000_000, 2_000_000, 5_000_000, 13_400_848, 25_000_000, 50_000_000):
        ctx = nes_b200.NesContext(counts=[D], seed=1234, table_len=1 << a.table_log2, device=local)
        ctx.set_theta(torch.randn(D) * 0.02)
        nes_b200.Adam(ctx, None, stepsize=0.001)
        for F in (1000, 2000, 5000, 10000, 20000):
            offs = ctx.sample_offsets(F, F, device=True)
            w = torch.randn(F, device=dev)
            if world > 1:
                # N GPUs: every rank sums all F rows for its slice of the parameter axis, then one all-reduce gives every rank the whole
                # vector (master.sharded_master_step); timed on the device as the max over ranks, the all-reduce included
                def k7():
                    g = ctx.gradient(offs, w, 0.01, out_device=dev, shard=(rank, world))
                    dist.all_reduce(g)
                tt = torch.tensor([time_ms(k7, 3)], dtype=torch.float64, device=dev)
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                ms = tt.item()
            else:
                ctx.last_gradient_ms()
                t = []
                for i in range(4):
                    ctx.gradient(offs, w, 0.01, want=False)
                    if i:
                        t.append(ctx.last_gradient_ms())
                ms = sum(t) / len(t)
            gb = (F * D * 2 + D * 4) / 1e9
            ms8 = time_ms(lambda: ctx.opt_update(None, 1e-7, want_ratio=False), 3)
            if rank == 0:
              print(json.dumps({'config': '5: K7 + K8 sweep', 'n_gpus': world, 'table_log2': a.table_log2, 'D': D, 'F': F, 'k7_ms': ms, 'k7_GBps': gb / (ms * 1e-3), 'k7_frac_of_measured_peak': gb / (ms * 1e-3) / hbm,
                              'k7_frac_of_8TBps': gb / (ms * 1e-3) / 8000.0, 'k8_adam_ms': ms8, 'k8_GBps_approx_56B_per_param': 56.0 * D / 1e9 / (ms8 * 1e-3)}))
        ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    ap = argparse.ArgumentParser()
    ap.add_argument('--config', type=int, required=True)
    ap.add_argument('--pairs', type=int, default=0)
    ap.add_argument('--wave', type=int, default=1024)
    ap.add_argument('--steps', type=int, default=2)
    ap.add_argument('--table-log2', type=int, default=32)
    a = ap.parse_args()
    if a.config == 1:
        sys.path.insert(0, ROOT)
        import bench
        print(json.dumps({'config': '1: mnist_es.json', **bench.mnist_extras(nes_b200, torch, torch.device('cuda', 0))}))
    elif a.config == 3:
        nes_config(a, 512, 512, 128, False, a.pairs or 5000, 0.01, '3: fc_caption NIC-NES Adam antithetic, population 5000')
    elif a.config == 4:
        nes_config(a, 128, 128, 512, True, a.pairs or 1250, 0.01, '4: bu_xent shapes, LayerNorm + affine, 512 rows')
    elif a.config == 5:
        sweep(a)
