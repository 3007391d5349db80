#!/usr/bin/env python
"""experiment: time and per-kernel-class breakdown of 1000 rollouts at the configs[1] shape in three variants -- sampled decode (Fitness.SAMPLE),
greedy on the shared batch, greedy with one batch per member (`single_batch: false`, the shipped mscoco_nes.json:21): python tools/decode_variants.py"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nes_b200  # noqa: E402

V, E, R, FD, T, n_m, rows = 9487, 512, 512, 2048, 16, 1000, 128
dev = torch.device('cuda:0')
ctx = nes_b200.NesContext(dict(vocab_size=V, input_encoding_size=E, rnn_size=R, fc_feat_size=FD), seed=1234, table_len=1 << 30, n_parents=1,
                          wave_members=n_m, rows=rows, device=0)
ctx.set_theta(nes_b200.FCModel(ctx, parent=0, rng_state=0).parameter_vector())
g = torch.Generator().manual_seed(1)
fc_shared = torch.randn(rows, FD, generator=g).to(dev)
signs = torch.tensor([1.0, -1.0], dtype=torch.float32, device=dev).repeat(n_m // 2)
offs = ctx.sample_offsets(7, n_m // 2, device=True).repeat_interleave(2)
u = torch.rand((n_m, rows, T), dtype=torch.float64, device=dev)
fc_each = torch.randn(n_m, rows, FD, generator=g).to(dev)
for mode in ('sampled', 'greedy', 'greedy_per_member_batches'):
    kw = dict(uniforms=u) if mode == 'sampled' else dict(precision=nes_b200.PREC_HYBRID)
    fc = fc_each if mode == 'greedy_per_member_batches' else fc_shared
    ctx.rollout(offs, signs, fc, 0.01, out_device=dev, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(2):
        ctx.rollout(offs, signs, fc, 0.01, out_device=dev, **kw)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 2
    ctx.profile(True)
    ctx.rollout(offs, signs, fc, 0.01, out_device=dev, **kw)
    torch.cuda.synchronize()
    pr = ctx.profile_read()
    ctx.profile(False)
    print(json.dumps({'mode': mode, 'ms_per_1000_rollouts': ms, 'kernel_ms': {k: round(v['ms'], 2) for k, v in pr.items()}, 'kernel_launches': {k: v['count'] for k, v in pr.items()}}))
