#!/usr/bin/env python
"""experiment: split the selection + survivor write-back phase of a NIC-ES generation into its parts (CUDA events)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch, nes_b200
V, E, R, FD = 9487, 512, 512, 2048
P, mu, el, rows = 1000, 50, 3, 128
ctx = nes_b200.NesContext(dict(vocab_size=V, input_encoding_size=E, rnn_size=R, fc_feat_size=FD), seed=1234, table_len=1 << 28, n_parents=2 * mu, wave_members=P, rows=rows)
nes_b200.FCModel(ctx, parent=0, rng_state=0)
for k in range(1, mu):
    ctx.evolve(int(ctx.sample_offsets(999, mu)[k]), 0.01, 1.0, parent=0, dst_parent=k, want_member=False, want_noise=False)
fc = torch.randn(rows, FD, generator=torch.Generator().manual_seed(0)).cuda()
ones = torch.ones(P, dtype=torch.float32, device=ctx.dev)
base, sigma = 0, 0.005
acc = np.zeros(4); n = 10
for i in range(n + 2):
    offs, pars = nes_b200.es_sample_generation(ctx, i, P, mu)
    ctx.rollout(offs, ones, fc, sigma, parents=pars + base, precision=nes_b200.PREC_HYBRID, out_device=ctx.dev)
    fit = ctx.logprob_fitness(P, rows, device=ctx.dev)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    ev[0].record()
    keep, _ = nes_b200.ESMaster.selection(fit, mu, el)
    ev[1].record()
    sel = torch.stack([keep, offs[keep], pars[keep].to(torch.int64)], 1).cpu().numpy()
    ev[2].record()
    nxt = mu - base
    m = el + len(sel)
    src = np.concatenate([base + np.arange(el), base + sel[:, 2]]).astype(np.int32)
    off = np.concatenate([np.zeros(el, dtype=np.int64), sel[:, 1]]).astype(np.int64)
    sg = np.concatenate([np.zeros(el), np.full(len(sel), sigma)]).astype(np.float32)
    ev[3].record()
    ctx.evolve_batch(src, off, sg, nxt + np.arange(m))
    ev[4].record()
    torch.cuda.synchronize()
    base = nxt
    if i >= 2:
        acc += [ev[k].elapsed_time(ev[k + 1]) for k in range(4)]
print('sort %.3f  gather+d2h %.3f  host arrays %.3f  evolve_batch (copies, kernel, row norms) %.3f ms' % tuple(acc / n))
