#!/usr/bin/env python
"""experiment: where one NIC-ES generation (bench.py's step) spends its time -- CUDA events around the phases of es_generation,
and the per-kernel-kind event profile inside the rollout; mean over n generations after warm-up"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, nes_b200
V, E, R, FD = 9487, 512, 512, 2048
P, mu, el, rows = 1000, 50, 3, 128
ctx = nes_b200.NesContext(dict(vocab_size=V, input_encoding_size=E, rnn_size=R, fc_feat_size=FD), seed=1234, table_len=1 << 28, n_parents=2 * mu, wave_members=P, rows=rows)
nes_b200.FCModel(ctx, parent=0, rng_state=0)
for k in range(1, mu):
    ctx.evolve(int(ctx.sample_offsets(999, mu)[k]), 0.01, 1.0, parent=0, dst_parent=k, want_member=False, want_noise=False)
fc = torch.randn(rows, FD, generator=torch.Generator().manual_seed(0)).cuda()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 12
prof = len(sys.argv) > 2 and sys.argv[2] == 'prof'
names = ['sample', 'rollout', 'fitness', 'select+materialise']
acc = [0.0] * 4
base = 0
ones = torch.ones(P, dtype=torch.float32, device=ctx.dev)
if prof:
    ctx.profile(True)
kinds = {}
for i in range(n + 3):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    ev[0].record()
    offs, pars = nes_b200.es_sample_generation(ctx, i, P, mu)
    ev[1].record()
    ctx.rollout(offs, ones, fc, 0.005, parents=pars + base, precision=nes_b200.PREC_HYBRID, out_device=ctx.dev)
    ev[2].record()
    fit = ctx.logprob_fitness(P, rows, device=ctx.dev)
    ev[3].record()
    base = nes_b200.es_select_and_materialise(ctx, fit, offs, pars, mu, el, base, 0.005)
    ev[4].record()
    torch.cuda.synchronize()
    if prof:
        pr = ctx.profile_read()
    if i >= 3:
        for k in range(4):
            acc[k] += ev[k].elapsed_time(ev[k + 1])
        if prof:
            for k, v in pr.items():
                kinds[k] = kinds.get(k, 0.0) + v['ms']
tot = sum(acc) / n
print('generation %.2f ms:' % tot, ' '.join('%s %.2f' % (nm, a / n) for nm, a in zip(names, acc)))
if prof:
    print('inside the rollout (event-bracketed kernels): ' + ' '.join('%s %.2f' % (k, v / n) for k, v in kinds.items()),
          '| sum %.2f of rollout %.2f' % (sum(kinds.values()) / n, acc[1] / n))
