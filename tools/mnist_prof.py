#!/usr/bin/env python
"""the MNIST population rollout alone (population 5000 x 64 images, configs[0] shapes) -- the program `ncu -k regex:k_mnist_rollout` captures"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nes_b200
from oracle import mnist as omnist

P, rows, mu = int(os.environ.get('MNIST_P', 5000)), 64, 10
ctx = nes_b200.NesContext(tensors=nes_b200.layout.mnist_tensors(), seed=4321, table_len=1 << 24, n_parents=mu)
for k in range(mu):
    ctx.set_theta(omnist.seeded_theta(k), k)
g = torch.Generator().manual_seed(2)
images = torch.randn(rows, 1, 28, 28, generator=g).cuda()
labels = torch.randint(0, 10, (rows,), generator=g).cuda()
offs = ctx.sample_offsets(77, P, device=True)
pars = ctx.sample_parents(77, P, mu, device=True)
ones = torch.ones(P, device='cuda')
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for _ in range(2):
    ctx.mnist_rollout(offs, ones, images, labels, 0.05, parents=pars, out_device='cuda')
torch.cuda.synchronize()
e0.record()
for _ in range(5):
    ctx.mnist_rollout(offs, ones, images, labels, 0.05, parents=pars, out_device='cuda')
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print('mnist rollout: %d members x %d images: %.3f ms, %.0f evals/s, %.1f TFLOP/s fp32' % (P, rows, ms, P / ms * 1e3, P * rows * 2 * 467200 / ms / 1e9))
