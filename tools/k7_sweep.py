#!/usr/bin/env python
"""experiment: K7 (k_gradient) at F = 5000, D = 13.4 M on a 2^32 table under the NES_K7_UNROLL switch (read at first use,
so one process per setting): prints ms and GB/s"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nes_b200
D, F = 13400848, 5000
ctx = nes_b200.NesContext(counts=[D], flat_granule=int(os.environ.get('K7_GRANULE', '2048')), seed=1234, table_len=1 << 32, device=0)
dev = torch.device('cuda', 0)
offs = ctx.sample_offsets(1, F, device=True)
w = torch.randn(F, device=dev)
ctx.last_gradient_ms()
t = []
for i in range(8):
    ctx.gradient(offs, w, 0.01, want=False)
    if i >= 3:
        t.append(ctx.last_gradient_ms())
ms = sum(t) / len(t)
gb = (F * D * 2 + D * 4) / 1e9
print('unroll %s: %.2f ms %.0f GB/s (%.3f of 8 TB/s)' % (os.environ.get('NES_K7_UNROLL', '16'), ms, gb / ms * 1e3, gb / ms * 1e3 / 8000))
