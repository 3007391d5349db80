#!/usr/bin/env python
"""experiment: K7 (k_gradient) at F = 5000, D = 13.4 M on a 2^32 table under the NES_K7 (vectors per thread x rows in flight) and
NES_K7_SORT switches (read at first use, so one process per setting): prints ms and GB/s for the rows as sampled and packed into 256 MB.
    for s in 1x16 2x4 4x2 4x4 8x1 8x2; do NES_K7=$s python tools/k7_sweep.py; done"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nes_b200
D, F = int(os.environ.get('K7_D', '13400848')), int(os.environ.get('K7_F', '5000'))
ctx = nes_b200.NesContext(counts=[D], flat_granule=int(os.environ.get('K7_GRANULE', '2048')), seed=1234, table_len=1 << 32, device=0)
dev = torch.device('cuda', 0)
offs = ctx.sample_offsets(1, F, device=True)
w = torch.randn(F, device=dev)
ctx.last_gradient_ms()
gb = (F * D * 2 + D * 4) / 1e9
# the rows as sampled (the library sorts them by offset unless NES_K7_SORT=0), a quarter of them (rows overlap 4x less), and all
# rows packed into the first 256 MB of the table (L2-resident: the L2 -> SM ceiling of the kernel)
packed = (torch.arange(F, device=dev, dtype=torch.int64) * ctx.G) % ((1 << 27) - ctx.D_pad) // ctx.G * ctx.G
F2 = int(os.environ.get('K7_F2', '1250'))
order = torch.argsort(offs)
for name, o, ww in (('sampled', offs, w), ('sampled', offs, w), ('pre-sorted', offs[order].contiguous(), w[order].contiguous()), ('F=%d' % F2, offs[:F2].contiguous(), w[:F2].contiguous()), ('packed-256MB', packed, w)):
    t = []
    for i in range(8):
        ctx.gradient(o, ww, 0.01, want=False)
        if i >= 3:
            t.append(ctx.last_gradient_ms())
    ms = sum(t) / len(t)
    gbn = (o.numel() * D * 2 + D * 4) / 1e9
    print('K7 %s sort %s rows %-12s: %.2f ms %.0f GB/s (%.3f of 8 TB/s)' % (os.environ.get('NES_K7', 'default'), os.environ.get('NES_K7_SORT', '1'), name, ms, gbn / ms * 1e3, gbn / ms * 1e3 / 8000))
