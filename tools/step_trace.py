#!/usr/bin/env python
"""experiment: per-step time of bench.py's generation over many consecutive steps (clock / power warm-up behaviour)"""
import os, sys, subprocess, threading, time
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
base = 0
n = int(sys.argv[1]) if len(sys.argv) > 1 else 40
ev = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
clk = []
def poll():
    while not done:
        try:
            out = subprocess.run(['nvidia-smi', '-i', '0', '--query-gpu=clocks.sm,power.draw,temperature.gpu,clocks.mem', '--format=csv,noheader,nounits'], capture_output=True, text=True).stdout.strip()
            clk.append((time.time(), out))
        except Exception:
            pass
        time.sleep(0.2)
done = False
th = threading.Thread(target=poll, daemon=True); th.start()
t0 = time.time()
ev[0].record()
for i in range(n):
    base, _, _, _ = nes_b200.es_generation(ctx, i, P, mu, el, fc, 0.005, base=base, precision=nes_b200.PREC_HYBRID)
    ev[i + 1].record()
torch.cuda.synchronize()
done = True
print('step ms:', ' '.join('%.1f' % ev[i].elapsed_time(ev[i + 1]) for i in range(n)))
print('clocks (t, sm MHz, W, C, mem MHz):', ' | '.join('%.1f %s' % (t - t0, o) for t, o in clk))
