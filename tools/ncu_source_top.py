#!/usr/bin/env python
"""experiment: top source lines of a kernel by warp-stall samples: python tools/ncu_source_top.py report.ncu-rep [n]
(reads `ncu -i X --page source --csv`; the report must have been taken with --import-source on from a -lineinfo build)"""
import csv
import subprocess
import sys
out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = None
data = []
for r in rows:
    if hdr is None:
        if any('Sampling' in c for c in r) and any(c == 'Source' for c in r):
            hdr = r
        continue
    if len(r) == len(hdr):
        data.append(r)
if hdr is None:
    print('no source table found'); print(out[:3000]); sys.exit(0)
ix = {h: i for i, h in enumerate(hdr)}
samp = [h for h in hdr if 'Sampling' in h and 'All' in h]
key = samp[0] if samp else [h for h in hdr if 'Sampling' in h][0]
inst = [h for h in hdr if h.startswith('# Instructions Executed') or h == 'Instructions Executed']
def num(x):
    try: return float(x.replace(',', ''))
    except Exception: return 0.0
tot = sum(num(r[ix[key]]) for r in data) or 1.0
print('columns:', hdr)
print('column:', key, 'total samples', tot)
for r in sorted(data, key=lambda r: -num(r[ix[key]]))[:n]:
    print('%6.2f%%  %s' % (100 * num(r[ix[key]]) / tot, r[ix['Source']].strip()[:170]))
