#!/usr/bin/env python
"""aggregate an ncu `--metrics gpu__time_duration.sum --csv` launch list by kernel name"""
import collections
import csv
import sys
rows = list(csv.reader(open(sys.argv[1])))
for i, r in enumerate(rows):
    if 'Kernel Name' in r:
        h = i
        break
hdr = rows[h]; ki = hdr.index('Kernel Name'); mi = hdr.index('Metric Value'); ui = hdr.index('Metric Unit')
agg = collections.OrderedDict()
for r in rows[h + 1:]:
    if len(r) <= mi:
        continue
    v = float(r[mi].replace(',', '')) / {'ns': 1e6, 'us': 1e3, 'ms': 1.0, 's': 1e-3}.get(r[ui], 1.0)
    a = agg.setdefault(r[ki][:64], [0, 0.0, []]); a[0] += 1; a[1] += v; a[2].append(v)
tot = sum(a[1] for a in agg.values())
print('total %.3f ms over %d launches' % (tot, sum(a[0] for a in agg.values())))
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print('%-66s n=%4d tot=%9.3f ms avg=%8.4f min=%8.4f max=%8.4f share=%.3f' % (k, a[0], a[1], a[1] / a[0], min(a[2]), max(a[2]), a[1] / tot))
