#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (count, total, mean, share)."""
import collections, csv, sys
hdr, agg = None, collections.defaultdict(lambda: [0, 0.0])
for r in csv.reader(open(sys.argv[1])):
    if hdr is None:
        if r and r[0] == 'ID':
            hdr = r
        continue
    d = dict(zip(hdr, r))
    if d.get('Metric Name') != 'gpu__time_duration.sum':
        continue
    v = float(d['Metric Value'].replace(',', ''))
    v = v / 1000 if d['Metric Unit'] == 'ns' else (v * 1000 if d['Metric Unit'] == 'ms' else v)
    k = d['Kernel Name'].split('(')[0][:70] + ' grid=' + d['Grid Size']
    agg[k][0] += 1
    agg[k][1] += v
tot = sum(v[1] for v in agg.values())
print(f"# {sys.argv[1]}: {sum(v[0] for v in agg.values())} launches, {tot:.1f} us total (cold-cache, serialised: compare shares)")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:100s} n={v[0]:5d} total_us={v[1]:10.1f} avg_us={v[1] / v[0]:8.2f} share={v[1] / tot:6.3f}")
