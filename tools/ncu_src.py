#!/usr/bin/env python
"""Top stall locations from `ncu --page source --csv` (SASS view): prints the N instructions with the most samples."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
hdr = rows[1]
idx = {h: i for i, h in enumerate(hdr)}
body = [(i, r) for i, r in enumerate(rows[2:]) if len(r) >= len(hdr) - 2]
def f(x):
    try: return float(x)
    except Exception: return 0.0
tot = sum(f(r[idx['# Samples']]) for _, r in body)
print('kernel:', rows[0][1][:100], ' total samples', tot, ' instructions', len(body))
for i, r in sorted(body, key=lambda t: -f(t[1][idx['# Samples']]))[:n]:
    print(f"{i:5d} samples={r[idx['# Samples']]:>6s} exec={r[idx['Instructions Executed']]:>7s} | {r[idx['Source']].strip()[:120]}")
