#!/usr/bin/env python
"""Summarise WB200_STEP_TRACE dumps: per lane, time per kernel kind inside the concurrent run and the lanes' overlap."""
import json, sys
from collections import defaultdict
for line in open(sys.argv[1]):
    d = json.loads(line)
    t0 = min(min(l["t_ns"]) for l in d["lanes"])
    lanes = []
    for l in d["lanes"]:  # split paired steps into their two halves
        subs = l.get("sub") or [0] * len(l["t_ns"])
        for sub in sorted(set(subs)):
            idx = [i for i, x in enumerate(subs) if x == sub]
            lanes.append({"lane": f"{l['lane']}.{sub}", "n": l["n"], "t_ns": [l["t_ns"][i] for i in idx], "names": [l["names"][i] for i in idx]})
    for l in lanes:
        t, names = l["t_ns"], l["names"]
        dur = defaultdict(float); cnt = defaultdict(int)
        for i in range(len(t) - 1):
            dur[names[i]] += (t[i + 1] - t[i]) / 1e3; cnt[names[i]] += 1
        tot = (t[-1] - t[0]) / 1e3
        print(f"lane {l['lane']} n={l['n']} start={(t[0]-t0)/1e3:9.1f}us span={tot:9.1f}us  " +
              "  ".join(f"{k}:{dur[k]/cnt[k]:.1f}x{cnt[k]}" for k in sorted(dur)))
    if len(sys.argv) > 2:
        l = d["lanes"][0]
        for i in range(int(sys.argv[2])):
            print(l["names"][i], (l["t_ns"][i] - t0) / 1e3)
