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
            lanes.append({"lane": f"{l['lane']}.{sub}", "n": l["n"], "t_ns": [l["t_ns"][i] for i in idx], "names": [l["names"][i] for i in idx],
                          "t_end_ns": [l["t_end_ns"][i] for i in idx] if "t_end_ns" in l else None})
    for l in lanes:
        t, names = l["t_ns"], l["names"]
        dur = defaultdict(float); cnt = defaultdict(int)
        for i in range(len(t) - 1):
            dur[names[i]] += (t[i + 1] - t[i]) / 1e3; cnt[names[i]] += 1
        tot = (t[-1] - t[0]) / 1e3
        print(f"lane {l['lane']} n={l['n']} start={(t[0]-t0)/1e3:9.1f}us span={tot:9.1f}us  " +
              "  ".join(f"{k}:{dur[k]/cnt[k]:.1f}x{cnt[k]}" for k in sorted(dur)))
        te = l.get("t_end_ns")
        if te:   # block 0's own run time (start stamp -> its exit) and what follows it until the next kernel's start stamp
            run = defaultdict(float); gap = defaultdict(float)
            for i in range(len(t) - 1):
                if te[i] >= t[i]:
                    run[names[i]] += (te[i] - t[i]) / 1e3; gap[names[i]] += (t[i + 1] - te[i]) / 1e3
            print("      block-0 run / gap to next start:  " + "  ".join(f"{k}:{run[k]/cnt[k]:.1f}/{gap[k]/cnt[k]:.1f}" for k in sorted(run)))
    if len(sys.argv) > 2:
        l = d["lanes"][0]
        for i in range(int(sys.argv[2])):
            print(l["names"][i], (l["t_ns"][i] - t0) / 1e3)
    # ---- concurrency of the K/V stream across lanes: a lane's cross interval = [its cross stamp, its next kernel's stamp)
    iv = []
    for l in lanes:
        t, names = l["t_ns"], l["names"]
        iv += [(t[i], t[i + 1]) for i in range(len(t) - 1) if names[i] == "cross"]
    if len(lanes) > 1 and iv:
        lo = max(l["t_ns"][0] for l in lanes); hi = min(l["t_ns"][-1] for l in lanes)   # window in which every lane is mid-step
        ev = sorted([(max(a, lo), 1) for a, b in iv if b > lo and a < hi] + [(min(b, hi), -1) for a, b in iv if b > lo and a < hi])
        occ = defaultdict(float); cur = 0; prev = lo
        for t, dlt in ev:
            occ[cur] += t - prev; prev = t; cur += dlt
        occ[cur] += hi - prev
        tot = max(1.0, float(hi - lo))
        print(f"  common window {tot/1e3:.1f}us; lanes inside cross_attn_step at once: " + "  ".join(f"{k}:{v/tot:.3f}" for k, v in sorted(occ.items())))
