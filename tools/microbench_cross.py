#!/usr/bin/env python
"""cross_attn_step (persistent) sweep: CTAs per SM x ring depth x rows.  One JSON line per case."""
import ctypes as C, json, os, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from whisper_rust_b200 import api
import torch  # noqa: F401
lib = api.load_library()
f = lib.whisper_b200_bench_kernel
f.restype = C.c_int
f.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
PEAK = 6552.0
try:
    PEAK = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]
except Exception:
    pass
for cps, stages in ((2, 5), (1, 5), (1, 6), (1, 7), (2, 4), (2, 3)):
    os.environ["WB200_XS_CPS"], os.environ["WB200_XS_STAGES"] = str(cps), str(stages)
    for n in (64, 128, 256):
        us, by = C.c_double(), C.c_double()
        r = f(b"cross_attn_step", n, 20, 1500, 8, 5, C.byref(us), C.byref(by))
        gbs = by.value / (us.value * 1e-6) / 1e9 if r == 0 and us.value > 0 else 0
        print(json.dumps({"kind": "cross_attn_step", "ctas_per_sm": cps, "stages": stages, "n": n, "rc": r, "us": round(us.value, 2),
                          "GBs": round(gbs, 1), "frac_hbm": round(gbs / PEAK, 3)}), flush=True)
