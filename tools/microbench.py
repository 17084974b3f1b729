#!/usr/bin/env python
"""Kernel micro-benchmarks on the GPU box (CUDA-event timed, graph-replayed): prints one JSON line per case."""
import ctypes as C, json, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from whisper_rust_b200 import api
import torch  # noqa: F401  (initialises CUDA)
lib = api.load_library()
f = lib.whisper_b200_bench_kernel
f.restype = C.c_int
f.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
PEAK = 6552.0
try:
    PEAK = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]
except Exception:
    pass
def run(kind, n, a, b, c, iters):
    us, by = C.c_double(), C.c_double()
    r = f(kind.encode(), n, a, b, c, iters, C.byref(us), C.byref(by))
    gbs = by.value / (us.value * 1e-6) / 1e9 if r == 0 and us.value > 0 else 0
    print(json.dumps({"kind": kind, "n": n, "a": a, "b": b, "c": c, "rc": r, "us": round(us.value, 3), "GBs": round(gbs, 1), "frac_hbm": round(gbs / PEAK, 3)}), flush=True)
for n in (1, 16, 64):
    run("cross_attn_step", n, 20, 1500, 8 if n > 16 else 32, 5)
for n in (16, 64):
    for (no, k) in ((1280, 1280), (3840, 1280), (5120, 1280), (1280, 5120)):
        run("gemm_wstream", n, no, k, 48, 5)
