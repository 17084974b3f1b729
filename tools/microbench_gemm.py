#!/usr/bin/env python
"""Encoder-GEMM micro-benchmark (CUDA-event timed): one JSON line per shape.  Env WB200_EPI / WB200_GEMM select kernel variants."""
import ctypes as C, json, os, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from whisper_rust_b200 import api
import torch  # noqa: F401
lib = api.load_library()
g = lib.whisper_b200_bench_gemm
g.restype = C.c_int
g.argtypes = [C.c_int] * 8 + [C.POINTER(C.c_double)]
TF_PEAK = 1400.0
try:
    TF_PEAK = json.load(open("MEASURED_PEAKS.json"))["bf16_tflops_sustained"]
except Exception:
    pass
def gemm(name, m, n, k, f16, flags, bn=0, iters=10):
    us = C.c_double()
    r = g(m, n, k, f16, flags, bn, 4, iters, C.byref(us))
    tf = 2.0 * m * n * k / (us.value * 1e-6) / 1e12 if r == 0 and us.value > 0 else 0
    print(json.dumps({"kind": "gemm_enc", "epi": os.environ.get("WB200_EPI", "0"), "name": name, "m": m, "n": n, "k": k, "f16": f16, "flags": flags,
                      "bn": bn, "rc": r, "us": round(us.value, 2), "TFs": round(tf, 1), "frac_tensor": round(tf / TF_PEAK, 3)}), flush=True)
Ws = [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "64").split(",")]
for W in Ws:
    R = W * 1500
    gemm("qk", R, 2560, 1280, 1, 2 | 4)
    gemm("o_proj", R, 1280, 1280, 0, 1 | 4)
    gemm("fc1", R, 5120, 1280, 1, 4)
    gemm("fc2", R, 1280, 5120, 0, 1 | 4)
    gemm("epi_only_f32", R, 1280, 64, 0, 1 | 4)
    gemm("epi_only_f16", R, 2560, 64, 1, 4)
