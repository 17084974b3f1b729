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
for n in (16, 64, 256):
    for (no, k) in ((1280, 1280), (3840, 1280), (5120, 1280), (1280, 5120)):
        run("gemm_wstream", n, no, k, 48, 5)
        run("gemm_wcluster", n, no, k, 48, 5)

# encoder GEMM shapes at large-v3, W windows of 1500 rows (tensor roofline)
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
    print(json.dumps({"kind": "gemm_enc", "name": name, "m": m, "n": n, "k": k, "f16": f16, "flags": flags, "bn": bn, "rc": r,
                      "us": round(us.value, 2), "TFs": round(tf, 1), "frac_tensor": round(tf / TF_PEAK, 3)}), flush=True)
for W in (8, 64):
    R = W * 1500
    for bn in (0, 256):   # 0 = persistent (auto), 256 = legacy one-tile-per-CTA kernel
        gemm("qk", R, 2560, 1280, 1, 2 | 4, bn)
        gemm("o_proj", R, 1280, 1280, 0, 1 | 4, bn)
        gemm("fc1", R, 5120, 1280, 1, 4, bn)
        gemm("fc2", R, 1280, 5120, 0, 1 | 4, bn)

# fused encoder attention, large-v3: FLOPs = 4 T^2 64 H W
fa = lib.whisper_b200_bench_enc_attn
fa.restype = C.c_int
fa.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double)]
for W, vt in ((8, 0), (64, 0), (64, 1)):
    us = C.c_double()
    r = fa(W, 20, 1500, 5, vt, C.byref(us))
    fl = 4.0 * 1500 * 1500 * 64 * 20 * W
    tf = fl / (us.value * 1e-6) / 1e12 if r == 0 and us.value > 0 else 0
    print(json.dumps({"kind": "enc_attn", "v_transposed": vt, "W": W, "H": 20, "T": 1500, "rc": r, "us": round(us.value, 1), "us_per_window": round(us.value / W, 1),
                      "TFs": round(tf, 1), "frac_tensor": round(tf / TF_PEAK, 3)}), flush=True)
