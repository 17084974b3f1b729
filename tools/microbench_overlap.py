#!/usr/bin/env python
"""Decode-lane concurrency probe: a PDL chain of cluster GEMMs timed alone and under a concurrent persistent cross attention."""
import ctypes as C, json, os, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from whisper_rust_b200 import api
import torch  # noqa: F401
lib = api.load_library()
f = lib.whisper_b200_bench_kernel
f.restype = C.c_int
f.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
cases = []
for n in (64, 128):
    for cps, stages in ((2, 5), (1, 5), (2, 3)):
        for nw in (48, 2):
            for cap in (192, 100, 60):
                cases.append((n, cps, stages, nw, cap, 0))
cases.append((128, 1, 5, 48, 100, 1))
cases.append((128, 2, 5, 48, 100, 1))
for n, cps, stages, nw, cap, noprio in cases:
    os.environ["WB200_XS_CPS"], os.environ["WB200_XS_STAGES"], os.environ["WB200_WC_SMEM_KB"] = str(cps), str(stages), str(cap)
    if noprio:
        os.environ["WB200_NO_PRIO"] = "1"
    else:
        os.environ.pop("WB200_NO_PRIO", None)
    us = (C.c_double * 2)()
    alone = C.c_double()
    r = f(b"overlap_gemm", n, 1280, 1280, nw, 5, us, C.byref(alone))
    print(json.dumps({"kind": "overlap_gemm", "n": n, "xs_ctas_per_sm": cps, "xs_stages": stages, "n_weights": nw, "wc_smem_kb": cap,
                      "no_prio": noprio, "rc": r, "gemm_us_alone": round(alone.value, 2), "gemm_us_with_cross": round(us[0], 2),
                      "cross_us_with_gemm": round(us[1], 2)}), flush=True)
