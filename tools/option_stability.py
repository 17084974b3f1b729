#!/usr/bin/env python
"""Runs tests/option_cases.py on the reference's two ISA builds (AVX2 and -march=native) and reports, per case, whether they agree
(ids, times, text) and how many tokens / segments each call produced.  Test infrastructure (oracle only).
Usage: python tools/option_stability.py [case ...]"""
import sys

sys.path.insert(0, "."); sys.path.insert(0, "tests")
import conftest as cf  # noqa: E402
import option_cases as oc  # noqa: E402
from whisper_rust_b200 import WhisperContext, WhisperContextParameters, api  # noqa: E402

libs = {}
for name, path in (("avx2", cf.ORACLE_LIB), ("native", cf.ORACLE_LIB.with_name("libwhisper_ref_native.so"))):
    l = cf.declare(api.load_library(path), cf._PROBE_SIGS)
    api.install_quiet_logging(l)
    libs[name] = l


def key(res):
    return [(r, lang, [(s.t0, s.t1, s.text, [t["id"] for t in s.tokens], [(t["t0"], t["t1"]) for t in s.tokens]) for s in segs]) for r, segs, lang in res]


stable = []
for name in (sys.argv[1:] or list(oc.CASES)):
    res = {}
    for b, lib in libs.items():
        ctx = WhisperContext.new_from_buffer_with_params(cf.model_blob(*oc.model_of(name)), WhisperContextParameters(use_gpu=False), lib=lib)
        res[b] = key(oc.run_case(name, lib, ctx))
        ctx.free()
    same = res["avx2"] == res["native"]
    shape = [(r, lang, len(segs), sum(len(s[3]) for s in segs)) for r, lang, segs in res["avx2"]]
    print("STABLE  " if same else "UNSTABLE", name, shape, flush=True)
    if same:
        stable.append(name)
print("STABLE =", stable)
