#!/usr/bin/env python
"""GPU diagnostic (test infrastructure): transcripts of the CUDA path vs the golden fixtures, token by token."""
import json, os, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
from conftest import GOLDEN, model_blob
from test_oracle_golden import FULL, RECIPE, RICH
from whisper_rust_b200 import *
from whisper_rust_b200 import api, synth
lib = api.load_library(); api.install_quiet_logging(lib)
cases = [("greedy_tiny_base_30.0s_seed3", "tiny", RECIPE, 30.0, 3), ("greedy_mini_rich_47.3s_seed5", "mini", RICH, 47.3, 5),
         ("greedy_mini_rich_30.0s_seed4", "mini", RICH, 30.0, 4), ("greedy_mini_rich_11.0s_seed3", "mini", RICH, 11.0, 3)]
for key, arch, recipe, seconds, seed in cases:
    ctx = WhisperContext.new_from_buffer_with_params(model_blob(arch, recipe), WhisperContextParameters(use_gpu=True), lib=lib)
    st = ctx.create_state()
    p = FullParams(lib, SamplingStrategy.GREEDY, best_of=1); p.set_n_threads(4).set_language("en").set_temperature_inc(0.0)
    st.full(p, synth.synth_audio(seconds, seed=seed))
    segs = st.segments()
    g = FULL[key]["segments"]
    ids = [(t["id"], t["p"]) for s in segs for t in s.tokens]
    gid = [(i, pp) for s in g for i, pp in zip(s["ids"], s["p"])]
    first = next((k for k, (a, b) in enumerate(zip(ids, gid)) if a[0] != b[0]), None)
    print(key, "mode", os.environ.get("WB200_NO_STEP_GRAPH", "step"), "segs", len(segs), "/", len(g), "tokens", len(ids), "/", len(gid), "first diff", first,
          "gold gap", FULL[key]["min_top2_logit_gap"])
    if first is not None:
        lo = max(0, first - 2)
        print("   ours", ids[lo:first + 3]); print("   gold", gid[lo:first + 3])
    else:
        d = max(abs(a[1] - b[1]) / max(b[1], 1e-9) for a, b in zip(ids, gid)) if ids else 0
        print("   max rel p diff", d)
    st.free(); ctx.free()
