#!/usr/bin/env python
"""How much do two ISA builds of the REFERENCE ITSELF differ (AVX2 vs -march=native: AVX-512/AMX kernels, other summation
order)?  Sets the floor for what any independent implementation can match (test infrastructure)."""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, conftest as cf
from conftest import model_blob, ref_embd_enc, rel_err
from whisper_rust_b200 import *
from whisper_rust_b200 import api, synth
libs = {}
for name, path in (("avx2", cf.ORACLE_LIB), ("native", cf.ORACLE_LIB.with_name("libwhisper_ref_native.so"))):
    l = cf.declare(api.load_library(path), cf._PROBE_SIGS); api.install_quiet_logging(l); libs[name] = l
for arch, recipe in (("tiny", synth.InitRecipe(te_std=0.2)), ("mini", synth.InitRecipe(te_std=0.2, pe_std=0.3, gain=3.0, ts_scale=1.6))):
    res = {}
    for name, lib in libs.items():
        c = WhisperContext.new_from_buffer_with_params(model_blob(arch, recipe), WhisperContextParameters(use_gpu=False), lib=lib)
        st = c.create_state(); st.pcm_to_mel(synth.synth_audio(30.0, seed=11), 4); st.encode(0, 8)
        e = ref_embd_enc(lib, st)
        st.decode([c.token_sot(), c.token_sot() + 1, c.token_transcribe()], 0, 8)
        lg = np.ctypeslib.as_array(lib.whisper_get_logits_from_state(st.ptr), shape=(3, c.n_vocab()))[2].copy()
        res[name] = (e, lg)
        p = FullParams(lib, SamplingStrategy.GREEDY, best_of=1); p.set_n_threads(8).set_language("en").set_temperature_inc(0.0)
        st.full(p, synth.synth_audio(47.3, seed=5))
        res[name + "_ids"] = [t["id"] for s in st.segments() for t in s.tokens]
        st.free(); c.free()
    (e0, l0), (e1, l1) = res["avx2"], res["native"]
    a, b = res["avx2_ids"], res["native_ids"]
    first = next((k for k, (x, y) in enumerate(zip(a, b)) if x != y), None)
    print(arch, "encoder rel", rel_err(e1, e0), "max abs", np.abs(e1 - e0).max(), "| logits rel", rel_err(l1, l0), "max abs", np.abs(l1 - l0).max(),
          "| 47.3 s transcript tokens", len(a), len(b), "first diff", first)
