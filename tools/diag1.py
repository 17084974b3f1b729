"""GPU diagnostic: where do the CUDA path and the oracle diverge? (test infrastructure, not product)"""
import sys, time
sys.path.insert(0, 'tests')
import numpy as np
import conftest as cf
from conftest import *
from whisper_rust_b200 import *
R = synth.InitRecipe(te_std=0.2) if (len(sys.argv) < 3 or sys.argv[2] != 'rich') else synth.InitRecipe(te_std=0.2, pe_std=0.3, gain=3.0, ts_scale=1.6)
ref = api.load_library(ORACLE_LIB); declare(ref, cf._PROBE_SIGS); api.install_quiet_logging(ref)
prod = api.load_library(); declare(prod, cf._B200_SIGS); api.install_quiet_logging(prod)
arch = sys.argv[1] if len(sys.argv) > 1 else 'tiny'
pcm = synth.synth_audio(30.0, seed=11)
cr = WhisperContext.new_from_buffer_with_params(model_blob(arch, R), WhisperContextParameters(use_gpu=False), lib=ref)
cp = WhisperContext.new_from_buffer_with_params(model_blob(arch, R), WhisperContextParameters(use_gpu=True), lib=prod)
sr, sp = cr.create_state(), cp.create_state()
sr.pcm_to_mel(pcm, 8); sp.pcm_to_mel(pcm, 8)
mr, _ = ref_mel(ref, sr); mp, _ = prod_mel(prod, sp)
d = np.abs(mr - mp)
print("mel: rel", rel_err(mp, mr), "max abs", d.max(), "n differing", int((d > 0).sum()), "of", d.size, "n>1e-6", int((d > 1e-6).sum()), "n>1e-5", int((d>1e-5).sum()))
idx = np.unravel_index(np.argmax(d), d.shape); print(" worst at", idx, mr[idx], mp[idx])
# shared mel -> encoder
sp.set_mel(mr[:, :])
sr.encode(0, 8); sp.encode(0, 8)
er = ref_embd_enc(ref, sr); ep = prod_embd_enc(prod, cp)
print("encoder (shared mel): rel", rel_err(ep, er), "max abs", np.abs(ep - er).max())
toks = [cr.token_sot(), cr.token_sot() + 1, cr.token_transcribe()]
sr.decode(toks, 0, 8); sp.decode(toks, 0, 8)
V = cr.n_vocab()
lr = np.ctypeslib.as_array(ref.whisper_get_logits_from_state(sr.ptr), shape=(3, V))[2].copy()
lp = np.ctypeslib.as_array(prod.whisper_get_logits_from_state(sp.ptr), shape=(3, V))[2].copy()
print("logits (own enc): rel", rel_err(lp, lr), "max abs", np.abs(lp - lr).max(), "std", lr.std(), "top2 gap", np.sort(lr)[-1] - np.sort(lr)[-2])
# timing
for name, st, lib in (("ref", sr, ref), ("b200", sp, prod)):
    t = time.time(); n = 3 if name == "b200" else 1
    for _ in range(n): st.encode(0, 8)
    print(name, "encode ms", (time.time() - t) / n * 1e3)
    t = time.time(); n = 20 if name == "b200" else 3
    for i in range(n): st.decode([1000 + i], 3 + i, 8)
    print(name, "decode ms/token", (time.time() - t) / n * 1e3)
