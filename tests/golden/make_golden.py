#!/usr/bin/env python
"""Generates the golden fixtures under tests/golden/ by running THE REFERENCE ITSELF: oracle/_ref/libwhisper_ref.so is
the reference's own CPU engine (whisper.cpp v1.7.5 + ggml-cpu) compiled in place from /root/reference by oracle/Makefile
(AVX2 ISA set).  The reference ships no golden vectors of its own for this path (SURVEY.md §4,
§8c), so these are "outputs of the reference run here" on frozen synthetic inputs:

    model  : whisper_rust_b200.synth.write_model(arch, InitRecipe(seed=1234, te_std=0.2))   (ggml file layout, Appendix B)
    audio  : whisper_rust_b200.synth.synth_audio(seconds, seed)

Run from the repo root in the build container (needs oracle/_ref):   python tests/golden/make_golden.py
The fixtures are small (sub-sampled tensors + sha256 of the full arrays) and are committed.
"""
import ctypes as C
import hashlib
import json
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import conftest as cf  # noqa: E402
from whisper_rust_b200 import (DtwModelPreset, DtwParameters, FullParams, SamplingStrategy, WhisperContext,  # noqa: E402
                               WhisperContextParameters, api, synth)

RECIPE = synth.InitRecipe(te_std=0.2)
# "rich" recipe: larger projection gain / positional embedding and boosted timestamp-token embeddings make the decoder emit
# varied, audio-dependent text tokens, timestamps and EOT, i.e. several segments and partial seeks per clip (mini arch only:
# the reference forces no_timestamps for 2-text-layer models, whisper.cpp:6956-6960)
RICH = synth.InitRecipe(te_std=0.2, pe_std=0.3, gain=3.0, ts_scale=1.6)
MEL_STRIDE = 8      # keep every 8th frame of the mel goldens
ENC_STRIDE = 10     # keep every 10th row of the encoder golden


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def seg_dump(segs):
    return [{"t0": s.t0, "t1": s.t1, "text": s.text.decode("utf-8", "replace"),
             "ids": [t["id"] for t in s.tokens], "tids": [t["tid"] for t in s.tokens],
             "p": [round(float(t["p"]), 7) for t in s.tokens], "plog": [round(float(t["plog"]), 6) for t in s.tokens],
             "t_dtw": [int(t["t_dtw"]) for t in s.tokens]} for s in segs]


def run_full(ref, arch, recipe, seconds, seed, strategy=SamplingStrategy.GREEDY, tinc=0.0, dtw=None, **kw):
    """One whisper_full_with_state call on the oracle; also records the smallest top-1/top-2 logit gap seen by the
    logits_filter_callback (wc:6226-6228) so that test cases with near-ties can be told apart from real mismatches."""
    c = WhisperContext.new_from_buffer_with_params(cf.model_blob(arch, recipe), WhisperContextParameters(use_gpu=False, dtw_parameters=dtw), lib=ref)
    st = c.create_state()
    p = FullParams(ref, strategy, **kw)
    p.set_n_threads(8).set_language("en").set_temperature_inc(tinc)
    V = c.n_vocab()
    gaps = []

    def on_logits(_ctx, _st, _tok, _n, logits, _ud):
        a = np.ctypeslib.as_array(logits, shape=(V,))
        top = np.partition(a[np.isfinite(a)], -2)[-2:]
        gaps.append(float(top[1] - top[0]))
    p.set_logits_filter_callback(on_logits)
    assert st.full(p, synth.synth_audio(seconds, seed=seed)) == 0
    out = {"segments": seg_dump(st.segments()), "min_top2_logit_gap": min(gaps) if gaps else None, "n_steps": len(gaps)}
    st.free(); c.free()
    return out


def main():
    ref = cf.declare(api.load_library(cf.ORACLE_LIB), cf._PROBE_SIGS)
    api.install_quiet_logging(ref)

    def ctx_of(arch, **kw):
        return WhisperContext.new_from_buffer_with_params(cf.model_blob(arch, RECIPE), WhisperContextParameters(use_gpu=False, **kw), lib=ref)

    # ---- M1: log-mel (wc:3186-3276)
    mel = {}
    ctx = ctx_of("micro")
    for seconds, seed in ((0.3, 7), (11.0, 7), (30.0, 7)):
        st = ctx.create_state()
        st.pcm_to_mel(synth.synth_audio(seconds, seed=seed), threads=4)
        m, n_org = cf.ref_mel(ref, st)
        key = f"s{seconds}_seed{seed}"
        mel[key + "_sub"] = m[:, ::MEL_STRIDE].copy()
        mel[key + "_meta"] = np.array([m.shape[0], m.shape[1], n_org], dtype=np.int64)
        mel[key + "_sha"] = np.frombuffer(bytes.fromhex(sha(m)), dtype=np.uint8)
        st.free()
    np.savez_compressed(HERE / "mel_micro.npz", **mel)

    # ---- E1-E3 + D1: encoder output and teacher-forced logits (wc:1994-2364, 2474-2852)
    enc = {}
    for arch in ("micro", "tiny"):
        c = ctx_of(arch)
        st = c.create_state()
        st.pcm_to_mel(synth.synth_audio(30.0, seed=11), threads=8)
        st.encode(0, threads=8)
        e = cf.ref_embd_enc(ref, st)
        toks = [c.token_sot(), c.token_sot() + 1, c.token_transcribe()]
        st.decode(toks, 0, threads=8)
        V = c.n_vocab()
        lg = np.ctypeslib.as_array(ref.whisper_get_logits_from_state(st.ptr), shape=(3, V))[2].copy()
        top = np.argsort(-lg, kind="stable")[:32]
        enc[f"{arch}_enc_sub"] = e[::ENC_STRIDE].copy()
        enc[f"{arch}_enc_shape"] = np.array(e.shape, dtype=np.int64)
        enc[f"{arch}_enc_sha"] = np.frombuffer(bytes.fromhex(sha(e)), dtype=np.uint8)
        enc[f"{arch}_logits_top_ids"] = top.astype(np.int32)
        enc[f"{arch}_logits_top_vals"] = lg[top]
        enc[f"{arch}_logits_sub"] = lg[::64].copy()
        enc[f"{arch}_logits_sha"] = np.frombuffer(bytes.fromhex(sha(lg)), dtype=np.uint8)
        st.free(); c.free()
    np.savez_compressed(HERE / "encoder_logits.npz", **enc)

    # ---- F1 + S1/S2: greedy transcripts (wc:6795-7711)
    # Every case is also run on the -march=native build of the reference (oracle/Makefile `native`: AVX-512/AMX kernels, another
    # f32 summation order).  "isa_stable" records whether the two builds of the reference produce the same transcript: on
    # random-init weights the logits are nearly flat and the reference's own builds diverge on about half of the long clips
    # (tools/ref_spread.py), so only isa_stable cases are meaningful token-parity targets for an independent implementation.
    nat_path = cf.ORACLE_LIB.with_name("libwhisper_ref_native.so")
    nat = None
    if nat_path.exists():
        nat = cf.declare(api.load_library(nat_path), cf._PROBE_SIGS)
        api.install_quiet_logging(nat)
    full = {}
    dtw2 = DtwParameters(preset=DtwModelPreset.N_TOP_MOST, n_top=2)
    B = SamplingStrategy.BEAM_SEARCH
    cases = {
        "greedy_micro_base_11.0s_seed3": dict(arch="micro", recipe=RECIPE, seconds=11.0, seed=3, best_of=1),
        "greedy_tiny_base_30.0s_seed3": dict(arch="tiny", recipe=RECIPE, seconds=30.0, seed=3, best_of=1),
        "greedy_mini_rich_11.0s_seed3": dict(arch="mini", recipe=RICH, seconds=11.0, seed=3, best_of=1),
        "greedy_mini_rich_30.0s_seed26": dict(arch="mini", recipe=RICH, seconds=30.0, seed=26, best_of=1),
        "greedy_mini_rich_47.3s_seed25": dict(arch="mini", recipe=RICH, seconds=47.3, seed=25, best_of=1),
        "greedy_mini_rich_47.3s_seed26": dict(arch="mini", recipe=RICH, seconds=47.3, seed=26, best_of=1),
        "greedy_mini_rich_47.3s_seed5": dict(arch="mini", recipe=RICH, seconds=47.3, seed=5, best_of=1),
        "greedy_mini_rich_65.0s_seed28": dict(arch="mini", recipe=RICH, seconds=65.0, seed=28, best_of=1),
        "greedy_mini_rich_0.3s_seed6": dict(arch="mini", recipe=RICH, seconds=0.3, seed=6, best_of=1),
        # temperature fallback ladder + best_of sampling (mt19937 streams, wc:6476, 6897); the base-recipe micro model repeats
        # one token, fails the entropy test at T=0 and walks the ladder
        "greedy_fallback_micro_base_11.0s_seed4": dict(arch="micro", recipe=RECIPE, seconds=11.0, seed=4, tinc=0.2, best_of=3),
        "greedy_fallback_mini_rich_47.3s_seed26": dict(arch="mini", recipe=RICH, seconds=47.3, seed=26, tinc=0.2, best_of=2),
        # S3 + X1/X2: beam search with DTW token timestamps (wc:6491-6564, 8647-8933)
        "beam5_dtw_mini_rich_30.0s_seed26": dict(arch="mini", recipe=RICH, seconds=30.0, seed=26, strategy=B, dtw=dtw2, beam_size=5),
        "beam5_dtw_mini_rich_11.0s_seed3": dict(arch="mini", recipe=RICH, seconds=11.0, seed=3, strategy=B, dtw=dtw2, beam_size=5),
        # greedy + DTW (deterministic sampler, alignment only)
        "greedy_dtw_mini_rich_30.0s_seed26": dict(arch="mini", recipe=RICH, seconds=30.0, seed=26, dtw=dtw2, best_of=1),
        "greedy_dtw_mini_rich_47.3s_seed25": dict(arch="mini", recipe=RICH, seconds=47.3, seed=25, dtw=dtw2, best_of=1),
    }
    for key, kw in cases.items():
        kw = dict(kw)
        arch, recipe, seconds, seed = kw.pop("arch"), kw.pop("recipe"), kw.pop("seconds"), kw.pop("seed")
        full[key] = run_full(ref, arch, recipe, seconds, seed, **kw)
        if nat is not None:
            other = run_full(nat, arch, recipe, seconds, seed, **kw)
            strip = lambda r: [(s["t0"], s["t1"], s["ids"], s["t_dtw"]) for s in r["segments"]]
            full[key]["isa_stable"] = strip(other) == strip(full[key])
    (HERE / "full_transcripts.json").write_text(json.dumps(full, indent=0))
    for k, v in full.items():
        print(f"{k}: {len(v['segments'])} segments, {sum(len(s['ids']) for s in v['segments'])} tokens, min top-2 gap "
              f"{v['min_top2_logit_gap']}, isa_stable {v.get('isa_stable')}")

    # ---- known answers: special tokens, language table, default params (SURVEY.md §4 "known-answer surface")
    ka = {"tokens": {}, "langs": [], "defaults": {}}
    for arch in ("micro", "large-v3"):
        cc = WhisperContext.new_from_buffer_with_params(synth.model_bytes(arch, RECIPE, with_tensors=False),
                                                        WhisperContextParameters(use_gpu=False), lib=ref)
        ka["tokens"][arch] = {n: getattr(ref, "whisper_token_" + n)(cc.ptr) for n in
                              ("eot", "sot", "solm", "prev", "nosp", "not", "beg", "translate", "transcribe")}
        ka["tokens"][arch]["lang_en"] = ref.whisper_token_lang(cc.ptr, 0)
        ka["tokens"][arch]["n_vocab"] = cc.n_vocab()
        ka["tokens"][arch]["multilingual"] = cc.is_multilingual()
        ka["tokens"][arch]["type"] = (cc.model_type_readable() or b"").decode()
        cc.free()
    for i in range(ref.whisper_lang_max_id() + 1):
        ka["langs"].append([ref.whisper_lang_str(i).decode(), ref.whisper_lang_str_full(i).decode()])
    for name, strat in (("greedy", 0), ("beam", 1)):
        fp = ref.whisper_full_default_params(strat)
        d = {}
        for fname, ftype in api.whisper_full_params._fields_:
            v = getattr(fp, fname)
            if isinstance(v, (bool, int, float)):
                d[fname] = v
            elif isinstance(v, bytes):
                d[fname] = v.decode()
        d["greedy.best_of"] = fp.greedy.best_of
        d["beam_search.beam_size"] = fp.beam_search.beam_size
        d["beam_search.patience"] = fp.beam_search.patience
        ka["defaults"][name] = d
    cp = ref.whisper_context_default_params()
    ka["defaults"]["context"] = {"use_gpu": cp.use_gpu, "flash_attn": cp.flash_attn, "gpu_device": cp.gpu_device,
                                 "dtw_token_timestamps": cp.dtw_token_timestamps, "dtw_aheads_preset": cp.dtw_aheads_preset,
                                 "dtw_n_top": cp.dtw_n_top, "dtw_mem_size": cp.dtw_mem_size}
    c = ctx_of("micro")
    toks = c.tokenize(" hello world t1234", 64)
    ka["tokenize"] = {" hello world t1234": toks}
    c.free()
    (HERE / "known_answers.json").write_text(json.dumps(ka, indent=0))
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
