"""Stage-level parity of the CUDA path against the oracle (the reference's ggml CPU engine), through the C ABI.

Tolerances are BASELINE.json's: mel 1e-4 relative, encoder output 1e-2 relative (we hold the f16-exact
path to 2e-3), logits compared on argmax + relative error.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import model_blob, prod_embd_enc, prod_mel, ref_embd_enc, ref_mel, rel_err
from whisper_rust_b200 import (FullParams, SamplingStrategy, WhisperContext, WhisperContextParameters, synth)

pytestmark = pytest.mark.gpu

RECIPE = synth.InitRecipe(te_std=0.2)


def _ctx(lib, arch, gpu, **kw):
    return WhisperContext.new_from_buffer_with_params(model_blob(arch, RECIPE), WhisperContextParameters(use_gpu=gpu, **kw), lib=lib)


@pytest.mark.parametrize("m,n,k,bn", [(128, 128, 64, 128), (128, 256, 128, 256), (1500, 384, 384, 0), (3, 1152, 384, 0),
                                      (77, 51865, 384, 0), (300, 200, 1500, 64), (1, 32, 4096, 32)])
def test_gemm_tcgen05_vs_torch(prod_lib, m, n, k, bn):
    import torch
    torch.manual_seed(m * 7 + n)
    kp = (k + 7) // 8 * 8
    a = torch.zeros(m, kp, device="cuda", dtype=torch.float16)
    b = torch.zeros(n, kp, device="cuda", dtype=torch.float16)
    a[:, :k] = torch.randn(m, k, device="cuda") * 0.5
    b[:, :k] = torch.randn(n, k, device="cuda") * 0.5
    bias = torch.randn(n, device="cuda", dtype=torch.float32)
    c = torch.full((m, n), float("nan"), device="cuda", dtype=torch.float32)
    torch.cuda.synchronize()
    r = prod_lib.whisper_b200_test_gemm(a.data_ptr(), b.data_ptr(), c.data_ptr(), m, n, k, kp, kp, n, 0, bias.data_ptr(), bn)
    assert r == 0
    torch.cuda.synchronize()
    ref = a[:, :k].float() @ b[:, :k].float().T + bias
    err = (c - ref).abs().max().item()
    assert torch.isfinite(c).all()
    assert err < 2e-3 * max(1.0, ref.abs().max().item()), err


# the persistent kernel (auto block_n, >= 148 tiles): TMA-store epilogue in row / per-head-panel layouts, f32 / f16,
# ragged M and N tails, and the in-place residual  c += A.B^T + bias  through TMA reduce-add (wc:2216-2220)
@pytest.mark.parametrize("m,n,k,f16,flags", [(12000, 1280, 1280, 0, 0), (12000, 1280, 1280, 0, 1), (6000, 2560, 1280, 1, 2),
                                              (4500, 5120, 1280, 1, 0), (3000, 1280, 5120, 0, 1), (20000, 384, 384, 1, 0),
                                              (19999, 392, 384, 0, 1), (9001, 1000, 192, 1, 0)])
def test_gemm_persistent_vs_torch(prod_lib, m, n, k, f16, flags):
    import torch
    torch.manual_seed(m + n + k)
    a = (torch.randn(m, k, device="cuda") * 0.5).half()
    b = (torch.randn(n, k, device="cuda") * 0.5).half()
    bias = torch.randn(n, device="cuda", dtype=torch.float32)
    ref = a.float() @ b.float().T + bias
    if flags & 1:
        c = torch.randn(m, n, device="cuda", dtype=torch.float32)
        ref = c + ref
    else:
        c = torch.full((m, n), float("nan"), device="cuda", dtype=torch.float16 if f16 else torch.float32)
    torch.cuda.synchronize()
    r = prod_lib.whisper_b200_test_gemm_ex(a.data_ptr(), b.data_ptr(), c.data_ptr(), m, n, k, k, k, n, f16, bias.data_ptr(), 0, flags)
    assert r == 0
    out = c.float()
    if flags & 2:  # [n/64][m][64] panels -> [m][n]
        out = c.view(n // 64, m, 64).permute(1, 0, 2).reshape(m, n).float()
    assert torch.isfinite(out).all()
    tol = (2e-3 if not f16 else 6e-3) * max(1.0, ref.abs().max().item())
    assert (out - ref).abs().max().item() < tol


# fused encoder attention (attn_enc.cu) against the reference's op sequence in torch: S = Q K^T (f32), softmax(S / 8) in f32,
# P rounded to f16, P V with f32 accumulation (wc:2183-2206).  T = 1500 is the real context; 300 / 129 / 77 are the ragged cases
# (exp_n_audio_ctx), 2 windows x 3 heads checks the panel addressing.
@pytest.mark.parametrize("vt", [0, 1])
@pytest.mark.parametrize("W,H,T", [(1, 1, 128), (2, 3, 1500), (1, 2, 300), (3, 1, 129), (1, 6, 77), (1, 1, 1000)])
def test_enc_attention_vs_torch(prod_lib, W, H, T, vt):
    import torch
    torch.manual_seed(W * 1000 + H * 10 + T)
    q = (torch.randn(W, H, T, 64, device="cuda") * 1.5).half()
    k = (torch.randn(W, H, T, 64, device="cuda") * 1.5).half()
    v = torch.randn(W, H, T, 64, device="cuda").half()
    out = torch.full((W * T, H * 64), float("nan"), device="cuda", dtype=torch.float16)
    torch.cuda.synchronize()
    assert prod_lib.whisper_b200_test_enc_attn(q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), W, H, T, vt) == 0
    s = q.float() @ k.float().transpose(-1, -2)
    p = torch.softmax(s * 0.125, dim=-1).half().float()
    ref = (p @ v.float()).permute(0, 2, 1, 3).reshape(W * T, H * 64)
    o = out.float()
    assert torch.isfinite(o).all()
    err = (o - ref).abs().max().item()
    rel = ((o - ref).norm() / ref.norm()).item()
    assert err < 4e-3 * max(1.0, ref.abs().max().item()) and rel < 2e-3, (err, rel)


@pytest.mark.parametrize("seconds", [0.3, 11.0, 30.0, 47.3])
def test_mel_matches_oracle(prod_lib, ref_lib, seconds):
    pcm = synth.synth_audio(seconds, seed=7)
    out = []
    for lib, gpu in ((ref_lib, False), (prod_lib, True)):
        ctx = _ctx(lib, "micro", gpu)
        st = ctx.create_state()
        st.pcm_to_mel(pcm, threads=4)
        out.append(ref_mel(lib, st) if not gpu else prod_mel(lib, st))
        assert st.n_len() == out[-1][1]
        st.free(); ctx.free()
    (mr, nr), (mp, np_) = out
    assert mr.shape == mp.shape and nr == np_
    assert rel_err(mp, mr) < 1e-4
    assert np.max(np.abs(mp - mr)) < 2e-3


@pytest.mark.parametrize("arch", ["micro", "tiny"])
def test_encoder_matches_oracle(prod_lib, ref_lib, arch):
    pcm = synth.synth_audio(30.0, seed=11)
    res = {}
    for name, lib, gpu in (("ref", ref_lib, False), ("b200", prod_lib, True)):
        ctx = _ctx(lib, arch, gpu)
        st = ctx.create_state()
        st.pcm_to_mel(pcm, threads=8)
        st.encode(0, threads=8)
        res[name] = ref_embd_enc(lib, st) if not gpu else prod_embd_enc(lib, ctx)
        st.decode([ctx.token_sot(), ctx.token_sot() + 1, ctx.token_transcribe()], 0, threads=8)
        V = ctx.n_vocab()
        p = lib.whisper_get_logits_from_state(st.ptr)
        res[name + "_logits"] = np.ctypeslib.as_array(p, shape=(3, V))[2].copy()
        st.free(); ctx.free()
    e = rel_err(res["b200"], res["ref"])
    print("encoder rel err", e, "max abs", np.max(np.abs(res["b200"] - res["ref"])))
    assert e < 2e-3
    le = rel_err(res["b200_logits"], res["ref_logits"])
    print("logits rel err", le, np.argmax(res["b200_logits"]), np.argmax(res["ref_logits"]))
    assert le < 5e-3
    assert np.argmax(res["b200_logits"]) == np.argmax(res["ref_logits"])


RICH = synth.InitRecipe(te_std=0.2, pe_std=0.3, gain=3.0, ts_scale=1.6)


def _run_full(lib, gpu, arch, pcm, strategy=SamplingStrategy.GREEDY, tinc=0.0, recipe=RECIPE, dtw=None, **kw):
    ctx = WhisperContext.new_from_buffer_with_params(model_blob(arch, recipe), WhisperContextParameters(use_gpu=gpu, dtw_parameters=dtw), lib=lib)
    st = ctx.create_state()
    p = FullParams(lib, strategy, **kw)
    p.set_n_threads(8).set_language("en").set_temperature_inc(tinc)
    assert st.full(p, pcm) == 0
    segs = st.segments()
    st.free(); ctx.free()
    return segs


def _same(a, b, p_rtol=5e-2):
    assert len(a) == len(b)
    for sa, sb in zip(a, b):
        assert (sa.t0, sa.t1, sa.text) == (sb.t0, sb.t1, sb.text)
        assert [t["id"] for t in sa.tokens] == [t["id"] for t in sb.tokens]
        ta, tb = [t["tid"] for t in sa.tokens], [t["tid"] for t in sb.tokens]
        assert all(x == y for x, y, t in zip(ta, tb, sa.tokens) if t["id"] >= 50364)
        assert sum(x != y for x, y in zip(ta, tb)) <= max(1, len(ta) // 50)   # arg-max over near-flat timestamp probabilities
        assert all(abs(x["t_dtw"] - y["t_dtw"]) <= 2 for x, y in zip(sa.tokens, sb.tokens))
        np.testing.assert_allclose([t["p"] for t in sa.tokens], [t["p"] for t in sb.tokens], rtol=p_rtol, atol=1e-4)


# Cases are chosen so that the oracle's smallest top-1/top-2 logit gap over the whole decode (recorded in
# tests/golden/full_transcripts.json) is >= 2e-3 AND the reference's AVX2 and native builds agree: below that the winner depends on fp32 summation order, which differs even
# between two ISA builds of the reference itself (DESIGN.md "parity").
@pytest.mark.parametrize("arch,recipe,seconds,seed", [("tiny", RECIPE, 30.0, 3), ("mini", RICH, 47.3, 26), ("mini", RICH, 30.0, 26),
                                                      ("mini", RICH, 11.0, 3), ("mini", RICH, 0.3, 6), ("mini", RICH, 65.0, 28)])
def test_full_greedy_tokens_match_oracle(prod_lib, ref_lib, arch, recipe, seconds, seed):
    pcm = synth.synth_audio(seconds, seed=seed)
    a = _run_full(ref_lib, False, arch, pcm, recipe=recipe, best_of=1)
    b = _run_full(prod_lib, True, arch, pcm, recipe=recipe, best_of=1)
    _same(a, b)


def test_logits_filter_callback_matches_oracle(prod_lib, ref_lib):
    """logits_filter_callback (wh:474-480) sees HOST logits after the static suppressions and before the timestamp rules
    (wc:6226-6228) and may mutate them: same callback on both engines -> same transcript, same number of calls."""
    pcm = synth.synth_audio(47.3, seed=26)   # two windows, ~230 tokens
    out = {}
    for name, lib, gpu in (("ref", ref_lib, False), ("b200", prod_lib, True)):
        calls = []

        def cb(ctx, state, tokens, n_tokens, logits, ud, calls=calls):
            calls.append((n_tokens, tokens[n_tokens - 1].id if n_tokens else -1))
            if n_tokens % 5 == 2:
                logits[1000 + n_tokens] += 40.0          # force a token
            for k in range(2000, 2064):
                logits[k] = -float("inf")                # suppress a range
            if n_tokens == 0:
                assert logits[50257] == -float("inf")    # suppress_blank already applied to <|eot|> at the first step

        ctx = WhisperContext.new_from_buffer_with_params(model_blob("mini", RICH), WhisperContextParameters(use_gpu=gpu), lib=lib)
        st = ctx.create_state()
        p = FullParams(lib, SamplingStrategy.GREEDY, best_of=1)
        p.set_n_threads(8).set_language("en").set_temperature_inc(0.0).set_logits_filter_callback(cb)
        assert st.full(p, pcm) == 0
        out[name] = (st.segments(), calls)
        st.free(); ctx.free()
    _same(out["ref"][0], out["b200"][0])
    assert out["ref"][1] == out["b200"][1] and len(out["ref"][1]) > 10
    ids = [t["id"] for s in out["b200"][0] for t in s.tokens]
    assert any(1000 <= i < 1300 for i in ids) and not any(2000 <= i < 2064 for i in ids)


@pytest.mark.parametrize("seconds,seed,max_len,split_on_word", [(30.0, 26, 0, False), (30.0, 26, 12, False), (47.3, 26, 20, True),
                                                                (11.0, 3, 1, False)])
def test_heuristic_token_timestamps_and_wrapping_match_oracle(prod_lib, ref_lib, seconds, seed, max_len, split_on_word):
    """`token_timestamps = true` (whisper_exp_compute_token_level_timestamps, wc:8390-8615: anchors from the timestamp-token
    statistics, voice-length split, signal-energy edges) and `max_len` / `split_on_word` (whisper_wrap_segment, wc:6049-6100):
    integer token times, voice lengths and the wrapped segments must be identical to the reference's."""
    pcm = synth.synth_audio(seconds, seed=seed)
    out = {}
    for name, lib, gpu in (("ref", ref_lib, False), ("b200", prod_lib, True)):
        ctx = WhisperContext.new_from_buffer_with_params(model_blob("mini", RICH), WhisperContextParameters(use_gpu=gpu), lib=lib)
        st = ctx.create_state()
        p = FullParams(lib, SamplingStrategy.GREEDY, best_of=1)
        p.set_n_threads(8).set_language("en").set_temperature_inc(0.0)
        p.set_token_timestamps(True).set_max_len(max_len).set_split_on_word(split_on_word)
        assert st.full(p, pcm) == 0
        out[name] = [(s.t0, s.t1, s.text, [(t["id"], t["t0"], t["t1"], t["vlen"]) for t in s.tokens]) for s in st.segments()]
        st.free(); ctx.free()
    assert len(out["ref"]) > 0 and any(t[1] >= 0 for s in out["ref"] for t in s[3])
    assert out["b200"] == out["ref"]
