"""Synthetic inputs for the transcription path: random-init Whisper models in the
reference's on-disk ggml layout, and deterministic 16 kHz mono audio.

There is no network, so no trained checkpoint exists here; BASELINE.json asks for
"random-init weights of the named architecture".  The file layout written here is the one
the reference loader parses (reader: sys/whisper.cpp/src/whisper.cpp:1513-1965, writer it
mirrors: sys/whisper.cpp/models/convert-pt-to-ggml.py:268-339):

    u32 magic 0x67676d6c | 11 x i32 hparams | i32 n_mel, i32 n_fft, f32 filters[n_mel*n_fft]
    | i32 n_tokens, { u32 len, bytes } * n_tokens
    | { i32 n_dims, i32 name_len, i32 ttype, i32 ne[n_dims] (reversed numpy shape), name, data } *

The initialisation recipe (seeds and standard deviations) is a frozen test-design
parameter: golden fixtures under tests/golden/ depend on it.
"""
from __future__ import annotations

import io
import struct
from dataclasses import dataclass

import numpy as np

GGML_MAGIC = 0x67676D6C
N_FFT_BINS = 201          # 1 + WHISPER_N_FFT/2   (whisper.h:33, whisper.cpp:3130)
SAMPLE_RATE = 16000
N_BASE_TOKENS = 50257     # tokens stored in the file; the loader synthesises the specials (whisper.cpp:1659-1690)


@dataclass(frozen=True)
class Arch:
    name: str
    n_vocab: int
    n_audio_ctx: int
    n_audio_state: int
    n_audio_head: int
    n_audio_layer: int
    n_text_ctx: int
    n_text_state: int
    n_text_head: int
    n_text_layer: int
    n_mels: int

    @property
    def hparams(self):
        return (self.n_vocab, self.n_audio_ctx, self.n_audio_state, self.n_audio_head, self.n_audio_layer,
                self.n_text_ctx, self.n_text_state, self.n_text_head, self.n_text_layer, self.n_mels)


def _arch(name, S, H, L, M=80, V=51865):
    return Arch(name, V, 1500, S, H, L, 448, S, H, L, M)


# SURVEY.md Appendix D (OpenAI Whisper dims; the reference infers the type from n_audio_layer, whisper.cpp:1543-1565)
ARCHS = {
    "tiny":     _arch("tiny", 384, 6, 4),
    "base":     _arch("base", 512, 8, 6),
    "small":    _arch("small", 768, 12, 12),
    "medium":   _arch("medium", 1024, 16, 24),
    "large-v3": _arch("large-v3", 1280, 20, 32, M=128, V=51866),
    # reduced shapes for fast CPU-side tests (same layer structure, D=64 heads)
    # (micro has 2 text layers, so the reference treats it as a first-release distilled model and forces no_timestamps,
    #  whisper.cpp:6956-6960; mini has 3 and keeps the timestamp logic live)
    # large-v3-turbo: 32 encoder / 4 decoder layers (n_audio_layer != n_text_layer); "miniturbo" is its reduced-shape stand-in
    "large-v3-turbo": Arch("large-v3-turbo", 51866, 1500, 1280, 20, 32, 448, 1280, 20, 4, 128),
    "miniturbo": Arch("miniturbo", 51866, 1500, 192, 3, 5, 448, 192, 3, 3, 128),
    "distil":   Arch("distil", 51865, 1500, 192, 3, 4, 448, 192, 3, 2, 80),
    "micro":    Arch("micro", 51865, 1500, 128, 2, 2, 448, 128, 2, 2, 80),
    "mini":     Arch("mini", 51865, 1500, 192, 3, 2, 448, 192, 3, 3, 80),
    # an English-only vocabulary (n_vocab 51864 = the *.en checkpoints): no language / task tokens in the prompt, unshifted special ids
    "minien":   Arch("minien", 51864, 1500, 192, 3, 2, 448, 192, 3, 3, 80),
    # every row length a multiple of 256: the smallest shape the K-quant block formats (q2_K ... q6_K) can hold
    "minik":    Arch("minik", 51865, 1500, 256, 4, 2, 448, 256, 4, 3, 80),
}


def mel_filterbank(n_mels: int, n_fft: int = 400, sr: int = SAMPLE_RATE) -> np.ndarray:
    """Deterministic Slaney-style triangular mel filterbank [n_mels, 201] (float32).

    The real mel_filters.npz is not in the reference tree (convert-pt-to-ggml.py:221 reads
    it from an OpenAI checkout), so the synthetic models carry this one.
    """
    n_bins = n_fft // 2 + 1
    fft_freqs = np.linspace(0.0, sr / 2.0, n_bins)

    def hz_to_mel(f):
        f = np.asarray(f, dtype=np.float64)
        mel = f / (200.0 / 3)
        log_t = f >= 1000.0
        mel = np.where(log_t, 15.0 + np.log(np.maximum(f, 1e-9) / 1000.0) / (np.log(6.4) / 27.0), mel)
        return mel

    def mel_to_hz(m):
        m = np.asarray(m, dtype=np.float64)
        f = m * (200.0 / 3)
        log_t = m >= 15.0
        f = np.where(log_t, 1000.0 * np.exp((np.log(6.4) / 27.0) * (m - 15.0)), f)
        return f

    mel_pts = mel_to_hz(np.linspace(hz_to_mel(0.0), hz_to_mel(sr / 2.0), n_mels + 2))
    fdiff = np.diff(mel_pts)
    ramps = mel_pts[:, None] - fft_freqs[None, :]
    lower = -ramps[:-2] / fdiff[:-1, None]
    upper = ramps[2:] / fdiff[1:, None]
    w = np.maximum(0.0, np.minimum(lower, upper))
    w *= (2.0 / (mel_pts[2:n_mels + 2] - mel_pts[:n_mels]))[:, None]
    return w.astype(np.float32)


def vocab_tokens():
    """50,257 distinct byte strings; ids < 256 are the single bytes, so " " (id 32) exists
    (the reference's suppress_blank does token_to_id.at(" "), whisper.cpp:6191)."""
    toks = [bytes([i]) for i in range(256)]
    toks[0] = b"<0>"  # avoid an embedded NUL: token text crosses the ABI as a C string
    toks += [b" t%d" % i if i % 3 == 0 else b"t%d" % i for i in range(256, N_BASE_TOKENS)]
    return toks


def tensor_plan(a: Arch):
    """(name, numpy shape, kind) in file order. kind: 'w' matrix (f16), 'b' bias, 'g' LN gamma,
    'beta' LN beta, 'pe' positional embedding, 'te' token embedding, 'cw' conv weight, 'cb' conv bias.
    Names: sys/whisper.cpp/src/whisper-arch.h:42-109."""
    S, T, M, V = a.n_audio_state, a.n_audio_ctx, a.n_mels, a.n_vocab
    plan = [
        ("encoder.positional_embedding", (T, S), "pe"),
        ("encoder.conv1.weight", (S, M, 3), "cw"), ("encoder.conv1.bias", (S, 1), "cb"),
        ("encoder.conv2.weight", (S, S, 3), "cw"), ("encoder.conv2.bias", (S, 1), "cb"),
        ("encoder.ln_post.weight", (S,), "g"), ("encoder.ln_post.bias", (S,), "beta"),
    ]

    def block(prefix, S, cross):
        out = []
        for ln in (["attn_ln", "cross_attn_ln", "mlp_ln"] if cross else ["attn_ln", "mlp_ln"]):
            out += [(f"{prefix}.{ln}.weight", (S,), "g"), (f"{prefix}.{ln}.bias", (S,), "beta")]
        for att in (["attn", "cross_attn"] if cross else ["attn"]):
            out += [(f"{prefix}.{att}.query.weight", (S, S), "w"), (f"{prefix}.{att}.query.bias", (S,), "b"),
                    (f"{prefix}.{att}.key.weight", (S, S), "w"),
                    (f"{prefix}.{att}.value.weight", (S, S), "w"), (f"{prefix}.{att}.value.bias", (S,), "b"),
                    (f"{prefix}.{att}.out.weight", (S, S), "w"), (f"{prefix}.{att}.out.bias", (S,), "b")]
        out += [(f"{prefix}.mlp.0.weight", (4 * S, S), "w"), (f"{prefix}.mlp.0.bias", (4 * S,), "b"),
                (f"{prefix}.mlp.2.weight", (S, 4 * S), "w"), (f"{prefix}.mlp.2.bias", (S,), "b")]
        return out

    for i in range(a.n_audio_layer):
        plan += block(f"encoder.blocks.{i}", S, cross=False)
    St = a.n_text_state
    plan += [("decoder.positional_embedding", (a.n_text_ctx, St), "pe"),
             ("decoder.token_embedding.weight", (V, St), "te"),
             ("decoder.ln.weight", (St,), "g"), ("decoder.ln.bias", (St,), "beta")]
    for i in range(a.n_text_layer):
        plan += block(f"decoder.blocks.{i}", St, cross=True)
    return plan


@dataclass(frozen=True)
class InitRecipe:
    """Frozen initialisation recipe. `te_std` widens the argmax margins of the
    tied-embedding logits; `gain` scales every projection matrix."""
    seed: int = 1234
    te_std: float = 0.05
    gain: float = 1.0
    bias_std: float = 0.02
    pe_std: float = 0.02
    ts_scale: float = 1.0   # multiplies the token-embedding rows of the timestamp tokens (ids >= 50364): > 1 makes the
                            # sampler emit timestamps, hence several segments / windows per clip (test-design parameter)


# ---- quantised checkpoints (examples/quantize: ggml_common_quantize_0 quantises the 2-D matrices, leaves the 3-D conv kernels f16 and
# the 1-D / positional tensors f32).  Block layouts: ggml/src/ggml-common.h; the arithmetic follows quantize_row_*_ref (ggml-quants.c).
FTYPE_TO_TTYPE = {2: 2, 3: 3, 7: 8, 8: 6, 9: 7,     # ggml_ftype -> ggml_type: q4_0, q4_1, q8_0, q5_0, q5_1
                  10: 10, 11: 11, 12: 12, 13: 13, 14: 14}   # q2_K, q3_K, q4_K, q5_K, q6_K (256-element super-blocks)
# This module packs q8_0 / q4_0 / q5_0 itself.  For the other types write_model takes the block codec from its caller:
#   quantizer(x: f32 array, ggml_type) -> bytes      dequantizer(raw: bytes, ggml_type, n) -> f32 array of n elements
# (the tests pass the reference's own ggml_quantize_chunk / dequantize_row_* through the oracle library).


def _quantize(x: np.ndarray, ttype: int) -> bytes:
    """x: f32, size a multiple of 32 -> the ggml block stream of type ttype (2 q4_0, 6 q5_0, 8 q8_0)"""
    b = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, 32)
    nb = b.shape[0]
    if ttype == 8:      # q8_0: d = amax / 127, q = round(x / d)   (quantize_row_q8_0_ref, ggml-quants.c:188-211)
        amax = np.abs(b).max(axis=1)
        d = (amax / np.float32(127)).astype(np.float32)
        inv = np.where(d != 0, np.float32(1) / np.where(d != 0, d, 1), 0).astype(np.float32)
        v = b * inv[:, None]
        q = (np.sign(v) * np.floor(np.abs(v) + np.float32(0.5))).astype(np.int8)      # roundf: half away from zero
        out = np.zeros((nb, 34), dtype=np.uint8)
        out[:, 0:2] = d.astype("<f2").view(np.uint8).reshape(nb, 2)
        out[:, 2:] = q.view(np.uint8)
        return out.tobytes()
    if ttype in (2, 6):  # q4_0 / q5_0: d = (value of largest magnitude) / -8 or -16   (ggml-quants.c:33-56, 99-141)
        levels = 8 if ttype == 2 else 16
        idx = np.abs(b).argmax(axis=1)
        mx = b[np.arange(nb), idx]
        d = (mx / np.float32(-levels)).astype(np.float32)
        inv = np.where(d != 0, np.float32(1) / np.where(d != 0, d, 1), 0).astype(np.float32)
        v = b * inv[:, None] + np.float32(levels + 0.5)
        xi = np.minimum(2 * levels - 1, v.astype(np.int8).astype(np.int32)).astype(np.uint8)   # MIN(15|31, (int8_t)(x*id + 8.5|16.5))
        lo, hi = xi[:, :16], xi[:, 16:]
        qs = ((lo & 0x0F) | ((hi & 0x0F) << 4)).astype(np.uint8)
        if ttype == 2:
            out = np.zeros((nb, 18), dtype=np.uint8)
            out[:, 0:2] = d.astype("<f2").view(np.uint8).reshape(nb, 2)
            out[:, 2:] = qs
            return out.tobytes()
        sh = np.arange(16, dtype=np.uint32)
        qh = (((lo.astype(np.uint32) & 0x10) >> 4) << sh).sum(axis=1, dtype=np.uint32) | \
             (((hi.astype(np.uint32) & 0x10) >> 4) << (sh + 16)).sum(axis=1, dtype=np.uint32)
        out = np.zeros((nb, 22), dtype=np.uint8)
        out[:, 0:2] = d.astype("<f2").view(np.uint8).reshape(nb, 2)
        out[:, 2:6] = qh.astype("<u4").view(np.uint8).reshape(nb, 4)
        out[:, 6:] = qs
        return out.tobytes()
    raise ValueError(f"synthetic writer: ggml type {ttype} not implemented")


def _dequantize(raw: bytes, ttype: int, n: int) -> np.ndarray:
    """inverse of _quantize, evaluated in f32 like dequantize_row_* (ggml-quants.c)"""
    bs = {2: 18, 6: 22, 8: 34}[ttype]
    b = np.frombuffer(raw, dtype=np.uint8).reshape(-1, bs)
    d = b[:, 0:2].copy().view("<f2").astype(np.float32)            # [nb][1]
    if ttype == 8:
        return (b[:, 2:].copy().view(np.int8).astype(np.float32) * d).reshape(n)
    qs = b[:, (2 if ttype == 2 else 6):].astype(np.int32)
    lo, hi = qs & 0x0F, qs >> 4
    if ttype == 6:
        qh = b[:, 2:6].copy().view("<u4").astype(np.uint32)        # [nb][1]
        sh = np.arange(16, dtype=np.uint32)
        lo = lo | (((qh >> sh) & 1) << 4).astype(np.int32)
        hi = hi | (((qh >> (sh + 16)) & 1) << 4).astype(np.int32)
    off = 8 if ttype == 2 else 16
    return (np.concatenate([lo - off, hi - off], axis=1).astype(np.float32) * d).reshape(n)


def write_model(arch: str | Arch, out, recipe: InitRecipe = InitRecipe(), ftype: int = 1, with_tensors: bool = True,
                expand_ftype: int | None = None, quantizer=None, dequantizer=None) -> int:
    """Write a random-init model to `out` (path or binary file object). Returns bytes written.

    with_tensors=False writes the reference's own test convention: a header-only file with
    zero tensor records, which its loader accepts (whisper.cpp:1959-1960)."""
    a = ARCHS[arch] if isinstance(arch, str) else arch
    own = isinstance(out, (str, bytes)) or hasattr(out, "__fspath__")
    f = open(out, "wb") if own else out
    n0 = f.tell()
    try:
        f.write(struct.pack("<I", GGML_MAGIC))
        f.write(struct.pack("<11i", *a.hparams, ftype))
        filt = mel_filterbank(a.n_mels)
        f.write(struct.pack("<2i", a.n_mels, N_FFT_BINS))
        f.write(filt.astype("<f4").tobytes())
        toks = vocab_tokens()
        f.write(struct.pack("<i", len(toks)))
        for t in toks:
            f.write(struct.pack("<I", len(t)))
            f.write(t)
        if with_tensors:
            rng = np.random.default_rng(recipe.seed)
            for name, shape, kind in tensor_plan(a):
                n = int(np.prod(shape))
                x = rng.standard_normal(n, dtype=np.float32).reshape(shape)
                if kind == "w":
                    x *= recipe.gain / np.sqrt(shape[1])
                elif kind == "cw":
                    x *= recipe.gain / np.sqrt(shape[1] * shape[2])
                elif kind in ("b", "cb", "beta"):
                    x *= recipe.bias_std
                elif kind == "g":
                    x = 1.0 + 0.02 * x
                elif kind == "pe":
                    x *= recipe.pe_std
                elif kind == "te":
                    x *= recipe.te_std
                    if recipe.ts_scale != 1.0:
                        x[N_BASE_TOKENS + 107:] *= recipe.ts_scale
                if expand_ftype is not None and kind in ("w", "te"):
                    # the f16 checkpoint a quantised one expands to: x = f16(dequantize(quantize(x))) (what the B200 loader keeps)
                    qt0 = FTYPE_TO_TTYPE[expand_ftype]
                    if quantizer is not None and dequantizer is not None:
                        x = np.asarray(dequantizer(quantizer(x, qt0), qt0, n), dtype=np.float32).reshape(shape)
                    else:
                        x = _dequantize(_quantize(x, qt0), qt0, n).reshape(shape)
                f16 = (ftype == 1 and kind in ("w", "cw", "te")) or (ftype >= 2 and kind == "cw")
                qt = FTYPE_TO_TTYPE[ftype] if ftype >= 2 and kind in ("w", "te") else None
                if qt is not None:
                    payload = quantizer(x, qt) if quantizer is not None else _quantize(x, qt)
                else:
                    payload = x.astype("<f2" if f16 else "<f4").tobytes()
                nm = name.encode()
                f.write(struct.pack("<3i", len(shape), len(nm), qt if qt is not None else (1 if f16 else 0)))
                f.write(struct.pack(f"<{len(shape)}i", *reversed(shape)))
                f.write(nm)
                f.write(payload)
        return f.tell() - n0
    finally:
        if own:
            f.close()


def model_bytes(arch, recipe: InitRecipe = InitRecipe(), **kw) -> bytes:
    bio = io.BytesIO()
    write_model(arch, bio, recipe, **kw)
    return bio.getvalue()


def synth_audio(seconds: float, seed: int = 12345, sr: int = SAMPLE_RATE) -> np.ndarray:
    """Deterministic speech-band test signal in [-1, 1]: a few AM/FM-modulated tones with
    syllable-rate envelopes plus low-level LCG white noise."""
    n = int(round(seconds * sr))
    t = np.arange(n, dtype=np.float64) / sr
    rng = np.random.default_rng(seed)
    x = np.zeros(n, dtype=np.float64)
    for k in range(6):
        f0 = rng.uniform(110.0, 2200.0)
        fm = rng.uniform(0.2, 1.5)
        am = rng.uniform(1.5, 6.0)
        ph = rng.uniform(0, 2 * np.pi, 3)
        env = 0.5 * (1.0 + np.sin(2 * np.pi * am * t + ph[0])) ** 2 / 2.0
        x += (0.6 / (k + 1)) * env * np.sin(2 * np.pi * f0 * t + 4.0 * np.sin(2 * np.pi * fm * t + ph[1]) + ph[2])
    # LCG noise (language-agnostic, reproducible without numpy's generator)
    s = np.uint64(seed * 2654435761 % (1 << 32))
    a, c = np.uint64(1664525), np.uint64(1013904223)
    lcg = np.empty(n, dtype=np.float64)
    state = int(s)
    # vectorised LCG: x_k = a^k x_0 + c (a^k - 1)/(a - 1) mod 2^32, done in blocks to stay exact in Python ints
    blk = 1 << 16
    ak = [1]; ck = [0]
    for _ in range(blk - 1):
        ak.append((ak[-1] * 1664525) & 0xFFFFFFFF)
        ck.append((ck[-1] * 1664525 + 1013904223) & 0xFFFFFFFF)
    ak = np.array(ak, dtype=np.uint64); ck = np.array(ck, dtype=np.uint64)
    for i0 in range(0, n, blk):
        m = min(blk, n - i0)
        vals = (ak[:m] * np.uint64(state) + ck[:m]) & np.uint64(0xFFFFFFFF)
        lcg[i0:i0 + m] = vals.astype(np.float64) / 2147483648.0 - 1.0
        state = int((int(ak[m - 1]) * state + int(ck[m - 1])) & 0xFFFFFFFF)
        state = (state * 1664525 + 1013904223) & 0xFFFFFFFF
    x += 0.02 * lcg
    x *= 0.8 / max(1e-9, np.max(np.abs(x)))
    return x.astype(np.float32)


# ---------------------------------------------------------------------------------------------------------------------------
# synthetic Silero-layout VAD checkpoint (reader: whisper.cpp:4775-5098; tensor names whisper-arch.h:181-197)
# ---------------------------------------------------------------------------------------------------------------------------
VAD_ENCODER = ((129, 128, 3), (128, 64, 3), (64, 64, 3), (64, 128, 3))   # (in channels, out channels, kernel) of the four conv layers


def write_vad_model(out, seed: int = 77, gain: float = 2.0, final_bias: float = -5.7, final_scale: float = 1.2) -> int:
    """Random-init VAD model in the ggml layout the reference loads.  The STFT basis is a real windowed Fourier basis (258 filters of
    256 taps: cosines then sines), the conv / LSTM weights are Gaussian; `final_bias` and the positive final-conv weights make quiet
    windows score below 0.5 and loud ones above, so that gated test audio produces several speech segments."""
    rng = np.random.default_rng(seed)
    own = isinstance(out, (str, bytes)) or hasattr(out, "__fspath__")
    f = open(out, "wb") if own else out
    n0 = f.tell()

    def tensor(name, arr, f16):
        a = np.ascontiguousarray(arr, dtype=np.float32)
        nm = name.encode()
        f.write(struct.pack("<3i", a.ndim, len(nm), 1 if f16 else 0))
        f.write(struct.pack(f"<{a.ndim}i", *reversed(a.shape)))
        f.write(nm)
        f.write(a.astype("<f2" if f16 else "<f4").tobytes())

    try:
        f.write(struct.pack("<I", GGML_MAGIC))
        t = b"silero-16k"
        f.write(struct.pack("<i", len(t))); f.write(t)
        f.write(struct.pack("<3i", 5, 1, 2))
        f.write(struct.pack("<2i", 512, 64))                      # n_window, n_context
        f.write(struct.pack("<i", len(VAD_ENCODER)))
        for ic, oc, k in VAD_ENCODER:
            f.write(struct.pack("<3i", ic, oc, k))
        f.write(struct.pack("<4i", 128, 128, 128, 1))             # lstm input / hidden, final conv in / out
        n = np.arange(256)
        win = 0.5 - 0.5 * np.cos(2 * np.pi * n / 256)
        k = np.arange(129)[:, None]
        basis = np.concatenate([np.cos(2 * np.pi * k * n / 256) * win, -np.sin(2 * np.pi * k * n / 256) * win]) / 16.0
        tensor("_model.stft.forward_basis_buffer", basis.reshape(258, 1, 256), True)
        for i, (ic, oc, kk) in enumerate(VAD_ENCODER):
            tensor(f"_model.encoder.{i}.reparam_conv.weight", rng.standard_normal((oc, ic, kk)) * gain / np.sqrt(ic * kk), True)
            tensor(f"_model.encoder.{i}.reparam_conv.bias", rng.standard_normal(oc) * 0.05, False)
        tensor("_model.decoder.rnn.weight_ih", rng.standard_normal((512, 128)) * gain / np.sqrt(128), False)
        tensor("_model.decoder.rnn.weight_hh", rng.standard_normal((512, 128)) * 0.5 / np.sqrt(128), False)
        tensor("_model.decoder.rnn.bias_ih", rng.standard_normal(512) * 0.1, False)
        tensor("_model.decoder.rnn.bias_hh", rng.standard_normal(512) * 0.1, False)
        tensor("_model.decoder.decoder.2.weight", np.abs(rng.standard_normal((1, 128))) * final_scale, True)
        tensor("_model.decoder.decoder.2.bias", np.array([final_bias]), False)
        return f.tell() - n0
    finally:
        if own:
            f.close()


def vad_model_bytes(**kw) -> bytes:
    bio = io.BytesIO()
    write_vad_model(bio, **kw)
    return bio.getvalue()


def gated_audio(seconds: float, seed: int = 12345, period: float = 4.0, duty: float = 0.55, floor: float = 0.002) -> np.ndarray:
    """synth_audio with alternating loud / near-silent stretches (speech-like gating for the VAD tests)"""
    a = synth_audio(seconds, seed=seed)
    t = np.arange(len(a)) / SAMPLE_RATE
    rng = np.random.default_rng(seed + 1)
    phase = rng.uniform(0, period)
    gate = (((t + phase) % period) < duty * period).astype(np.float32)
    return (a * (floor + (1.0 - floor) * gate)).astype(np.float32)
