"""Heuristic token timestamps (token_timestamps = true) and segment wrapping (max_len, split_on_word) without a device: the product's
host logic (csrc/full.cu: heuristic_token_timestamps_core, wrap_last_segment_core) against the reference's static
whisper_exp_compute_token_level_timestamps / whisper_wrap_segment (whisper.cpp:8390-8615, 6049-6100; reached through
oracle/ref_probe.cpp) on the same synthetic segments: anchors from the timestamp statistics, proportional splitting of the runs between
anchors, the energy-edge refinement, and the carry of t_beg / t_last / tid_last from segment to segment.  Integer / float host
arithmetic in the reference's own types: every field of every token must be equal.  (tests/test_gpu_stages.py and the option cases run the
same code end to end on the GPU with real sampler statistics.)"""
import ctypes as C

import numpy as np
import pytest

from whisper_rust_b200 import WhisperContext, WhisperContextParameters, synth

TOKEN = np.dtype([("id", "<i4"), ("tid", "<i4"), ("p", "<f4"), ("plog", "<f4"), ("pt", "<f4"), ("ptsum", "<f4"), ("t0", "<i8"), ("t1", "<i8"),
                  ("t_dtw", "<i8"), ("vlen", "<f4")], align=True)
assert TOKEN.itemsize == 56                                                         # whisper_token_data (wh:131-147)
FIELDS = [n for n in TOKEN.names]
V, I, F = C.c_void_p, C.c_int, C.c_float
TAIL = [V, I, V, V, V, V, I, F, F, I, I, V, I, V, V, V, I]


@pytest.fixture(scope="module")
def sides(prod_lib, ref_lib):
    prod_lib.whisper_b200_test_token_timestamps.restype = I
    prod_lib.whisper_b200_test_token_timestamps.argtypes = [V, I, I, I] + TAIL
    ref_lib.ref_probe_token_timestamps.restype = I
    ref_lib.ref_probe_token_timestamps.argtypes = [V, V] + TAIL
    ref_lib.whisper_token_to_str.restype = C.c_char_p
    ref_lib.whisper_token_to_str.argtypes = [V, I]
    blob = synth.model_bytes("micro", synth.InitRecipe(), with_tensors=False)
    cr = WhisperContext.new_from_buffer_with_params(blob, WhisperContextParameters(use_gpu=False), lib=ref_lib)
    st = cr.create_state()
    nv, eot, beg = cr.n_vocab(), cr.token_eot(), cr.token_beg()
    texts = [ref_lib.whisper_token_to_str(cr.ptr, i) or b"" for i in range(nv)]
    arr = (C.c_char_p * nv)(*texts)

    def run(energy, tokens, seg_n, seg_t0, seg_t1, thold_pt, thold_ptsum, max_len, split_on_word):
        energy = np.ascontiguousarray(energy, dtype=np.float32)
        seg_n = np.ascontiguousarray(seg_n, dtype=np.int32)
        seg_t0 = np.ascontiguousarray(seg_t0, dtype=np.int64)
        seg_t1 = np.ascontiguousarray(seg_t1, dtype=np.int64)
        cap_s, cap_t = 64 * len(seg_n) + 64, len(tokens) * 64 + 64                  # wrapping re-lists the tail of a segment per piece
        out = []
        for side in ("ref", "prod"):
            ot = np.zeros(cap_t, dtype=TOKEN)
            o0, o1, on = np.zeros(cap_s, dtype=np.int64), np.zeros(cap_s, dtype=np.int64), np.zeros(cap_s, dtype=np.int32)
            tail = (energy.ctypes.data, len(energy), tokens.ctypes.data, seg_n.ctypes.data, seg_t0.ctypes.data, seg_t1.ctypes.data, len(seg_n),
                    thold_pt, thold_ptsum, max_len, int(split_on_word), ot.ctypes.data, cap_t, o0.ctypes.data, o1.ctypes.data, on.ctypes.data, cap_s)
            if side == "ref":
                n = ref_lib.ref_probe_token_timestamps(cr.ptr, st.ptr, *tail)
            else:
                n = prod_lib.whisper_b200_test_token_timestamps(arr, nv, eot, beg, *tail)
            assert n >= 0
            out.append((n, o0[:n].copy(), o1[:n].copy(), on[:n].copy(), ot[:int(on[:n].sum())].copy()))
        return out

    yield run, eot, beg
    st.free(); cr.free()


def _clip_energy(rng, seconds):
    """what the engine keeps per clip: a moving average of |signal| (wc:8359-8388), here of gated noise"""
    n = int(seconds * 16000)
    env = (np.sin(np.arange(n) / 16000.0 * 2 * np.pi * rng.uniform(0.3, 2.0) + rng.uniform(0, 6)) > rng.uniform(-0.6, 0.6)).astype(np.float32)
    x = np.abs(rng.standard_normal(n).astype(np.float32)) * (0.02 + env)
    k = 32
    return np.convolve(x, np.ones(2 * k + 1, dtype=np.float32) / (2 * k + 1), mode="same").astype(np.float32)


def _segments(rng, seconds, eot, beg):
    """consecutive segments over the clip; each: [opening timestamp token?] text tokens [closing timestamp token?]"""
    toks, seg_n, seg_t0, seg_t1 = [], [], [], []
    t, t_end = 0, int(seconds * 100)
    while t < t_end - 20 and len(seg_n) < 12:
        dur = int(rng.integers(20, 900))
        a, b = t, min(t_end, t + dur)
        n_text = int(rng.integers(0, 30))
        ids = []
        if rng.random() < 0.6:
            ids.append(beg + (0 if (not seg_n and rng.random() < 0.7) else int(rng.integers(0, 1500))))
        ids += [int(rng.integers(1, eot)) if rng.random() > 0.3 else int(rng.integers(256, 5000)) for _ in range(n_text)]
        if rng.random() < 0.7:
            ids.append(beg + int(rng.integers(0, 1500)))
        if not ids:
            ids = [int(rng.integers(1, eot))]
        seg = np.zeros(len(ids), dtype=TOKEN)
        seg["id"] = ids
        # the sampler's best timestamp token per step: mostly advancing through the segment, sometimes stuck or jumping back
        base = beg + (a % 3000) // 2
        steps = np.cumsum(rng.integers(0, max(2, (b - a) // max(1, len(ids))), size=len(ids)))
        tid = base + steps + np.where(rng.random(len(ids)) < 0.15, -rng.integers(0, 40, size=len(ids)), 0)
        seg["tid"] = np.maximum(tid, beg)
        seg["p"] = rng.random(len(ids)).astype(np.float32)
        seg["plog"] = np.log(np.maximum(seg["p"], 1e-6)).astype(np.float32)
        seg["pt"] = (rng.random(len(ids)) ** 2).astype(np.float32)
        seg["ptsum"] = (rng.random(len(ids)) ** 2).astype(np.float32)
        seg["t0"], seg["t1"], seg["t_dtw"] = -1, -1, -1
        toks.append(seg); seg_n.append(len(ids)); seg_t0.append(a); seg_t1.append(b)
        t = b if rng.random() > 0.2 else b + int(rng.integers(0, 200))              # sometimes a gap (a seek past silence)
    return np.concatenate(toks), seg_n, seg_t0, seg_t1


STATS = dict(cases=0, tokens=0, distinct_times=0, pieces=0)


@pytest.mark.parametrize("seed", range(40))
def test_token_times_and_wrapping_equal_the_reference(sides, seed):
    run, eot, beg = sides
    rng = np.random.default_rng(1000 + seed)
    seconds = float(rng.choice([3.0, 11.0, 30.0]))
    energy = _clip_energy(rng, seconds)
    tokens, seg_n, seg_t0, seg_t1 = _segments(rng, seconds, eot, beg)
    thold_pt, thold_ptsum = [(0.01, 0.01), (0.3, 0.3), (0.0, 0.0), (0.9, 0.1)][seed % 4]
    for max_len, split in ((0, False), (int(rng.integers(1, 40)), False), (int(rng.integers(1, 40)), True)):
        (na, a0, a1, an, at), (nb, b0, b1, bn, bt) = run(energy, tokens, seg_n, seg_t0, seg_t1, thold_pt, thold_ptsum, max_len, split)
        ctx = (seed, max_len, split)
        assert na == nb, ctx
        np.testing.assert_array_equal(an, bn, err_msg=str(ctx))
        np.testing.assert_array_equal(a0, b0, err_msg=str(ctx))
        np.testing.assert_array_equal(a1, b1, err_msg=str(ctx))
        for f in FIELDS:
            x, y = at[f], bt[f]
            if x.dtype.kind == "f":
                x, y = x.view(np.uint32), y.view(np.uint32)
            np.testing.assert_array_equal(x, y, err_msg=f"{ctx} field {f}")
        if max_len == 0:
            assert na == len(seg_n)
            STATS["cases"] += 1
            STATS["tokens"] += len(at)
            STATS["distinct_times"] += len(np.unique(at["t0"]))
        else:
            STATS["pieces"] += nb - len(seg_n)


def test_the_fuzz_cases_were_not_trivial():
    if STATS["cases"] < 40:
        pytest.skip("runs after the 40 fuzz cases")
    # two thousand tokens, most of them given a time of their own, and the wrap producing over a thousand extra segments
    assert STATS["tokens"] > 2000 and STATS["distinct_times"] > STATS["tokens"] // 3 and STATS["pieces"] > 300, STATS


def test_single_token_and_empty_energy(sides):
    run, eot, beg = sides
    one = np.zeros(1, dtype=TOKEN)
    one["id"], one["tid"], one["t0"], one["t1"] = 300, beg + 5, -1, -1
    (na, a0, a1, an, at), (nb, b0, b1, bn, bt) = run(np.ones(16000, dtype=np.float32), one, [1], [10], [90], 0.01, 0.01, 0, False)
    assert na == nb == 1 and at["t0"][0] == bt["t0"][0] == 10 and at["t1"][0] == bt["t1"][0] == 90
    # no signal data: both sides leave the tokens as they are
    two = np.zeros(2, dtype=TOKEN)
    two["id"], two["t0"], two["t1"] = [300, 301], -1, -1
    (na, _, _, _, at), (nb, _, _, _, bt) = run(np.zeros(0, dtype=np.float32), two, [2], [0], [100], 0.01, 0.01, 0, False)
    assert na == nb == 1 and (at["t0"] == -1).all() and (bt["t0"] == -1).all()
