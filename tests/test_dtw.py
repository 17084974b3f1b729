"""X2: dtw_and_backtrace (wc:8647-8731) is integer/index work: given the same cost matrix the device wavefront + backtrace must
return the reference's path bit for bit, ties included (tie rule wc:8670-8679: diagonal only if strictly below both, else "up"
only if strictly below both, else "left").

* CPU (`not gpu`): a plain-Python restatement of the reference loop is pinned against the reference itself (ref_probe_dtw_path).
* GPU: the device kernels against the reference on random, tie-heavy and real (oracle alignment-head) matrices; X1's float
  post-processing (token-axis norm eps 1e-9, median filter 7, head mean, negate: wc:8864-8877) against the reference's ggml graph.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import model_blob
from whisper_rust_b200 import (DtwModelPreset, DtwParameters, FullParams, SamplingStrategy, WhisperContext, WhisperContextParameters,
                               synth)

RICH = synth.InitRecipe(te_std=0.2, pe_std=0.3, gain=3.0, ts_scale=1.6)


def dtw_restated(x):
    """x[N][M] -> (token indices, time indices): statement-for-statement semantics of wc:8647-8731 (test restatement)."""
    N, M = x.shape
    cost = np.full((N + 1, M + 1), np.inf, dtype=np.float32)
    trace = np.full((N + 1, M + 1), -1, dtype=np.int32)
    cost[0, 0] = 0.0
    for j in range(1, M + 1):
        for i in range(1, N + 1):
            c0, c1, c2 = cost[i - 1, j - 1], cost[i - 1, j], cost[i, j - 1]
            if c0 < c1 and c0 < c2:
                c, t = c0, 0
            elif c1 < c0 and c1 < c2:
                c, t = c1, 1
            else:
                c, t = c2, 2
            cost[i, j] = np.float32(x[i - 1, j - 1]) + c
            trace[i, j] = t
    trace[0, :] = 2
    trace[:, 0] = 1
    i, j, ti, tj = N, M, [], []
    while i > 0 or j > 0:
        ti.append(i - 1); tj.append(j - 1)
        t = trace[i, j]
        if t == 0:
            i -= 1; j -= 1
        elif t == 1:
            i -= 1
        else:
            j -= 1
    return np.array(ti[::-1], dtype=np.int32), np.array(tj[::-1], dtype=np.int32)


def _declare(lib, ref):
    if ref:
        lib.ref_probe_dtw_path.restype = C.c_int
        lib.ref_probe_dtw_path.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
        lib.ref_probe_dtw.restype = C.c_int
        lib.ref_probe_dtw.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
    else:
        lib.whisper_b200_test_dtw_path.restype = C.c_int
        lib.whisper_b200_test_dtw_path.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
        lib.whisper_b200_test_dtw.restype = C.c_int
        lib.whisper_b200_test_dtw.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
    return lib


def ref_path(ref, x):
    """the reference's dtw_and_backtrace on x[N][M]"""
    N, M = x.shape
    xt = np.ascontiguousarray(x.T, dtype=np.float32)   # ggml order: token index contiguous
    out = np.zeros(2 * (N + M), dtype=np.int32)
    L = _declare(ref, True).ref_probe_dtw_path(xt.ctypes.data, N, M, out.ctypes.data, out.size)
    return out[:L].copy(), out[L:2 * L].copy()


def prod_path(lib, x):
    N, M = x.shape
    xc = np.ascontiguousarray(x, dtype=np.float32)
    out = np.zeros(2 * (N + M), dtype=np.int32)
    L = _declare(lib, False).whisper_b200_test_dtw_path(xc.ctypes.data, N, M, out.ctypes.data, out.size)
    assert L > 0
    return out[:L].copy(), out[L:2 * L].copy()


def cost_cases():
    rng = np.random.default_rng(99)
    cases = {"ones_3x4": np.ones((3, 4), dtype=np.float32), "zeros_3x4": np.zeros((3, 4), dtype=np.float32), "zeros_1x9": np.zeros((1, 9), dtype=np.float32)}
    # (M = 1 is outside the reference's domain: dtw_and_backtrace asserts ggml_n_dims(x) == 2, wc:8648)
    for (N, M) in ((2, 8), (7, 40), (66, 750), (224, 1500), (120, 333)):
        cases[f"gauss_{N}x{M}"] = -rng.standard_normal((N, M)).astype(np.float32)
        cases[f"ties_{N}x{M}"] = -rng.integers(0, 3, size=(N, M)).astype(np.float32)          # heavy ties: the rule decides everything
        q = -(rng.integers(0, 64, size=(N, M)) / 64.0).astype(np.float32)                       # exactly representable sums -> equal paths costs
        cases[f"dyadic_{N}x{M}"] = q
    return cases


SMALL = ["ones_3x4", "zeros_1x9", "gauss_2x8", "ties_2x8", "gauss_7x40", "ties_7x40", "dyadic_7x40"]


@pytest.mark.parametrize("key", SMALL + ["ties_66x750"])
def test_restatement_equals_reference_dtw(ref_lib, key):
    x = cost_cases()[key]
    a, b = ref_path(ref_lib, x)
    c, d = dtw_restated(x)
    np.testing.assert_array_equal(a, c)
    np.testing.assert_array_equal(b, d)


def test_tie_rule_known_answer(ref_lib):
    """all-equal costs: every comparison ties, so the rule always answers "left" until the first row pins "up" (wc:8670-8695)"""
    a, b = ref_path(ref_lib, np.zeros((3, 4), dtype=np.float32))
    assert a.tolist() == [0, 1, 2, 2, 2, 2] and b.tolist() == [0, 0, 0, 1, 2, 3]
    a, b = ref_path(ref_lib, np.ones((3, 4), dtype=np.float32))   # positive costs: the diagonal is strictly cheapest
    assert a.tolist() == [0, 1, 2, 2] and b.tolist() == [0, 1, 2, 3]


@pytest.mark.gpu
@pytest.mark.parametrize("key", list(cost_cases().keys()))
def test_device_dtw_path_is_bit_identical(prod_lib, ref_lib, key):
    x = cost_cases()[key]
    a, b = ref_path(ref_lib, x)
    c, d = prod_path(prod_lib, x)
    np.testing.assert_array_equal(a, c)
    np.testing.assert_array_equal(b, d)


def _oracle_qk(ref, seconds, seed):
    """alignment-head softmax rows of the LAST window of a DTW run on the oracle: data[head][audio_ctx][token] (wc:8842-8852)"""
    ctx = WhisperContext.new_from_buffer_with_params(model_blob("mini", RICH),
                                                     WhisperContextParameters(use_gpu=False, dtw_parameters=DtwParameters(preset=DtwModelPreset.N_TOP_MOST, n_top=2)), lib=ref)
    st = ctx.create_state()
    p = FullParams(ref, SamplingStrategy.GREEDY, best_of=1)
    p.set_n_threads(8).set_language("en").set_temperature_inc(0.0)
    assert st.full(p, synth.synth_audio(seconds, seed=seed)) == 0
    n = C.c_int64()
    ptr = ref.ref_probe_aheads_qks(st.ptr, C.byref(n))
    H, T = 2 * 3, 1500                                   # n_top=2 layers x 3 heads (mini)
    n_tok = n.value // (H * T)
    data = np.ctypeslib.as_array(ptr, shape=(H, T, n_tok)).copy()
    st.free(); ctx.free()
    return data


@pytest.mark.gpu
@pytest.mark.parametrize("seconds,seed,n_audio", [(30.0, 26, 1500), (47.3, 25, 865), (11.0, 3, 550), (30.0, 26, 8)])
def test_device_alignment_chain_vs_reference(prod_lib, ref_lib, seconds, seed, n_audio):
    """X1+X2 on the oracle's own alignment-head tensor: cost matrix to float tolerance, the path from the REFERENCE's cost matrix
    bit-identical, and the path from the device's own cost matrix identical as well."""
    data = _oracle_qk(ref_lib, seconds, seed)
    H, T, n_tok = data.shape
    sot_len, N = 2, n_tok - 3
    assert N > 0
    _declare(ref_lib, True)
    cost_ref = np.zeros((n_audio, N), dtype=np.float32)
    out = np.zeros(2 * (N + n_audio), dtype=np.int32)
    L = ref_lib.ref_probe_dtw(data.ctypes.data, n_tok, T, H, n_audio, sot_len, 7, cost_ref.ctypes.data, out.ctypes.data, out.size)
    ra, rb = out[:L].copy(), out[L:2 * L].copy()
    qk = np.ascontiguousarray(data.transpose(0, 2, 1))   # [H][n_tok][T]: the product's capture layout
    cost_dev = np.zeros((n_tok, n_audio), dtype=np.float32)
    out2 = np.zeros(2 * (N + n_audio), dtype=np.int32)
    L2 = _declare(prod_lib, False).whisper_b200_test_dtw(qk.ctypes.data, H, n_tok, T, n_audio, 7, sot_len, cost_dev.ctypes.data,
                                                         out2.ctypes.data, out2.size)
    assert L2 > 0
    cd = cost_dev[sot_len:sot_len + N]                    # rows kept (wc:8882)
    err = np.max(np.abs(cd - cost_ref.T))
    print("alignment cost max abs err", err, "max |cost|", np.max(np.abs(cost_ref)))
    assert err <= 2e-6 * max(1.0, float(np.max(np.abs(cost_ref))))
    # the integer part, strictly: same matrix in -> same path out
    pa, pb = prod_path(prod_lib, np.ascontiguousarray(cost_ref.T))
    np.testing.assert_array_equal(pa, ra)
    np.testing.assert_array_equal(pb, rb)
    # and end to end from the QK rows
    np.testing.assert_array_equal(out2[:L2], ra)
    np.testing.assert_array_equal(out2[L2:2 * L2], rb)


def test_alignment_head_preset_tables_equal_the_reference(prod_lib, ref_lib):
    """DtwModelPreset::{TinyEn .. LargeV3Turbo} (rs/whisper_ctx.rs DtwModelPreset -> whisper_alignment_heads_preset, wh:96-112): the
    (text layer, head) tables of whisper.cpp:417-443, entry for entry and in order (the slot order of the captured QK rows follows it)"""
    import ctypes as C
    import numpy as np
    for lib, name in ((prod_lib, "whisper_b200_test_aheads_preset"), (ref_lib, "ref_probe_aheads_preset")):
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = C.c_int, [C.c_int, C.c_void_p, C.c_int]
    n_tables = 0
    for preset in range(-1, 20):
        a, b = np.full((64, 2), -7, dtype=np.int32), np.full((64, 2), -7, dtype=np.int32)
        na = ref_lib.ref_probe_aheads_preset(preset, a.ctypes.data, 64)
        nb = prod_lib.whisper_b200_test_aheads_preset(preset, b.ctypes.data, 64)
        assert na == nb, preset
        np.testing.assert_array_equal(a, b, err_msg=f"preset {preset}")
        n_tables += int(na > 0)
    assert n_tables == 12
