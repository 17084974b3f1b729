"""K1: the unified self-attention KV cache's cell bookkeeping (csrc/engine.cu: KvCache::{find_slot, cell_max, seq_rm, seq_cp, clear})
against the reference's static whisper_kv_cache_* functions (whisper.cpp:1049-1167; reached through oracle/ref_probe.cpp) WITHOUT a
device: random operation scripts and scripts shaped like whisper_full's own use (prompt into sequence 0, copies to the decoders, one
cell per decoder and step, the beam-search reshuffle through the upper half of the sequence ids, wc:7243-7291).  Which cell a token
lands in and which cells a sequence attends to follow from this state alone, so it must be identical after every operation."""
import ctypes as C

import numpy as np
import pytest


@pytest.fixture(scope="module")
def run(prod_lib, ref_lib):
    for lib, name in ((prod_lib, "whisper_b200_test_kv_script"), (ref_lib, "ref_probe_kv_script")):
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]

    def both(size, ops):
        ops = np.ascontiguousarray([list(o) + [0] * (5 - len(o)) for o in ops], dtype=np.int32).reshape(-1, 5)
        out = []
        for fn in (ref_lib.ref_probe_kv_script, prod_lib.whisper_b200_test_kv_script):
            per_op = np.full((len(ops), 3), -99, dtype=np.int32)
            cells = np.full((size, 2), -99, dtype=np.int32)
            assert fn(size, ops.ctypes.data, len(ops), per_op.ctypes.data, cells.ctypes.data) == 0
            out.append((per_op, cells))
        return out

    return both


def _beam_script(rng, size, n_dec, n_steps):
    ops = [(0, 0, 0, 0, 0)]
    n_prompt = int(rng.integers(1, max(2, size // 3)))
    ops.append((1, n_prompt, 0, 0, 0))
    for j in range(1, n_dec):
        ops.append((3, 0, j, -1, -1))
    pos = n_prompt
    for _ in range(n_steps):
        for j in range(n_dec):
            ops.append((1, 1, j, pos, 0))
        pos += 1
        if rng.random() < 0.5:                                   # the beam reshuffle: park, drop, re-assign
            src = rng.integers(0, n_dec, size=n_dec)
            for j in range(n_dec):
                ops.append((3, j, n_dec + j, -1, -1))
            for j in range(n_dec):
                ops.append((2, j, -1, -1))
            for j in range(n_dec):
                ops.append((3, n_dec + int(src[j]), j, -1, -1))
            for j in range(n_dec):
                ops.append((2, n_dec + j, -1, -1))
        if rng.random() < 0.1:
            ops.append((4, 0, 0, 0, 0))
        if rng.random() < 0.05:                                  # a failed decoder rolls back to the prompt (wc:7078: seq_rm(j, n_prompt, -1))
            ops.append((2, int(rng.integers(0, n_dec)), n_prompt, -1))
    return ops


def _random_script(rng, size, n_ops):
    ops = []
    for _ in range(n_ops):
        k = rng.choice([0, 1, 1, 1, 1, 2, 2, 3, 3, 4], p=None)
        if k == 0 and rng.random() < 0.8:
            k = 1                                                # clears are rare
        if k == 1:
            ops.append((1, int(rng.integers(0, max(2, size // 4))) if rng.random() < 0.9 else int(rng.integers(size - 2, size + 3)),
                        int(rng.integers(0, 10)), int(rng.integers(0, 400)), 0))
        elif k == 2:
            ops.append((2, int(rng.integers(-1, 10)), int(rng.integers(-1, 300)), int(rng.integers(-1, 500)), 0))
        elif k == 3:
            ops.append((3, int(rng.integers(0, 10)), int(rng.integers(0, 10)), int(rng.integers(-1, 300)), int(rng.integers(-1, 500))))
        else:
            ops.append((int(k), 0, 0, 0, 0))
    return ops


@pytest.mark.parametrize("seed", range(24))
def test_kv_cell_state_equals_the_reference(run, seed):
    rng = np.random.default_rng(300 + seed)
    size = int(rng.choice([8, 33, 64, 224, 448, 512]))
    if seed % 2 == 0:
        n_dec = int(rng.integers(1, 6))
        ops = _beam_script(rng, size, n_dec, int(rng.integers(5, 120)))
    else:
        ops = _random_script(rng, size, 400)
    (pa, ca), (pb, cb) = run(size, ops)
    bad = np.nonzero((pa != pb).any(axis=1))[0]
    assert len(bad) == 0, f"first differing op {bad[0]}: {ops[bad[0]]} ref {pa[bad[0]]} got {pb[bad[0]]}"
    np.testing.assert_array_equal(ca, cb)
    assert (ca[:, 0] >= 0).any() or seed % 2                     # the beam scripts leave a populated cache behind
