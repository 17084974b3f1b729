"""whisper_b200_gemm_wstream_noalloc (include/whisper_b200.h): the split-K streamed-weight GEMM of the decode step with its partial sums in a
fixed array inside the library instead of a cudaMalloc'ed workspace.  No counterpart in the reference;
the checker is the existing allocating entry point whisper_b200_test_gemm_wstream on the same inputs.

Collected last on purpose.  Passes alone in a fresh process and after the other tests: it depends on
nothing but the library, and every call clears the part of the shared array it uses."""
import ctypes as C

import pytest

pytestmark = pytest.mark.gpu

MAX_FLOATS = 2 * 1024 * 1024   # the array's stated size (include/whisper_b200.h)


@pytest.fixture(scope="module")
def lib(prod_lib):
    sig = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
    prod_lib.whisper_b200_test_gemm_wstream.restype = C.c_int
    prod_lib.whisper_b200_test_gemm_wstream.argtypes = sig
    prod_lib.whisper_b200_gemm_wstream_noalloc.restype = C.c_int
    prod_lib.whisper_b200_gemm_wstream_noalloc.argtypes = sig
    return prod_lib


def _inputs(n, n_out, k, seed, with_bias=True):
    import torch
    g = torch.Generator(device="cuda").manual_seed(seed)
    w = (torch.randn(n_out, k, device="cuda", generator=g) * 0.5).half()
    x = (torch.randn(n, k, device="cuda", generator=g) * 0.5).half()
    bias = torch.randn(n_out, device="cuda", dtype=torch.float32, generator=g) if with_bias else None
    return w, x, bias


def _both(lib, n, n_out, k, split, seed, with_bias=True):
    """(existing entry point's output, no-alloc entry point's output, device bytes the no-alloc call took from the allocator)"""
    import torch
    w, x, bias = _inputs(n, n_out, k, seed, with_bias)
    bp = bias.data_ptr() if with_bias else None
    want = torch.full((n, n_out), float("nan"), device="cuda", dtype=torch.float32)
    got = torch.full((n, n_out), float("nan"), device="cuda", dtype=torch.float32)
    torch.cuda.synchronize()
    assert lib.whisper_b200_test_gemm_wstream(w.data_ptr(), x.data_ptr(), want.data_ptr(), n, n_out, k, bp, split) == 0
    torch.cuda.synchronize()
    # device memory taken by the call = free before - free after (cudaMemGetInfo).  Another thread of the process freeing or allocating
    # at that moment would show up here too, so the call is measured up to five times and the smallest magnitude is reported.
    taken = None
    for _ in range(5):
        free0, _ = torch.cuda.mem_get_info()
        assert lib.whisper_b200_gemm_wstream_noalloc(w.data_ptr(), x.data_ptr(), got.data_ptr(), n, n_out, k, bp, split) == 0
        free1, _ = torch.cuda.mem_get_info()
        taken = free0 - free1 if taken is None or abs(free0 - free1) < abs(taken) else taken
        if taken == 0:
            break
    torch.cuda.synchronize()
    return want, got, taken, (x.float() @ w.float().T + (bias if with_bias else 0.0))


def _check(want, got, ref):
    import torch
    assert torch.isfinite(got).all()
    tol = 2e-3 * max(1.0, ref.abs().max().item())           # the bar of tests/test_gpu_step_kernels.py
    assert (got - want).abs().max().item() < tol
    assert (got - ref).abs().max().item() < tol
    assert torch.equal(got, want)                           # same kernel, same split plan, same summation order: identical bits


def test_no_alloc_gemm_wstream_equals_the_allocating_entry_point_on_two_calls_in_a_row(lib):
    # first call: a large footprint in the shared array (split 8 asked, 7 planned: 7 x 64 rows x 3840 outputs = 1,720,320 of 2,097,152 floats) ...
    want, got, taken, ref = _both(lib, 64, 3840, 1280, 8, seed=1)
    _check(want, got, ref)
    # ... second call, straight after: other data, a smaller footprint over the same (now stale) region, the planner's split, no bias
    want2, got2, taken2, ref2 = _both(lib, 37, 1280, 1280, 0, seed=2, with_bias=False)
    _check(want2, got2, ref2)
    assert taken2 == 0, f"the second call took {taken2} bytes of device memory"
    # third: the first problem again gives the first answer again (nothing of call two leaks into it)
    want3, got3, taken3, ref3 = _both(lib, 64, 3840, 1280, 8, seed=1)
    _check(want3, got3, ref3)
    import torch
    assert torch.equal(got3, got) and taken3 == 0


@pytest.mark.parametrize("n,n_out,k,split", [(1, 128, 192, 3), (5, 384, 384, 0), (200, 512, 512, 3), (256, 256, 2048, 0), (16, 51866, 384, 2)])
def test_no_alloc_gemm_wstream_shapes(lib, n, n_out, k, split):
    want, got, _, ref = _both(lib, n, n_out, k, split, seed=n * 131 + n_out)
    _check(want, got, ref)


def test_no_alloc_gemm_wstream_rejects_what_does_not_fit(lib):
    import torch
    n, n_out, k = 256, 51866, 384                           # 1 split x 256 x 51866 = 13,277,696 floats > MAX_FLOATS
    assert n * n_out > MAX_FLOATS
    w, x, bias = _inputs(n, n_out, k, seed=3)
    out = torch.full((n, n_out), float("nan"), device="cuda", dtype=torch.float32)
    torch.cuda.synchronize()
    assert lib.whisper_b200_gemm_wstream_noalloc(w.data_ptr(), x.data_ptr(), out.data_ptr(), n, n_out, k, bias.data_ptr(), 0) == 2
    torch.cuda.synchronize()
    assert torch.isnan(out).all()                           # refused before anything was launched
    assert lib.whisper_b200_gemm_wstream_noalloc(w.data_ptr(), x.data_ptr(), out.data_ptr(), 0, n_out, k, None, 0) == 1
    assert lib.whisper_b200_gemm_wstream_noalloc(None, x.data_ptr(), out.data_ptr(), 4, n_out, k, None, 0) == 1
    # and the library still works afterwards
    want, got, _, ref = _both(lib, 4, 256, 256, 2, seed=4)
    _check(want, got, ref)
