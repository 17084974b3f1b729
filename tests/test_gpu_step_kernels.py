"""The decode-step kernels (whisper_rust_b200/csrc/step.cu) against plain torch f32 references, through the C ABI."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib(prod_lib):
    prod_lib.whisper_b200_test_gemm_wstream.restype = C.c_int
    prod_lib.whisper_b200_test_gemm_wstream.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
    prod_lib.whisper_b200_test_gemm_wcluster.restype = C.c_int
    prod_lib.whisper_b200_test_gemm_wcluster.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                                         C.c_float, C.c_int]
    prod_lib.whisper_b200_test_cross_attn_step.restype = C.c_int
    prod_lib.whisper_b200_test_cross_attn_step.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
    return prod_lib


@pytest.mark.parametrize("n,n_out,k,split", [(1, 128, 64, 1), (16, 1280, 1280, 0), (64, 3840, 1280, 0), (5, 384, 384, 0), (64, 1280, 5120, 0),
                                             (33, 51866, 384, 1), (200, 512, 512, 3), (256, 256, 2048, 0), (7, 192, 192, 1)])
def test_gemm_wstream_vs_torch(lib, n, n_out, k, split):
    import torch
    torch.manual_seed(n * 131 + n_out)
    w = (torch.randn(n_out, k, device="cuda") * 0.5).half()
    x = (torch.randn(n, k, device="cuda") * 0.5).half()
    bias = torch.randn(n_out, device="cuda", dtype=torch.float32)
    out = torch.full((n, n_out), float("nan"), device="cuda", dtype=torch.float32)
    torch.cuda.synchronize()
    assert lib.whisper_b200_test_gemm_wstream(w.data_ptr(), x.data_ptr(), out.data_ptr(), n, n_out, k, bias.data_ptr(), split) == 0
    torch.cuda.synchronize()
    ref = x.float() @ w.float().T + bias
    assert torch.isfinite(out).all()
    err = (out - ref).abs().max().item()
    assert err < 2e-3 * max(1.0, ref.abs().max().item()), err


# the cluster split-K GEMM (partials reduced through distributed shared memory): f16 output with the column-range scale
# (the decoder's Q|K|V: wc:2558-2583) and the in-place f32 residual (wc:2657-2664), over ragged token counts and every
# cluster size the plan produces (tiles 1..40, k-blocks 3..80)
@pytest.mark.parametrize("n,n_out,k", [(1, 128, 192), (16, 1280, 1280), (64, 3840, 1280), (5, 384, 384), (64, 1280, 5120),
                                        (200, 512, 512), (256, 1280, 1280), (256, 5120, 1280), (255, 1152, 384), (37, 1536, 384)])
@pytest.mark.parametrize("mode", [0, 1])
def test_gemm_wcluster_vs_torch(lib, n, n_out, k, mode):
    import torch
    torch.manual_seed(n * 131 + n_out + mode)
    w = (torch.randn(n_out, k, device="cuda") * 0.5).half()
    x = (torch.randn(n, k, device="cuda") * 0.5).half()
    bias = torch.randn(n_out, device="cuda", dtype=torch.float32)
    ref = x.float() @ w.float().T + bias
    scale_f1 = (n_out // 3) // 4 * 4 * 2
    if mode == 0:
        out = torch.full((n, n_out), float("nan"), device="cuda", dtype=torch.float16)
        ref[:, :scale_f1] *= 0.35355338
    else:
        out = torch.randn(n, n_out, device="cuda", dtype=torch.float32)
        ref = ref + out
    torch.cuda.synchronize()
    assert lib.whisper_b200_test_gemm_wcluster(w.data_ptr(), x.data_ptr(), out.data_ptr(), n, n_out, k, bias.data_ptr(), mode,
                                               0.35355338, scale_f1) == 0
    torch.cuda.synchronize()
    o = out.float()
    assert torch.isfinite(o).all()
    err = (o - ref).abs().max().item()
    assert err < (6e-3 if mode == 0 else 2e-3) * max(1.0, ref.abs().max().item()), err


@pytest.mark.parametrize("n,H,T", [(1, 2, 1500), (16, 20, 1500), (3, 6, 92), (4, 3, 128), (2, 8, 777)])
def test_cross_attn_step_vs_torch(lib, n, H, T):
    import torch
    torch.manual_seed(n + H + T)
    S = H * 64
    q = (torch.randn(n, S, device="cuda") * 0.7).half()
    kv = (torch.randn(2 * H, T, 64, device="cuda") * 0.7).half()     # K heads then V heads (layer 0)
    out = torch.zeros(n, S, device="cuda", dtype=torch.float16)
    torch.cuda.synchronize()
    assert lib.whisper_b200_test_cross_attn_step(q.data_ptr(), kv.data_ptr(), out.data_ptr(), n, H, T) == 0
    torch.cuda.synchronize()
    scale = 64.0 ** -0.25
    qh = q.float().view(n, H, 64)
    k, v = kv[:H].float(), kv[H:].float()
    sc = torch.einsum("nhd,htd->nht", qh, k) * scale
    p = torch.softmax(sc, dim=-1).half().float()
    ref = torch.einsum("nht,htd->nhd", p, v).reshape(n, S)
    err = (out.float() - ref).abs().max().item()
    assert err < 3e-3 * max(1.0, ref.abs().max().item()), err
    # and against the general-path kernel (same rounding points)
    out2 = torch.zeros_like(out)
    assert lib.whisper_b200_test_cross_attn(q.data_ptr(), kv.data_ptr(), out2.data_ptr(), n, H, T) == 0
    torch.cuda.synchronize()
    assert (out.float() - out2.float()).abs().max().item() < 2e-3
