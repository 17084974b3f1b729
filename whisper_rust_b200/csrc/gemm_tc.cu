// tcgen05/TMEM GEMM (sm_100a).  See gemm_tc.cuh for the contract.
//
// Kernel anatomy (one 128 x BLOCK_N output tile per CTA, 192 threads):
//   warp 0      : TMA producer  — cp.async.bulk.tensor 3-D loads of the A [128 x 64] and B [BLOCK_N x 64]
//                 f16 k-blocks into a kStages-deep shared-memory ring (128-byte swizzle), mbarrier tx-count.
//   warp 1      : TMEM allocator + MMA issuer — one elected lane issues 4 x tcgen05.mma (K=16 each) per
//                 k-block, tcgen05.commit releases the ring slot; a final commit signals the epilogue.
//   warps 2..5  : epilogue — tcgen05.ld 32 lanes x 32 columns at a time, fused bias / scale / GELU /
//                 residual, f16 or f32 store in one of four layouts.
//
// Replaces, for the Whisper graphs, the reference's ggml MUL_MAT + ADD + SCALE + GELU + CPY node chains
// (sys/whisper.cpp/src/whisper.cpp:2131-2148,2212-2254,2321-2356,2558-2606,2657-2835) with the
// rounding points of its CPU path: f16 operands, f32 accumulate (ggml-cpu.c:1269-1461).
#include "gemm_tc.cuh"
#include "prof.h"

#include <atomic>
#include <map>
#include <mutex>
#include <tuple>

namespace wb {

static constexpr int BLOCK_M = 128;
static constexpr int BLOCK_K = 64;  // 64 f16 = 128 bytes = one swizzle row
static constexpr int GEMM_THREADS = 192;

__host__ __device__ constexpr int gemm_stages(int block_n) { return block_n >= 256 ? 4 : (block_n >= 128 ? 6 : 8); }
__host__ __device__ constexpr int gemm_stage_bytes(int block_n) { return (BLOCK_M + block_n) * BLOCK_K * 2; }
__host__ __device__ constexpr int gemm_smem_bytes(int block_n) {
    return gemm_stages(block_n) * gemm_stage_bytes(block_n) + 1024 /*align slack*/ + 256 /*barriers*/;
}

__device__ __forceinline__ float gelu_ref_f16table(float x, const __half * __restrict__ table) {
    // ggml_vec_gelu_f32 with GGML_GELU_FP16 (ggml-cpu/vec.h:571-585)
    if (x <= -10.0f) return 0.0f;
    if (x >= 10.0f) return x;
    const __half h = __float2half_rn(x);
    return __half2float(__ldg(table + __half_as_ushort(h)));
}

template <int BLOCK_N>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_f16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, int M, int N, int K,
                        int a_bcast, int b_bcast, GemmTaps taps, GemmEpilogue epi) {
    constexpr int kStages     = gemm_stages(BLOCK_N);
    constexpr int kStageBytes = gemm_stage_bytes(BLOCK_N);
    constexpr int kABytes     = BLOCK_M * BLOCK_K * 2;

    extern __shared__ uint8_t smem_raw[];
    uint8_t * smem = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);

    uint64_t * full_bar  = (uint64_t *) (smem + kStages * kStageBytes);
    uint64_t * empty_bar = full_bar + kStages;
    uint64_t * tmem_full = empty_bar + kStages;
    uint32_t * tmem_slot = (uint32_t *) (tmem_full + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    const int n0    = blockIdx.x * BLOCK_N;
    const int m0    = blockIdx.y * BLOCK_M;
    const int batch = blockIdx.z;

    const int kb_per_tap = (K + BLOCK_K - 1) / BLOCK_K;  // K = reduction length of one tap
    const int num_kb     = kb_per_tap * taps.taps;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(full_bar + s, 1);
            mbar_init(empty_bar + s, 1);
        }
        mbar_init(tmem_full, 1);
        mbar_fence_init();
    }
    if (warp == 1) {
        tmem_alloc<BLOCK_N>(tmem_slot);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            const int ba = a_bcast ? 0 : batch;
            const int bb = b_bcast ? 0 : batch;
            for (int kb = 0; kb < num_kb; ++kb) {
                const int tap = kb / kb_per_tap;
                const int kk  = (kb - tap * kb_per_tap) * BLOCK_K;
                mbar_wait(empty_bar + stage, phase ^ 1);
                uint8_t * sa = smem + stage * kStageBytes;
                uint8_t * sb = sa + kABytes;
                mbar_expect_tx(full_bar + stage, kStageBytes);
                tma_load_4d(sa, &tmap_a, full_bar + stage, kk, taps.p[tap], m0 + taps.r[tap], ba);
                tma_load_4d(sb, &tmap_b, full_bar + stage, tap * taps.k_tap + kk, 0, n0, bb);
                if (++stage == kStages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            constexpr uint32_t idesc = umma_idesc_f16(BLOCK_M, BLOCK_N);
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(full_bar + stage, phase);
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + stage * kStageBytes);
                const uint32_t sb = sa + kABytes;
                const uint64_t da = umma_desc_k_sw128(sa);
                const uint64_t db = umma_desc_k_sw128(sb);
#pragma unroll
                for (int k = 0; k < BLOCK_K / 16; ++k) {
                    // advance 16 f16 = 32 bytes along K inside the swizzled row: +2 in the (addr >> 4) field
                    umma_f16_ss(tmem_base, da + (uint64_t) (2 * k), db + (uint64_t) (2 * k), idesc, (kb | k) != 0);
                }
                umma_commit(empty_bar + stage);  // frees the ring slot when these MMAs retire
                if (++stage == kStages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            umma_commit(tmem_full);  // accumulator complete
        }
    } else {
        // ---------------- epilogue: warps 2..5, TMEM lane group = warp % 4 ----------------
        const int grp = warp & 3;
        const int m   = m0 + grp * 32 + lane;
        const bool group_live = (m0 + grp * 32) < M;  // warp-uniform

        mbar_wait(tmem_full, 0);
        tc_fence_after();

        if (group_live) {
            const bool row_ok = m < M;
            const int64_t ob  = (int64_t) (batch / epi.batch_div) * epi.out_batch_stride + (int64_t) (batch % epi.batch_div) * epi.out_batch_stride2;
            const float * bias = epi.bias ? epi.bias + (int64_t) batch * epi.bias_batch_stride : nullptr;
            int64_t mcol = m;  // row coordinate used by the transposed layouts
            if (epi.rows_per_panel > 0) mcol = (int64_t) (m / epi.rows_per_panel) * epi.panel_stride + (m % epi.rows_per_panel);
            const float * res_row = nullptr;
            if (epi.res != nullptr) {
                const int mr = epi.res_row_mod > 0 ? (m % epi.res_row_mod) : m;
                res_row = epi.res + (int64_t) batch * epi.res_batch_stride + (int64_t) mr * epi.ldr;
            }
#pragma unroll 1
            for (int c = 0; c < BLOCK_N / 32; ++c) {
                const int nc = n0 + c * 32;
                if (nc >= N) break;  // warp-uniform
                uint32_t r[32];
                tmem_ld_32x32(tmem_base + ((uint32_t) (grp * 32) << 16) + (uint32_t) (c * 32), r);
                tmem_ld_wait();

                float v[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int n = nc + j;
                    float x = __uint_as_float(r[j]);
                    if (n < N) {
                        if (bias != nullptr) x = __fadd_rn(x, __ldg(bias + n));
                        if (n >= epi.scale_n0 && n < epi.scale_n1) x = __fmul_rn(x, epi.scale);
                        if (epi.gelu_table != nullptr) x = gelu_ref_f16table(x, epi.gelu_table);
                        if (res_row != nullptr && row_ok) x = __fadd_rn(x, res_row[n]);
                    }
                    v[j] = x;
                }

                const bool full_chunk = (nc + 32 <= N);
                if (epi.out_mode == OUT_ROW || epi.out_mode == OUT_HEAD) {
                    if (row_ok) {
                        int64_t off;
                        if (epi.out_mode == OUT_ROW) {
                            off = ob + (int64_t) m * epi.ldo + nc;
                        } else {
                            off = ob + (int64_t) (nc >> 6) * epi.head_stride + (int64_t) m * 64 + (nc & 63);
                        }
                        if (epi.out_f16) {
                            __half * o = (__half *) epi.out + off;
                            if (full_chunk && ((off & 7) == 0)) {
#pragma unroll
                                for (int q = 0; q < 4; ++q) {
                                    uint4 pk;
                                    __half2 h0 = __floats2half2_rn(v[q * 8 + 0], v[q * 8 + 1]);
                                    __half2 h1 = __floats2half2_rn(v[q * 8 + 2], v[q * 8 + 3]);
                                    __half2 h2 = __floats2half2_rn(v[q * 8 + 4], v[q * 8 + 5]);
                                    __half2 h3 = __floats2half2_rn(v[q * 8 + 6], v[q * 8 + 7]);
                                    pk.x = *(uint32_t *) &h0;
                                    pk.y = *(uint32_t *) &h1;
                                    pk.z = *(uint32_t *) &h2;
                                    pk.w = *(uint32_t *) &h3;
                                    *(uint4 *) (o + q * 8) = pk;
                                }
                            } else {
#pragma unroll
                                for (int j = 0; j < 32; ++j)
                                    if (nc + j < N) o[j] = __float2half_rn(v[j]);
                            }
                        } else {
                            float * o = (float *) epi.out + off;
                            if (full_chunk && ((off & 3) == 0)) {
#pragma unroll
                                for (int q = 0; q < 8; ++q)
                                    *(float4 *) (o + q * 4) = make_float4(v[q * 4 + 0], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
                            } else {
#pragma unroll
                                for (int j = 0; j < 32; ++j)
                                    if (nc + j < N) o[j] = v[j];
                            }
                        }
                    }
                } else {
                    // transposed layouts: for a fixed column the 32 lanes hold 32 consecutive m -> coalesced
                    if (row_ok) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const int n = nc + j;
                            if (n < N) {
                                int64_t off;
                                if (epi.out_mode == OUT_COL) {
                                    off = ob + (int64_t) n * epi.ldo + mcol;
                                } else {
                                    off = ob + (int64_t) (n >> 6) * epi.head_stride + (int64_t) (n & 63) * epi.ldo + mcol;
                                }
                                if (epi.out_f16) ((__half *) epi.out)[off] = __float2half_rn(v[j]);
                                else ((float *) epi.out)[off] = v[j];
                            }
                        }
                    }
                }
            }
        }
        tc_fence_before();
    }

    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<BLOCK_N>(tmem_base);
    }
}

// ---------------------------------------------------------------------------------------------
// host side: tensor-map construction (driver entry point fetched through the runtime) and launch
// ---------------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
    static PFN_encodeTiled fn = nullptr;
    static std::once_flag once;
    std::call_once(once, []() {
        void * p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess) {
            fn = (PFN_encodeTiled) p;
        }
    });
    return fn;
}

static bool make_tmap(CUtensorMap * out, const GemmOperand & op, int K, int box_rows) {
    PFN_encodeTiled enc = get_encode_fn();
    WB_CHECK(enc != nullptr, "cuTensorMapEncodeTiled entry point unavailable\n");
    WB_CHECK(((uintptr_t) op.ptr & 15) == 0, "operand base must be 16-byte aligned\n");
    WB_CHECK((op.row_stride * 2) % 16 == 0, "operand row stride must be a multiple of 16 bytes (got %lld elements)\n",
             (long long) op.row_stride);
    const int64_t nb = op.batch > 0 ? op.batch : 1;
    int64_t bstride = op.batch_stride > 0 ? op.batch_stride : op.rows * op.row_stride;
    if ((bstride * 2) % 16 != 0) {
        WB_CHECK(nb == 1, "operand batch stride must be a multiple of 16 bytes\n");
        bstride = pad_to(bstride, 8);
    }
    const int64_t np = op.phases > 1 ? op.phases : 1;
    // dims {k, phase, rows, batch}; with phases the "row" stride is phases*row_stride apart per phase pair
    const int64_t pstride = np > 1 ? op.phase_stride : op.row_stride;
    WB_CHECK((pstride * 2) % 16 == 0, "phase stride must be a multiple of 16 bytes\n");
    cuuint64_t dims[4]    = {(cuuint64_t) K, (cuuint64_t) np, (cuuint64_t) op.rows, (cuuint64_t) nb};
    cuuint64_t strides[3] = {(cuuint64_t) pstride * 2, (cuuint64_t) op.row_stride * 2, (cuuint64_t) bstride * 2};
    cuuint32_t box[4]     = {(cuuint32_t) BLOCK_K, 1, (cuuint32_t) box_rows, 1};
    cuuint32_t estr[4]    = {1, 1, 1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, (void *) op.ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    WB_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed: %d (K=%d rows=%lld stride=%lld batch=%lld)\n", (int) r, K,
             (long long) op.rows, (long long) op.row_stride, (long long) nb);
    return true;
}

// plain 2-D K-major f16 matrix [rows][inner] (row stride in elements), 128-byte swizzle, box {box_inner (<= 64), box_rows (<= 256)}
bool tmap_encode_2d(CUtensorMap * out, const void * ptr, uint64_t inner, uint64_t rows, uint64_t row_stride, uint32_t box_inner,
                    uint32_t box_rows) {
    PFN_encodeTiled enc = get_encode_fn();
    WB_CHECK(enc != nullptr, "cuTensorMapEncodeTiled entry point unavailable\n");
    WB_CHECK(((uintptr_t) ptr & 15) == 0 && (row_stride * 2) % 16 == 0, "tensor map operand must be 16-byte aligned\n");
    WB_CHECK(box_inner * 2 <= 128 && box_rows <= 256 && box_rows >= 1, "bad TMA box %u x %u\n", box_inner, box_rows);
    cuuint64_t dims[2]    = {(cuuint64_t) inner, (cuuint64_t) rows};
    cuuint64_t strides[1] = {(cuuint64_t) row_stride * 2};
    cuuint32_t box[2]     = {box_inner, box_rows};
    cuuint32_t estr[2]    = {1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (void *) ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    WB_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(2d) failed: %d (inner=%llu rows=%llu stride=%llu box=%ux%u)\n", (int) r,
             (unsigned long long) inner, (unsigned long long) rows, (unsigned long long) row_stride, box_inner, box_rows);
    return true;
}

static std::atomic<uint64_t> g_gemm_launches{0};
void gemm_launch_count_add(uint64_t n) { g_gemm_launches.fetch_add(n); }
uint64_t gemm_launch_count() { return g_gemm_launches.load(); }

template <int BN>
static bool launch_gemm(const CUtensorMap & ta, const CUtensorMap & tb, int M, int N, int K, int batch, int a_bcast, int b_bcast,
                        const GemmTaps & taps, const GemmEpilogue & epi, cudaStream_t stream) {
    static bool attr_set = false;
    constexpr int smem = gemm_smem_bytes(BN);
    if (!attr_set) {
        WB_CUDA_OK(cudaFuncSetAttribute(gemm_f16_tcgen05_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_set = true;
    }
    dim3 grid(ceil_div(N, BN), ceil_div(M, BLOCK_M), batch);
    ProfScope prof_scope(g_prof_gemm_cat, stream,
                         (double) batch * ((double) M * K * taps.taps * 2 + (double) N * K * taps.taps * 2 * (b_bcast ? 1.0 / batch : 1.0) +
                                          (double) M * N * (epi.out_f16 ? 2 : 4)),
                         2.0 * batch * M * N * (double) K * taps.taps);
    gemm_f16_tcgen05_kernel<BN><<<grid, GEMM_THREADS, smem, stream>>>(ta, tb, M, N, K, a_bcast, b_bcast, taps, epi);
    WB_CUDA_OK(cudaGetLastError());
    g_gemm_launches.fetch_add(1);
    return true;
}

bool gemm_f16_tc(const GemmOperand & A, const GemmOperand & B, int M, int N, int K, int batch, const GemmEpilogue & epi_in, int block_n,
                 cudaStream_t stream, const GemmTaps * taps_in) {
    GemmEpilogue epi = epi_in;
    if (epi.batch_div < 1) epi.batch_div = 1;
    GemmTaps taps;
    if (taps_in) taps = *taps_in;
    else taps = GemmTaps{1, {0, 0, 0}, {0, 0, 0}, 0};
    const int a_bcast = (A.batch <= 1 && batch > 1) ? 1 : 0;
    const int b_bcast = (B.batch <= 1 && batch > 1) ? 1 : 0;
    const int Kb = taps.taps * (taps.taps > 1 ? taps.k_tap : K);  // extent of B's k dimension
    WB_CHECK(M > 0 && N > 0 && K > 0 && batch > 0, "bad GEMM shape %d %d %d %d\n", M, N, K, batch);
    if (block_n == 0) {
        // enough CTAs to cover the 148 SMs for weight streaming at small M; wide tiles for big M
        const int64_t mt = ceil_div(M, BLOCK_M) * (int64_t) batch;
        if (mt * ceil_div(N, 256) >= 148) block_n = 256;
        else if (mt * ceil_div(N, 128) >= 148) block_n = 128;
        else if (mt * ceil_div(N, 64) >= 148) block_n = 64;
        else block_n = 32;
    }
    CUtensorMap ta, tb;
    if (!make_tmap(&ta, A, K, BLOCK_M)) return false;
    if (!make_tmap(&tb, B, Kb, block_n)) return false;
    switch (block_n) {
        case 32: return launch_gemm<32>(ta, tb, M, N, K, batch, a_bcast, b_bcast, taps, epi, stream);
        case 64: return launch_gemm<64>(ta, tb, M, N, K, batch, a_bcast, b_bcast, taps, epi, stream);
        case 128: return launch_gemm<128>(ta, tb, M, N, K, batch, a_bcast, b_bcast, taps, epi, stream);
        case 256: return launch_gemm<256>(ta, tb, M, N, K, batch, a_bcast, b_bcast, taps, epi, stream);
        default: WB_CHECK(false, "unsupported block_n %d\n", block_n);
    }
    return false;
}

}  // namespace wb
