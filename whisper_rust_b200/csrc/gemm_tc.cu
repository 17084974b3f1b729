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

#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
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
// Persistent variant for the large encoder GEMMs (one CTA per SM, static round-robin tile scheduler).
//   * TMEM holds TWO accumulators (2 x BLOCK_N columns): the MMA warp starts tile i+1 while the epilogue
//     warps drain tile i  (tmem_full / tmem_empty mbarriers).
//   * The epilogue never touches global memory through the LSU: each warp transposes its 32-row slab
//     chunk by chunk through a 128-byte-swizzled shared-memory buffer and hands it to TMA
//     (cp.async.bulk.tensor store).  The residual GEMMs (x += W.y + b: wc:2216-2220, 2251-2254) use
//     cp.reduce.async.bulk.tensor .add: the f32 addition is done at the L2, x is never loaded by the SM.
//     fl(x + (acc + b)) == fl((acc + b) + x): same rounding as the reference's ggml_add.
// ---------------------------------------------------------------------------------------------
struct PersistEpi {
    const float * bias;
    int64_t       bias_batch_stride;
    float         scale;
    int           scale_n0, scale_n1;
    int           head_mode;   // 0: coords (n, m, 0, b)   1: per-head panels, coords (0, m, n / 64, b)
    int           accumulate;  // f32 only: out += value
};

__host__ __device__ constexpr int pgemm_stages(int block_n) { return block_n >= 256 ? 4 : 6; }
static constexpr int kEpiBufBytes = 32 * 128;               // 32 rows x 128 B (32 f32 or 64 f16 columns)
static constexpr int kEpiBytes    = 4 * 2 * kEpiBufBytes;   // 4 warps x 2 buffers
__host__ __device__ constexpr int pgemm_smem_bytes(int block_n) {
    return pgemm_stages(block_n) * gemm_stage_bytes(block_n) + kEpiBytes + 1024 + 256;
}

template <int BLOCK_N, bool kF16>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_f16_persist_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                        const __grid_constant__ CUtensorMap tmap_c, int M, int N, int K, int batch, int a_bcast, int b_bcast,
                        PersistEpi epi) {
    constexpr int kStages     = pgemm_stages(BLOCK_N);
    constexpr int kStageBytes = gemm_stage_bytes(BLOCK_N);
    constexpr int kABytes     = BLOCK_M * BLOCK_K * 2;
    constexpr int kChunkCols  = kF16 ? 64 : 32;  // columns per 128-byte staging row

    extern __shared__ uint8_t smem_raw[];
    uint8_t * smem = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);
    uint8_t * epi_smem = smem + kStages * kStageBytes;

    uint64_t * full_bar   = (uint64_t *) (epi_smem + kEpiBytes);
    uint64_t * empty_bar  = full_bar + kStages;
    uint64_t * tmem_full  = empty_bar + kStages;
    uint64_t * tmem_empty = tmem_full + 2;
    uint32_t * tmem_slot  = (uint32_t *) (tmem_empty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    const int m_tiles = (M + BLOCK_M - 1) / BLOCK_M;
    const int n_tiles = (N + BLOCK_N - 1) / BLOCK_N;
    const int total   = m_tiles * n_tiles * batch;
    const int num_kb  = (K + BLOCK_K - 1) / BLOCK_K;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
        tma_prefetch_desc(&tmap_c);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(full_bar + s, 1);
            mbar_init(empty_bar + s, 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(tmem_full + a, 1);
            mbar_init(tmem_empty + a, 4);
        }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc<2 * BLOCK_N>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < total; t += gridDim.x) {
                const int n0 = (t % n_tiles) * BLOCK_N;
                const int m0 = ((t / n_tiles) % m_tiles) * BLOCK_M;
                const int b  = t / (n_tiles * m_tiles);
                const int ba = a_bcast ? 0 : b, bb = b_bcast ? 0 : b;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(empty_bar + stage, phase ^ 1);
                    uint8_t * sa = smem + stage * kStageBytes;
                    mbar_expect_tx(full_bar + stage, kStageBytes);
                    tma_load_4d(sa, &tmap_a, full_bar + stage, kb * BLOCK_K, 0, m0, ba);
                    tma_load_4d(sa + kABytes, &tmap_b, full_bar + stage, kb * BLOCK_K, 0, n0, bb);
                    if (++stage == kStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            constexpr uint32_t idesc = umma_idesc_f16(BLOCK_M, BLOCK_N);
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (int t = blockIdx.x; t < total; t += gridDim.x) {
                mbar_wait(tmem_empty + acc, acc_phase ^ 1);  // epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t) (acc * BLOCK_N);
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(full_bar + stage, phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * kStageBytes);
                    const uint64_t da = umma_desc_k_sw128(sa);
                    const uint64_t db = umma_desc_k_sw128(sa + kABytes);
#pragma unroll
                    for (int k = 0; k < BLOCK_K / 16; ++k) {
                        umma_f16_ss(d_tmem, da + (uint64_t) (2 * k), db + (uint64_t) (2 * k), idesc, (kb | k) != 0);
                    }
                    umma_commit(empty_bar + stage);
                    if (++stage == kStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                umma_commit(tmem_full + acc);
                if ((acc ^= 1) == 0) acc_phase ^= 1;
            }
        }
    } else {
        // ---------------- epilogue warps 2..5: TMEM lane quadrant = warp % 4 ----------------
        const int grp = warp & 3;
        uint8_t * my_buf  = epi_smem + grp * 2 * kEpiBufBytes;
        const uint32_t row_off = (uint32_t) lane * 128u;
        const uint32_t sw = (uint32_t) (lane & 7);
        int acc = 0, buf = 0;
        uint32_t acc_phase = 0;
        for (int t = blockIdx.x; t < total; t += gridDim.x) {
            const int n0 = (t % n_tiles) * BLOCK_N;
            const int m0 = ((t / n_tiles) % m_tiles) * BLOCK_M;
            const int b  = t / (n_tiles * m_tiles);
            const int mw = m0 + grp * 32;  // first row of this warp's slab

            const float * bias = epi.bias ? epi.bias + (int64_t) b * epi.bias_batch_stride : nullptr;

            mbar_wait(tmem_full + acc, acc_phase);
            tc_fence_after();
            if (mw < M) {
                const uint32_t t_acc = tmem_base + ((uint32_t) (grp * 32) << 16) + (uint32_t) (acc * BLOCK_N);
#pragma unroll 1
                for (int c = 0; c < BLOCK_N / kChunkCols; ++c) {
                    const int nc = n0 + c * kChunkCols;
                    if (nc >= N) break;
                    uint32_t r[kChunkCols];
                    {
                        uint32_t (&r0)[32] = *reinterpret_cast<uint32_t (*)[32]>(&r[0]);
                        tmem_ld_32x32(t_acc + (uint32_t) (c * kChunkCols), r0);
                        if constexpr (kF16) {
                            uint32_t (&r1)[32] = *reinterpret_cast<uint32_t (*)[32]>(&r[32]);
                            tmem_ld_32x32(t_acc + (uint32_t) (c * kChunkCols + 32), r1);
                        }
                    }
                    // bias of this chunk's columns: the same address in every lane (one broadcast transaction per load, L1-resident)
                    float4 bsrc[kChunkCols / 4];
                    if (bias != nullptr && nc + kChunkCols <= N) {
#pragma unroll
                        for (int q = 0; q < kChunkCols / 4; ++q) bsrc[q] = __ldg((const float4 *) (bias + nc) + q);
                    } else {
#pragma unroll
                        for (int q = 0; q < kChunkCols / 4; ++q) {
                            bsrc[q].x = (bias != nullptr && nc + q * 4 + 0 < N) ? __ldg(bias + nc + q * 4 + 0) : 0.0f;
                            bsrc[q].y = (bias != nullptr && nc + q * 4 + 1 < N) ? __ldg(bias + nc + q * 4 + 1) : 0.0f;
                            bsrc[q].z = (bias != nullptr && nc + q * 4 + 2 < N) ? __ldg(bias + nc + q * 4 + 2) : 0.0f;
                            bsrc[q].w = (bias != nullptr && nc + q * 4 + 3 < N) ? __ldg(bias + nc + q * 4 + 3) : 0.0f;
                        }
                    }
                    // column-range scale: the range is aligned to the chunk width (checked on the host)
                    const float sc = (nc >= epi.scale_n0 && nc < epi.scale_n1) ? epi.scale : 1.0f;
                    // the TMA store issued two chunks ago from this buffer must have finished reading it
                    if (lane == 0) bulk_wait_group_read<1>();
                    tmem_ld_wait();
                    __syncwarp();
                    uint8_t * dst = my_buf + buf * kEpiBufBytes + row_off;
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        uint4 pk;
                        if constexpr (kF16) {
                            const float4 b0 = bsrc[q * 2], b1 = bsrc[q * 2 + 1];  // same address in every lane: broadcast
                            const float v0 = __fmul_rn(__fadd_rn(__uint_as_float(r[q * 8 + 0]), b0.x), sc);
                            const float v1 = __fmul_rn(__fadd_rn(__uint_as_float(r[q * 8 + 1]), b0.y), sc);
                            const float v2 = __fmul_rn(__fadd_rn(__uint_as_float(r[q * 8 + 2]), b0.z), sc);
                            const float v3 = __fmul_rn(__fadd_rn(__uint_as_float(r[q * 8 + 3]), b0.w), sc);
                            const float v4 = __fmul_rn(__fadd_rn(__uint_as_float(r[q * 8 + 4]), b1.x), sc);
                            const float v5 = __fmul_rn(__fadd_rn(__uint_as_float(r[q * 8 + 5]), b1.y), sc);
                            const float v6 = __fmul_rn(__fadd_rn(__uint_as_float(r[q * 8 + 6]), b1.z), sc);
                            const float v7 = __fmul_rn(__fadd_rn(__uint_as_float(r[q * 8 + 7]), b1.w), sc);
                            __half2 h0 = __floats2half2_rn(v0, v1), h1 = __floats2half2_rn(v2, v3);
                            __half2 h2 = __floats2half2_rn(v4, v5), h3 = __floats2half2_rn(v6, v7);
                            pk.x = *(uint32_t *) &h0; pk.y = *(uint32_t *) &h1; pk.z = *(uint32_t *) &h2; pk.w = *(uint32_t *) &h3;
                        } else {
                            const float4 b0 = bsrc[q];
                            pk.x = __float_as_uint(__fmul_rn(__fadd_rn(__uint_as_float(r[q * 4 + 0]), b0.x), sc));
                            pk.y = __float_as_uint(__fmul_rn(__fadd_rn(__uint_as_float(r[q * 4 + 1]), b0.y), sc));
                            pk.z = __float_as_uint(__fmul_rn(__fadd_rn(__uint_as_float(r[q * 4 + 2]), b0.z), sc));
                            pk.w = __float_as_uint(__fmul_rn(__fadd_rn(__uint_as_float(r[q * 4 + 3]), b0.w), sc));
                        }
                        *(uint4 *) (dst + (((uint32_t) q ^ sw) << 4)) = pk;
                    }
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        const void * src = my_buf + buf * kEpiBufBytes;
                        const int c0 = epi.head_mode ? 0 : nc, c2 = epi.head_mode ? (nc >> 6) : 0;
                        if (epi.accumulate) tma_reduce_add_4d(&tmap_c, src, c0, mw, c2, b);
                        else tma_store_4d(&tmap_c, src, c0, mw, c2, b);
                        bulk_commit_group();
                    }
                    buf ^= 1;
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tmem_empty + acc);
            if ((acc ^= 1) == 0) acc_phase ^= 1;
        }
        if (lane == 0) bulk_wait_group<0>();
        tc_fence_before();
    }

    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<2 * BLOCK_N>(tmem_base);
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
    static PerDeviceOnce attr_set;
    constexpr int smem = gemm_smem_bytes(BN);
    if (attr_set.first()) {
        WB_CUDA_OK(cudaFuncSetAttribute(gemm_f16_tcgen05_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
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


// ---- persistent path: output tensor map + launch -------------------------------------------------
static int sm_count() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    }
    return n;
}

static int gemm_mode_env() {  // WB200_GEMM=legacy forces the one-tile-per-CTA kernel (A/B measurements)
    static int mode = -1;
    if (mode < 0) {
        const char * e = getenv("WB200_GEMM");
        mode = (e && strcmp(e, "legacy") == 0) ? 1 : 0;
    }
    return mode;
}

// 4-D view of the output for TMA stores: row mode {N, M, 1, batch}; head mode {64, M, N/64, batch}; box = 32 rows x 128 bytes
static bool make_tmap_out(CUtensorMap * out, const GemmEpilogue & e, int M, int N, int batch) {
    PFN_encodeTiled enc = get_encode_fn();
    WB_CHECK(enc != nullptr, "cuTensorMapEncodeTiled entry point unavailable\n");
    const int es = e.out_f16 ? 2 : 4;
    const bool head = e.out_mode == OUT_HEAD;
    const int64_t bstride = (batch > 1 || e.out_batch_stride > 0) && e.out_batch_stride > 0 ? e.out_batch_stride
                                                                                             : (head ? (int64_t) (N / 64) * e.head_stride : (int64_t) M * e.ldo);
    cuuint64_t dims[4];
    cuuint64_t strides[3];
    if (head) {
        dims[0] = 64; dims[1] = (cuuint64_t) M; dims[2] = (cuuint64_t) (N / 64); dims[3] = (cuuint64_t) batch;
        strides[0] = 64 * es; strides[1] = (cuuint64_t) e.head_stride * es; strides[2] = (cuuint64_t) bstride * es;
    } else {
        dims[0] = (cuuint64_t) N; dims[1] = (cuuint64_t) M; dims[2] = 1; dims[3] = (cuuint64_t) batch;
        strides[0] = (cuuint64_t) e.ldo * es; strides[1] = (cuuint64_t) e.ldo * es * (cuuint64_t) M; strides[2] = (cuuint64_t) bstride * es;
    }
    cuuint32_t box[4]  = {(cuuint32_t) (128 / es), 32, 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(out, e.out_f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, e.out, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    WB_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(out) failed: %d (M=%d N=%d batch=%d ldo=%lld head=%d)\n", (int) r, M, N, batch,
             (long long) e.ldo, (int) head);
    return true;
}

// can this GEMM run on the persistent kernel?
static bool persist_eligible(const GemmEpilogue & e, const GemmTaps & taps, int M, int N, int batch) {
    if (gemm_mode_env() == 1) return false;
    if (taps.taps != 1 || e.gelu_table != nullptr || e.batch_div != 1 || e.rows_per_panel != 0) return false;
    if (e.out_mode != OUT_ROW && e.out_mode != OUT_HEAD) return false;
    const int es = e.out_f16 ? 2 : 4;
    if (((uintptr_t) e.out & 15) != 0) return false;
    if (e.out_mode == OUT_ROW && (((e.ldo * es) & 15) != 0 || e.ldo < N)) return false;
    if (e.out_mode == OUT_HEAD && (!e.out_f16 || (N & 63) != 0 || ((e.head_stride * es) & 15) != 0)) return false;
    if (batch > 1 && ((e.out_batch_stride * es) & 15) != 0) return false;
    if (e.bias != nullptr && ((((uintptr_t) e.bias) & 15) != 0 || (batch > 1 && (e.bias_batch_stride & 3) != 0))) return false;
    if (e.scale_n1 > e.scale_n0 && ((e.scale_n0 | e.scale_n1) & 63) != 0) return false;  // the epilogue scales whole 64-column chunks
    if (e.res != nullptr) {  // only the in-place residual  out += ...
        if (e.res != (const float *) e.out || e.out_f16 || e.out_mode != OUT_ROW || e.ldr != e.ldo || e.res_row_mod != 0) return false;
        if (batch > 1 && e.res_batch_stride != e.out_batch_stride) return false;
    }
    // worth it only when there are at least as many 128 x 128 tiles as SMs
    return (int64_t) ceil_div(M, BLOCK_M) * ceil_div(N, 128) * batch >= sm_count();
}

template <int BN, bool F16>
static bool launch_gemm_persist(const CUtensorMap & ta, const CUtensorMap & tb, const CUtensorMap & tc, int M, int N, int K, int batch,
                                int a_bcast, int b_bcast, const GemmEpilogue & e, cudaStream_t stream) {
    static PerDeviceOnce attr_set;
    constexpr int smem = pgemm_smem_bytes(BN);
    if (attr_set.first()) {
        WB_CUDA_OK(cudaFuncSetAttribute(gemm_f16_persist_kernel<BN, F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    }
    PersistEpi pe{};
    pe.bias = e.bias; pe.bias_batch_stride = e.bias_batch_stride;
    pe.scale = e.scale; pe.scale_n0 = e.scale_n0; pe.scale_n1 = e.scale_n1;
    pe.head_mode = e.out_mode == OUT_HEAD ? 1 : 0;
    pe.accumulate = e.res != nullptr ? 1 : 0;
    const int64_t tiles = (int64_t) ceil_div(M, BLOCK_M) * ceil_div(N, BN) * batch;
    const int grid = (int) std::min<int64_t>(tiles, sm_count());
    ProfScope prof_scope(g_prof_gemm_cat, stream,
                         (double) batch * ((double) M * K * 2 * (a_bcast ? 1.0 / batch : 1.0) + (double) N * K * 2 * (b_bcast ? 1.0 / batch : 1.0) +
                                          (double) M * N * (e.out_f16 ? 2 : 4)),
                         2.0 * batch * M * N * (double) K);
    gemm_f16_persist_kernel<BN, F16><<<grid, GEMM_THREADS, smem, stream>>>(ta, tb, tc, M, N, K, batch, a_bcast, b_bcast, pe);
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
    if (block_n == 0 && persist_eligible(epi, taps, M, N, batch)) {
        const int bn = ((int64_t) ceil_div(M, BLOCK_M) * ceil_div(N, 256) * batch >= 2 * sm_count() || N % 256 == 0) ? 256 : 128;
        CUtensorMap ta, tb, tc;
        if (!make_tmap(&ta, A, K, BLOCK_M)) return false;
        if (!make_tmap(&tb, B, K, bn)) return false;
        if (!make_tmap_out(&tc, epi, M, N, batch)) return false;
        if (bn == 256) {
            return epi.out_f16 ? launch_gemm_persist<256, true>(ta, tb, tc, M, N, K, batch, a_bcast, b_bcast, epi, stream)
                               : launch_gemm_persist<256, false>(ta, tb, tc, M, N, K, batch, a_bcast, b_bcast, epi, stream);
        }
        return epi.out_f16 ? launch_gemm_persist<128, true>(ta, tb, tc, M, N, K, batch, a_bcast, b_bcast, epi, stream)
                           : launch_gemm_persist<128, false>(ta, tb, tc, M, N, K, batch, a_bcast, b_bcast, epi, stream);
    }
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
