// Fused encoder self-attention for sm_100a:  out = softmax(Q K^T * scale) V  per (window, head), T ~ 1500, D = 64.
//
// Replaces the reference's materialised  KQ = mul_mat(K, Q); soft_max_ext(KQ, scale); mul_mat(V, KQ_soft_max)
// (sys/whisper.cpp/src/whisper.cpp:2183-2206), which on a GPU costs three passes over a [H][T][T] score tensor.
// Here the scores never leave the SM: S lives in TMEM, P in shared memory.
//
// One CTA = one (window, head, 256 query rows).  320 threads:
//   warp 0     TMA producer: Q (2 x 128 rows) once, then K tiles [128 kv x 64] and V^T tiles [64 x 128 kv] through mbarrier rings
//   warp 1     tcgen05 issuer: S_g = Q_g K_j^T (128x128x64) into TMEM, O_g += P_g V_j (128x64x128) from shared-memory P
//   warps 2-9  two softmax groups (g = 0, 1), one thread per query row: tcgen05.ld of the S row, exp2, f16 pack into the
//              128-byte-swizzled P tile the next MMA consumes
// Two passes over the K tiles keep the reference's structure (max over the whole row first, wc softmax = ggml ops.cpp:4731-4827)
// without a rescaling step: pass A computes the row maxima only (no exponentials), pass B computes p = exp((s - max) * scale),
// accumulates sum(p) in f32 and O += f16(p) V on the tensor cores; the epilogue writes f16(O / sum).
// Rounding points as in the reference's CPU path: Q, K, V f16; S f32; P f16 before P.V; f32 accumulate.  P is rounded before
// the division by sum(p) instead of after it.
#include "kernels.cuh"
#include "gemm_tc.cuh"
#include "prof.h"

#include <atomic>

namespace wb {

static constexpr int AT_THREADS   = 320;
static constexpr int AT_KV        = 128;             // kv rows per tile
static constexpr int AT_NK        = 3;               // K ring depth
static constexpr int AT_NV        = 3;               // V ring depth
static constexpr int AT_TILE      = 128 * 64 * 2;    // 16 KB: a [128 x 64] f16 tile, 128-B rows
static constexpr int AT_Q_BYTES   = 2 * AT_TILE;     // two query groups
static constexpr int AT_P_BYTES   = 2 * 2 * AT_TILE; // per group: two 64-kv halves of [128 q x 64 kv]
static constexpr int AT_SMEM      = AT_Q_BYTES + AT_P_BYTES + (AT_NK + AT_NV) * AT_TILE + 1024 + 256;
static constexpr uint32_t AT_TMEM_COLS = 512;        // S0 [0,128) S1 [128,256) O0 [256,320) O1 [320,384)

__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// kVRow = true : V in the same per-head [t][64] panels as Q and K (panel 2H + h): the P.V MMA reads it as an MN-major B operand
// kVRow = false: V^T panels [H][W][64][Tpad] (K-major B operand)
// kOnePass = true : one pass over the K tiles with a running reference maximum per row.  p = exp2((s - m_run) * scale) where m_run
//                   only moves when a tile's maximum exceeds it by more than 2^8 (then O and sum(p) of that row are rescaled in
//                   TMEM by exp2(m_old - m_new)); softmax is shift-invariant, so O / sum(p) is the same quantity as with the exact
//                   row maximum, P keeps f16's relative precision (p <= 256), and the second Q.K^T pass disappears.
// kOnePass = false: the two-pass form described above (WB200_ATTN_TWOPASS=1).
template <bool kVRow, bool kOnePass>
__global__ void __launch_bounds__(AT_THREADS, 1)
enc_attn_kernel(const __grid_constant__ CUtensorMap tmap_qk, const __grid_constant__ CUtensorMap tmap_vt, __half * __restrict__ out,
                int T, int R, int H, int W, int S, float scale_log2e) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t * smem = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);
    uint8_t * q_smem = smem;
    uint8_t * p_smem = q_smem + AT_Q_BYTES;
    uint8_t * k_smem = p_smem + AT_P_BYTES;
    uint8_t * v_smem = k_smem + AT_NK * AT_TILE;
    uint64_t * bars    = (uint64_t *) (v_smem + AT_NV * AT_TILE);
    uint64_t * q_full  = bars;                 // 1
    uint64_t * k_full  = q_full + 1;           // AT_NK
    uint64_t * k_empty = k_full + AT_NK;       // AT_NK
    uint64_t * v_full  = k_empty + AT_NK;      // AT_NV
    uint64_t * v_empty = v_full + AT_NV;       // AT_NV
    uint64_t * s_full  = v_empty + AT_NV;      // 2
    uint64_t * s_empty = s_full + 2;           // 2
    uint64_t * p_full  = s_empty + 2;          // 2
    uint64_t * p_empty = p_full + 2;           // 2
    uint64_t * o_full  = p_empty + 2;          // 2
    uint32_t * tmem_slot = (uint32_t *) (o_full + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * 256;  // first query row (within the window) of this CTA
    const int h  = blockIdx.y;
    const int w  = blockIdx.z;
    const int nkv = (T + AT_KV - 1) / AT_KV;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_qk);
        tma_prefetch_desc(&tmap_vt);
        mbar_init(q_full, 1);
        for (int i = 0; i < AT_NK; ++i) { mbar_init(k_full + i, 1); mbar_init(k_empty + i, 1); }
        for (int i = 0; i < AT_NV; ++i) { mbar_init(v_full + i, 1); mbar_init(v_empty + i, 1); }
        for (int g = 0; g < 2; ++g) {
            mbar_init(s_full + g, 1);
            mbar_init(s_empty + g, 4);
            mbar_init(p_full + g, 4);
            mbar_init(p_empty + g, 1);
            mbar_init(o_full + g, 1);
        }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc<AT_TMEM_COLS>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ------------------------------------------- TMA producer -------------------------------------------
        if (elect_one()) {
            const int q_row = h * R + w * T + q0;
            const int k_row = (H + h) * R + w * T;
            const int v_row = (h * W + w) * 64;
            mbar_expect_tx(q_full, AT_Q_BYTES);
            tma_load_2d(q_smem, &tmap_qk, q_full, 0, q_row);
            tma_load_2d(q_smem + AT_TILE, &tmap_qk, q_full, 0, q_row + 128);
            int kn = 0;  // K tiles issued so far (pass A then pass B)
            auto load_k = [&](int j) {
                const int st = kn % AT_NK;
                mbar_wait(k_empty + st, ((kn / AT_NK) & 1) ^ 1);
                mbar_expect_tx(k_full + st, AT_TILE);
                tma_load_2d(k_smem + st * AT_TILE, &tmap_qk, k_full + st, 0, k_row + j * AT_KV);
                ++kn;
            };
            if constexpr (!kOnePass)
                for (int j = 0; j < nkv; ++j) load_k(j);
            for (int j = 0; j < nkv; ++j) {
                load_k(j);
                const int st = j % AT_NV;
                mbar_wait(v_empty + st, ((j / AT_NV) & 1) ^ 1);
                mbar_expect_tx(v_full + st, AT_TILE);
                if constexpr (kVRow) {
                    tma_load_2d(v_smem + st * AT_TILE, &tmap_qk, v_full + st, 0, (2 * H + h) * R + w * T + j * AT_KV);  // [128 kv][64 d]
                } else {
                    tma_load_2d(v_smem + st * AT_TILE, &tmap_vt, v_full + st, j * AT_KV, v_row);                   // kv [0,64) of the tile
                    tma_load_2d(v_smem + st * AT_TILE + AT_TILE / 2, &tmap_vt, v_full + st, j * AT_KV + 64, v_row);  // kv [64,128)
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------- MMA issuer ---------------------------------------------
        if (elect_one()) {
            constexpr uint32_t idesc_s = umma_idesc_f16(128, 128);
            // bit 16 of the instruction descriptor: B operand is MN-major (rows of V are kv = the reduction index)
            constexpr uint32_t idesc_o = umma_idesc_f16(128, 64) | (kVRow ? (1u << 16) : 0u);
            const uint32_t q_addr = smem_u32(q_smem), p_addr = smem_u32(p_smem), k_addr = smem_u32(k_smem), v_addr = smem_u32(v_smem);
            mbar_wait(q_full, 0);
            int sn = 0;  // S tiles issued so far (pass A: 0..nkv-1, pass B: nkv..2nkv-1)
            auto issue_s = [&]() {
                const int st = sn % AT_NK;
                mbar_wait(k_full + st, (sn / AT_NK) & 1);
                const uint64_t dk = umma_desc_k_sw128(k_addr + st * AT_TILE);
                for (int g = 0; g < 2; ++g) {
                    mbar_wait(s_empty + g, (sn & 1) ^ 1);
                    tc_fence_after();
                    const uint64_t dq = umma_desc_k_sw128(q_addr + g * AT_TILE);
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_f16_ss(tmem_base + g * 128, dq + (uint64_t) (2 * k), dk + (uint64_t) (2 * k), idesc_s, k != 0);
                    umma_commit(s_full + g);
                }
                umma_commit(k_empty + st);
                ++sn;
            };
            if constexpr (!kOnePass)
                for (int j = 0; j < nkv; ++j) issue_s();  // pass A
            issue_s();                                    // pass B, tile 0
            for (int j = 0; j < nkv; ++j) {
                if (j + 1 < nkv) issue_s();
                const int st = j % AT_NV;
                mbar_wait(v_full + st, (j / AT_NV) & 1);
                for (int g = 0; g < 2; ++g) {
                    mbar_wait(p_full + g, j & 1);
                    tc_fence_after();
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                        const uint64_t dp = umma_desc_k_sw128(p_addr + (g * 2 + hh) * AT_TILE);
                        // K-major V^T: 32 bytes per 16-kv step inside the 128-byte row; MN-major V: 16 rows x 128 bytes per step
                        const uint64_t dv = umma_desc_k_sw128(v_addr + st * AT_TILE + hh * (AT_TILE / 2));
                        constexpr uint64_t vstep = kVRow ? (16 * 128) >> 4 : 2;
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_f16_ss(tmem_base + 256 + g * 64, dp + (uint64_t) (2 * k), dv + vstep * (uint64_t) k, idesc_o, (j | hh | k) != 0);
                    }
                    umma_commit(p_empty + g);
                }
                umma_commit(v_empty + st);
            }
            umma_commit(o_full + 0);
            umma_commit(o_full + 1);
        }
    } else {
        // ------------------------------------------- softmax groups -----------------------------------------
        const int g    = (warp - 2) >> 2;
        const int quad = warp & 3;                  // TMEM lane quadrant this warp may read
        const int row  = quad * 32 + lane;          // row inside the group's 128-row tile
        const uint32_t t_s = tmem_base + ((uint32_t) (quad * 32) << 16) + (uint32_t) (g * 128);
        const uint32_t t_o = tmem_base + ((uint32_t) (quad * 32) << 16) + (uint32_t) (256 + g * 64);
        uint32_t r[4][32];

        auto load_s = [&](int n) {
            mbar_wait(s_full + g, n & 1);
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < 4; ++c) tmem_ld_32x32(t_s + c * 32, r[c]);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(s_empty + g);
        };

        // columns beyond the window only exist in the last kv tile: they are set to -inf once there (exp2 -> 0, max unaffected), so the
        // per-element work of every tile carries no bounds predicate
        auto mask_tail = [&](int valid) {
            if (valid < AT_KV) {
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (c * 32 + i >= valid) r[c][i] = 0xff800000u;
            }
        };

        // pass A: row maximum over the valid kv columns
        float mx = -INFINITY;
        if constexpr (!kOnePass) {
            for (int j = 0; j < nkv; ++j) {
                load_s(j);
                mask_tail(T - j * AT_KV);  // columns >= valid are beyond this window
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int i = 0; i < 32; ++i) mx = fmaxf(mx, __uint_as_float(r[c][i]));
            }
        }
        float m_scaled = mx * scale_log2e;  // one-pass: the running reference maximum, set at tile 0

        // pass B: p = exp2(s * scale * log2e - m_scaled), P tile in shared memory, running sum
        float lsum = 0.0f;
        uint8_t * p_row = p_smem + (g * 2) * AT_TILE + row * 128;
        const uint32_t sw = (uint32_t) (row & 7);
        for (int j = 0; j < nkv; ++j) {
            load_s(kOnePass ? j : nkv + j);
            mask_tail(T - j * AT_KV);
            mbar_wait(p_empty + g, (j & 1) ^ 1);  // the P.V MMAs of tile j-1 have read this buffer (and O is quiescent)
            if constexpr (kOnePass) {
                float tm4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};  // four independent chains: the compare latency overlaps
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int i = 0; i < 32; ++i) tm4[i & 3] = fmaxf(tm4[i & 3], __uint_as_float(r[c][i]));
                const float tm = fmaxf(fmaxf(tm4[0], tm4[1]), fmaxf(tm4[2], tm4[3]));
                const float tms = tm * scale_log2e;
                if (j == 0) {
                    m_scaled = tms;
                } else {
                    const bool grow = tms > m_scaled + 8.0f;
                    if (__any_sync(0xffffffffu, grow)) {  // rare: the row's scale moves by more than 2^8
                        const float m_new = grow ? tms : m_scaled;
                        const float f = fast_exp2(m_scaled - m_new);  // exactly 1 for the rows that keep their reference
                        m_scaled = m_new;
                        lsum *= f;
                        tc_fence_after();
#pragma unroll 1
                        for (int c0 = 0; c0 < 64; c0 += 16) {
                            uint32_t o[16];
                            tmem_ld_32x16(t_o + (uint32_t) c0, o);
                            tmem_ld_wait();
#pragma unroll
                            for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * f);
                            tmem_st_32x16(t_o + (uint32_t) c0, o);
                        }
                        tmem_st_wait();
                        tc_fence_before();
                    }
                }
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float pv[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float p = fast_exp2(fmaf(__uint_as_float(r[c][q * 8 + i]), scale_log2e, -m_scaled));
                        lsum += p;  // one chain, in column order: the sum's rounding is part of the pinned numerics (goldens)
                        pv[i] = p;
                    }
                    uint4 pk;
                    __half2 h0 = __floats2half2_rn(pv[0], pv[1]);
                    __half2 h1 = __floats2half2_rn(pv[2], pv[3]);
                    __half2 h2 = __floats2half2_rn(pv[4], pv[5]);
                    __half2 h3 = __floats2half2_rn(pv[6], pv[7]);
                    pk.x = *(uint32_t *) &h0; pk.y = *(uint32_t *) &h1; pk.z = *(uint32_t *) &h2; pk.w = *(uint32_t *) &h3;
                    const int c16 = c * 4 + q;  // 16-byte chunk 0..15 of the 128-kv row: half = c16 / 8, chunk-in-row = c16 % 8
                    *(uint4 *) (p_row + (c16 >> 3) * AT_TILE + ((((uint32_t) c16 & 7) ^ sw) << 4)) = pk;
                }
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_full + g);
        }

        // epilogue: O / sum -> f16, merged-head layout [t][S]
        mbar_wait(o_full + g, 0);
        tc_fence_after();
        tmem_ld_32x32(t_o, r[0]);
        tmem_ld_32x32(t_o + 32, r[1]);
        tmem_ld_wait();
        tc_fence_before();
        const int qrow = q0 + g * 128 + row;
        if (qrow < T) {
            const float inv = 1.0f / lsum;
            __half * o = out + ((int64_t) w * T + qrow) * S + h * 64;
#pragma unroll
            for (int c = 0; c < 2; ++c)
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    uint4 pk;
                    __half2 h0 = __floats2half2_rn(__uint_as_float(r[c][q * 8 + 0]) * inv, __uint_as_float(r[c][q * 8 + 1]) * inv);
                    __half2 h1 = __floats2half2_rn(__uint_as_float(r[c][q * 8 + 2]) * inv, __uint_as_float(r[c][q * 8 + 3]) * inv);
                    __half2 h2 = __floats2half2_rn(__uint_as_float(r[c][q * 8 + 4]) * inv, __uint_as_float(r[c][q * 8 + 5]) * inv);
                    __half2 h3 = __floats2half2_rn(__uint_as_float(r[c][q * 8 + 6]) * inv, __uint_as_float(r[c][q * 8 + 7]) * inv);
                    pk.x = *(uint32_t *) &h0; pk.y = *(uint32_t *) &h1; pk.z = *(uint32_t *) &h2; pk.w = *(uint32_t *) &h3;
                    *(uint4 *) (o + c * 32 + q * 8) = pk;
                }
        }
    }

    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<AT_TMEM_COLS>(tmem_base);
    }
}

static std::atomic<uint64_t> g_attn_launches{0};
uint64_t enc_attn_launch_count() { return g_attn_launches.load(); }

// qk: f16 per-head panels [2H][R][64] (Q heads then K heads, R = W*T rows) followed, when vT == nullptr, by the V heads
// ([3H][R][64]); otherwise vT: f16 [H][W][64][Tpad]; out: f16 [R][S]
bool enc_attention(const __half * qk, const __half * vT, __half * out, int W, int T, int Tpad, int H, int S, float scale, cudaStream_t stream) {
    static PerDeviceOnce attr_set;
    if (attr_set.first()) {
        WB_CUDA_OK(cudaFuncSetAttribute(enc_attn_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
        WB_CUDA_OK(cudaFuncSetAttribute(enc_attn_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
        WB_CUDA_OK(cudaFuncSetAttribute(enc_attn_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
        WB_CUDA_OK(cudaFuncSetAttribute(enc_attn_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
    }
    const int R = W * T;
    const bool vrow = vT == nullptr;
    CUtensorMap tq, tv;
    if (!tmap_encode_2d(&tq, qk, 64, (uint64_t) (vrow ? 3 : 2) * H * R, 64, 64, 128)) return false;
    if (vrow) tv = tq;
    else if (!tmap_encode_2d(&tv, vT, (uint64_t) Tpad, (uint64_t) H * W * 64, (uint64_t) Tpad, 64, 64)) return false;
    dim3 grid(ceil_div(T, 256), H, W);
    const double flops = 4.0 * (double) T * T * 64 * H * W;
    ProfScope prof_scope(PC_ENC_SOFTMAX, stream, (double) R * S * 2 * 4, flops);
    static const bool two_pass = getenv("WB200_ATTN_TWOPASS") != nullptr;
    const float sl2 = scale * 1.4426950408889634f;
    if (two_pass) {
        if (vrow) enc_attn_kernel<true, false><<<grid, AT_THREADS, AT_SMEM, stream>>>(tq, tv, out, T, R, H, W, S, sl2);
        else enc_attn_kernel<false, false><<<grid, AT_THREADS, AT_SMEM, stream>>>(tq, tv, out, T, R, H, W, S, sl2);
    } else {
        if (vrow) enc_attn_kernel<true, true><<<grid, AT_THREADS, AT_SMEM, stream>>>(tq, tv, out, T, R, H, W, S, sl2);
        else enc_attn_kernel<false, true><<<grid, AT_THREADS, AT_SMEM, stream>>>(tq, tv, out, T, R, H, W, S, sl2);
    }
    WB_CUDA_OK(cudaGetLastError());
    g_attn_launches.fetch_add(1);
    return true;
}

}  // namespace wb
