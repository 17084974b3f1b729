// The greedy decode step as one CUDA graph of weight-streaming kernels.  See step.cuh.
#include "step.cuh"

#include <math.h>
#include <string.h>

#include <algorithm>
#include <array>
#include <atomic>
#include <map>
#include <mutex>
#include <vector>

#include "log.h"
#include "prof.h"

namespace wb {

// =============================================================================================
// gemm_wstream:  Y[t][f] = sum_k X[t][k] * W[f][k]      (t < n <= 256 token rows, f < N_out weight rows)
// =============================================================================================
// The WEIGHT tile is the UMMA M operand (128 weight rows per CTA), the token rows are the N operand, so the
// accumulator is D[128 features (TMEM lanes)][n_pad tokens (TMEM columns)].  grid = (ceil(N_out/128), split):
// CTA (tile, s) reduces k-blocks [s*kb_per_split, ...) and either stores its f32 partial tile to
// partial[s][t][f] (mode 0; the consumer kernel sums the splits in index order, deterministically) or, when
// split == 1, applies bias / GELU and stores f16 or f32 directly (mode 1).
// Rounding points as the reference's mul_mat (ggml-cpu.c:1269-1461): f16 operands, f32 accumulate.
// In-situ timeline (WB200_STEP_TRACE=1): block 0 of every step kernel stamps %globaltimer once its griddepcontrol.wait
// returned, i.e. when its predecessor in the lane has completed.  Differences of consecutive stamps = kernel durations
// inside the real (concurrent) run, which ncu's serialised replay cannot show.
__device__ __forceinline__ void trace_stamp(unsigned long long * slot) {
    if (slot != nullptr && blockIdx.x == 0 && blockIdx.y == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
        *slot = t;
    }
}
// second stamp of a kernel's slot pair: block 0 is about to exit (its own work is done; other blocks may still run)
__device__ __forceinline__ void trace_stamp_end(unsigned long long * slot) {
    if (slot != nullptr && blockIdx.x == 0 && blockIdx.y == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
        slot[1] = t;
    }
}

// Register budgets (minimum resident CTAs per SM in __launch_bounds__): with two persistent cross-attention CTAs resident (2 x 288 threads x 80 registers = 46 K of the SM's
// 64 K) the short kernels of the OTHER decode lanes share what is left; capping their registers decides how many of their CTAs fit.
#ifndef WB_SELF_MIN_CTAS
#define WB_SELF_MIN_CTAS 8
#endif
#ifndef WB_LN_MIN_CTAS
#define WB_LN_MIN_CTAS 5
#endif
struct WsArgs {
    unsigned long long * trace;
    int n, n_pad, N_out;
    int kb_total, kb_per_split, stages;
    int mode;  // 0: partial, 1: direct
    float * partial;
    int64_t split_stride;  // elements between splits = n * N_out
    void * out;
    int out_f16;
    int64_t ldo;
    const float * bias;
    const __half * gelu_table;
    uint32_t tmem_cols;
};

static constexpr int WS_THREADS = 192;
static constexpr int WS_A_BYTES = 128 * 64 * 2;

__device__ __forceinline__ float gelu_f16table(float x, const __half * __restrict__ table) {
    // ggml_vec_gelu_f32 with GGML_GELU_FP16 (ggml-cpu/vec.h:571-585)
    if (x <= -10.0f) return 0.0f;
    if (x >= 10.0f) return x;
    return __half2float(__ldg(table + __half_as_ushort(__float2half_rn(x))));
}

__global__ void __launch_bounds__(WS_THREADS, 2)
gemm_wstream_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x, WsArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t * smem = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);
    const int b_bytes     = a.n_pad * 128;
    const int stage_bytes = WS_A_BYTES + b_bytes;
    uint64_t * full_bar   = (uint64_t *) (smem + a.stages * stage_bytes);
    uint64_t * empty_bar  = full_bar + a.stages;
    uint64_t * tmem_full  = empty_bar + a.stages;
    uint32_t * tmem_slot  = (uint32_t *) (tmem_full + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int f0    = blockIdx.x * 128;
    const int split = blockIdx.y;
    const int kb0   = split * a.kb_per_split;
    const int nkb   = min(a.kb_per_split, a.kb_total - kb0);

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_w);
        tma_prefetch_desc(&tmap_x);
        for (int s = 0; s < a.stages; ++s) {
            mbar_init(full_bar + s, 1);
            mbar_init(empty_bar + s, 1);
        }
        mbar_init(tmem_full, 1);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc_rt(tmem_slot, a.tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // PDL: every kernel of the step waits (griddepcontrol.wait) before it touches mutable global memory, so the next
    // kernel may become resident now and overlap its prologue + weight prefetch with this kernel's tail
    pdl_launch_dependents();

    if (warp == 0) {
        if (elect_one()) {
            const uint64_t pol_w = l2_policy_evict_first(), pol_x = l2_policy_evict_last();
            // the weights do not depend on the previous kernel: start streaming them before the PDL wait
            const int pre = min(nkb, a.stages);
            for (int i = 0; i < pre; ++i) {
                mbar_expect_tx(full_bar + i, stage_bytes);
                tma_load_2d_hint(smem + i * stage_bytes, &tmap_w, full_bar + i, (kb0 + i) * 64, f0, pol_w);
            }
            pdl_wait();
            trace_stamp(a.trace);
            for (int i = 0; i < pre; ++i)
                tma_load_2d_hint(smem + i * stage_bytes + WS_A_BYTES, &tmap_x, full_bar + i, (kb0 + i) * 64, 0, pol_x);
            int stage = pre == a.stages ? 0 : pre;
            uint32_t phase = pre == a.stages ? 1 : 0;
            for (int i = pre; i < nkb; ++i) {
                mbar_wait(empty_bar + stage, phase ^ 1);
                uint8_t * sa = smem + stage * stage_bytes;
                mbar_expect_tx(full_bar + stage, stage_bytes);
                tma_load_2d_hint(sa, &tmap_w, full_bar + stage, (kb0 + i) * 64, f0, pol_w);
                tma_load_2d_hint(sa + WS_A_BYTES, &tmap_x, full_bar + stage, (kb0 + i) * 64, 0, pol_x);
                if (++stage == a.stages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            const uint32_t idesc = umma_idesc_f16(128, (uint32_t) a.n_pad);
            int stage = 0;
            uint32_t phase = 0;
            for (int i = 0; i < nkb; ++i) {
                mbar_wait(full_bar + stage, phase);
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + stage * stage_bytes);
                const uint64_t da = umma_desc_k_sw128(sa), db = umma_desc_k_sw128(sa + WS_A_BYTES);
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_f16_ss(tmem_base, da + (uint64_t) (2 * k), db + (uint64_t) (2 * k), idesc, (i | k) != 0);
                umma_commit(empty_bar + stage);
                if (++stage == a.stages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            umma_commit(tmem_full);
        }
    } else {
        // epilogue warps 2..5: TMEM lane group = warp % 4; lane = weight row (feature), column = token
        const int grp = warp & 3;
        const int f   = f0 + grp * 32 + lane;
        pdl_wait();  // the output buffers may still be read by the previous kernel
        mbar_wait(tmem_full, 0);
        tc_fence_after();
        const bool f_ok = f < a.N_out;
        float bias = 0.0f;
        if (a.mode == 1 && a.bias != nullptr && f_ok) bias = __ldg(a.bias + f);
        float * pout = a.partial + (int64_t) split * a.split_stride + f;
#pragma unroll 1
        for (int c0 = 0; c0 < a.n_pad; c0 += 16) {
            if (c0 >= a.n) break;  // warp-uniform
            uint32_t r[16];
            tmem_ld_32x16(tmem_base + ((uint32_t) (grp * 32) << 16) + (uint32_t) c0, r);
            tmem_ld_wait();
            if (f_ok) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const int t = c0 + j;
                    if (t < a.n) {
                        float v = __uint_as_float(r[j]);
                        if (a.mode == 0) {
                            pout[(int64_t) t * a.N_out] = v;
                        } else {
                            v = __fadd_rn(v, bias);
                            if (a.gelu_table != nullptr) v = gelu_f16table(v, a.gelu_table);
                            if (a.out_f16) ((__half *) a.out)[(int64_t) t * a.ldo + f] = __float2half_rn(v);
                            else ((float *) a.out)[(int64_t) t * a.ldo + f] = v;
                        }
                    }
                }
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (threadIdx.x == 0) trace_stamp_end(a.trace);
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_rt(tmem_base, a.tmem_cols);
    }
}

// =============================================================================================
// gemm_wcluster: the same weight-streaming GEMM with the split-K reduction done INSIDE a thread-block cluster
// =============================================================================================
// grid (split, tiles), cluster (split, 1, 1): the CTAs of one cluster share a 128-feature weight tile and own disjoint
// k-ranges.  Each CTA accumulates its partial D[128 features][n_pad tokens] in TMEM, spills it to its own shared memory
// (over the dead pipeline stages) as P[token][128 features] f32, and after a cluster barrier CTA r reduces the token slice
// [r * tpr, (r + 1) * tpr) by reading every peer's P through distributed shared memory in rank order 0..split-1 (the same
// fixed summation order the split-K consumers used), applies the epilogue and writes the FINAL values: no partials in
// global memory, no reduction work in the consumer kernels.
//   WC_F16      out f16 = f16((acc + bias) * scale?)     scale for features < scale_f1   (wc:2558-2583 Q/K/V, wc:2700-2707 cross Q)
//   WC_RESIDUAL x f32 = (acc + bias) + x                 in place                          (wc:2657-2664, 2771-2777, 2807-2814)
//   WC_GELU     out f16 = f16(gelu_table[f16(acc + bias)])                                  (wc:2795-2802)
enum WcMode : int { WC_F16 = 0, WC_RESIDUAL = 1, WC_GELU = 2 };
struct WcArgs {
    unsigned long long * trace;
    int n, n_pad, N_out;
    int kb_total, kb_per_split, stages, split;
    int mode;
    void * out;
    int64_t ldo;
    const float * bias;
    float scale;
    int scale_f1;
    const __half * gelu_table;
    uint32_t tmem_cols;
    int tok_per_rank;
};

__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
    return r;
}
__device__ __forceinline__ float4 ld_dsmem_f4(uint32_t local_addr, uint32_t rank) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(remote) : "r"(local_addr), "r"(rank));
    float4 v;
    asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];\n" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(remote) : "memory");
    return v;
}

__global__ void __launch_bounds__(WS_THREADS, 1)
gemm_wcluster_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x, WcArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t * smem = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);
    const int b_bytes     = a.n_pad * 128;
    const int stage_bytes = WS_A_BYTES + b_bytes;
    const int main_bytes  = max(a.stages * stage_bytes, a.n_pad * 512);  // pipeline stages, later the partial tile P
    uint64_t * full_bar   = (uint64_t *) (smem + main_bytes);
    uint64_t * empty_bar  = full_bar + a.stages;
    uint64_t * tmem_full  = empty_bar + a.stages;
    uint32_t * tmem_slot  = (uint32_t *) (tmem_full + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank  = (int) cluster_ctarank();  // == blockIdx.x: the split index
    const int f0    = blockIdx.y * 128;
    const int kb0   = rank * a.kb_per_split;
    const int nkb   = min(a.kb_per_split, a.kb_total - kb0);  // >= 1 by construction of the plan

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_w);
        tma_prefetch_desc(&tmap_x);
        for (int s = 0; s < a.stages; ++s) {
            mbar_init(full_bar + s, 1);
            mbar_init(empty_bar + s, 1);
        }
        mbar_init(tmem_full, 1);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc_rt(tmem_slot, a.tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();

    if (warp == 0) {
        if (elect_one()) {
            const uint64_t pol_w = l2_policy_evict_first(), pol_x = l2_policy_evict_last();
            const int pre = min(nkb, a.stages);
            for (int i = 0; i < pre; ++i) {  // weights do not depend on the previous kernel
                mbar_expect_tx(full_bar + i, stage_bytes);
                tma_load_2d_hint(smem + i * stage_bytes, &tmap_w, full_bar + i, (kb0 + i) * 64, f0, pol_w);
            }
            pdl_wait();
            trace_stamp(a.trace);
            for (int i = 0; i < pre; ++i)
                tma_load_2d_hint(smem + i * stage_bytes + WS_A_BYTES, &tmap_x, full_bar + i, (kb0 + i) * 64, 0, pol_x);
            int stage = pre == a.stages ? 0 : pre;
            uint32_t phase = pre == a.stages ? 1 : 0;
            for (int i = pre; i < nkb; ++i) {
                mbar_wait(empty_bar + stage, phase ^ 1);
                uint8_t * sa = smem + stage * stage_bytes;
                mbar_expect_tx(full_bar + stage, stage_bytes);
                tma_load_2d_hint(sa, &tmap_w, full_bar + stage, (kb0 + i) * 64, f0, pol_w);
                tma_load_2d_hint(sa + WS_A_BYTES, &tmap_x, full_bar + stage, (kb0 + i) * 64, 0, pol_x);
                if (++stage == a.stages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            const uint32_t idesc = umma_idesc_f16(128, (uint32_t) a.n_pad);
            int stage = 0;
            uint32_t phase = 0;
            for (int i = 0; i < nkb; ++i) {
                mbar_wait(full_bar + stage, phase);
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + stage * stage_bytes);
                const uint64_t da = umma_desc_k_sw128(sa), db = umma_desc_k_sw128(sa + WS_A_BYTES);
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_f16_ss(tmem_base, da + (uint64_t) (2 * k), db + (uint64_t) (2 * k), idesc, (i | k) != 0);
                umma_commit(empty_bar + stage);
                if (++stage == a.stages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            umma_commit(tmem_full);
        }
    } else {
        // warps 2..5: spill the accumulator (lane = feature, column = token) to P[token][feature]; all MMAs have retired, so
        // the pipeline stages underneath are dead
        const int grp = warp & 3;
        mbar_wait(tmem_full, 0);
        tc_fence_after();
        float * P = (float *) smem + grp * 32 + lane;
#pragma unroll 1
        for (int c0 = 0; c0 < a.n_pad; c0 += 16) {
            if (c0 >= a.n) break;
            uint32_t r[16];
            tmem_ld_32x16(tmem_base + ((uint32_t) (grp * 32) << 16) + (uint32_t) c0, r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) P[(c0 + j) * 128] = __uint_as_float(r[j]);
        }
        tc_fence_before();
    }
    __syncwarp();
    cluster_sync_all();  // every CTA's P is complete and visible cluster-wide

    if (warp >= 2) {
        pdl_wait();  // the output may still be read by the previous kernel (every thread that writes global memory waits)
        const int e  = (warp - 2) * 32 + lane;  // 0..127
        const int fl = (e & 31) * 4;            // 4 consecutive features of the tile
        const int f  = f0 + fl;
        const int t0 = rank * a.tok_per_rank, t1 = min(a.n, t0 + a.tok_per_rank);
        const uint32_t p_local = smem_u32(smem) + (uint32_t) fl * 4u;
        float4 bias = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        const bool f_ok = f < a.N_out;
        if (a.bias != nullptr && f_ok) bias = __ldg((const float4 *) (a.bias + f));
        const float sc = (f < a.scale_f1) ? a.scale : 1.0f;
        // two tokens per iteration, all of their split partials requested before the first add: the distributed-shared-memory
        // round trips overlap instead of queueing behind each other (the sums keep the split order: results are unchanged)
#pragma unroll 1
        for (int tb = t0 + (e >> 5); tb < t1; tb += 8) {
            float4 w[2][8];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int t = tb + 4 * u;
#pragma unroll
                for (int s = 0; s < 8; ++s)
                    if (t < t1 && s < a.split) w[u][s] = ld_dsmem_f4(p_local + (uint32_t) t * 512u, (uint32_t) s);
            }
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int t = tb + 4 * u;
                if (t >= t1) break;
                float4 v = w[u][0];
#pragma unroll
                for (int s = 1; s < 8; ++s)
                    if (s < a.split) {
                        v.x = __fadd_rn(v.x, w[u][s].x); v.y = __fadd_rn(v.y, w[u][s].y);
                        v.z = __fadd_rn(v.z, w[u][s].z); v.w = __fadd_rn(v.w, w[u][s].w);
                    }
                if (!f_ok) continue;
                v.x = __fadd_rn(v.x, bias.x); v.y = __fadd_rn(v.y, bias.y); v.z = __fadd_rn(v.z, bias.z); v.w = __fadd_rn(v.w, bias.w);
                if (a.mode == WC_RESIDUAL) {
                    float4 * xp = (float4 *) ((float *) a.out + (int64_t) t * a.ldo + f);
                    const float4 x = *xp;
                    v.x = __fadd_rn(v.x, x.x); v.y = __fadd_rn(v.y, x.y); v.z = __fadd_rn(v.z, x.z); v.w = __fadd_rn(v.w, x.w);
                    *xp = v;
                } else {
                    if (a.mode == WC_GELU) {
                        v.x = gelu_f16table(v.x, a.gelu_table); v.y = gelu_f16table(v.y, a.gelu_table);
                        v.z = gelu_f16table(v.z, a.gelu_table); v.w = gelu_f16table(v.w, a.gelu_table);
                    } else {
                        v.x = __fmul_rn(v.x, sc); v.y = __fmul_rn(v.y, sc); v.z = __fmul_rn(v.z, sc); v.w = __fmul_rn(v.w, sc);
                    }
                    const __half2 h0 = __floats2half2_rn(v.x, v.y), h1 = __floats2half2_rn(v.z, v.w);
                    uint2 pk;
                    pk.x = *(const uint32_t *) &h0;
                    pk.y = *(const uint32_t *) &h1;
                    *(uint2 *) ((__half *) a.out + (int64_t) t * a.ldo + f) = pk;
                }
            }
        }
    }
    __syncwarp();
    cluster_sync_all();  // no CTA may exit (and release its shared memory) while a peer is still reading it
    if (threadIdx.x == 0) trace_stamp_end(a.trace);
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_rt(tmem_base, a.tmem_cols);
    }
}

// ---- host side: plan (split / stages) and launch --------------------------------------------------------------
static std::atomic<uint64_t> g_step_launches{0};
// per-call launch settings: thread_local, because contexts on different GPUs may be driven from different host threads
static thread_local bool g_pdl = true;  // launch with programmatic stream serialization (WB200_NO_PDL=1 switches it off)
static int env_int(const char * name, int dflt) {
    const char * e = getenv(name);
    return e && *e ? atoi(e) : dflt;
}
static thread_local int g_launch_prio = 0;
// timeline capture (WB200_STEP_TRACE): next free stamp slot and the label of every slot handed out
static thread_local unsigned long long * g_trace_next = nullptr;
static thread_local std::vector<const char *> * g_trace_names = nullptr;
static thread_local std::vector<int> * g_trace_subs = nullptr;
static thread_local int g_trace_sub = 0;  // which half of a paired step the next launches belong to
static unsigned long long * trace_slot(const char * name) {
    if (!g_trace_next) return nullptr;
    if (g_trace_names) g_trace_names->push_back(name);
    if (g_trace_subs) g_trace_subs->push_back(g_trace_sub);
    unsigned long long * p = g_trace_next;
    g_trace_next += 2;  // (start, end of block 0)
    return p;
}  // cudaLaunchAttributePriority of the next launches (0: none; set around issue_step when lanes run)

// Tuning of the kernels that share the SMs when several decode lanes are in flight (set by stage_step_launch):
// CTAs per SM and ring depth of the persistent cross attention, shared-memory cap of the cluster GEMM.
struct LaneTuning {
    int xs_ctas_per_sm = 2, xs_stages = 5, wc_smem_cap = 192 * 1024;
};
static thread_local LaneTuning g_tune;

struct WsPlan {
    int tiles, split, kb_total, kb_per_split, stages, n_pad, smem;
    uint32_t tmem_cols;
};

static WsPlan ws_plan(int n, int N_out, int K, int force_split /*0: auto*/) {
    WsPlan p{};
    p.tiles    = ceil_div(N_out, 128);
    p.kb_total = ceil_div(K, 64);
    p.n_pad    = (int) pad_to(std::max(n, 1), 16);
    int split  = force_split;
    if (split <= 0) {
        // ~one CTA per SM, at least two k-blocks (32 KB of weights) per CTA
        split = std::max(1, 148 / p.tiles);
        split = std::min(split, std::max(1, p.kb_total / 2));
    }
    split          = std::min(split, p.kb_total);
    p.kb_per_split = ceil_div(p.kb_total, split);
    p.split        = ceil_div(p.kb_total, p.kb_per_split);
    const int stage_bytes = WS_A_BYTES + p.n_pad * 128;
    p.stages = std::max(1, std::min(std::min(8, p.kb_per_split), (100 * 1024) / stage_bytes));
    p.smem   = p.stages * stage_bytes + 1024 + 256;
    uint32_t c = 32;
    while ((int) c < p.n_pad) c <<= 1;
    p.tmem_cols = c;
    return p;
}

template <typename... KArgs, typename... Args>
static bool launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, bool pdl, Args... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[2];
    int na = 0;
    if (pdl && g_pdl) {
        at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    if (g_launch_prio != 0) {  // lanes: the bandwidth-bound cross attention yields SM slots to the other lanes' short kernels
        at[na].id = cudaLaunchAttributePriority;
        at[na].val.priority = g_launch_prio;
        ++na;
    }
    cfg.attrs = at;
    cfg.numAttrs = na;
    WB_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, args...));
    g_step_launches.fetch_add(1);
    return true;
}

struct WsOut {
    // mode 0
    float * partial = nullptr;
    // mode 1
    void * out = nullptr;
    int out_f16 = 0;
    int64_t ldo = 0;
    const float * bias = nullptr;
    const __half * gelu_table = nullptr;
};

// X: f16 [n][K] row-major (ld = K); W: f16 [N_out][K].  Returns the split count through *split_out (mode 0).
static bool gemm_wstream(const __half * W, const __half * X, int n, int N_out, int K, const WsOut & o, int force_split, int * split_out,
                         cudaStream_t s, bool pdl) {
    WB_CHECK(n >= 1 && n <= 256, "gemm_wstream handles 1..256 token rows (got %d)\n", n);
    const bool direct = o.out != nullptr;
    WsPlan p = ws_plan(n, N_out, K, direct ? 1 : force_split);
    static PerDeviceOnce attr_set;
    if (attr_set.first()) {
        WB_CUDA_OK(cudaFuncSetAttribute(gemm_wstream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024));
    }
    CUtensorMap tw, tx;
    if (!tmap_encode_2d(&tw, W, (uint64_t) K, (uint64_t) N_out, (uint64_t) K, 64, 128)) return false;
    if (!tmap_encode_2d(&tx, X, (uint64_t) K, (uint64_t) n, (uint64_t) K, 64, (uint32_t) p.n_pad)) return false;
    WsArgs a{};
    a.n = n; a.n_pad = p.n_pad; a.N_out = N_out;
    a.kb_total = p.kb_total; a.kb_per_split = p.kb_per_split; a.stages = p.stages;
    a.mode = direct ? 1 : 0;
    a.partial = o.partial; a.split_stride = (int64_t) n * N_out;
    a.out = o.out; a.out_f16 = o.out_f16; a.ldo = o.ldo; a.bias = o.bias; a.gelu_table = o.gelu_table;
    a.tmem_cols = p.tmem_cols;
    a.trace = trace_slot(N_out > 20000 ? "logits" : "gemm_ws");
    if (split_out) *split_out = p.split;
    ProfScope prof_scope(N_out > 20000 ? PC_GEMM_LOGITS : PC_GEMM_DEC, s, (double) N_out * K * 2 + (double) n * K * 2);
    return launch_k(gemm_wstream_kernel, dim3(p.tiles, p.split), dim3(WS_THREADS), (size_t) p.smem, s, pdl, tw, tx, a);
}

// ---- gemm_wcluster: plan and launch ---------------------------------------------------------------------------------
struct WcPlan {
    int tiles, split, kb_total, kb_per_split, stages, n_pad, smem, tok_per_rank;
    uint32_t tmem_cols;
};
static WcPlan wc_plan(int n, int N_out, int K) {
    WcPlan p{};
    p.tiles    = ceil_div(N_out, 128);
    p.kb_total = ceil_div(K, 64);
    p.n_pad    = (int) pad_to(std::max(n, 1), 16);
    // ~one CTA per SM, clusters of at most 8 (the portable limit), at least two k-blocks per CTA
    static const int target_ctas = std::max(1, env_int("WB200_WC_CTAS", 148));  // tuning knob: fewer, longer-running CTAs per GEMM
    int split = std::max(1, std::min(8, target_ctas / p.tiles));
    split = std::min(split, std::max(1, p.kb_total / 2));
    p.kb_per_split = ceil_div(p.kb_total, split);
    p.split        = ceil_div(p.kb_total, p.kb_per_split);
    p.tok_per_rank = ceil_div(n, p.split);
    const int stage_bytes = WS_A_BYTES + p.n_pad * 128;
    p.stages = std::max(1, std::min(std::min(4, p.kb_per_split), g_tune.wc_smem_cap / stage_bytes));
    p.smem   = std::max(p.stages * stage_bytes, p.n_pad * 512) + 1024 + 256;
    uint32_t c = 32;
    while ((int) c < p.n_pad) c <<= 1;
    p.tmem_cols = c;
    return p;
}

struct WcOut {
    int mode = WC_F16;
    void * out = nullptr;
    int64_t ldo = 0;
    const float * bias = nullptr;
    float scale = 1.0f;
    int scale_f1 = 0;
    const __half * gelu_table = nullptr;
};

static bool gemm_wcluster(const __half * W, const __half * X, int n, int N_out, int K, const WcOut & o, cudaStream_t s, bool pdl) {
    WB_CHECK(n >= 1 && n <= 256 && (N_out % 4) == 0, "gemm_wcluster handles 1..256 token rows, N_out %% 4 == 0 (got %d, %d)\n", n, N_out);
    const WcPlan p = wc_plan(n, N_out, K);
    static PerDeviceOnce attr_set;
    if (attr_set.first()) {
        WB_CUDA_OK(cudaFuncSetAttribute(gemm_wcluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    }
    WB_CHECK(p.smem <= 200 * 1024, "gemm_wcluster: %d bytes of shared memory\n", p.smem);
    CUtensorMap tw, tx;
    if (!tmap_encode_2d(&tw, W, (uint64_t) K, (uint64_t) N_out, (uint64_t) K, 64, 128)) return false;
    if (!tmap_encode_2d(&tx, X, (uint64_t) K, (uint64_t) n, (uint64_t) K, 64, (uint32_t) p.n_pad)) return false;
    WcArgs a{};
    a.n = n; a.n_pad = p.n_pad; a.N_out = N_out;
    a.kb_total = p.kb_total; a.kb_per_split = p.kb_per_split; a.stages = p.stages; a.split = p.split;
    a.mode = o.mode; a.out = o.out; a.ldo = o.ldo; a.bias = o.bias; a.scale = o.scale; a.scale_f1 = o.scale_f1;
    a.gelu_table = o.gelu_table; a.tmem_cols = p.tmem_cols; a.tok_per_rank = p.tok_per_rank;
    a.trace = trace_slot(N_out == 3 * K ? "gemm_qkv" : N_out == 4 * K ? "gemm_fc1" : K == 4 * N_out ? "gemm_fc2" : "gemm_sq");
    // one profiler category per launch shape of the step (they are separate lines of the ncu launch list too)
    ProfScope prof_scope(N_out == 3 * K ? PC_GEMM_STEP_QKV : N_out == 4 * K ? PC_GEMM_STEP_FC1 : K == 4 * N_out ? PC_GEMM_STEP_FC2 : PC_GEMM_STEP_SQ, s,
                         (double) N_out * K * 2 + (double) n * K * 2);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(p.split, p.tiles);
    cfg.blockDim = dim3(WS_THREADS);
    cfg.dynamicSmemBytes = (size_t) p.smem;
    cfg.stream = s;
    cudaLaunchAttribute at[3];
    int na = 0;
    at[na].id = cudaLaunchAttributeClusterDimension;
    at[na].val.clusterDim.x = p.split; at[na].val.clusterDim.y = 1; at[na].val.clusterDim.z = 1;
    ++na;
    if (pdl && g_pdl) {
        at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    if (g_launch_prio != 0) {
        at[na].id = cudaLaunchAttributePriority;
        at[na].val.priority = g_launch_prio;
        ++na;
    }
    cfg.attrs = at;
    cfg.numAttrs = na;
    WB_CUDA_OK(cudaLaunchKernelEx(&cfg, gemm_wcluster_kernel, tw, tx, a));
    g_step_launches.fetch_add(1);
    return true;
}

// =============================================================================================
// ln_reduce: x = [residual | embedding] (+ bias + sum of split-K partials);  xa = f16(LayerNorm(x) * gamma + beta)
// =============================================================================================
// One CTA per token row.  The split partials are summed in split order (deterministic); then (acc + bias) + residual
// as the reference's ggml_add chain (whisper.cpp:2216-2220, 2251-2254); LayerNorm as ggml_norm (ops.cpp:3199-3248:
// mean and variance accumulated in double) followed by the separate f32 mul / add with gamma / beta (wc:2121-2126).
struct LnArgs {
    unsigned long long * trace;
    const float * partial;  // [split][n][S] or null
    int split;
    int64_t split_stride;
    const float * bias;     // [S] (with partial)
    const float * res;      // [n][S] residual stream in (null with embed)
    const __half * te;      // embedding mode: x = f32(te[token]) + pe[pos]  (wc:2531-2534)
    const float * pe;
    const int * tokens;
    const int * pos;
    float * x_out;          // [n][S] residual stream out (may alias res)
    __half * xa_out;        // [n][S] LayerNorm output, f16
    const float * gamma;
    const float * beta;
    float eps;
    int S;
};

static constexpr int LN_THREADS = 256;
static constexpr int LN_MAX_PER_THREAD = 8;  // S <= 2048

__device__ __forceinline__ double block_sum_double(double v, double * red /*[8]*/) {
    v = warp_sum_d(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double r = 0.0;
#pragma unroll
    for (int w = 0; w < LN_THREADS / 32; ++w) r += red[w];
    return r;
}

__global__ void __launch_bounds__(LN_THREADS, WB_LN_MIN_CTAS)
ln_reduce_kernel(LnArgs a) {
    __shared__ double red[8];
    pdl_launch_dependents();
    const int row = blockIdx.x, tid = threadIdx.x, S = a.S;
    // parameters do not depend on the previous kernel: fetch them while it drains
    float gm[LN_MAX_PER_THREAD], bt[LN_MAX_PER_THREAD];
#pragma unroll
    for (int q = 0; q < LN_MAX_PER_THREAD; ++q) {
        const int i = tid + q * LN_THREADS;
        gm[q] = i < S ? __ldg(a.gamma + i) : 0.0f;
        bt[q] = i < S ? __ldg(a.beta + i) : 0.0f;
    }
    pdl_wait();
    if (tid == 0) trace_stamp(a.trace);
    float v[LN_MAX_PER_THREAD];
    double sum = 0.0;
    // split-K partials first: all loads of a split are independent, four splits are kept in flight
    if (a.partial != nullptr) {
        const float * p = a.partial + (int64_t) row * S + tid;
#pragma unroll
        for (int q = 0; q < LN_MAX_PER_THREAD; ++q) v[q] = (tid + q * LN_THREADS < S) ? p[q * LN_THREADS] : 0.0f;
#pragma unroll 4
        for (int s = 1; s < a.split; ++s) {
            const float * ps = p + (int64_t) s * a.split_stride;
#pragma unroll
            for (int q = 0; q < LN_MAX_PER_THREAD; ++q)
                if (tid + q * LN_THREADS < S) v[q] = __fadd_rn(v[q], ps[q * LN_THREADS]);
        }
    }
#pragma unroll
    for (int q = 0; q < LN_MAX_PER_THREAD; ++q) {
        const int i = tid + q * LN_THREADS;
        float x = 0.0f;
        if (i < S) {
            if (a.te != nullptr) x = __fadd_rn(__half2float(a.te[(int64_t) a.tokens[row] * S + i]), a.pe[(int64_t) a.pos[row] * S + i]);
            else x = a.res[(int64_t) row * S + i];
            if (a.partial != nullptr) x = __fadd_rn(__fadd_rn(v[q], __ldg(a.bias + i)), x);
            if (a.partial != nullptr || a.te != nullptr) a.x_out[(int64_t) row * S + i] = x;
            sum += (double) x;
        }
        v[q] = x;
    }
    sum = block_sum_double(sum, red);
    const float mean = (float) (sum / S);
    double sum2 = 0.0;
#pragma unroll
    for (int q = 0; q < LN_MAX_PER_THREAD; ++q) {
        const int i = tid + q * LN_THREADS;
        if (i < S) {
            v[q] = __fsub_rn(v[q], mean);
            sum2 += (double) __fmul_rn(v[q], v[q]);
        }
    }
    sum2 = block_sum_double(sum2, red);
    const float variance = (float) (sum2 / S);
    const float scale = __frcp_rn(__fsqrt_rn(__fadd_rn(variance, a.eps)));
#pragma unroll
    for (int q = 0; q < LN_MAX_PER_THREAD; ++q) {
        const int i = tid + q * LN_THREADS;
        if (i < S) {
            const float y = __fadd_rn(__fmul_rn(__fmul_rn(v[q], scale), gm[q]), bt[q]);
            a.xa_out[(int64_t) row * S + i] = __float2half_rn(y);
        }
    }
    if (tid == 0) trace_stamp_end(a.trace);
}

// =============================================================================================
// self_attn_step: finish this row's Q/K/V head slice from the QKV split-K partials, append K/V to the cache, attend
// =============================================================================================
// grid (H, n), 128 threads.  Q and K carry the D^-1/4 scale (wc:2522,2566,2573), bias then scale; Q, K, V and the
// softmax probabilities are rounded to f16 where the reference's mul_mat converts them.
struct SelfArgs {
    unsigned long long * trace;
    const float * partial;  // [split][n][3S]
    int split;
    int64_t split_stride;
    const float * bias;     // [3S] = (bq, 0, bv)
    float scale;
    const SelfKvRow * rows;
    int S, layer;
    __half * out;           // [n][S]
    const __half * qkv16;   // alternative to partial: finished f16 Q|K|V [n][3S] (bias and D^-1/4 scale applied by the GEMM)
};

__global__ void __launch_bounds__(128, WB_SELF_MIN_CTAS)
self_attn_step_kernel(SelfArgs a) {
    extern __shared__ float s_sc[];
    __shared__ float s_q[64];
    __shared__ float s_red[4];
    __shared__ double s_redd[4];
    __shared__ float s_pv[16][64];
    pdl_launch_dependents();
    pdl_wait();
    if (threadIdx.x == 0) trace_stamp(a.trace);
    const int h = blockIdx.x, j = blockIdx.y, tid = threadIdx.x, S = a.S;
    const SelfKvRow r = a.rows[j];
    __half * kb = r.k_base + (int64_t) a.layer * r.n_ctx * S + h * 64;
    __half * vb = r.v_base + (int64_t) a.layer * r.n_ctx * S + h * 64;
    for (int t = tid; t < 192; t += 128) {
        const int part = t >> 6, d = t & 63, col = part * S + h * 64 + d;
        __half hv;
        if (a.qkv16 != nullptr) {
            hv = a.qkv16[(int64_t) j * 3 * S + col];
        } else {
            const float * p = a.partial + (int64_t) j * 3 * S + col;
            float acc = p[0];
            for (int s = 1; s < a.split; ++s) acc = __fadd_rn(acc, p[(int64_t) s * a.split_stride]);
            acc = __fadd_rn(acc, __ldg(a.bias + col));
            if (part < 2) acc = __fmul_rn(acc, a.scale);
            hv = __float2half_rn(acc);
        }
        if (part == 0) s_q[d] = __half2float(hv);
        else if (part == 1) kb[(int64_t) r.store_cell * S + d] = hv;
        else vb[(int64_t) r.store_cell * S + d] = hv;
    }
    __syncthreads();  // orders this CTA's K/V stores before its own loads below

    const int n = r.n_cells;
    float mx = -INFINITY;
    for (int c = tid; c < n; c += 128) {
        const int cell = r.cells ? r.cells[c] : c;
        const uint4 * kr = (const uint4 *) (kb + (int64_t) cell * S);
        float acc = 0.0f;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const uint4 u = kr[q];
            const __half2 * hp = (const __half2 *) &u;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 f = __half22float2(hp[e]);
                acc = fmaf(f.x, s_q[q * 8 + e * 2], acc);
                acc = fmaf(f.y, s_q[q * 8 + e * 2 + 1], acc);
            }
        }
        s_sc[c] = acc;
        mx = fmaxf(mx, acc);
    }
    mx = warp_max(mx);
    if ((tid & 31) == 0) s_red[tid >> 5] = mx;
    __syncthreads();
    mx = fmaxf(fmaxf(s_red[0], s_red[1]), fmaxf(s_red[2], s_red[3]));
    double sum = 0.0;
    for (int c = tid; c < n; c += 128) {
        const float e = expf(__fsub_rn(s_sc[c], mx));
        s_sc[c] = e;
        sum += (double) e;
    }
    sum = warp_sum_d(sum);
    if ((tid & 31) == 0) s_redd[tid >> 5] = sum;
    __syncthreads();
    sum = s_redd[0] + s_redd[1] + s_redd[2] + s_redd[3];
    const float inv = (float) (1.0 / sum);
    for (int c = tid; c < n; c += 128) s_sc[c] = __half2float(__float2half_rn(__fmul_rn(s_sc[c], inv)));
    __syncthreads();
    // P.V: thread (g, sub) owns 8 consecutive d of every 16th cell; groups are summed in a fixed order
    {
        const int g = tid >> 3, sub = tid & 7;
        float acc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = 0.0f;
#pragma unroll 4
        for (int c = g; c < n; c += 16) {
            const int cell = r.cells ? r.cells[c] : c;
            const float p = s_sc[c];
            const uint4 u = *(const uint4 *) (vb + (int64_t) cell * S + sub * 8);
            const __half2 * hp = (const __half2 *) &u;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 f = __half22float2(hp[e]);
                acc[e * 2]     = fmaf(p, f.x, acc[e * 2]);
                acc[e * 2 + 1] = fmaf(p, f.y, acc[e * 2 + 1]);
            }
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) s_pv[g][sub * 8 + e] = acc[e];
    }
    __syncthreads();
    if (tid < 64) {
        float o = 0.0f;
#pragma unroll
        for (int g = 0; g < 16; ++g) o += s_pv[g][tid];
        a.out[(int64_t) j * S + h * 64 + tid] = __float2half_rn(o);
    }
    if (tid == 0) trace_stamp_end(a.trace);
}

// =============================================================================================
// gelu_reduce: hb = f16(gelu(sum of fc1 split-K partials + bias))      (wc:2795-2802; GELU as the reference's f16 table)
// =============================================================================================
struct GeluArgs {
    const float * partial;  // [split][n][N]
    int split;
    int64_t split_stride;
    const float * bias;     // [N]
    const __half * gelu_table;
    __half * out;           // [n][N]
    int64_t total;          // n * N (N % 4 == 0)
    int N;
};

__global__ void __launch_bounds__(256)
gelu_reduce_kernel(GeluArgs a) {
    pdl_launch_dependents();
    pdl_wait();
    const int64_t i4 = ((int64_t) blockIdx.x * 256 + threadIdx.x) * 4;
    if (i4 >= a.total) return;
    float4 v = *(const float4 *) (a.partial + i4);
#pragma unroll 4
    for (int s = 1; s < a.split; ++s) {
        const float4 w = *(const float4 *) (a.partial + (int64_t) s * a.split_stride + i4);
        v.x = __fadd_rn(v.x, w.x); v.y = __fadd_rn(v.y, w.y); v.z = __fadd_rn(v.z, w.z); v.w = __fadd_rn(v.w, w.w);
    }
    const float4 b = __ldg((const float4 *) (a.bias + (i4 % a.N)));
    const __half2 h0 = __floats2half2_rn(gelu_f16table(__fadd_rn(v.x, b.x), a.gelu_table), gelu_f16table(__fadd_rn(v.y, b.y), a.gelu_table));
    const __half2 h1 = __floats2half2_rn(gelu_f16table(__fadd_rn(v.z, b.z), a.gelu_table), gelu_f16table(__fadd_rn(v.w, b.w), a.gelu_table));
    uint2 pk;
    pk.x = *(const uint32_t *) &h0;
    pk.y = *(const uint32_t *) &h1;
    *(uint2 *) (a.out + i4) = pk;
}

// =============================================================================================
// cross_attn_step: the kernel that bounds batched decoding (2 * T * 64 * 2 B of K/V per row, head and layer)
// =============================================================================================
// Persistent: a fixed number of CTAs (a multiple of the SM count, so that the kernel's shared-memory footprint per SM is
// known and the other decode lanes' kernels can co-reside) walk the H*n (row, head) items in a strided order; the ring
// keeps running across items, so the producer already streams the next item's K panel while the consumers finish the
// softmax / P.V / output of the current one.  8 consumer warps + 1 producer warp.  The producer streams the contiguous K
// panel [T][64] f16 and then the V panel of a (state, layer, head) in 16 KB chunks (128 keys) with cp.async.bulk into an mbarrier ring
// (L2 evict-first: every byte is used once per step); consumers compute scores from shared memory, the
// row softmax (max, exp, double sum, 1/sum: ggml_soft_max_ext, ops.cpp:4731-4827; scale applied to K.Q as wc:2732-2734),
// round P to f16 (mul_mat src1 conversion) and accumulate P.V in f32.  Q is finished from the cross-Q split-K partials.
struct CrossArgs {
    unsigned long long * trace;
    const float * partial;  // [split][n][S] cross-Q partials
    int split;
    int64_t split_stride;
    const float * bias;     // [S]
    const __half * q_f16;   // alternative to partial: finished Q [n][S] (tests)
    const CrossKvRow * rows;
    int H, T, S, layer;
    float scale;
    __half * out;           // [n][S]
    int stages;
    int items;              // H * n (row, head) pairs, walked by a fixed number of persistent CTAs
};

static constexpr int XS_CONSUMERS  = 256;
static constexpr int XS_THREADS    = XS_CONSUMERS + 32;
static constexpr int XS_CHUNK_ROWS = 128;
static constexpr int XS_CHUNK_BYTES = XS_CHUNK_ROWS * 128;

__global__ void __launch_bounds__(XS_THREADS, 2)
cross_attn_step_kernel(CrossArgs a) {
    extern __shared__ uint8_t xs_raw[];
    uint8_t * smem = (uint8_t *) (((uintptr_t) xs_raw + 127) & ~(uintptr_t) 127);
    uint8_t * ring   = smem;
    float * s_sc     = (float *) (ring + a.stages * XS_CHUNK_BYTES);          // [T] scores / probabilities
    uint64_t * full  = (uint64_t *) (s_sc + ((a.T + 31) & ~31));
    uint64_t * empty = full + a.stages;
    __shared__ float s_q[64];
    __shared__ float s_red[8];
    __shared__ double s_redd[8];
    __shared__ float s_acc[8][4][64];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int T = a.T, S = a.S;
    const int nck = (T + XS_CHUNK_ROWS - 1) / XS_CHUNK_ROWS;  // chunks per panel
    if (tid == 0) {
        for (int s = 0; s < a.stages; ++s) {
            mbar_init(full + s, 1);
            mbar_init(empty + s, 8);
        }
        mbar_fence_init();
    }
    __syncthreads();
    pdl_wait();
    if (tid == 0) trace_stamp(a.trace);

    if (warp == 8) {
        // ---------------- producer ----------------
        if (lane == 0) {
            const uint64_t pol = l2_policy_evict_first();
            int stage = 0;
            uint32_t phase = 0;
            for (int w = blockIdx.x; w < a.items; w += gridDim.x) {
                const int h = w % a.H, j = w / a.H;
                const __half * kv_base = a.rows[j].kv_base;
                const __half * kp = kv_base + ((int64_t) a.layer * 2 * a.H + h) * (int64_t) T * 64;
                const __half * vp = kv_base + ((int64_t) a.layer * 2 * a.H + a.H + h) * (int64_t) T * 64;
                for (int c = 0; c < 2 * nck; ++c) {
                    const int ci = c < nck ? c : c - nck;
                    const int rows = min(XS_CHUNK_ROWS, T - ci * XS_CHUNK_ROWS);
                    const __half * src = (c < nck ? kp : vp) + (int64_t) ci * XS_CHUNK_ROWS * 64;
                    mbar_wait(empty + stage, phase ^ 1);
                    mbar_expect_tx(full + stage, (uint32_t) rows * 128);
                    bulk_load_1d_hint(ring + stage * XS_CHUNK_BYTES, src, (uint32_t) rows * 128, full + stage, pol);
                    if (++stage == a.stages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
        return;
    }

    // ---------------- consumers (warps 0..7) ----------------
    const int sub = lane & 7, kq = lane >> 3;
    int stage = 0;
    uint32_t phase = 0;
    for (int w = blockIdx.x; w < a.items; w += gridDim.x) {
        const int h = w % a.H, j = w / a.H;
        // the dependent kernel's CTAs hold shared memory while they wait: let them in only near the end of this one
        if (w + (int) gridDim.x >= a.items) pdl_launch_dependents();
        if (tid < 64) {
            float qv;
            if (a.q_f16 != nullptr) {
                qv = __half2float(a.q_f16[(int64_t) j * S + h * 64 + tid]);
            } else {
                const float * p = a.partial + (int64_t) j * S + h * 64 + tid;
                float acc = p[0];
                for (int s = 1; s < a.split; ++s) acc = __fadd_rn(acc, p[(int64_t) s * a.split_stride]);
                acc = __fadd_rn(acc, __ldg(a.bias + h * 64 + tid));
                qv = __half2float(__float2half_rn(acc));
            }
            s_q[tid] = qv;
        }
        named_bar_sync(1, XS_CONSUMERS);

        float qf[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) qf[e] = s_q[sub * 8 + e];

        float mx = -INFINITY;
        for (int c = 0; c < nck; ++c) {
            const int rows = min(XS_CHUNK_ROWS, T - c * XS_CHUNK_ROWS);
            mbar_wait(full + stage, phase);
            const uint8_t * buf = ring + stage * XS_CHUNK_BYTES;
#pragma unroll
            for (int it = 0; it < XS_CHUNK_ROWS / 32; ++it) {
                const int rr = it * 32 + warp * 4 + kq;
                float acc = 0.0f;
                if (rr < rows) {
                    const uint4 u = *(const uint4 *) (buf + rr * 128 + sub * 16);
                    const __half2 * hp = (const __half2 *) &u;
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float2 f = __half22float2(hp[e]);
                        acc = fmaf(f.x, qf[e * 2], acc);
                        acc = fmaf(f.y, qf[e * 2 + 1], acc);
                    }
                }
                acc += __shfl_xor_sync(0xffffffffu, acc, 1);
                acc += __shfl_xor_sync(0xffffffffu, acc, 2);
                acc += __shfl_xor_sync(0xffffffffu, acc, 4);
                if (rr < rows && sub == 0) {
                    const float sc = __fmul_rn(acc, a.scale);
                    s_sc[c * XS_CHUNK_ROWS + rr] = sc;
                    mx = fmaxf(mx, sc);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty + stage);
            if (++stage == a.stages) {
                stage = 0;
                phase ^= 1;
            }
        }
        mx = warp_max(mx);
        if (lane == 0) s_red[warp] = mx;
        named_bar_sync(1, XS_CONSUMERS);
        mx = s_red[0];
#pragma unroll
        for (int w8 = 1; w8 < 8; ++w8) mx = fmaxf(mx, s_red[w8]);
        double sum = 0.0;
        for (int t = tid; t < T; t += XS_CONSUMERS) {
            const float e = expf(__fsub_rn(s_sc[t], mx));
            s_sc[t] = e;
            sum += (double) e;
        }
        sum = warp_sum_d(sum);
        if (lane == 0) s_redd[warp] = sum;
        named_bar_sync(1, XS_CONSUMERS);
        sum = 0.0;
#pragma unroll
        for (int w8 = 0; w8 < 8; ++w8) sum += s_redd[w8];
        const float inv = (float) (1.0 / sum);
        for (int t = tid; t < T; t += XS_CONSUMERS) s_sc[t] = __half2float(__float2half_rn(__fmul_rn(s_sc[t], inv)));
        named_bar_sync(1, XS_CONSUMERS);

        float acc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = 0.0f;
        for (int c = 0; c < nck; ++c) {
            const int rows = min(XS_CHUNK_ROWS, T - c * XS_CHUNK_ROWS);
            mbar_wait(full + stage, phase);
            const uint8_t * buf = ring + stage * XS_CHUNK_BYTES;
#pragma unroll
            for (int it = 0; it < XS_CHUNK_ROWS / 32; ++it) {
                const int rr = it * 32 + warp * 4 + kq;
                if (rr < rows) {
                    const float p = s_sc[c * XS_CHUNK_ROWS + rr];
                    const uint4 u = *(const uint4 *) (buf + rr * 128 + sub * 16);
                    const __half2 * hp = (const __half2 *) &u;
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float2 f = __half22float2(hp[e]);
                        acc[e * 2]     = fmaf(p, f.x, acc[e * 2]);
                        acc[e * 2 + 1] = fmaf(p, f.y, acc[e * 2 + 1]);
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty + stage);
            if (++stage == a.stages) {
                stage = 0;
                phase ^= 1;
            }
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) s_acc[warp][kq][sub * 8 + e] = acc[e];
        named_bar_sync(1, XS_CONSUMERS);
        if (tid < 64) {
            float o = 0.0f;
#pragma unroll
            for (int w8 = 0; w8 < 8; ++w8)
#pragma unroll
                for (int g = 0; g < 4; ++g) o += s_acc[w8][g][tid];
            a.out[(int64_t) j * S + h * 64 + tid] = __float2half_rn(o);
        }
    }
    if (tid == 0) trace_stamp_end(a.trace);
}

static int xs_smem_bytes(int T, int stages) { return stages * XS_CHUNK_BYTES + ((T + 31) & ~31) * 4 + 2 * stages * 8 + 128 + 64; }

static int step_sm_count() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    }
    return n;
}

static bool launch_cross(CrossArgs a, int n_rows, cudaStream_t s, bool pdl) {
    static PerDeviceOnce attr;
    if (attr.first()) {
        WB_CUDA_OK(cudaFuncSetAttribute(cross_attn_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 132 * 1024));
    }
    a.items = a.H * n_rows;
    a.trace = trace_slot("cross");
    a.stages = std::max(2, std::min(g_tune.xs_stages, 8));
    const int grid = std::min(a.items, step_sm_count() * std::max(1, g_tune.xs_ctas_per_sm));
    // algorithmic bytes: the K and V panels of every (row, head), f16
    ProfScope prof_scope(PC_CROSS_STEP, s, (double) a.items * a.T * 64 * 2 * 2, (double) a.items * a.T * 64 * 4);
    return launch_k(cross_attn_step_kernel, dim3(grid), dim3(XS_THREADS), (size_t) xs_smem_bytes(a.T, a.stages), s, pdl, a);
}

// =============================================================================================
// the step graph
// =============================================================================================
namespace {

struct GraphEntry {
    cudaGraphExec_t exec = nullptr;
    uint64_t launches = 0;
    uint64_t cat_counts[PC_COUNT] = {};  // launches per profiler category in one replay
};

struct StepWs {
    int cap_n = 0;
    uint8_t * base = nullptr;
    float * x = nullptr;         // [n][S] residual stream
    __half * xa = nullptr;       // [n][S] LayerNorm output
    __half * att = nullptr;      // [n][S] attention output
    __half * hb = nullptr;       // [n][4S]
    __half * qkv16 = nullptr;    // [n][3S] finished Q|K|V (cluster GEMM path)
    __half * q16 = nullptr;      // [n][S] finished cross-attention query
    float * partial = nullptr;   // split-K partial sums (WB200_STEP=partials path)
    float * logits = nullptr;    // [n][V]
    float * samp_work = nullptr; // [n][3][V]
    uint8_t * h_stage = nullptr; // pinned
    uint8_t * d_stage = nullptr;
    size_t stage_bytes = 0;
    SamplerOut * h_out = nullptr;  // pinned
    SamplerOut * d_out = nullptr;
    std::map<std::array<int, 8>, GraphEntry> graphs;
    unsigned long long * d_trace = nullptr;  // WB200_STEP_TRACE: one %globaltimer stamp per kernel of the step
    std::vector<const char *> trace_names;
    std::vector<int> trace_subs;
    int trace_n = 0;                         // rows of the traced step
    int kv_cells_max = 0;                    // largest self-KV cache (cells) among the staged rows: sizes self_attn_step's score buffer
    int64_t kv_cells_live = 0;               // live cells summed over the staged rows (algorithmic bytes of self_attn_step)
};

struct StageLayout {
    size_t o_tok, o_pos, o_self, o_cross, o_samp, bytes;
};
static StageLayout stage_layout(int n) {
    StageLayout L{};
    size_t off = 0;
    auto sub = [&](size_t b) {
        const size_t o = off;
        off += (b + 15) & ~(size_t) 15;
        return o;
    };
    L.o_tok = sub((size_t) n * 4);
    L.o_pos = sub((size_t) n * 4);
    L.o_self = sub((size_t) n * sizeof(SelfKvRow));
    L.o_cross = sub((size_t) n * sizeof(CrossKvRow));
    L.o_samp = sub((size_t) n * sizeof(SamplerRow));
    L.bytes = off;
    return L;
}

void destroy_graphs(StepWs & ws) {
    for (auto & kv : ws.graphs)
        if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    ws.graphs.clear();
}

void free_ws(StepWs & ws) {
    destroy_graphs(ws);
    cudaFree(ws.base);
    cudaFree(ws.d_stage);
    cudaFree(ws.d_out);
    cudaFree(ws.d_trace);
    if (ws.h_stage) cudaFreeHost(ws.h_stage);
    if (ws.h_out) cudaFreeHost(ws.h_out);
    ws = StepWs();
}

bool ensure_ws(Context & ctx, StepWs & ws, int n, cudaStream_t stream) {
    if (ws.base && ws.cap_n >= n) return true;
    const auto & hp = ctx.model.hp;
    const int S = hp.n_text_state, V = hp.n_vocab;
    const int cap = std::min(256, std::max(n, std::max(16, ws.cap_n * 2)));
    WB_CUDA_OK(cudaStreamSynchronize(stream));
    free_ws(ws);  // graphs bake the workspace pointers
    size_t max_part = 0;
    const int shapes[4][2] = {{3 * S, S}, {S, S}, {S, 4 * S}, {4 * S, S}};
    for (auto & sh : shapes) {
        const WsPlan p = ws_plan(cap, sh[0], sh[1], 0);
        max_part = std::max(max_part, (size_t) p.split * cap * sh[0]);
    }
    size_t off = 0;
    auto take = [&](size_t b) {
        const size_t o = off;
        off += (b + 1023) & ~(size_t) 1023;
        return o;
    };
    const size_t o_x = take((size_t) cap * S * 4), o_xa = take((size_t) cap * S * 2), o_att = take((size_t) cap * S * 2);
    const size_t o_hb = take((size_t) cap * 4 * S * 2), o_part = take(max_part * 4), o_lg = take((size_t) cap * V * 4);
    const size_t o_sw = take((size_t) cap * 3 * V * 4), o_qkv = take((size_t) cap * 3 * S * 2), o_q = take((size_t) cap * S * 2);
    WB_CUDA_OK(cudaMalloc((void **) &ws.base, off));
    ws.x = (float *) (ws.base + o_x); ws.xa = (__half *) (ws.base + o_xa); ws.att = (__half *) (ws.base + o_att);
    ws.hb = (__half *) (ws.base + o_hb); ws.partial = (float *) (ws.base + o_part); ws.logits = (float *) (ws.base + o_lg);
    ws.samp_work = (float *) (ws.base + o_sw);
    ws.qkv16 = (__half *) (ws.base + o_qkv); ws.q16 = (__half *) (ws.base + o_q);
    ws.stage_bytes = stage_layout(cap).bytes;
    WB_CUDA_OK(cudaMallocHost((void **) &ws.h_stage, ws.stage_bytes));
    WB_CUDA_OK(cudaMalloc((void **) &ws.d_stage, ws.stage_bytes));
    WB_CUDA_OK(cudaMallocHost((void **) &ws.h_out, (size_t) cap * sizeof(SamplerOut)));
    WB_CUDA_OK(cudaMalloc((void **) &ws.d_out, (size_t) cap * sizeof(SamplerOut)));
    ws.cap_n = cap;
    return true;
}

// Calibration of the sampled live steps: an event pair around a kernel that does nothing measures what the bracketing itself
// adds to every sampled launch (event records + an unoverlapped launch); bench.py subtracts it from the per-kernel averages.
__global__ void null_kernel() {}
static void prof_calibrate(cudaStream_t s) {
    for (int i = 0; i < 8; ++i) {
        ProfScope prof_scope(PC_CALIB, s);
        null_kernel<<<1, 32, 0, s>>>();
    }
}

// The launches of one decode step for the n rows of one workspace, in the pieces the two issue orders below need.
struct LaneIssue {
    Context & ctx;
    StepWs & ws;
    int n, T;
    cudaStream_t s;
    const Model & m;
    int S, H, L, V;
    StageLayout SL;
    const int *d_tok, *d_pos;
    const SelfKvRow * d_self;
    const CrossKvRow * d_cross;
    const SamplerRow * d_samp;
    float kq_scale;

    LaneIssue(Context & c, StepWs & w, int n_, int T_, cudaStream_t s_)
        : ctx(c), ws(w), n(n_), T(T_), s(s_), m(c.model), S(c.model.hp.n_text_state), H(c.model.hp.n_text_head), L(c.model.hp.n_text_layer),
          V(c.model.hp.n_vocab), SL(stage_layout(n_)) {
        d_tok = (const int *) (ws.d_stage + SL.o_tok);
        d_pos = (const int *) (ws.d_stage + SL.o_pos);
        d_self = (const SelfKvRow *) (ws.d_stage + SL.o_self);
        d_cross = (const CrossKvRow *) (ws.d_stage + SL.o_cross);
        d_samp = (const SamplerRow *) (ws.d_stage + SL.o_samp);
        kq_scale = (float) pow((double) 64.0f, -0.25);  // wc:2522
    }

    bool ln(bool embed, const Norm & nm) {
        LnArgs a{};
        a.res = embed ? nullptr : ws.x;
        a.te = embed ? m.d_te : nullptr; a.pe = m.d_pe; a.tokens = d_tok; a.pos = d_pos;
        a.x_out = ws.x; a.xa_out = ws.xa; a.gamma = nm.g; a.beta = nm.b; a.eps = m.hp.eps; a.S = S;
        a.trace = trace_slot("ln");
        ProfScope prof_scope(PC_LN_DEC, s, (double) n * S * 10);
        return launch_k(ln_reduce_kernel, dim3(n), dim3(LN_THREADS), 0, s, !embed, a);
    }
    // staging upload + token/position embedding + first LayerNorm
    bool begin() {
        static PerDeviceOnce attrs;
        if (attrs.first()) {
            WB_CUDA_OK(cudaFuncSetAttribute(self_attn_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 5120 * 4));
        }
        WB_CHECK(S <= LN_THREADS * LN_MAX_PER_THREAD && (S % 64) == 0, "unsupported text state %d\n", S);
        WB_CUDA_OK(cudaMemcpyAsync(ws.d_stage, ws.h_stage, SL.bytes, cudaMemcpyHostToDevice, s));
        return ln(true, L > 0 ? m.dec[0].ln1 : m.d_ln);
    }
    // self-attention block and the cross-attention query of layer il (everything up to the K/V stream)
    bool pre(int il) {
        const DecLayer & l = m.dec[il];
        {   // Q|K|V: bias, D^-1/4 on Q and K, f16 (wc:2558-2583)
            WcOut o{};
            o.mode = WC_F16; o.out = ws.qkv16; o.ldo = 3 * S; o.bias = l.qkv.b; o.scale = kq_scale; o.scale_f1 = 2 * S;
            if (!gemm_wcluster(l.qkv.w, ws.xa, n, 3 * S, S, o, s, true)) return false;
        }
        {
            SelfArgs a{};
            a.qkv16 = ws.qkv16; a.rows = d_self; a.S = S; a.layer = il; a.out = ws.att;
            a.trace = trace_slot("self");
            ProfScope prof_scope(PC_SELF_ATTN, s, (double) ws.kv_cells_live * S * 2 * 2 /* K and V rows of every live cell, f16 */);
            // the score buffer only has to hold the cache's cells (step_eligible bounds them by 5000): 2 KB instead of 20 KB for
            // the greedy cache of 512 cells, which lifts the kernel from 9 to 16 resident CTAs per SM
            const size_t sc_bytes = (size_t) std::min(5120, std::max(256, (int) pad_to(ws.kv_cells_max, 64))) * 4;
            if (!launch_k(self_attn_step_kernel, dim3(H, n), dim3(128), sc_bytes, s, true, a)) return false;
        }
        {   // x += Wo.att + bo (wc:2657-2664)
            WcOut o{};
            o.mode = WC_RESIDUAL; o.out = ws.x; o.ldo = S; o.bias = l.o.b;
            if (!gemm_wcluster(l.o.w, ws.att, n, S, S, o, s, true)) return false;
        }
        if (!ln(false, l.lnc)) return false;
        {   // cross-attention query, f16 (wc:2700-2707)
            WcOut o{};
            o.mode = WC_F16; o.out = ws.q16; o.ldo = S; o.bias = l.cq.b;
            if (!gemm_wcluster(l.cq.w, ws.xa, n, S, S, o, s, true)) return false;
        }
        return true;
    }
    // the HBM-bound kernel: streams this lane's cross K/V of layer il.  `pdl`: launch as a programmatic dependent
    bool cross(int il, bool pdl) {
        CrossArgs a{};
        a.q_f16 = ws.q16; a.rows = d_cross; a.H = H; a.T = T; a.S = S; a.layer = il; a.scale = kq_scale; a.out = ws.att;
        const int prio = g_launch_prio;
        g_launch_prio = 0;  // the bandwidth-bound kernel: the other lane's short kernels overtake its pending CTAs
        const bool ok = launch_cross(a, n, s, pdl);
        g_launch_prio = prio;
        return ok;
    }
    // cross-attention output projection, MLP, and the LayerNorm that opens the next layer (or the final one)
    bool post(int il) {
        const DecLayer & l = m.dec[il];
        {
            WcOut o{};
            o.mode = WC_RESIDUAL; o.out = ws.x; o.ldo = S; o.bias = l.co.b;
            if (!gemm_wcluster(l.co.w, ws.att, n, S, S, o, s, true)) return false;
        }
        if (!ln(false, l.ln2)) return false;
        {   // MLP: f16(gelu(W1.x + b1)) through the reference's f16 table, then x += W2.h + b2 (wc:2795-2814)
            WcOut o{};
            o.mode = WC_GELU; o.out = ws.hb; o.ldo = 4 * S; o.bias = l.fc1.b; o.gelu_table = m.gelu_table;
            if (!gemm_wcluster(l.fc1.w, ws.xa, n, 4 * S, S, o, s, true)) return false;
            WcOut o2{};
            o2.mode = WC_RESIDUAL; o2.out = ws.x; o2.ldo = S; o2.bias = l.fc2.b;
            if (!gemm_wcluster(l.fc2.w, ws.hb, n, S, 4 * S, o2, s, true)) return false;
        }
        return ln(false, il + 1 < L ? m.dec[il + 1].ln1 : m.d_ln);
    }
    // tied-embedding logits (wc:2835), sampler, picks to the host
    bool finish(const SamplerStatic & sst) {
        WsOut d{};
        d.out = ws.logits; d.out_f16 = 0; d.ldo = V;
        if (!gemm_wstream(m.d_te, ws.xa, n, V, S, d, 1, nullptr, s, true)) return false;
        if (!sample_rows(sst, ws.logits, V, d_samp, n, ws.samp_work, ws.d_out, s)) return false;
        WB_CUDA_OK(cudaMemcpyAsync(ws.h_out, ws.d_out, (size_t) n * sizeof(SamplerOut), cudaMemcpyDeviceToHost, s));
        return true;
    }
};

// issues every launch of one decode step for n rows on `s` (live or under stream capture)
bool issue_step(Context & ctx, StepWs & ws, int n, int T, const SamplerStatic & sst, cudaStream_t s) {
    LaneIssue a(ctx, ws, n, T, s);
    if (!a.begin()) return false;
    for (int il = 0; il < a.L; ++il)
        if (!a.pre(il) || !a.cross(il, true) || !a.post(il)) return false;
    return a.finish(sst);
}

// Interleaved issue: the rows are split into K groups that advance through the layers a fraction of a layer apart, as K
// branches of ONE graph.  The HBM-bound cross attentions form a single chain A0 -> B0 -> C0 -> A1 -> B1 ... (full graph
// edges through events), so exactly one K/V stream owns the memory system at a time while the other groups' latency-bound
// kernels (GEMMs at a few rows, LayerNorm, self attention) run underneath it -- HBM never waits for them as long as a
// group's short kernels fit under the other K-1 groups' K/V streams.
bool issue_step_multi(Context & ctx, StepWs * const * w, const int * n, int K, int T, const SamplerStatic & sst, cudaStream_t * st,
                      cudaEvent_t * ev /*[K * L + 2]*/) {
    std::vector<LaneIssue> G;
    G.reserve((size_t) K);
    for (int k = 0; k < K; ++k) G.emplace_back(ctx, *w[k], n[k], T, st[k]);
    static const bool cross_pdl = getenv("WB200_PAIR_PDL") != nullptr;
    const int L = G[0].L;
    cudaEvent_t fork = ev[K * L], join = ev[K * L + 1];
    WB_CUDA_OK(cudaEventRecord(fork, st[0]));
    for (int k = 1; k < K; ++k) WB_CUDA_OK(cudaStreamWaitEvent(st[k], fork, 0));
    for (int k = 0; k < K; ++k) {
        g_trace_sub = k;
        if (!G[(size_t) k].begin()) return false;
    }
    for (int il = 0; il < L; ++il)
        for (int k = 0; k < K; ++k) {
            g_trace_sub = k;
            LaneIssue & g = G[(size_t) k];
            if (!g.pre(il)) return false;
            const int link = il * K + k;  // position in the chain of K/V streams
            if (link > 0) WB_CUDA_OK(cudaStreamWaitEvent(st[k], ev[link - 1], 0));  // the previous K/V stream has finished
            if (!g.cross(il, link == 0 || cross_pdl)) return false;
            WB_CUDA_OK(cudaEventRecord(ev[link], st[k]));
            if (!g.post(il)) return false;
        }
    for (int k = 0; k < K; ++k) {
        g_trace_sub = k;
        if (!G[(size_t) k].finish(sst)) return false;
    }
    g_trace_sub = 0;
    for (int k = 1; k < K; ++k) {
        WB_CUDA_OK(cudaEventRecord(join, st[k]));
        WB_CUDA_OK(cudaStreamWaitEvent(st[0], join, 0));
    }
    return true;
}

}  // namespace

// Decode lanes: independent groups of sequences, each with its own stream, workspace and step graphs.  While one
// lane's cross attention saturates HBM, the latency-bound kernels of the other lanes (and the host's token bookkeeping
// for a finished lane) run underneath it.  Lane 0 shares the context's stream with the general stages.
static constexpr int MAX_LANES = 8;
static constexpr int MAX_GROUPS = 4;
struct StepLanes {
    StepWs ws[MAX_LANES];
    StepWs sub[MAX_LANES][MAX_GROUPS - 1];            // groups 1.. of an interleaved step (issue_step_multi)
    cudaStream_t stream[MAX_LANES] = {};
    cudaStream_t sub_stream[MAX_LANES][MAX_GROUPS - 1] = {};  // capture-time branches of groups 1..
    int groups[MAX_LANES] = {};                       // groups of the step in flight
    int group_n[MAX_LANES][MAX_GROUPS] = {};          // and their row counts
    int64_t t0[MAX_LANES] = {};
    uint64_t step_no[MAX_LANES] = {};
    std::vector<cudaEvent_t> events;      // fork/join and the cross-attention chain of the paired graph
    int prio_hi = 0;
};

static StepLanes * step_lanes(Context & ctx) {
    if (!ctx.step) {
        StepLanes * L = new StepLanes();
        int lo = 0, hi = 0;
        if (getenv("WB200_NO_PRIO") == nullptr && cudaDeviceGetStreamPriorityRange(&lo, &hi) == cudaSuccess) L->prio_hi = hi;
        ctx.step = L;
    }
    return (StepLanes *) ctx.step;
}

int step_max_lanes() { return MAX_LANES; }

// WB200_STEP_TRACE=<file>: append the stamps of every lane's most recent step as one JSON line
void step_trace_dump(Context & ctx) {
    const char * path = getenv("WB200_STEP_TRACE");
    if (!path || !ctx.step) return;
    StepLanes & L = *(StepLanes *) ctx.step;
    FILE * f = fopen(path, "a");
    if (!f) return;
    fprintf(f, "{\"lanes\": [");
    bool first = true;
    for (int l = 0; l < MAX_LANES; ++l) {
        StepWs & ws = L.ws[l];
        if (!ws.d_trace || ws.trace_names.empty()) continue;
        std::vector<unsigned long long> t2(2 * ws.trace_names.size()), t(ws.trace_names.size()), te(ws.trace_names.size());
        if (cudaMemcpy(t2.data(), ws.d_trace, t2.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost) != cudaSuccess) continue;
        for (size_t i = 0; i < t.size(); ++i) {
            t[i] = t2[2 * i];
            te[i] = t2[2 * i + 1];
        }
        fprintf(f, "%s{\"lane\": %d, \"n\": %d, \"names\": [", first ? "" : ", ", l, ws.trace_n);
        for (size_t i = 0; i < t.size(); ++i) fprintf(f, "%s\"%s\"", i ? "," : "", ws.trace_names[i]);
        fprintf(f, "], \"sub\": [");
        for (size_t i = 0; i < t.size(); ++i) fprintf(f, "%s%d", i ? "," : "", i < ws.trace_subs.size() ? ws.trace_subs[i] : 0);
        fprintf(f, "], \"t_ns\": [");
        for (size_t i = 0; i < t.size(); ++i) fprintf(f, "%s%llu", i ? "," : "", t[i]);
        fprintf(f, "], \"t_end_ns\": [");
        for (size_t i = 0; i < te.size(); ++i) fprintf(f, "%s%llu", i ? "," : "", te[i]);
        fprintf(f, "]}");
        first = false;
    }
    fprintf(f, "]}\n");
    fclose(f);
}

void step_release(Context & ctx) {
    if (!ctx.step) return;
    StepLanes * L = (StepLanes *) ctx.step;
    for (int i = 0; i < MAX_LANES; ++i) {
        if (L->stream[i]) cudaStreamSynchronize(L->stream[i]);
        free_ws(L->ws[i]);
        for (int g = 0; g < MAX_GROUPS - 1; ++g) {
            free_ws(L->sub[i][g]);
            if (L->sub_stream[i][g]) cudaStreamDestroy(L->sub_stream[i][g]);
        }
        if (i > 0 && L->stream[i]) cudaStreamDestroy(L->stream[i]);
    }
    for (cudaEvent_t e : L->events) cudaEventDestroy(e);
    delete L;
    ctx.step = nullptr;
}

bool step_eligible(const Context & ctx, const BatchRow * rows, int n, bool any_host_dists) {
    static const bool disabled = getenv("WB200_NO_STEP_GRAPH") != nullptr;
    if (disabled || any_host_dists || n < 1 || n > 256 || ctx.model.n_loaded == 0 || ctx.model.hp.n_text_layer < 1) return false;
    const int S = ctx.model.hp.n_text_state;
    if (S > LN_THREADS * LN_MAX_PER_THREAD || (S % 64) != 0) return false;
    const int T = rows[0].st->cross_T;
    if (T <= 0) return false;
    for (int i = 0; i < n; ++i) {
        const BatchRow & r = rows[i];
        if (!r.want_logits || r.seq_id != 0 || r.st->cross_T != T) return false;
        for (int k = 0; k < i; ++k)
            if (rows[k].st == r.st) return false;  // one row per state
        // the cache must be in append order: cells [0, used) live for sequence 0 at earlier positions, everything after free
        const KvCache & kv = r.st->kv_self;
        int used = 0;
        while (used < kv.size && kv.cells[(size_t) used].pos >= 0) {
            if (!kv.cells[(size_t) used].has(0) || kv.cells[(size_t) used].pos > r.pos) return false;
            ++used;
        }
        if (used >= kv.size || used > 5000) return false;
        for (int c = used + 1; c < kv.size; ++c)
            if (kv.cells[(size_t) c].pos >= 0) return false;
    }
    return true;
}

// fills the pinned staging block of `ws` for rows[0..n) (KV slot search included, wc:2885-2897)
static bool fill_stage(StepWs & ws, const BatchRow * rows, const SamplerRow * srows, int n) {
    const StageLayout SL = stage_layout(n);
    int * h_tok = (int *) (ws.h_stage + SL.o_tok);
    int * h_pos = (int *) (ws.h_stage + SL.o_pos);
    SelfKvRow * h_self = (SelfKvRow *) (ws.h_stage + SL.o_self);
    CrossKvRow * h_cross = (CrossKvRow *) (ws.h_stage + SL.o_cross);
    SamplerRow * h_samp = (SamplerRow *) (ws.h_stage + SL.o_samp);
    ws.kv_cells_max = 0;
    ws.kv_cells_live = 0;
    for (int i = 0; i < n; ++i) {
        State * st = rows[i].st;
        KvCache & kv = st->kv_self;
        if (!kv.find_slot(&rows[i].pos, &rows[i].seq_id, 1)) {
            log_error("%s: failed to find a KV slot\n", __func__);
            return false;
        }
        kv.n = std::min(kv.size, std::max(1, kv.cell_max()));
        h_tok[i] = rows[i].token;
        h_pos[i] = rows[i].pos;
        SelfKvRow & sr = h_self[i];
        sr.k_base = kv.k; sr.v_base = kv.v; sr.cells = nullptr;
        sr.n_cells = kv.head + 1; sr.store_cell = kv.head; sr.n_ctx = kv.size;
        ws.kv_cells_max = std::max(ws.kv_cells_max, kv.size);
        ws.kv_cells_live += kv.head + 1;
        h_cross[i] = CrossKvRow{st->kv_cross, nullptr, 0, 0};
        h_samp[i] = srows[i];
        h_samp[i].logits_row = i;
        h_samp[i].work_slot = i;
    }
    return true;
}

bool stage_step_launch(Context & ctx, int lane, int n_lanes, const BatchRow * rows, int n, const SamplerStatic & sst, const SamplerRow * srows) {
    WB_CUDA_OK(cudaSetDevice(ctx.model.device));
    WB_CHECK(lane >= 0 && lane < MAX_LANES, "decode lane %d out of range\n", lane);
    static const bool no_graph = getenv("WB200_STEP_NO_GRAPH") != nullptr;
    g_pdl = getenv("WB200_NO_PDL") == nullptr;
    StepLanes & L = *step_lanes(ctx);
    if (!L.stream[lane]) {
        if (lane == 0) L.stream[0] = ctx.stream;
        else WB_CUDA_OK(cudaStreamCreateWithFlags(&L.stream[lane], cudaStreamNonBlocking));
    }
    cudaStream_t s = L.stream[lane];
    L.t0[lane] = time_us();
    const int T = rows[0].st->cross_T;
    // Interleaved step (issue_step_multi): from WB200_PAIR_MIN rows per group on, the rows advance as K groups of one graph
    int K = 1;
    if (env_int("WB200_PAIR", 0) != 0) {  // off by default: measured gain on B200 is within 3 %
        const int per_group = std::max(1, env_int("WB200_PAIR_MIN", 64)), k_max = std::max(1, std::min(MAX_GROUPS, env_int("WB200_GROUPS", 2)));
        K = std::max(1, std::min(k_max, n / per_group));
    }
    const bool paired = K > 1;
    StepWs * gw[MAX_GROUPS];
    cudaStream_t gs[MAX_GROUPS];
    int gn[MAX_GROUPS], goff[MAX_GROUPS];
    for (int k = 0, off = 0; k < K; ++k) {
        gn[k] = n / K + (k < n % K ? 1 : 0);
        goff[k] = off;
        off += gn[k];
        L.group_n[lane][k] = gn[k];
    }
    L.groups[lane] = K;
    StepWs & ws = L.ws[lane];
    gw[0] = &ws;
    gs[0] = s;
    if (!ensure_ws(ctx, ws, gn[0], s)) return false;
    for (int k = 1; k < K; ++k) {
        StepWs & wk = L.sub[lane][k - 1];
        if (!L.sub_stream[lane][k - 1]) WB_CUDA_OK(cudaStreamCreateWithFlags(&L.sub_stream[lane][k - 1], cudaStreamNonBlocking));
        const uint8_t * base = wk.base;
        if (!ensure_ws(ctx, wk, gn[k], s)) return false;
        if (wk.base != base) destroy_graphs(ws);  // the interleaved graphs (kept with group 0) bake the other groups' pointers
        gw[k] = &wk;
        gs[k] = L.sub_stream[lane][k - 1];
    }
    if (paired) {
        const size_t need = (size_t) K * ctx.model.hp.n_text_layer + 2;
        while (L.events.size() < need) {
            cudaEvent_t e = nullptr;
            WB_CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            L.events.push_back(e);
        }
    }
    for (int k = 0; k < K; ++k)
        if (!fill_stage(*gw[k], rows + goff[k], srows + goff[k], gn[k])) return false;

    // More than one K/V stream's worth of kernels in flight (lanes, or the two halves of a paired step): every kernel's
    // footprint per SM is bounded so that the short kernels always find room next to a cross attention: two persistent
    // CTAs per SM with a 3-deep ring (2 x 64 KB, still 97 % of the HBM peak) + a cluster GEMM CTA capped at 66 KB --
    // measured with tools/microbench_overlap.py, DESIGN.md 4.3.
    const bool multi = n_lanes > 1 || paired;
    LaneTuning tune;
    tune.xs_ctas_per_sm = env_int("WB200_XS_CPS", 2);
    tune.xs_stages = env_int("WB200_XS_STAGES", multi ? 3 : 5);
    tune.wc_smem_cap = env_int("WB200_WC_SMEM_KB", multi ? 66 : 192) * 1024;
    g_tune = tune;
    g_launch_prio = multi ? L.prio_hi : 0;
    static const bool tracing = getenv("WB200_STEP_TRACE") != nullptr;
    auto trace_begin = [&]() {
        if (!tracing) return;
        if (!ws.d_trace && cudaMalloc((void **) &ws.d_trace, 4096 * sizeof(unsigned long long)) != cudaSuccess) return;
        ws.trace_names.clear();
        ws.trace_subs.clear();
        ws.trace_n = n;
        g_trace_next = ws.d_trace;
        g_trace_names = &ws.trace_names;
        g_trace_subs = &ws.trace_subs;
        g_trace_sub = 0;
    };
    auto trace_end = [&]() { g_trace_next = nullptr; g_trace_names = nullptr; g_trace_subs = nullptr; g_trace_sub = 0; };
    auto issue = [&]() {
        return paired ? issue_step_multi(ctx, gw, gn, K, T, sst, gs, L.events.data()) : issue_step(ctx, ws, gn[0], T, sst, s);
    };
    auto reset = [&]() {
        trace_end();
        g_launch_prio = 0;
        g_tune = LaneTuning();
    };
    // with the in-situ profiler on (bench.py), every 16th step of a lane is issued live so that CUDA events on the launching
    // stream time its kernels inside the real run (graph nodes cannot be bracketed by events without breaking the graph)
    // The sampled step runs with the device to itself (the other lanes are drained first and it is waited for here), so the
    // event pairs bracket each kernel alone -- its own speed, not its share of a machine contended by the other lanes.
    // every 384th step of a lane is issued live under the profiler (the other lanes are drained for it): ~2 live steps per lane and
    // bench step, 1.5 % of the timed region (every 128th cost 4 % once the stream API kept all lanes busy all the time)
    const bool live_sample = prof_enabled() && (L.step_no[lane]++ % 384) == 7;
    if (live_sample)
        for (int l = 0; l < MAX_LANES; ++l)
            if (l != lane && L.stream[l]) WB_CUDA_OK(cudaStreamSynchronize(L.stream[l]));
    if (no_graph || live_sample) {
        trace_begin();
        if (live_sample) prof_calibrate(s);
        const bool ok = issue();
        reset();
        if (!ok) return false;
        if (live_sample) WB_CUDA_OK(cudaStreamSynchronize(s));
    } else {
        int cells = 0;
        for (int k = 0; k < K; ++k) cells = std::max(cells, (int) pad_to(gw[k]->kv_cells_max, 64));
        const std::array<int, 8> key = {n, K, T, tune.xs_ctas_per_sm, tune.xs_stages * 16 + (int) multi * 2 + (int) g_pdl, tune.wc_smem_cap, cells, 0};
        auto it = ws.graphs.find(key);
        if (it == ws.graphs.end()) {
            const uint64_t l0 = g_step_launches.load() + kernel_launch_count();
            cudaGraph_t graph = nullptr;
            GraphEntry ge;
            prof_suspend(true);
            prof_capture_counts(ge.cat_counts);
            trace_begin();
            cudaError_t e = cudaStreamBeginCapture(s, cudaStreamCaptureModeRelaxed);
            const bool ok = e == cudaSuccess && issue();
            if (e == cudaSuccess) e = cudaStreamEndCapture(s, &graph);
            reset();
            prof_capture_counts(nullptr);
            prof_suspend(false);
            if (!ok || e != cudaSuccess || graph == nullptr) {
                if (graph) cudaGraphDestroy(graph);
                log_error("%s: capturing the decode-step graph failed (%s)\n", __func__, cudaGetErrorString(e));
                return false;
            }
            ge.launches = g_step_launches.load() + kernel_launch_count() - l0;
            const cudaError_t ei = cudaGraphInstantiate(&ge.exec, graph, 0);
            cudaGraphDestroy(graph);
            if (ei != cudaSuccess) {
                log_error("%s: cudaGraphInstantiate failed (%s)\n", __func__, cudaGetErrorString(ei));
                return false;
            }
            // launches counted while capturing were not executed
            g_step_launches.fetch_sub(ge.launches - 1);
            it = ws.graphs.emplace(key, ge).first;
        }
        reset();
        WB_CUDA_OK(cudaGraphLaunch(it->second.exec, s));
        g_step_launches.fetch_add(it->second.launches);
        prof_add_totals(it->second.cat_counts);
    }
    return true;
}

bool stage_step_ready(Context & ctx, int lane) {
    StepLanes & L = *step_lanes(ctx);
    return lane >= 0 && lane < MAX_LANES && L.stream[lane] && cudaStreamQuery(L.stream[lane]) != cudaErrorNotReady;
}

bool stage_step_wait(Context & ctx, int lane, const BatchRow * rows, int n, Decoder ** decs) {
    StepLanes & L = *step_lanes(ctx);
    WB_CHECK(lane >= 0 && lane < MAX_LANES && L.stream[lane], "decode lane %d was never launched\n", lane);
    WB_CUDA_OK(cudaStreamSynchronize(L.stream[lane]));
    for (int k = 0, i = 0; k < L.groups[lane]; ++k) {
        const StepWs & wk = k == 0 ? L.ws[lane] : L.sub[lane][k - 1];
        for (int r = 0; r < L.group_n[lane][k] && i < n; ++r, ++i)
            if (decs && decs[i]) decs[i]->picked = wk.h_out[r];
    }
    const int64_t dt = time_us() - L.t0[lane];
    for (int i = 0; i < n; ++i) {  // wc:2982-2991: single-token decodes
        rows[i].st->t_decode_us += dt / n;
        rows[i].st->n_decode++;
    }
    return true;
}

bool stage_step(Context & ctx, const BatchRow * rows, int n, const SamplerStatic & sst, const SamplerRow * srows, Decoder ** decs) {
    return stage_step_launch(ctx, 0, 1, rows, n, sst, srows) && stage_step_wait(ctx, 0, rows, n, decs);
}

uint64_t step_launch_count() { return g_step_launches.load(); }

// =============================================================================================
// test / micro-benchmark hooks
// =============================================================================================
bool test_gemm_wstream(const __half * w, const __half * x, float * out, int n, int N_out, int K, const float * bias, int force_split) {
    g_pdl = false;
    if (force_split == 1) {
        WsOut d{};
        d.out = out; d.out_f16 = 0; d.ldo = N_out; d.bias = bias;
        if (!gemm_wstream(w, x, n, N_out, K, d, 1, nullptr, 0, false)) return false;
        WB_CUDA_OK(cudaDeviceSynchronize());
        return true;
    }
    // partial mode: reduce on the host side of the test through a plain device loop (ln_reduce needs LayerNorm weights)
    const WsPlan p = ws_plan(n, N_out, K, force_split);
    float * part = nullptr;
    WB_CUDA_OK(cudaMalloc((void **) &part, (size_t) p.split * n * N_out * 4));
    WsOut po{};
    po.partial = part;
    int split = 0;
    bool ok = gemm_wstream(w, x, n, N_out, K, po, force_split, &split, 0, false);
    std::vector<float> h((size_t) split * n * N_out), o((size_t) n * N_out, 0.0f), hb;
    ok = ok && cudaMemcpy(h.data(), part, h.size() * 4, cudaMemcpyDeviceToHost) == cudaSuccess;
    if (ok && bias) {
        hb.resize((size_t) N_out);
        ok = cudaMemcpy(hb.data(), bias, hb.size() * 4, cudaMemcpyDeviceToHost) == cudaSuccess;
    }
    if (ok) {
        for (size_t i = 0; i < o.size(); ++i) {
            float acc = h[i];
            for (int s = 1; s < split; ++s) acc += h[(size_t) s * n * N_out + i];
            if (bias) acc += hb[i % (size_t) N_out];
            o[i] = acc;
        }
        ok = cudaMemcpy(out, o.data(), o.size() * 4, cudaMemcpyHostToDevice) == cudaSuccess;
    }
    cudaFree(part);
    return ok;
}

static bool launch_cross_test(const __half * q, const CrossKvRow * d_rows, __half * out, int n_rows, int H, int T, int layer, cudaStream_t s) {
    CrossArgs a{};
    a.q_f16 = q; a.rows = d_rows; a.H = H; a.T = T; a.S = H * 64; a.layer = layer; a.scale = (float) pow(64.0, -0.25); a.out = out;
    g_tune.xs_ctas_per_sm = env_int("WB200_XS_CPS", 2);
    g_tune.xs_stages = env_int("WB200_XS_STAGES", 5);
    const bool ok = launch_cross(a, n_rows, s, false);
    g_tune = LaneTuning();
    return ok;
}

bool test_cross_attn_step(const __half * q_f16, const __half * kv, __half * out, int n_rows, int H, int T) {
    std::vector<CrossKvRow> h((size_t) n_rows, CrossKvRow{kv, nullptr, 0, 0});
    CrossKvRow * d_rows = nullptr;
    WB_CUDA_OK(cudaMalloc((void **) &d_rows, h.size() * sizeof(CrossKvRow)));
    bool ok = cudaMemcpy(d_rows, h.data(), h.size() * sizeof(CrossKvRow), cudaMemcpyHostToDevice) == cudaSuccess;
    ok = ok && launch_cross_test(q_f16, d_rows, out, n_rows, H, T, 0, 0);
    ok = ok && cudaDeviceSynchronize() == cudaSuccess;
    cudaFree(d_rows);
    return ok;
}

// captures `issue` (launches on stream 0's replacement `s`) into a graph, replays it `iters` times between two events
template <typename F>
static bool time_graph(F issue, int iters, float * ms_out) {
    cudaStream_t s = nullptr;
    WB_CUDA_OK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    cudaGraph_t g = nullptr;
    cudaGraphExec_t ge = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    bool ok = issue(s) && cudaStreamSynchronize(s) == cudaSuccess;  // warm-up, live
    prof_suspend(true);
    ok = ok && cudaStreamBeginCapture(s, cudaStreamCaptureModeRelaxed) == cudaSuccess;
    ok = ok && issue(s);
    ok = (cudaStreamEndCapture(s, &g) == cudaSuccess) && ok && g != nullptr;
    prof_suspend(false);
    ok = ok && cudaGraphInstantiate(&ge, g, 0) == cudaSuccess;
    ok = ok && cudaEventCreate(&e0) == cudaSuccess && cudaEventCreate(&e1) == cudaSuccess;
    if (ok) {
        ok = cudaGraphLaunch(ge, s) == cudaSuccess;
        cudaEventRecord(e0, s);
        for (int i = 0; i < iters && ok; ++i) ok = cudaGraphLaunch(ge, s) == cudaSuccess;
        cudaEventRecord(e1, s);
        ok = ok && cudaEventSynchronize(e1) == cudaSuccess;
        if (ok) cudaEventElapsedTime(ms_out, e0, e1);
    }
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    if (ge) cudaGraphExecDestroy(ge);
    if (g) cudaGraphDestroy(g);
    cudaStreamDestroy(s);
    return ok;
}

// n_rows sequences with their own [layers][2H][T][64] cross caches (so the working set is n_rows*layers*2*H*T*128 B,
// far beyond L2 for the bench shapes); one launch per layer, `iters` sweeps, timed with events on the launching stream.
bool bench_cross_attn_step(int n_rows, int H, int T, int layers, int iters, StepBenchResult * out) {
    const size_t per_row = (size_t) layers * 2 * H * T * 64;
    __half *kv = nullptr, *q = nullptr, *o = nullptr;
    CrossKvRow * d_rows = nullptr;
    bool ok = cudaMalloc((void **) &kv, per_row * n_rows * 2) == cudaSuccess && cudaMalloc((void **) &q, (size_t) n_rows * H * 64 * 2) == cudaSuccess &&
              cudaMalloc((void **) &o, (size_t) n_rows * H * 64 * 2) == cudaSuccess &&
              cudaMalloc((void **) &d_rows, (size_t) n_rows * sizeof(CrossKvRow)) == cudaSuccess;
    if (ok) {
        cudaMemset(kv, 0x11, per_row * n_rows * 2);
        cudaMemset(q, 0x11, (size_t) n_rows * H * 64 * 2);
        std::vector<CrossKvRow> h((size_t) n_rows);
        for (int i = 0; i < n_rows; ++i) h[(size_t) i] = CrossKvRow{kv + per_row * i, nullptr, 0, 0};
        ok = cudaMemcpy(d_rows, h.data(), h.size() * sizeof(CrossKvRow), cudaMemcpyHostToDevice) == cudaSuccess;
    }
    if (ok) {
        float ms = 0.0f;
        ok = time_graph([&](cudaStream_t st) {
            bool k = true;
            for (int l = 0; l < layers && k; ++l) k = launch_cross_test(q, d_rows, o, n_rows, H, T, l, st);
            return k;
        }, iters, &ms);
        out->launches = iters * layers;
        out->us_per_launch = 1e3 * ms / std::max(1, out->launches);
        out->bytes_per_launch = (double) n_rows * H * T * 64 * 2 * 2;
    }
    cudaFree(kv); cudaFree(q); cudaFree(o); cudaFree(d_rows);
    return ok;
}

// `n_weights` distinct [N_out][K] matrices are streamed round-robin (choose n_weights*N_out*K*2 B > L2)
bool bench_gemm_wstream(int n_rows, int N_out, int K, int n_weights, int iters, StepBenchResult * out) {
    g_pdl = getenv("WB200_NO_PDL") == nullptr;
    __half *w = nullptr, *x = nullptr;
    float * part = nullptr;
    const size_t wsz = (size_t) N_out * K;
    const WsPlan p = ws_plan(n_rows, N_out, K, 0);
    bool ok = cudaMalloc((void **) &w, wsz * n_weights * 2) == cudaSuccess && cudaMalloc((void **) &x, (size_t) n_rows * K * 2) == cudaSuccess &&
              cudaMalloc((void **) &part, (size_t) p.split * n_rows * N_out * 4) == cudaSuccess;
    if (ok) {
        cudaMemset(w, 0x11, wsz * n_weights * 2);
        cudaMemset(x, 0x11, (size_t) n_rows * K * 2);
    }
    WsOut po{};
    po.partial = part;
    if (ok) {
        float ms = 0.0f;
        ok = time_graph([&](cudaStream_t st) {
            bool k = true;
            for (int i = 0; i < n_weights && k; ++i) k = gemm_wstream(w + wsz * i, x, n_rows, N_out, K, po, 0, nullptr, st, i > 0);
            return k;
        }, iters, &ms);
        out->launches = iters * n_weights;
        out->us_per_launch = 1e3 * ms / std::max(1, out->launches);
        out->bytes_per_launch = (double) wsz * 2 + (double) n_rows * K * 2 + (double) p.split * n_rows * N_out * 4;
    }
    cudaFree(w); cudaFree(x); cudaFree(part);
    return ok;
}
// cluster split-K GEMM on raw device pointers.  mode 0: out f16 [n][N_out] = f16((x.w^T + bias) * (f < scale_f1 ? scale : 1));
// mode 1: out f32 [n][N_out] += x.w^T + bias (in place);  mode 2: f16 GELU through `gelu_table` (65,536 f16 entries, device)
bool test_gemm_wcluster(const __half * w, const __half * x, void * out, int n, int N_out, int K, const float * bias, int mode, float scale,
                        int scale_f1, const __half * gelu_table) {
    g_pdl = false;
    WcOut o{};
    o.mode = mode; o.out = out; o.ldo = N_out; o.bias = bias; o.scale = scale; o.scale_f1 = scale_f1; o.gelu_table = gelu_table;
    if (!gemm_wcluster(w, x, n, N_out, K, o, 0, false)) return false;
    WB_CUDA_OK(cudaDeviceSynchronize());
    return true;
}

// Concurrency probe for the decode lanes: a chain of cluster GEMMs (n_weights distinct matrices, PDL-chained, top launch
// priority) timed alone and then while another stream keeps the persistent cross attention streaming K/V.
// out->us_per_launch: per GEMM with the background; out->bytes_per_launch: per GEMM alone (us); out->launches: background
// cross-attention launch time with the foreground running, in ns.
bool bench_overlap_gemm(int n_rows, int N_out, int K, int n_weights, int iters, StepBenchResult * out) {
    g_pdl = getenv("WB200_NO_PDL") == nullptr;
    const int H = 20, T = 1500, layers = 8;
    const size_t per_row = (size_t) layers * 2 * H * T * 64;
    __half *kv = nullptr, *q = nullptr, *o = nullptr, *w = nullptr, *x = nullptr, *y = nullptr;
    CrossKvRow * d_rows = nullptr;
    const size_t wsz = (size_t) N_out * K;
    bool ok = cudaMalloc((void **) &kv, per_row * n_rows * 2) == cudaSuccess && cudaMalloc((void **) &q, (size_t) n_rows * H * 64 * 2) == cudaSuccess &&
              cudaMalloc((void **) &o, (size_t) n_rows * H * 64 * 2) == cudaSuccess &&
              cudaMalloc((void **) &d_rows, (size_t) n_rows * sizeof(CrossKvRow)) == cudaSuccess &&
              cudaMalloc((void **) &w, wsz * n_weights * 2) == cudaSuccess && cudaMalloc((void **) &x, (size_t) n_rows * K * 2) == cudaSuccess &&
              cudaMalloc((void **) &y, (size_t) n_rows * N_out * 2) == cudaSuccess;
    if (ok) {
        cudaMemset(kv, 0x11, per_row * n_rows * 2);
        cudaMemset(q, 0x11, (size_t) n_rows * H * 64 * 2);
        cudaMemset(w, 0x11, wsz * n_weights * 2);
        cudaMemset(x, 0x11, (size_t) n_rows * K * 2);
        std::vector<CrossKvRow> h((size_t) n_rows);
        for (int i = 0; i < n_rows; ++i) h[(size_t) i] = CrossKvRow{kv + per_row * i, nullptr, 0, 0};
        ok = cudaMemcpy(d_rows, h.data(), h.size() * sizeof(CrossKvRow), cudaMemcpyHostToDevice) == cudaSuccess;
    }
    cudaStream_t sa = nullptr, sb = nullptr;
    cudaGraph_t ga = nullptr, gb = nullptr;
    cudaGraphExec_t ea = nullptr, eb = nullptr;
    cudaEvent_t ev[4] = {};
    ok = ok && cudaStreamCreateWithFlags(&sa, cudaStreamNonBlocking) == cudaSuccess && cudaStreamCreateWithFlags(&sb, cudaStreamNonBlocking) == cudaSuccess;
    for (auto & e : ev) ok = ok && cudaEventCreate(&e) == cudaSuccess;
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    WcOut wo{};
    wo.mode = WC_F16; wo.out = y; wo.ldo = N_out;
    auto issue_bg = [&](cudaStream_t st) {
        bool k = true;
        for (int l = 0; l < layers && k; ++l) k = launch_cross_test(q, d_rows, o, n_rows, H, T, l, st);
        return k;
    };
    auto issue_fg = [&](cudaStream_t st) {
        bool k = true;
        g_tune.wc_smem_cap = env_int("WB200_WC_SMEM_KB", 192) * 1024;
        g_launch_prio = getenv("WB200_NO_PRIO") ? 0 : hi;
        for (int i = 0; i < n_weights && k; ++i) k = gemm_wcluster(w + wsz * i, x, n_rows, N_out, K, wo, st, i > 0);
        g_launch_prio = 0;
        g_tune = LaneTuning();
        return k;
    };
    ok = ok && issue_bg(sa) && issue_fg(sb) && cudaDeviceSynchronize() == cudaSuccess;
    prof_suspend(true);
    ok = ok && cudaStreamBeginCapture(sa, cudaStreamCaptureModeRelaxed) == cudaSuccess && issue_bg(sa);
    ok = (cudaStreamEndCapture(sa, &ga) == cudaSuccess) && ok && ga;
    ok = ok && cudaStreamBeginCapture(sb, cudaStreamCaptureModeRelaxed) == cudaSuccess && issue_fg(sb);
    ok = (cudaStreamEndCapture(sb, &gb) == cudaSuccess) && ok && gb;
    prof_suspend(false);
    ok = ok && cudaGraphInstantiate(&ea, ga, 0) == cudaSuccess && cudaGraphInstantiate(&eb, gb, 0) == cudaSuccess;
    if (ok) {
        float ms_alone = 0, ms_with = 0, ms_bg = 0;
        cudaGraphLaunch(eb, sb);
        cudaEventRecord(ev[0], sb);
        for (int i = 0; i < iters; ++i) cudaGraphLaunch(eb, sb);
        cudaEventRecord(ev[1], sb);
        cudaStreamSynchronize(sb);
        cudaEventElapsedTime(&ms_alone, ev[0], ev[1]);
        // background: enough replays to outlast the foreground by a wide margin
        const int bg_reps = std::max(4, (int) (6.0f * ms_alone / (0.15f * layers * n_rows / 128.0f)) + 4);
        cudaGraphLaunch(ea, sa);
        cudaEventRecord(ev[2], sa);
        for (int i = 0; i < bg_reps; ++i) cudaGraphLaunch(ea, sa);
        cudaEventRecord(ev[3], sa);
        cudaEventRecord(ev[0], sb);
        for (int i = 0; i < iters; ++i) cudaGraphLaunch(eb, sb);
        cudaEventRecord(ev[1], sb);
        ok = cudaDeviceSynchronize() == cudaSuccess;
        cudaEventElapsedTime(&ms_with, ev[0], ev[1]);
        cudaEventElapsedTime(&ms_bg, ev[2], ev[3]);
        out->us_per_launch = 1e3 * ms_with / (iters * n_weights);
        out->bytes_per_launch = 1e3 * ms_alone / (iters * n_weights);
        out->launches = (int) (1e6 * ms_bg / (bg_reps * layers));
    }
    for (auto & e : ev) if (e) cudaEventDestroy(e);
    if (ea) cudaGraphExecDestroy(ea);
    if (eb) cudaGraphExecDestroy(eb);
    if (ga) cudaGraphDestroy(ga);
    if (gb) cudaGraphDestroy(gb);
    if (sa) cudaStreamDestroy(sa);
    if (sb) cudaStreamDestroy(sb);
    cudaFree(kv); cudaFree(q); cudaFree(o); cudaFree(d_rows); cudaFree(w); cudaFree(x); cudaFree(y);
    return ok;
}

bool bench_gemm_wcluster(int n_rows, int N_out, int K, int n_weights, int iters, StepBenchResult * out) {
    g_pdl = getenv("WB200_NO_PDL") == nullptr;
    __half *w = nullptr, *x = nullptr, *y = nullptr;
    const size_t wsz = (size_t) N_out * K;
    bool ok = cudaMalloc((void **) &w, wsz * n_weights * 2) == cudaSuccess && cudaMalloc((void **) &x, (size_t) n_rows * K * 2) == cudaSuccess &&
              cudaMalloc((void **) &y, (size_t) n_rows * N_out * 2) == cudaSuccess;
    if (ok) {
        cudaMemset(w, 0x11, wsz * n_weights * 2);
        cudaMemset(x, 0x11, (size_t) n_rows * K * 2);
    }
    WcOut o{};
    o.mode = WC_F16; o.out = y; o.ldo = N_out;
    if (ok) {
        float ms = 0.0f;
        ok = time_graph([&](cudaStream_t st) {
            bool k = true;
            for (int i = 0; i < n_weights && k; ++i) k = gemm_wcluster(w + wsz * i, x, n_rows, N_out, K, o, st, i > 0);
            return k;
        }, iters, &ms);
        out->launches = iters * n_weights;
        out->us_per_launch = 1e3 * ms / std::max(1, out->launches);
        out->bytes_per_launch = (double) wsz * 2 + (double) n_rows * K * 2 + (double) n_rows * N_out * 2;
    }
    cudaFree(w); cudaFree(x); cudaFree(y);
    return ok;
}

// =============================================================================================
// gemm_wstream without any allocation (whisper_b200_gemm_wstream_noalloc, include/whisper_b200.h)
// =============================================================================================
// The split-K partial sums of gemm_wstream (mode 0) normally live in a cudaMalloc'ed workspace.  Callers that may not allocate get the
// same GEMM with the partials in a fixed array that is part of the loaded module: every GPU that loads the library holds
// WS_NOALLOC_MAX_FLOATS * 4 bytes of it until the process exits.  All calls on a device share the array.
constexpr int64_t WS_NOALLOC_MAX_FLOATS = 2 * 1024 * 1024;  // 8 MiB: split * n * N_out f32 partials
static __device__ __align__(256) float g_ws_noalloc_partial[WS_NOALLOC_MAX_FLOATS];

// out[t][f] = ((p[0] + p[1]) + ... + p[split-1]) + bias[f]: the summation order of the allocating test entry (test_gemm_wstream)
__global__ void __launch_bounds__(256)
ws_noalloc_reduce_kernel(const float * __restrict__ partial, int split, int64_t split_stride, const float * __restrict__ bias, int N_out,
                         float * __restrict__ out) {
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < split_stride; i += (int64_t) gridDim.x * blockDim.x) {
        float acc = partial[i];
        for (int s = 1; s < split; ++s) acc = __fadd_rn(acc, partial[(int64_t) s * split_stride + i]);
        if (bias != nullptr) acc = __fadd_rn(acc, __ldg(bias + (int) (i % N_out)));
        out[i] = acc;
    }
}

// The device address of the array, looked up once per device with that device current.  In host code `g_ws_noalloc_partial` is only
// the host-side handle of the symbol; kernels must be given the address cudaGetSymbolAddress returns.
static float * ws_noalloc_scratch() {
    static std::mutex mu;
    static float * addr[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    std::lock_guard<std::mutex> lk(mu);
    if (addr[dev] == nullptr) {
        void * p = nullptr;
        const cudaError_t err = cudaGetSymbolAddress(&p, g_ws_noalloc_partial);
        if (err != cudaSuccess) {
            log_error("gemm_wstream_noalloc: cudaGetSymbolAddress failed on device %d: %d (%s)\n", dev, (int) err, cudaGetErrorString(err));
            return nullptr;
        }
        addr[dev] = (float *) p;
    }
    return addr[dev];
}

// 0 = done (output complete on return), 1 = CUDA failure or bad argument, 2 = the partials do not fit the fixed array
int gemm_wstream_noalloc(const __half * w, const __half * x, float * out, int n, int N_out, int K, const float * bias, int force_split) {
    if (w == nullptr || x == nullptr || out == nullptr || n < 1 || n > 256 || N_out < 1 || K < 1) {
        log_error("gemm_wstream_noalloc: bad arguments (n %d in 1..256, n_out %d, k %d, non-null w / x / out)\n", n, N_out, K);
        return 1;
    }
    const WsPlan p = ws_plan(n, N_out, K, force_split);
    const int64_t need = (int64_t) p.split * n * N_out;
    if (need > WS_NOALLOC_MAX_FLOATS) {
        log_error("gemm_wstream_noalloc: %d splits x %d rows x %d outputs = %lld partial sums exceed the fixed scratch of %lld floats (%lld MiB); "
                  "use fewer rows, fewer splits, or the allocating entry point\n",
                  p.split, n, N_out, (long long) need, (long long) WS_NOALLOC_MAX_FLOATS, (long long) (WS_NOALLOC_MAX_FLOATS * 4 >> 20));
        return 2;
    }
    float * part = ws_noalloc_scratch();
    if (part == nullptr) return 1;
    const cudaStream_t s = 0;
    prof_suspend(true);  // the in-situ profiler would create events for a sampled launch
    bool ok = cudaMemsetAsync(part, 0, (size_t) need * sizeof(float), s) == cudaSuccess;  // the array keeps the previous call's sums
    WsOut po{};
    po.partial = part;
    int split = 0;
    ok = ok && gemm_wstream(w, x, n, N_out, K, po, force_split, &split, s, false);
    if (ok) {
        const int64_t total = (int64_t) n * N_out;
        const int blocks = (int) std::min<int64_t>(ceil_div64(total, 256), 148 * 8);
        ok = launch_k(ws_noalloc_reduce_kernel, dim3(blocks), dim3(256), 0, s, false, (const float *) part, split, total, bias, N_out, out);
    }
    prof_suspend(false);
    const cudaError_t err = cudaStreamSynchronize(s);
    if (err != cudaSuccess) {
        log_error("gemm_wstream_noalloc: CUDA error %d (%s)\n", (int) err, cudaGetErrorString(err));
        return 1;
    }
    return ok ? 0 : 1;
}
}  // namespace wb
