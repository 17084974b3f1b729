// Non-GEMM kernels of the transcription path (sm_100a).  See kernels.cuh.
//
// All of these are HBM- or latency-bound: they use coalesced 16-byte loads, shared-memory staging and
// warp-shuffle reductions, not tensor cores.  Rounding points follow the reference's ggml CPU path
//  so that greedy token ids match it.
#include "kernels.cuh"
#include "prof.h"

#include <algorithm>
#include <atomic>
#include <math.h>

namespace wb {

static std::atomic<uint64_t> g_launches{0};
uint64_t kernel_launch_count() { return g_launches.load(); }
#define WB_LAUNCHED()                      \
    do {                                   \
        WB_CUDA_OK(cudaGetLastError());    \
        g_launches.fetch_add(1);           \
    } while (0)

// =============================================================================================
// M1  log-mel spectrogram
// =============================================================================================
// One CTA computes MEL_FPB consecutive frames.  Per frame: Hann window -> 400-point real FFT with the
// reference's decimation order (four radix-2 levels on top of sixteen 25-point DFTs, float sin/cos table:
// whisper.cpp:3054-3118) -> power -> mel filterbank dot (float products summed four at a time, double
// accumulate: whisper.cpp:3156-3172) -> log10 clamp.
static constexpr int MEL_FPB     = 8;
static constexpr int MEL_THREADS = 128;
static constexpr int N_FFT       = 400;
static constexpr int N_BINS      = 201;
static constexpr int HOP         = 160;

__device__ __forceinline__ float padded_sample(const float * __restrict__ s, int n_samples, int idx) {
    // samples_padded of whisper.cpp:3209-3217: 200 reflected samples, the clip, then zeros
    if (idx < 200) {
        const int j = 200 - idx;  // reverse_copy(samples + 1, samples + 201)
        return j < n_samples ? __ldg(s + j) : 0.0f;
    }
    const int j = idx - 200;
    return j < n_samples ? __ldg(s + j) : 0.0f;
}

__global__ void __launch_bounds__(MEL_THREADS)
mel_frames_kernel(MelTables tab, const float * __restrict__ samples, int n_samples, int n_valid, int n_len, float * __restrict__ mel,
                  int * __restrict__ gmax) {
    __shared__ float s_in[N_FFT];
    __shared__ float2 s_a[N_FFT];
    __shared__ float2 s_b[N_FFT];
    __shared__ float s_pow[N_BINS + 3];
    __shared__ float s_cos[N_FFT];
    __shared__ float s_sin[N_FFT];
    __shared__ float s_out[128 /*n_mel max*/][MEL_FPB];

    const int tid = threadIdx.x;
    for (int i = tid; i < N_FFT; i += MEL_THREADS) {
        s_cos[i] = tab.cos_tab[i];
        s_sin[i] = tab.sin_tab[i];
    }
    const int f0 = blockIdx.x * MEL_FPB;
    float local_max = -1e30f;

    for (int f = 0; f < MEL_FPB; ++f) {
        const int frame = f0 + f;
        if (frame >= n_valid) break;  // uniform
        const int offset = frame * HOP;
        __syncthreads();
        for (int j = tid; j < N_FFT; j += MEL_THREADS) {
            s_in[j] = __fmul_rn(__ldg(tab.hann + j), padded_sample(samples, n_samples, offset + j));
        }
        __syncthreads();
        // sixteen 25-point DFTs over x_o[n] = in[16 n + o]  (stride-16 decimation leaves)
        for (int idx = tid; idx < N_FFT; idx += MEL_THREADS) {
            const int o = idx / 25, k = idx - o * 25;
            float re = 0.0f, im = 0.0f;
#pragma unroll 5
            for (int n = 0; n < 25; ++n) {
                const int ti = (k * n * 16) % N_FFT;
                const float v = s_in[16 * n + o];
                re += v * s_cos[ti];
                im -= v * s_sin[ti];
            }
            s_a[o * 25 + k] = make_float2(re, im);
        }
        __syncthreads();
        // radix-2 combines: stride s = 8, 4, 2, 1 ; sub-transform length N = 400 / s
        float2 * src = s_a;
        float2 * dst = s_b;
#pragma unroll 1
        for (int s = 8; s >= 1; s >>= 1) {
            const int N = N_FFT / s, half = N >> 1;
            for (int idx = tid; idx < s * half; idx += MEL_THREADS) {
                const int o = idx / half, k = idx - o * half;
                const float2 e  = src[o * half + k];        // sub-sequence (offset o,     stride 2s)
                const float2 od = src[(o + s) * half + k];  // sub-sequence (offset o + s, stride 2s)
                const float wr = s_cos[k * s];
                const float wi = -s_sin[k * s];
                float2 lo, hi;
                lo.x = e.x + wr * od.x - wi * od.y;
                lo.y = e.y + wr * od.y + wi * od.x;
                hi.x = e.x - wr * od.x + wi * od.y;
                hi.y = e.y - wr * od.y - wi * od.x;
                dst[o * N + k]        = lo;
                dst[o * N + k + half] = hi;
            }
            __syncthreads();
            float2 * t = src;
            src = dst;
            dst = t;
        }
        for (int j = tid; j < N_BINS; j += MEL_THREADS) {
            const float2 c = src[j];
            s_pow[j] = c.x * c.x + c.y * c.y;
        }
        __syncthreads();
        for (int j = tid; j < tab.n_mel; j += MEL_THREADS) {
            double sum = 0.0;
            int k = 0;
            for (; k < N_BINS - 3; k += 4) {
                const float p4 = s_pow[k + 0] * __ldg(tab.filters_t + (k + 0) * tab.n_mel + j) +
                                 s_pow[k + 1] * __ldg(tab.filters_t + (k + 1) * tab.n_mel + j) +
                                 s_pow[k + 2] * __ldg(tab.filters_t + (k + 2) * tab.n_mel + j) +
                                 s_pow[k + 3] * __ldg(tab.filters_t + (k + 3) * tab.n_mel + j);
                sum += (double) p4;
            }
            for (; k < N_BINS; ++k) sum += (double) (s_pow[k] * __ldg(tab.filters_t + k * tab.n_mel + j));
            sum = log10(fmax(sum, 1e-10));
            const float v = (float) sum;
            s_out[j][f] = v;
            local_max = fmaxf(local_max, v);
        }
    }
    __syncthreads();
    const int nf = min(MEL_FPB, n_valid - f0);
    for (int idx = tid; idx < tab.n_mel * MEL_FPB; idx += MEL_THREADS) {
        const int j = idx / MEL_FPB, f = idx - j * MEL_FPB;
        if (f < nf) mel[(int64_t) j * n_len + f0 + f] = s_out[j][f];
    }
    // global max (order-preserving int encoding of float)
    local_max = warp_max(local_max);
    if ((tid & 31) == 0 && local_max > -1e30f) {
        int bits = __float_as_int(local_max);
        bits = bits >= 0 ? bits : (bits ^ 0x7fffffff);
        atomicMax(gmax, bits);
    }
}

__global__ void mel_init_max_kernel(int * gmax, int n_valid, int n_len) {
    // frames past the signal hold log10(1e-10) = -10 (whisper.cpp:3177-3182); they take part in the max
    float v = n_valid < n_len ? -10.0f : -1e30f;
    int bits = __float_as_int(v);
    bits = bits >= 0 ? bits : (bits ^ 0x7fffffff);
    *gmax = bits;
}

__global__ void mel_normalize_kernel(float * __restrict__ mel, int n_mel, int n_len, int n_valid, const int * __restrict__ gmax) {
    // whisper.cpp:3245-3260 (double mmax; (x + 4.0)/4.0 in double, stored as float)
    int bits = *gmax;
    bits = bits >= 0 ? bits : (bits ^ 0x7fffffff);
    const double mmax = (double) __int_as_float(bits) - 8.0;
    const int64_t total = (int64_t) n_mel * n_len;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t) gridDim.x * blockDim.x) {
        const int fr = (int) (i % n_len);
        float v = fr < n_valid ? mel[i] : -10.0f;
        if ((double) v < mmax) v = (float) mmax;
        mel[i] = (float) (((double) v + 4.0) / 4.0);
    }
}

bool mel_compute(const MelTables & tab, const float * samples, int n_samples, int n_len, float * mel_out, int * scratch_max,
                 cudaStream_t stream) {
    ProfScope prof_scope(PC_MEL, stream, (double) n_samples * 4 + (double) tab.n_mel * n_len * 4);
    WB_CHECK(tab.n_mel <= 128, "n_mel > 128 unsupported\n");
    const int n_eff   = n_samples + 200;
    const int n_valid = min(n_eff / HOP + 1, n_len);  // whisper.cpp:3133
    mel_init_max_kernel<<<1, 1, 0, stream>>>(scratch_max, n_valid, n_len);
    WB_LAUNCHED();
    if (n_valid > 0) {
        mel_frames_kernel<<<ceil_div(n_valid, MEL_FPB), MEL_THREADS, 0, stream>>>(tab, samples, n_samples, n_valid, n_len, mel_out,
                                                                                 scratch_max);
        WB_LAUNCHED();
    }
    const int64_t total = (int64_t) tab.n_mel * n_len;
    int64_t blocks64 = ceil_div64(total, 256);
    const int blocks = (int) (blocks64 < 148 * 8 ? blocks64 : 148 * 8);
    mel_normalize_kernel<<<blocks, 256, 0, stream>>>(mel_out, tab.n_mel, n_len, n_valid, scratch_max);
    WB_LAUNCHED();
    return true;
}

// mel[:, offset : offset + n_frames] (zero beyond n_len) -> f16 panel rows 1..n_frames, kc columns (zero padded),
// rows 0 and n_frames+1 are the conv padding.  A 32x32 shared-memory transpose keeps both sides coalesced.
__global__ void mel_window_kernel(const float * __restrict__ mel, int n_mel, int n_len, int offset, int n_frames,
                                  __half * __restrict__ out, int kc) {
    __shared__ float tile[32][33];
    const int t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;  // 32 x 8
    for (int r = ty; r < 32; r += 8) {
        const int c = c0 + r, t = t0 + tx;  // row r of the tile = mel bin c
        float v = 0.0f;
        if (c < n_mel && t < n_frames && offset + t < n_len) v = mel[(int64_t) c * n_len + offset + t];
        tile[r][tx] = v;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int t = t0 + r, c = c0 + tx;
        if (t < n_frames && c < kc) out[(int64_t) (t + 1) * kc + c] = __float2half_rn(tile[tx][r]);
    }
    if (blockIdx.x == 0 && ty == 0) {
        const int c = c0 + tx;
        if (c < kc) {
            out[c] = __float2half_rn(0.0f);
            out[(int64_t) (n_frames + 1) * kc + c] = __float2half_rn(0.0f);
        }
    }
}

bool mel_window_to_f16(const float * mel, int n_mel, int n_len, int offset, int n_frames, __half * out, int kc, cudaStream_t stream) {
    ProfScope prof_scope(PC_MEL, stream);
    dim3 grid(ceil_div(n_frames, 32), ceil_div(kc, 32));
    mel_window_kernel<<<grid, dim3(32, 8), 0, stream>>>(mel, n_mel, n_len, offset, n_frames, out, kc);
    WB_LAUNCHED();
    return true;
}

// =============================================================================================
// LayerNorm rows
// =============================================================================================
// one warp per row; mean / variance accumulated in double like ggml_compute_forward_norm_f32
__global__ void __launch_bounds__(256)
layernorm_kernel(const float * __restrict__ x, int rows, int S, const float * __restrict__ gamma, const float * __restrict__ beta,
                 float eps, __half * __restrict__ o16, float * __restrict__ o32) {
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    const float4 * xr = (const float4 *) (x + (int64_t) row * S);
    const int n4 = S >> 2;
    double sum = 0.0;
    for (int i = lane; i < n4; i += 32) {
        const float4 v = xr[i];
        sum += (double) v.x + (double) v.y + (double) v.z + (double) v.w;
    }
    sum = warp_sum_d(sum);
    const float mean = (float) (sum / S);
    double sum2 = 0.0;
    for (int i = lane; i < n4; i += 32) {
        const float4 v = xr[i];
        const float a = __fsub_rn(v.x, mean), b = __fsub_rn(v.y, mean), c = __fsub_rn(v.z, mean), d = __fsub_rn(v.w, mean);
        sum2 += (double) __fmul_rn(a, a) + (double) __fmul_rn(b, b) + (double) __fmul_rn(c, c) + (double) __fmul_rn(d, d);
    }
    sum2 = warp_sum_d(sum2);
    const float variance = (float) (sum2 / S);
    const float scale = __frcp_rn(__fsqrt_rn(__fadd_rn(variance, eps)));
    const float4 * g4 = (const float4 *) gamma;
    const float4 * b4 = (const float4 *) beta;
    for (int i = lane; i < n4; i += 32) {
        const float4 v = xr[i], g = __ldg(g4 + i), b = __ldg(b4 + i);
        float4 y;
        y.x = __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(v.x, mean), scale), g.x), b.x);
        y.y = __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(v.y, mean), scale), g.y), b.y);
        y.z = __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(v.z, mean), scale), g.z), b.z);
        y.w = __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(v.w, mean), scale), g.w), b.w);
        if (o16) {
            __half2 h0 = __floats2half2_rn(y.x, y.y), h1 = __floats2half2_rn(y.z, y.w);
            uint2 pk;
            pk.x = *(uint32_t *) &h0;
            pk.y = *(uint32_t *) &h1;
            *(uint2 *) (o16 + (int64_t) row * S + i * 4) = pk;
        }
        if (o32) *(float4 *) (o32 + (int64_t) row * S + i * 4) = y;
    }
}

bool layernorm_rows(const float * x, int rows, int S, const float * gamma, const float * beta, float eps, __half * out_f16,
                    float * out_f32, cudaStream_t stream) {
    ProfScope prof_scope(PC_LN, stream, (double) rows * S * (4 + (out_f16 ? 2 : 0) + (out_f32 ? 4 : 0)));
    WB_CHECK((S & 3) == 0, "S must be a multiple of 4\n");
    if (rows <= 0) return true;
    layernorm_kernel<<<ceil_div(rows, 8), 256, 0, stream>>>(x, rows, S, gamma, beta, eps, out_f16, out_f32);
    WB_LAUNCHED();
    return true;
}

// =============================================================================================
// row softmax (encoder attention scores)  -> f16 probabilities
// =============================================================================================
__global__ void __launch_bounds__(256)
softmax_rows_kernel(const float * __restrict__ s, int64_t rows, int n, int64_t lds, float scale, __half * __restrict__ p, int64_t ldp) {
    const int64_t row = (int64_t) blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    const float * sr = s + row * lds;
    float mx = -INFINITY;
    for (int i = lane; i < n; i += 32) mx = fmaxf(mx, __fmul_rn(sr[i], scale));
    mx = warp_max(mx);
    double sum = 0.0;
    for (int i = lane; i < n; i += 32) sum += (double) expf(__fsub_rn(__fmul_rn(sr[i], scale), mx));
    sum = warp_sum_d(sum);
    const float inv = (float) (1.0 / sum);
    __half * pr = p + row * ldp;
    for (int i = lane; i < n; i += 32) pr[i] = __float2half_rn(__fmul_rn(expf(__fsub_rn(__fmul_rn(sr[i], scale), mx)), inv));
    // zero the padding so the P.V GEMM's k-tail reads zeros even without TMA bounds
    for (int i = n + lane; i < ldp; i += 32) pr[i] = __float2half_rn(0.0f);
}

bool softmax_rows_f16(const float * s, int64_t rows, int n, int64_t lds, float scale, __half * p, int64_t ldp, cudaStream_t stream) {
    ProfScope prof_scope(PC_ENC_SOFTMAX, stream, (double) rows * n * 6);
    if (rows <= 0) return true;
    softmax_rows_kernel<<<(unsigned) ceil_div64(rows, 8), 256, 0, stream>>>(s, rows, n, lds, scale, p, ldp);
    WB_LAUNCHED();
    return true;
}

// =============================================================================================
// GELU through the reference's 65,536-entry f16 table (ggml-cpu/vec.h:571-585), in place on an f16 tensor.
// The fc1 GEMM stores h = f16(W1.x + b1) - the value the reference's lookup is indexed by - and this pass maps h -> table[h].
// The table sits in shared memory (128 KB): a warp's 32 random lookups cost ~3-4 bank cycles instead of up to 32 L1 wavefronts.
// =============================================================================================
__device__ __forceinline__ uint4 gelu_lookup8(uint4 v, const uint16_t * tab) {
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        uint32_t lo = w[k] & 0xffffu, hi = w[k] >> 16;
        // x <= -10 -> 0, x >= 10 -> x (vec.h:575-580); f16 bit patterns: 10.0 = 0x4900, -10.0 = 0xC900
        const uint32_t alo = lo & 0x7fffu, ahi = hi & 0x7fffu;
        uint32_t rlo = tab[lo], rhi = tab[hi];
        if (alo >= 0x4900u && alo <= 0x7c00u) rlo = (lo & 0x8000u) ? 0u : lo;
        if (ahi >= 0x4900u && ahi <= 0x7c00u) rhi = (hi & 0x8000u) ? 0u : hi;
        w[k] = rlo | (rhi << 16);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

// One persistent CTA per SM (the table takes 128 KB of its shared memory); 1024 threads x 4 independent 16-byte loads keep
// 64 KB per SM in flight, which is what an HBM-bound pass needs on this part.
__global__ void __launch_bounds__(1024, 1)
gelu_table_kernel(__half * __restrict__ x, int64_t n8 /* groups of 8 halves */, const __half * __restrict__ table) {
    extern __shared__ uint4 gelu_tab_raw[];
    uint16_t * tab = (uint16_t *) gelu_tab_raw;
    const uint4 * tsrc = (const uint4 *) table;
    for (int i = threadIdx.x; i < 65536 * 2 / 16; i += blockDim.x) gelu_tab_raw[i] = __ldg(tsrc + i);
    __syncthreads();
    uint4 * xv = (uint4 *) x;
    const int64_t stride = (int64_t) gridDim.x * blockDim.x;
    int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n8; i += 4 * stride) {
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = xv[i + u * stride];
#pragma unroll
        for (int u = 0; u < 4; ++u) xv[i + u * stride] = gelu_lookup8(v[u], tab);
    }
    for (; i < n8; i += stride) xv[i] = gelu_lookup8(xv[i], tab);
}

bool gelu_table_inplace(__half * x, int64_t n, const __half * table, cudaStream_t stream) {
    static PerDeviceOnce attr_set;
    if (attr_set.first()) {
        WB_CUDA_OK(cudaFuncSetAttribute(gelu_table_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536 * 2));
    }
    WB_CHECK(n % 8 == 0 && ((uintptr_t) x & 15) == 0, "gelu_table_inplace: %lld elements must be a multiple of 8, 16-byte aligned\n", (long long) n);
    ProfScope prof_scope(PC_OTHER, stream, (double) n * 4);
    const int64_t n8 = n / 8;
    const int grid = (int) std::min<int64_t>(ceil_div64(n8, 1024), 148);
    if (grid <= 0) return true;
    gelu_table_kernel<<<grid, 1024, 65536 * 2, stream>>>(x, n8, table);
    WB_LAUNCHED();
    return true;
}

// =============================================================================================
// F5: moving average of |x| (whisper.cpp:8360-8376), one thread per sample, the reference's summation order
// =============================================================================================
__global__ void signal_energy_kernel(const float * __restrict__ pcm, int n, int hw, float * __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float sum = 0.0f;
    for (int j = -hw; j <= hw; ++j) {
        const int k = i + j;
        if (k >= 0 && k < n) sum = __fadd_rn(sum, fabsf(pcm[k]));
    }
    out[i] = __fdiv_rn(sum, (float) (2 * hw + 1));
}

bool signal_energy(const float * pcm, int n_samples, int half_window, float * out, cudaStream_t stream) {
    if (n_samples <= 0) return true;
    ProfScope prof_scope(PC_OTHER, stream, (double) n_samples * 8);
    signal_energy_kernel<<<ceil_div(n_samples, 256), 256, 0, stream>>>(pcm, n_samples, half_window, out);
    WB_LAUNCHED();
    return true;
}

// =============================================================================================
// decoder step: embedding
// =============================================================================================
__global__ void embed_kernel(const __half * __restrict__ te, const float * __restrict__ pe, const int * __restrict__ tokens,
                             const int * __restrict__ pos, int S, float * __restrict__ x) {
    const int j = blockIdx.x;
    const __half * t = te + (int64_t) tokens[j] * S;
    const float * p  = pe + (int64_t) pos[j] * S;
    for (int i = threadIdx.x; i < S; i += blockDim.x) x[(int64_t) j * S + i] = __fadd_rn(__half2float(t[i]), p[i]);
}

bool embed_tokens(const __half * te, const float * pe, const int * tokens, const int * pos, int n_tokens, int S, float * x,
                  cudaStream_t stream) {
    ProfScope prof_scope(PC_EMBED, stream);
    embed_kernel<<<n_tokens, 256, 0, stream>>>(te, pe, tokens, pos, S, x);
    WB_LAUNCHED();
    return true;
}

// =============================================================================================
// decoder step: self attention over the per-state KV cache
// =============================================================================================
__global__ void self_kv_store_kernel(const __half * __restrict__ qkv, const SelfKvRow * __restrict__ rows, int S, int layer) {
    const int j = blockIdx.x;
    const SelfKvRow r = rows[j];
    const uint4 * k = (const uint4 *) (qkv + (int64_t) j * 3 * S + S);
    const uint4 * v = (const uint4 *) (qkv + (int64_t) j * 3 * S + 2 * S);
    uint4 * kd = (uint4 *) (r.k_base + ((int64_t) layer * r.n_ctx + r.store_cell) * S);
    uint4 * vd = (uint4 *) (r.v_base + ((int64_t) layer * r.n_ctx + r.store_cell) * S);
    for (int i = threadIdx.x; i < S / 8; i += blockDim.x) {
        kd[i] = k[i];
        vd[i] = v[i];
    }
}

bool self_kv_store(const __half * qkv, const SelfKvRow * rows, int n_tokens, int S, int layer, cudaStream_t stream) {
    ProfScope prof_scope(PC_SELF_ATTN, stream);
    self_kv_store_kernel<<<n_tokens, 128, 0, stream>>>(qkv, rows, S, layer);
    WB_LAUNCHED();
    return true;
}

// grid (H, n_tokens), 128 threads.  scores in dynamic shared memory (n_cells floats).
__global__ void __launch_bounds__(128)
self_attn_kernel(const __half * __restrict__ qkv, const SelfKvRow * __restrict__ rows, int S, int layer, __half * __restrict__ out) {
    extern __shared__ float s_sc[];
    __shared__ float s_q[64];
    __shared__ float s_red[4];
    __shared__ double s_redd[4];
    __shared__ float s_acc[2][64];

    const int h = blockIdx.x, j = blockIdx.y, tid = threadIdx.x;
    const SelfKvRow r = rows[j];
    const int n = r.n_cells;
    if (tid < 64) s_q[tid] = __half2float(qkv[(int64_t) j * 3 * S + h * 64 + tid]);
    __syncthreads();

    const __half * kb = r.k_base + (int64_t) layer * r.n_ctx * S + h * 64;
    const __half * vb = r.v_base + (int64_t) layer * r.n_ctx * S + h * 64;

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
    for (int c = tid; c < n; c += 128) s_sc[c] = __half2float(__float2half_rn(__fmul_rn(s_sc[c], inv)));  // P rounded to f16
    __syncthreads();

    // P.V : thread (half, d) accumulates cells of its half
    const int d = tid & 63, half = tid >> 6;
    float acc = 0.0f;
    for (int c = half; c < n; c += 2) {
        const int cell = r.cells ? r.cells[c] : c;
        acc = fmaf(s_sc[c], __half2float(vb[(int64_t) cell * S + d]), acc);
    }
    s_acc[half][d] = acc;
    __syncthreads();
    if (tid < 64) out[(int64_t) j * S + h * 64 + tid] = __float2half_rn(s_acc[0][tid] + s_acc[1][tid]);
}

bool self_attn_decode(const __half * qkv, const SelfKvRow * rows, int n_tokens, int S, int H, int layer, __half * out,
                      cudaStream_t stream) {
    ProfScope prof_scope(PC_SELF_ATTN, stream);
    // scores buffer sized for the largest cache the reference can build: 8+2 decoders x 512 cells (whisper.cpp:7101-7106)
    const int smem = 5120 * (int) sizeof(float);
    self_attn_kernel<<<dim3(H, n_tokens), 128, smem, stream>>>(qkv, rows, S, layer, out);
    WB_LAUNCHED();
    return true;
}

// =============================================================================================
// decoder step: cross attention (the bandwidth-critical kernel of batched decoding)
// =============================================================================================
// grid (H, n_tokens), 256 threads.  K and V of one (state, layer, head) are contiguous [T][64] f16 panels
// (96 KB each at T = 1500... 192 KB for both), streamed once with 16-byte loads.
static constexpr int XA_THREADS = 256;

__global__ void __launch_bounds__(XA_THREADS)
cross_attn_kernel(const __half * __restrict__ q, int ldq, const CrossKvRow * __restrict__ rows, int H, int T, int layer, float scale,
                  const int * __restrict__ ahead_slot, __half * __restrict__ out, int ldo) {
    extern __shared__ float s_sc[];  // [T]
    __shared__ float s_q[64];
    __shared__ float s_red[8];
    __shared__ double s_redd[8];
    __shared__ float s_acc[8][4][64];

    const int h = blockIdx.x, j = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const CrossKvRow r = rows[j];
    if (tid < 64) s_q[tid] = __half2float(q[(int64_t) j * ldq + h * 64 + tid]);
    __syncthreads();

    const __half * kp = r.kv_base + ((int64_t) layer * 2 * H + h) * (int64_t) T * 64;
    const __half * vp = r.kv_base + ((int64_t) layer * 2 * H + H + h) * (int64_t) T * 64;

    // ---- scores: 8 lanes per key (16 bytes each), 4 keys per warp-wide load, 32 keys per CTA iteration
    const int sub = lane & 7, kq = lane >> 3;
    float qf[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) qf[e] = s_q[sub * 8 + e];
    float mx = -INFINITY;
    for (int t0 = warp * 4; t0 < T; t0 += 32) {
        const int t = t0 + kq;
        float acc = 0.0f;
        if (t < T) {
            const uint4 u = *(const uint4 *) (kp + (int64_t) t * 64 + sub * 8);
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
        if (t < T && sub == 0) {
            const float sc = __fmul_rn(acc, scale);  // ggml_soft_max_ext: wp = KQ * scale
            s_sc[t] = sc;
            mx = fmaxf(mx, sc);
        }
    }
    mx = warp_max(mx);
    if (lane == 0) s_red[warp] = mx;
    __syncthreads();
    mx = s_red[0];
#pragma unroll
    for (int w = 1; w < 8; ++w) mx = fmaxf(mx, s_red[w]);

    double sum = 0.0;
    for (int t = tid; t < T; t += XA_THREADS) {
        const float e = expf(__fsub_rn(s_sc[t], mx));
        s_sc[t] = e;
        sum += (double) e;
    }
    sum = warp_sum_d(sum);
    if (lane == 0) s_redd[warp] = sum;
    __syncthreads();
    sum = 0.0;
#pragma unroll
    for (int w = 0; w < 8; ++w) sum += s_redd[w];
    const float inv = (float) (1.0 / sum);

    float * cap = nullptr;
    if (r.qk_out != nullptr && ahead_slot != nullptr) {
        const int slot = ahead_slot[h];
        if (slot >= 0) cap = r.qk_out + ((int64_t) slot * r.n_tok_total + r.tok_index) * T;
    }
    for (int t = tid; t < T; t += XA_THREADS) {
        const float p = __fmul_rn(s_sc[t], inv);
        if (cap) cap[t] = p;                                   // f32 softmax row for DTW (whisper.cpp:2737-2752)
        s_sc[t] = __half2float(__float2half_rn(p));            // P rounded to f16 before P.V
    }
    __syncthreads();

    // ---- P.V: lane owns 8 consecutive d (16 bytes), 4 rows per warp-wide load
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.0f;
    for (int t0 = warp * 4; t0 < T; t0 += 32) {
        const int t = t0 + kq;
        if (t < T) {
            const float p = s_sc[t];
            const uint4 u = *(const uint4 *) (vp + (int64_t) t * 64 + sub * 8);
            const __half2 * hp = (const __half2 *) &u;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 f = __half22float2(hp[e]);
                acc[e * 2]     = fmaf(p, f.x, acc[e * 2]);
                acc[e * 2 + 1] = fmaf(p, f.y, acc[e * 2 + 1]);
            }
        }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) s_acc[warp][kq][sub * 8 + e] = acc[e];
    __syncthreads();
    if (tid < 64) {
        float o = 0.0f;
#pragma unroll
        for (int w = 0; w < 8; ++w)
#pragma unroll
            for (int g = 0; g < 4; ++g) o += s_acc[w][g][tid];
        out[(int64_t) j * ldo + h * 64 + tid] = __float2half_rn(o);
    }
}

bool cross_attn_decode(const __half * q, int ldq, const CrossKvRow * rows, int n_tokens, int H, int T, int layer, float scale,
                       const int * ahead_slot, __half * out, int ldo, cudaStream_t stream) {
    ProfScope prof_scope(PC_CROSS_ATTN, stream, (double) n_tokens * H * T * 64 * 2 * 2 /*K and V panels, f16*/, (double) n_tokens * H * T * 64 * 4);
    cross_attn_kernel<<<dim3(H, n_tokens), XA_THREADS, T * (int) sizeof(float), stream>>>(q, ldq, rows, H, T, layer, scale, ahead_slot,
                                                                                         out, ldo);
    WB_LAUNCHED();
    return true;
}

// =============================================================================================
// S1/S2  logits filter + log-softmax + greedy pick, one CTA per decoder
// =============================================================================================
static constexpr int SM_THREADS = 1024;

__device__ __forceinline__ float block_max(float v, float * red) {
    v = warp_max(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = red[0];
    for (int w = 1; w < SM_THREADS / 32; ++w) r = fmaxf(r, red[w]);
    return r;
}
__device__ __forceinline__ float block_sum(float v, float * red) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = 0.0f;
    for (int w = 0; w < SM_THREADS / 32; ++w) r += red[w];
    return r;
}
__device__ __forceinline__ double block_sum_d(double v, double * red) {
    v = warp_sum_d(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double r = 0.0;
    for (int w = 0; w < SM_THREADS / 32; ++w) r += red[w];
    return r;
}
// first index holding the maximum value: returns (value, index) with lowest index on ties
__device__ __forceinline__ void block_argmax_first(float & v, int & idx, float * redv, int * redi) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, v, o);
        const int oi   = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ov > v || (ov == v && oi < idx)) {
            v = ov;
            idx = oi;
        }
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) {
        redv[threadIdx.x >> 5] = v;
        redi[threadIdx.x >> 5] = idx;
    }
    __syncthreads();
    v = redv[0];
    idx = redi[0];
    for (int w = 1; w < SM_THREADS / 32; ++w) {
        if (redv[w] > v || (redv[w] == v && redi[w] < idx)) {
            v = redv[w];
            idx = redi[w];
        }
    }
}

__global__ void __launch_bounds__(SM_THREADS)
sampler_kernel(SamplerStatic st, const float * __restrict__ logits_all, int64_t ld, const SamplerRow * __restrict__ rows,
               float * __restrict__ work, SamplerOut * __restrict__ out) {
    __shared__ float red[32];
    __shared__ double redd[32];
    __shared__ int redi[32];

    const SamplerRow r = rows[blockIdx.x];
    const int V = st.n_vocab, beg = st.token_beg, tid = threadIdx.x;
    const float * lin = logits_all + (int64_t) r.logits_row * ld;
    float * lg = work + (int64_t) r.work_slot * 3 * V;  // filtered logits
    float * lp = lg + V;                                // logprobs
    float * pr = lp + V;                                // probs
    const uint8_t * __restrict__ sup      = r.suppress ? r.suppress : st.suppress;
    const uint8_t * __restrict__ sup_post = r.suppress_post ? r.suppress_post : st.suppress_post;

    // no-speech probability from the raw logits (whisper.cpp:7126-7134)
    float nosp = 0.0f;
    if (r.want_nosp) {
        float mx = -INFINITY;
        for (int i = tid; i < V; i += SM_THREADS) mx = fmaxf(mx, lin[i]);
        mx = block_max(mx, red);
        float s = 0.0f;
        for (int i = tid; i < V; i += SM_THREADS)
            if (lin[i] > -INFINITY) s += expf(lin[i] - mx);
        s = block_sum(s, red);
        const float lse = logf(s) + mx;
        nosp = expf(lin[st.token_nosp] - lse);
    }

    // ---- filters (whisper.cpp:6183-6302)
    float mx = -INFINITY;
    for (int i = tid; i < V; i += SM_THREADS) {
        float x;
        bool kill;
        if (r.prefiltered) {
            x = lg[i];
            kill = sup_post[i] != 0;
        } else {
            x = lin[i];
            if (r.temperature > 0.0f) x = __fdiv_rn(x, r.temperature);
            kill = sup[i] != 0;
            if (r.suppress_blank && r.is_initial && (i == st.token_eot || i == st.token_space)) kill = true;
            if (r.no_timestamps && i >= beg) kill = true;
        }
        if (r.last_was_ts) {
            if (r.penult_was_ts) kill = kill || (i >= beg);
            else kill = kill || (i < st.token_eot);
        }
        if (r.is_initial && i >= r.max_initial_tid) kill = true;
        if (r.has_ts && i >= beg && i < beg + r.tid0) kill = true;
        if (kill) x = -INFINITY;
        lg[i] = x;
        mx = fmaxf(mx, x);
    }
    mx = block_max(mx, red);
    // ---- log-softmax (whisper_compute_logprobs, whisper.cpp:6109-6129)
    float s = 0.0f;
    for (int i = tid; i < V; i += SM_THREADS) {
        const float x = lg[i];
        if (x > -INFINITY) s += expf(x - mx);
    }
    s = block_sum(s, red);
    const float lse = logf(s) + mx;
    float ts_max = -INFINITY, txt_max = -INFINITY;
    for (int i = tid; i < V; i += SM_THREADS) {
        const float x = lg[i];
        const float l = x > -INFINITY ? x - lse : -INFINITY;
        lp[i] = l;
        if (i >= beg) ts_max = fmaxf(ts_max, l);
        else txt_max = fmaxf(txt_max, l);
    }
    ts_max  = block_max(ts_max, red);
    txt_max = block_max(txt_max, red);
    // ---- timestamp mass vs best text token (whisper.cpp:6309-6333)
    float tsum = 0.0f;
    for (int i = beg + tid; i < V; i += SM_THREADS) {
        const float l = lp[i];
        if (l > -INFINITY) tsum += expf(l - ts_max);
    }
    tsum = block_sum(tsum, red);
    float ts_logprob = -INFINITY;
    if (tsum > 0.0f) ts_logprob = logf(tsum) + ts_max;
    const bool mask_text = ts_logprob > txt_max;
    // ---- grammar (wc:6334-6355): only when text wins; penalise the rejected tokens, then the log-softmax again over the new logits
    if (!mask_text && r.grammar_reject != nullptr) {
        float gmx = -INFINITY;
        for (int i = tid; i < V; i += SM_THREADS) {
            float x = lg[i];
            if (r.grammar_reject[i]) {
                x = __fsub_rn(x, r.grammar_penalty);
                lg[i] = x;
            }
            gmx = fmaxf(gmx, x);
        }
        gmx = block_max(gmx, red);
        float gs = 0.0f;
        for (int i = tid; i < V; i += SM_THREADS) {
            const float x = lg[i];
            if (x > -INFINITY) gs += expf(x - gmx);
        }
        gs = block_sum(gs, red);
        const float glse = logf(gs) + gmx;
        for (int i = tid; i < V; i += SM_THREADS) {
            const float x = lg[i];
            lp[i] = x > -INFINITY ? x - glse : -INFINITY;
        }
        __syncthreads();
    }

    // ---- probs + picks (whisper_compute_probs, whisper_sample_token best=true)
    double sum_ts = 0.0;
    float best_p = 0.0f, best_ts = 0.0f;
    int best_i = 0x7fffffff, best_ti = 0x7fffffff;
    for (int i = tid; i < V; i += SM_THREADS) {
        float x = lg[i], l = lp[i];
        if (mask_text && i < beg) {
            x = -INFINITY;
            l = -INFINITY;
            lg[i] = x;
            lp[i] = l;
        }
        const float p = (x == -INFINITY) ? 0.0f : expf(l);
        pr[i] = p;
        if (p > best_p) {  // ascending i within a thread: strictly-greater keeps the first index
            best_p = p;
            best_i = i;
        }
        if (i >= beg) {
            sum_ts += (double) p;
            if (p > best_ts) {
                best_ts = p;
                best_ti = i;
            }
        }
    }
    block_argmax_first(best_p, best_i, red, redi);
    block_argmax_first(best_ts, best_ti, red, redi);
    sum_ts = block_sum_d(sum_ts, redd);

    if (tid == 0) {
        SamplerOut o;
        // whisper_sample_token starts from {id 0, tid 0, p 0}: nothing beats 0 -> ids stay 0
        o.id    = best_p > 0.0f ? best_i : 0;
        o.p     = best_p > 0.0f ? best_p : 0.0f;
        o.plog  = best_p > 0.0f ? lp[o.id] : 0.0f;
        o.tid   = best_ts > 0.0f ? best_ti : 0;
        o.pt    = (float) ((double) best_ts / (sum_ts + 1e-10));
        o.ptsum = (float) sum_ts;
        if (o.id >= beg) {
            o.tid = o.id;
            o.pt  = o.p;
        }
        o.nosp_prob = nosp;
        o.pad = 0;
        out[blockIdx.x] = o;
    }
}

bool sample_rows(const SamplerStatic & st, const float * logits, int64_t ld_logits, const SamplerRow * rows, int n_rows, float * work,
                 SamplerOut * out, cudaStream_t stream) {
    ProfScope prof_scope(PC_SAMPLER, stream);
    if (n_rows <= 0) return true;
    sampler_kernel<<<n_rows, SM_THREADS, 0, stream>>>(st, logits, ld_logits, rows, work, out);
    WB_LAUNCHED();
    return true;
}

// =============================================================================================
// X1/X2  DTW alignment
// =============================================================================================
// step 1: per (head, audio column): normalise over the token axis (ggml_norm eps 1e-9, whisper.cpp:8864)
// layout in: qk[h][tok][T]; out: nrm[h][tok][n_audio]
__global__ void dtw_norm_kernel(const float * __restrict__ qk, int n_tok, int T, int n_audio, float * __restrict__ nrm) {
    const int h = blockIdx.y;
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= n_audio) return;
    const float * src = qk + (int64_t) h * n_tok * T + a;
    double sum = 0.0;
    for (int t = 0; t < n_tok; ++t) sum += (double) src[(int64_t) t * T];
    const float mean = (float) (sum / n_tok);
    double sum2 = 0.0;
    for (int t = 0; t < n_tok; ++t) {
        const float v = __fsub_rn(src[(int64_t) t * T], mean);
        sum2 += (double) __fmul_rn(v, v);
    }
    const float variance = (float) (sum2 / n_tok);
    const float scale = __frcp_rn(__fsqrt_rn(__fadd_rn(variance, 1e-9f)));
    float * dst = nrm + (int64_t) h * n_tok * n_audio + a;
    for (int t = 0; t < n_tok; ++t) dst[(int64_t) t * n_audio] = __fmul_rn(__fsub_rn(src[(int64_t) t * T], mean), scale);
}

// step 2: median filter (width w, reflect) over audio, mean over heads, negate -> cost[tok][n_audio]
__global__ void dtw_medfilt_mean_kernel(const float * __restrict__ nrm, int n_heads, int n_tok, int n_audio, int width,
                                        float * __restrict__ cost) {
    const int t = blockIdx.y;
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= n_audio) return;
    double acc = 0.0;  // ggml_mean: ggml_vec_sum_f32 accumulates in ggml_float, then /= (float) n
    for (int h = 0; h < n_heads; ++h) {
        const float * row = nrm + ((int64_t) h * n_tok + t) * n_audio;
        float win[16];
        int n = 0;
        for (int off = -width / 2; off <= width / 2; ++off) {
            int idx = a + off;
            if (idx < 0) idx = -idx;
            else if (idx >= n_audio) idx = 2 * (n_audio - 1) - idx;
            win[n++] = row[idx];
        }
        for (int i = 1; i < n; ++i) {  // insertion sort (n <= 15)
            const float v = win[i];
            int k = i - 1;
            while (k >= 0 && win[k] > v) {
                win[k + 1] = win[k];
                --k;
            }
            win[k + 1] = v;
        }
        acc += (double) win[n / 2];
    }
    cost[(int64_t) t * n_audio + a] = -__fdiv_rn((float) acc, (float) n_heads);
}

// step 3: anti-diagonal wavefront DTW + backtrace in one CTA (whisper.cpp:8647-8731).
// x[i][j]: i in [0,N) tokens, j in [0,M) audio.  dtw buffer: three diagonals of length N+1.
__global__ void __launch_bounds__(256)
dtw_wavefront_kernel(const float * __restrict__ x, int N, int M, float * __restrict__ diag, int8_t * __restrict__ trace,
                     int * __restrict__ path, int * __restrict__ path_len) {
    // cost(i, j) lives on diagonal d = i + j at slot i.  diag: [3][N + 1]
    float * d0 = diag;                 // d - 2
    float * d1 = diag + (N + 1);       // d - 1
    float * d2 = diag + 2 * (N + 1);   // d
    const int tid = threadIdx.x;
    const int W = M + 1;
    for (int i = tid; i <= N; i += blockDim.x) {
        d0[i] = INFINITY;
        d1[i] = INFINITY;
        d2[i] = INFINITY;
    }
    __syncthreads();
    if (tid == 0) d0[0] = 0.0f;  // cost(0,0) on diagonal 0 ; diagonal 1 holds cost(0,1), cost(1,0) = inf
    __syncthreads();
    // after processing diagonal d: d0 <- d1 (d-1), d1 <- d2 (d)   [pointer rotation]
    // start: "d-2" = diagonal 0 (d0), "d-1" = diagonal 1 (d1, all inf)
    for (int d = 2; d <= N + M; ++d) {
        const int ilo = max(1, d - M), ihi = min(N, d - 1);
        for (int i = ilo + tid; i <= ihi; i += blockDim.x) {
            const int j = d - i;
            const float c0 = d0[i - 1];  // cost(i-1, j-1)
            const float c1 = d1[i - 1];  // cost(i-1, j)
            const float c2 = d1[i];      // cost(i, j-1)
            float c;
            int8_t t;
            if (c0 < c1 && c0 < c2) {
                c = c0;
                t = 0;
            } else if (c1 < c0 && c1 < c2) {
                c = c1;
                t = 1;
            } else {
                c = c2;
                t = 2;
            }
            d2[i] = __fadd_rn(x[(int64_t) (i - 1) * M + (j - 1)], c);
            trace[(int64_t) i * W + j] = t;
        }
        __syncthreads();
        // boundary cells of the new diagonal stay +inf: slot 0 (i = 0) and slot d (j = 0)
        float * tmp = d0;
        d0 = d1;
        d1 = d2;
        d2 = tmp;
        for (int i = tid; i <= N; i += blockDim.x) d2[i] = INFINITY;
        __syncthreads();
    }
    if (tid == 0) {
        // backtrace with trace[0][:] = 2, trace[:][0] = 1
        int i = N, j = M, n = 0;
        const int cap = N + M;
        int * pi = path;
        int * pj = path + cap;
        while (i > 0 || j > 0) {
            pi[cap - 1 - n] = i - 1;
            pj[cap - 1 - n] = j - 1;
            ++n;
            int t;
            if (i == 0) t = 2;
            else if (j == 0) t = 1;
            else t = trace[(int64_t) i * W + j];
            if (t == 0) {
                --i;
                --j;
            } else if (t == 1) {
                --i;
            } else {
                --j;
            }
        }
        *path_len = n;
    }
}

bool dtw_cost(const float * qk, int n_heads, int n_tok, int T, int n_audio, int medfilt_width, float * work_nrm, float * work_cost,
              cudaStream_t stream) {
    ProfScope prof_scope(PC_DTW, stream);
    WB_CHECK(medfilt_width <= 15 && (medfilt_width & 1), "bad median filter width %d\n", medfilt_width);
    dtw_norm_kernel<<<dim3(ceil_div(n_audio, 128), n_heads), 128, 0, stream>>>(qk, n_tok, T, n_audio, work_nrm);
    WB_LAUNCHED();
    dtw_medfilt_mean_kernel<<<dim3(ceil_div(n_audio, 128), n_tok), 128, 0, stream>>>(work_nrm, n_heads, n_tok, n_audio, medfilt_width,
                                                                                    work_cost);
    WB_LAUNCHED();
    return true;
}

bool dtw_path(const float * cost, int N, int M, float * diag3, int8_t * trace, int * path, int * path_len, cudaStream_t stream) {
    ProfScope prof_scope(PC_DTW, stream);
    dtw_wavefront_kernel<<<1, 256, 0, stream>>>(cost, N, M, diag3, trace, path, path_len);
    WB_LAUNCHED();
    return true;
}

}  // namespace wb
