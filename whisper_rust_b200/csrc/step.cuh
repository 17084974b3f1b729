// The greedy decode step as ONE CUDA graph of weight-streaming kernels (sm_100a).
//
// A decode step of the token loop (whisper.cpp:7403-7477: one new token per live sequence) is HBM-bound: it reads
// every decoder weight once (amortised over the n sequences of the batch) and, per sequence, that state's
// cross-attention K/V (2*T*S*L_t f16).  The general decoder driver (engine.cu stage_decode) launches ~13 kernels per
// layer with host work in between; this path replaces it, for batches of one row per state, with
//   * gemm_wstream: tcgen05 GEMM with the WEIGHT tile as the M=128 operand and the n token rows as the N operand
//     (TMA for both, f32 accumulate in TMEM), split-K across CTAs so that ~148 CTAs stream disjoint weight slabs;
//     partial sums go to an L2-resident f32 workspace and are reduced in a fixed order by the consumer kernel
//   * ln_reduce: split-K reduce + bias + residual + LayerNorm (double accumulate, like ggml_norm) -> f32 x, f16 LN(x)
//   * self_attn_step / cross_attn_step: reduce their own Q(K,V) partials, append K/V to the cache, attend.
//     cross_attn_step streams the K and V panels with cp.async.bulk into an mbarrier ring (it is the kernel that
//     bounds batched decoding: 2*T*64*2 B per (row, head, layer))
//   * the logits GEMM (same gemm_wstream, direct f32 epilogue) and the sampler kernel
// captured once per batch size into a CUDA graph: one graph launch + one stream sync per generated token.
#pragma once

#include "engine.h"

namespace wb {

// rows: exactly one row per state (greedy, temperature 0), all flagged for logits, every KV cache in append order.
bool step_eligible(const Context & ctx, const BatchRow * rows, int n, bool any_host_dists);
// decoder forward + sampler for one token per row; results land in decs[i]->picked.
bool stage_step(Context & ctx, const BatchRow * rows, int n, const SamplerStatic & sst, const SamplerRow * srows, Decoder ** decs);
// the same, split for decode lanes: launch on lane `lane`'s stream without waiting / wait for it and collect the picks.
// Lanes hold disjoint sets of states; a lane must be waited for before its next launch.
bool stage_step_launch(Context & ctx, int lane, int n_lanes, const BatchRow * rows, int n, const SamplerStatic & sst, const SamplerRow * srows);
bool stage_step_ready(Context & ctx, int lane);  // true once the lane's step has finished (non-blocking)
bool stage_step_wait(Context & ctx, int lane, const BatchRow * rows, int n, Decoder ** decs);
int step_max_lanes();
void step_trace_dump(Context & ctx);  // WB200_STEP_TRACE=<file>: in-situ kernel timeline of the last step of every lane
void step_release(Context & ctx);
uint64_t step_launch_count();

// micro-benchmark hooks (bench.py / tests): average device time of one launch, CUDA events on the launching stream
struct StepBenchResult {
    double us_per_launch;
    double bytes_per_launch;  // algorithmic bytes
    int launches;
};
bool bench_cross_attn_step(int n_rows, int H, int T, int layers, int iters, StepBenchResult * out);
bool bench_gemm_wstream(int n_rows, int N_out, int K, int n_weights, int iters, StepBenchResult * out);

// test hooks on raw device pointers
bool test_gemm_wstream(const __half * w, const __half * x, float * out, int n, int N_out, int K, const float * bias, int force_split);
bool test_gemm_wcluster(const __half * w, const __half * x, void * out, int n, int N_out, int K, const float * bias, int mode, float scale,
                        int scale_f1, const __half * gelu_table);
bool bench_gemm_wcluster(int n_rows, int N_out, int K, int n_weights, int iters, StepBenchResult * out);
bool bench_overlap_gemm(int n_rows, int N_out, int K, int n_weights, int iters, StepBenchResult * out);
bool test_cross_attn_step(const __half * q_f16, const __half * kv, __half * out, int n_rows, int H, int T);
// gemm_wstream with its split-K partial sums in a fixed module-scope device array: no cudaMalloc / cudaFree on any call
// (0 = done, 1 = failure, 2 = beyond the array's size).  step.cu, end of file.
int gemm_wstream_noalloc(const __half * w, const __half * x, float * out, int n, int N_out, int K, const float * bias, int force_split);

}  // namespace wb
