// Hand-written sm_100a kernels of the transcription path that are not GEMMs (HBM/latency-bound work).
#pragma once

#include "common.cuh"

namespace wb {

// ------------------------------------------------------------------------------------------
// M1: log-mel front end  (reference: whisper.cpp:3054-3276)
// ------------------------------------------------------------------------------------------
struct MelTables {
    float * sin_tab;   // [400]  sinf(2*pi*i/400)   (host-computed exactly like whisper.cpp:3031-3037)
    float * cos_tab;   // [400]
    float * hann;      // [400]  periodic Hann       (whisper.cpp:3039-3047)
    float * filters_t; // [201][n_mel] transposed filterbank
    int     n_mel;
};

// samples: device f32 [n_samples]; mel_raw/mel_out: device f32 [n_mel][n_len]; scratch_max: device int[1]
bool mel_compute(const MelTables & tab, const float * samples, int n_samples, int n_len, float * mel_out, int * scratch_max,
                 cudaStream_t stream);

static inline int mel_n_len(int n_samples) { return (n_samples + 480000) / 160; }           // whisper.cpp:3222
static inline int mel_n_len_org(int n_samples) { return 1 + (n_samples + 200 - 400) / 160; } // whisper.cpp:3224

// mel window (whisper.cpp:2406-2420) -> zero-padded, time-major f16 panel [3000 + 2][kc] for the conv1 implicit GEMM
bool mel_window_to_f16(const float * mel, int n_mel, int n_len, int offset, int n_frames /*2*n_ctx*/, __half * out, int kc,
                       cudaStream_t stream);

// ------------------------------------------------------------------------------------------
// LayerNorm (ggml_norm + mul + add: ggml-cpu/ops.cpp:3199-3248, whisper.cpp:2121-2126) -> f16 (+ optional f32 copy)
// ------------------------------------------------------------------------------------------
bool layernorm_rows(const float * x, int rows, int S, const float * gamma, const float * beta, float eps, __half * out_f16,
                    float * out_f32, cudaStream_t stream);

// row softmax with scale (ggml_soft_max_ext, ops.cpp:4731-4827): p = softmax(s*scale) -> f16
bool softmax_rows_f16(const float * s, int64_t rows, int n, int64_t lds, float scale, __half * p, int64_t ldp, cudaStream_t stream);

// in place x[i] = gelu_table[x[i]] with the reference's clamps (x <= -10 -> 0, x >= 10 -> x): ggml-cpu/vec.h:571-585
bool gelu_table_inplace(__half * x, int64_t n, const __half * table, cudaStream_t stream);

// fused encoder self-attention (attn_enc.cu): qk f16 per-head panels [2H][R][64] (Q heads then K heads, R = W*T), vT f16
// [H][W][64][Tpad], out f16 [R][S]; softmax(Q K^T * scale) V   (whisper.cpp:2183-2206)
bool enc_attention(const __half * qk, const __half * vT, __half * out, int W, int T, int Tpad, int H, int S, float scale,
                   cudaStream_t stream);
uint64_t enc_attn_launch_count();

// ------------------------------------------------------------------------------------------
// decoder step kernels
// ------------------------------------------------------------------------------------------
// x[j] = f32(te[token_j]) + pe[pos_j]     (whisper.cpp:2531-2534)
bool embed_tokens(const __half * te, const float * pe, const int * tokens, const int * pos, int n_tokens, int S, float * x,
                  cudaStream_t stream);

// Per-token view of the self-attention KV cache (one entry per batch row).
struct SelfKvRow {
    __half *    k_base;      // this state's K cache for layer 0: [L][n_ctx][S]
    __half *    v_base;
    const int * cells;       // ascending cell indices this row attends to (null: identity 0..n_cells-1)
    int         n_cells;     // number of attended cells (includes this row's own cell)
    int         store_cell;  // cell this row's K/V are written to
    int         n_ctx;       // cells per layer in this cache
};

// write this step's K/V rows (from the fused QKV buffer [n_tokens][3S]) into their cache cells
bool self_kv_store(const __half * qkv, const SelfKvRow * rows, int n_tokens, int S, int layer, cudaStream_t stream);
// out[j][h*64..] = softmax(q_j . K^T) V   over the row's cells   (whisper.cpp:2611-2651)
bool self_attn_decode(const __half * qkv, const SelfKvRow * rows, int n_tokens, int S, int H, int layer, __half * out,
                      cudaStream_t stream);

// Cross attention against the encoder K/V of each row's state (whisper.cpp:2715-2758).
struct CrossKvRow {
    const __half * kv_base;  // this state's cross cache: [L][2H][T][64]  (K heads then V heads)
    float *        qk_out;   // null, or f32 [n_aheads][n_tok_total][T] capture buffer (DTW alignment heads)
    int            tok_index;    // row index of this token inside qk_out
    int            n_tok_total;
};
bool cross_attn_decode(const __half * q, int ldq, const CrossKvRow * rows, int n_tokens, int H, int T, int layer, float scale,
                       const int * ahead_slot /*[H] slot index or -1, device, may be null*/, __half * out, int ldo,
                       cudaStream_t stream);

// ------------------------------------------------------------------------------------------
// S1/S2: logits filter + log-softmax + greedy pick  (whisper.cpp:6149-6417, 6432-6489)
// ------------------------------------------------------------------------------------------
struct SamplerStatic {
    const uint8_t * suppress;  // [V] 1 = always -inf (not/sot/nosp/solm/translate/transcribe/prev/langs/nst/regex)
    const uint8_t * suppress_post;  // [V] the subset applied after logits_filter_callback (regex, non-speech tokens; wc:6230-6262)
    int n_vocab, token_eot, token_beg, token_space, token_nosp;
};
struct SamplerRow {
    int   logits_row;      // row in the logits matrix
    float temperature;
    int   is_initial, suppress_blank, no_timestamps;
    int   last_was_ts, penult_was_ts;
    int   max_initial_tid; // first suppressed timestamp index when is_initial (token_beg + tid0 + 1), or n_vocab
    int   has_ts, tid0;    // monotone rule: suppress [beg, beg + tid0)
    int   want_nosp;       // also report softmax(raw logits)[nosp]
    int   work_slot;       // which [3][V] work area (logits, logprobs, probs) to use
    int   prefiltered;     // 1: the work area already holds the logits after temperature, the static suppressions of wc:6187-6224 and
                           // logits_filter_callback (done on the host); continue from the rules that follow the callback
    int   pad0;
    const uint8_t * suppress;       // per-row masks (jobs with different suppress rules may share a launch); null: SamplerStatic's
    const uint8_t * suppress_post;
    const uint8_t * grammar_reject; // null, or [V]: tokens the row's grammar rejects; applied when text beats the timestamps (wc:6334-6355)
    float grammar_penalty;          // subtracted from the rejected tokens' logits (whisper_full_params.grammar_penalty)
    int   pad1;
};
struct SamplerOut {
    int   id, tid;
    float p, plog, pt, ptsum;
    float nosp_prob;
    int   pad;
};
bool sample_rows(const SamplerStatic & st, const float * logits, int64_t ld_logits, const SamplerRow * rows, int n_rows, float * work,
                 SamplerOut * out, cudaStream_t stream);

// ------------------------------------------------------------------------------------------
// F5: signal energy for the token-timestamp heuristic (get_signal_energy, whisper.cpp:8360-8376)
// ------------------------------------------------------------------------------------------
// out[i] = (sum over j in [-hw, hw] of |pcm[i+j]|, samples outside the signal skipped, added in ascending j) / (2 hw + 1); f32 like the reference
bool signal_energy(const float * pcm, int n_samples, int half_window, float * out, cudaStream_t stream);

// ------------------------------------------------------------------------------------------
// X1/X2: DTW token timestamps  (whisper.cpp:8647-8933)
// ------------------------------------------------------------------------------------------
// qk: f32 [n_heads][n_tok][T] softmax rows; uses the first n_audio columns.
// dtw_cost: token-axis normalisation (eps 1e-9) -> median filter over audio (reflect) -> mean over heads -> negate
//           work_nrm: f32 [n_heads][n_tok][n_audio], cost: f32 [n_tok][n_audio]
bool dtw_cost(const float * qk, int n_heads, int n_tok, int T, int n_audio, int medfilt_width, float * work_nrm, float * cost,
              cudaStream_t stream);
// dtw_path: anti-diagonal wavefront DTW over x = cost rows [N][M] (row stride M) + backtrace.
//           diag3: f32 [3][N+1], trace: int8 [(N+1)][(M+1)], path: int32 [2][N+M] right-aligned, path_len: int32[1]
bool dtw_path(const float * cost, int N, int M, float * diag3, int8_t * trace, int * path, int * path_len, cudaStream_t stream);

uint64_t kernel_launch_count();

}  // namespace wb
