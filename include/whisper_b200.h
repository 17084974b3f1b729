/*
 * whisper_b200.h — additive extensions of libwhisper_b200.so.  Nothing here changes the whisper.h ABI;
 * a caller that only knows whisper.h never needs this header.
 *
 *  - whisper_b200_full_batch: explicit cross-state batching of the transcription path (SURVEY.md §8b
 *    "proposed additive extension"): N states of one context advance in lock-step, so the encoder runs
 *    over N windows per launch and every decoder step streams the weights once for all N sequences.
 *    Semantics per state are exactly those of whisper_full_with_state (reference: wc:6795-7711).
 *  - stage probes used by the parity tests (the reference keeps these values in whisper_state:
 *    mel wc:896, embd_enc wc:912, kv_cross wc:891) and launch counters used by bench.py.
 */
#ifndef WHISPER_B200_H
#define WHISPER_B200_H

#include "whisper.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Runs whisper_full_with_state for n states of `ctx` as one batch. samples[i] / n_samples[i] is the PCM of state i.
 * Returns 0, or the first non-zero per-state code of whisper_full_with_state. */
WHISPER_API int whisper_b200_full_batch(struct whisper_context * ctx, struct whisper_state ** states, int n_states,
                                        struct whisper_full_params params, const float * const * samples, const int * n_samples);

/* Harness for the unchanged API's threading contract: one host thread per state, each calling whisper_full_with_state(ctx, states[i],
 * params, samples[i], n_samples[i]) (what a Rust caller's thread pool does; rs/whisper_state.rs:19-21).  Returns the first non-zero code. */
WHISPER_API int whisper_b200_full_threads(struct whisper_context * ctx, struct whisper_state ** states, int n_states,
                                          struct whisper_full_params params, const float * const * samples, const int * n_samples);

/* The same harness as a stream: n_states worker threads (one state each) drain a shared queue of rounds * n_states clips, samples[k] /
 * n_samples[k]; a thread takes the next clip as soon as its previous whisper_full_with_state call returns (no barrier between rounds; the
 * scheduler admits it while other clips are still decoding).  read_back != 0: every call's result is read through the public accessors;
 * *bytes_read (optional) receives the bytes read; last_item[i] (optional) the queue index of the last clip states[i] transcribed. */
WHISPER_API int whisper_b200_full_stream(struct whisper_context * ctx, struct whisper_state ** states, int n_states,
                                         struct whisper_full_params params, const float * const * samples, const int * n_samples,
                                         int rounds, int read_back, uint64_t * bytes_read, int * last_item);

/* Batched stage entry points (same contracts as whisper_encode_with_state / whisper_decode_with_state, wc:3933-3971). */
WHISPER_API int whisper_b200_encode_batch(struct whisper_context * ctx, struct whisper_state ** states, const int * offsets, int n_states);
/* one new token per state at position n_past[i]; logits readable through whisper_get_logits_from_state */
WHISPER_API int whisper_b200_decode_batch(struct whisper_context * ctx, struct whisper_state ** states, const whisper_token * tokens,
                                          const int * n_past, int n_states);

/* Probes: copy a stage output to host memory.  Return the element count (or -1); copy only when cap is large enough. */
WHISPER_API int64_t whisper_b200_get_mel(struct whisper_state * state, float * dst, int64_t cap, int * n_len, int * n_len_org, int * n_mel);
WHISPER_API int64_t whisper_b200_get_embd_enc(struct whisper_context * ctx, int window, float * dst, int64_t cap);   /* [n_ctx][n_state] */
/* Grammar engine in isolation (whisper.cpp:5784-5890): start on `rules`, accept the given tokens, write the reject mask of the next step
 * to mask_out[n_vocab]; returns the number of live grammar stacks */
WHISPER_API int whisper_b200_test_grammar(struct whisper_context * ctx, const whisper_grammar_element ** rules, size_t n_rules, size_t i_start_rule,
                                          const whisper_token * accept, int n_accept, uint8_t * mask_out);
/* the same automaton on a caller-supplied vocabulary (token_texts[n_vocab], NUL-terminated strings; ids >= token_eot are never
 * rejected, wc:5830-5840): host only, needs neither a context nor a device */
WHISPER_API int whisper_b200_test_grammar_vocab(const char * const * token_texts, int n_vocab, int token_eot,
                                                const whisper_grammar_element ** rules, size_t n_rules, size_t i_start_rule,
                                                const whisper_token * accept, int n_accept, uint8_t * mask_out);
/* the tokenizer (replaces wc:3288-3336) on a caller-supplied vocabulary, n_texts strings per call: host only.  tokens[k * n_max_tokens ...)
 * and counts[k] are what whisper_tokenize(ctx, texts[k], ..., n_max_tokens) writes and returns.  Returns 0. */
WHISPER_API int whisper_b200_test_tokenize_vocab(const char * const * token_texts, int n_vocab, const char * const * texts, int n_texts,
                                                 whisper_token * tokens, int n_max_tokens, int * counts);
/* VAD host-logic test hook: replace the probabilities of a VAD context (so that whisper_vad_segments_from_probs can be checked against
 * the reference on arbitrary sequences, also without a device) */
WHISPER_API int whisper_b200_test_vad_set_probs(struct whisper_vad_context * vctx, const float * probs, int n);
/* E1 alone (wc:1994-2054): conv stem of the window at mel_offset, without the positional embedding = whisper_state::embd_conv; f32 [T][S] */
WHISPER_API int64_t whisper_b200_test_conv_stem(struct whisper_context * ctx, struct whisper_state * state, int mel_offset, float * dst, int64_t cap);
WHISPER_API int64_t whisper_b200_get_kv_cross(struct whisper_state * state, uint16_t * dst, int64_t cap);            /* f16 [L][2H][T][64] */
WHISPER_API int64_t whisper_b200_get_aheads_qk(struct whisper_state * state, float * dst, int64_t cap);              /* [n_heads][n_tok][T] */
WHISPER_API void whisper_b200_get_counters(struct whisper_state * state, int32_t * counts7, int64_t * us6);          /* wc:868-881 */

/* In-situ kernel timing: CUDA events on the launching stream around the first `max_samples_per_category` launches of each
 * kernel category (0 switches it off).  whisper_b200_profile_read(0, ...) snapshots every category (device sync) and
 * returns the category count; then read categories 1..count-1. */
WHISPER_API void whisper_b200_profile(int max_samples_per_category);
WHISPER_API int whisper_b200_profile_read(int category, char * name, int name_cap, double * ms_sum, uint64_t * n_sampled, uint64_t * n_total,
                                          double * bytes_sampled /* algorithmic bytes of the sampled launches */, double * flops_sampled);

/* number of kernels this library has launched in this process (all contexts) */
WHISPER_API uint64_t whisper_b200_kernel_launches(void);
WHISPER_API void whisper_b200_device_sync(struct whisper_context * ctx);

/* Kernel-level test/bench hooks on raw DEVICE pointers (plain pointers and sizes; used by tests/ and bench.py). */
/* C[m][n] (f32 or f16) = sum_k A[m][k] * B[n][k], f16 row-major operands, optional f32 bias[n]. Returns 0 on success. */
WHISPER_API int whisper_b200_test_gemm(const void * a_f16, const void * b_f16, void * c, int m, int n, int k, int lda, int ldb, int ldc,
                                       int c_is_f16, const float * bias, int block_n);
/* as whisper_b200_test_gemm; flags: 1 = accumulate (f32: c += A.B^T + bias, done by TMA reduce-add), 2 = per-head f16 panels
 * c[n/64][m][64].  Synchronises the device before returning. */
WHISPER_API int whisper_b200_test_gemm_ex(const void * a_f16, const void * b_f16, void * c, int m, int n, int k, int lda, int ldb, int ldc,
                                          int c_is_f16, const float * bias, int block_n, int flags);
/* encoder-GEMM micro-benchmark on scratch buffers (flags as above, 4 = with bias; block_n 0 = auto/persistent) */
WHISPER_API int whisper_b200_bench_gemm(int m, int n, int k, int c_is_f16, int flags, int block_n, int n_weights, int iters,
                                        double * us_per_launch);
/* fused encoder self-attention softmax(Q K^T / 8) V on device buffers: q, k, v f16 [W][H][T][64], out f16 [W*T][H*64] */
/* v_transposed: 0 = V in [t][64] panels like Q and K (MN-major MMA operand), 1 = V^T panels [64][t] */
WHISPER_API int whisper_b200_test_enc_attn(const void * q_f16, const void * k_f16, const void * v_f16, void * out_f16, int n_windows, int n_head,
                                           int n_ctx, int v_transposed);
WHISPER_API int whisper_b200_bench_enc_attn(int n_windows, int n_head, int n_ctx, int iters, int v_transposed, double * us_per_launch);
/* log-mel of device PCM -> device mel [n_mel][n_len] using the context's filterbank */
WHISPER_API int whisper_b200_test_mel(struct whisper_context * ctx, const float * d_pcm, int n_samples, float * d_mel);
/* the decode-step cross-attention kernel on device buffers: q f16 [n][S], kv f16 [2H][T][64] per row (same kv for all rows) */
WHISPER_API int whisper_b200_test_cross_attn(const void * q_f16, const void * kv_f16, void * out_f16, int n_rows, int n_head, int n_ctx);

/* decode-step kernels (step.cu): weight-streaming tcgen05 GEMM out[n][n_out] = x[n][k] . w[n_out][k]^T (+bias), n <= 256;
 * force_split: 0 auto split-K, 1 direct epilogue, >1 that many k-splits */
WHISPER_API int whisper_b200_test_gemm_wstream(const void * w_f16, const void * x_f16, float * out, int n, int n_out, int k, const float * bias,
                                               int force_split);
/* decode-step GEMM with the split-K reduction inside a thread-block cluster (distributed shared memory), n <= 256:
 * mode 0: out f16 [n][n_out] = f16((x.w^T + bias) * (feature < scale_f1 ? scale : 1));  mode 1: out f32 [n][n_out] += x.w^T + bias */
WHISPER_API int whisper_b200_test_gemm_wcluster(const void * w_f16, const void * x_f16, void * out, int n, int n_out, int k, const float * bias,
                                                int mode, float scale, int scale_f1);
/* Host logic of whisper_vad without a device (tests; replaces wc:6643-6790 and map_processed_to_original_time wc:7882-7925): segments
 * from the context's current probabilities, then the sample filter (out[cap]) and the time-mapping table (table[2 * table_cap] as
 * (processed, original) centisecond pairs, *n_table entries); map_in[n_map] -> map_out through that table.  Returns the number of
 * filtered samples or -1. */
WHISPER_API int whisper_b200_test_vad_filter(struct whisper_vad_context * vctx, struct whisper_vad_params params, const float * samples,
                                             int n_samples, float * out, int cap, int64_t * table, int table_cap, int * n_table,
                                             const int64_t * map_in, int64_t * map_out, int n_map);
/* Host logic without a device (tests; replaces whisper_exp_compute_token_level_timestamps wc:8390-8615 and whisper_wrap_segment
 * wc:6049-6100): heuristic token times and segment wrapping applied to caller-supplied segments in order, as whisper_full applies them
 * to each new segment (wc:7615-7630; max_len <= 0: no wrapping).  Segment s owns seg_n_tokens[s] consecutive entries of tokens[];
 * energy[n_energy] is the smoothed |signal| of the clip (wc:8359-8388).  The resulting segments come back flattened the same way.
 * Returns their number, or -1. */
WHISPER_API int whisper_b200_test_token_timestamps(const char * const * token_texts, int n_vocab, int token_eot, int token_beg,
                                                   const float * energy, int n_energy, const whisper_token_data * tokens,
                                                   const int * seg_n_tokens, const int64_t * seg_t0, const int64_t * seg_t1, int n_segments,
                                                   float thold_pt, float thold_ptsum, int max_len, int split_on_word,
                                                   whisper_token_data * out_tokens, int out_cap_tokens, int64_t * out_t0, int64_t * out_t1,
                                                   int * out_n_tokens, int out_cap_segments);
/* Host only (tests): the unified self-attention KV cache's cell bookkeeping (replaces whisper_kv_cache_{find_slot, cell_max, seq_rm,
 * seq_cp, clear}, wc:1049-1167) on a cache of `size` cells, driven by n_ops operations of five ints each: {0} clear, {1, n_tokens, seq,
 * first pos} find_slot, {2, seq, p0, p1} seq_rm, {3, src, dst, p0, p1} seq_cp, {4} cell_max.  per_op[3 i ..] = {return value, head,
 * cell_max}; cells_out[2 c ..] = {pos, sequence-id bit mask}.  Returns 0. */
WHISPER_API int whisper_b200_test_kv_script(int size, const int * ops, int n_ops, int * per_op, int * cells_out);
/* Host only (tests): the constant tables the loader uploads: gelu_f16[65536] (replaces ggml_table_gelu_f16, G/ggml-cpu/ggml-cpu.c:3515),
 * sin400 / cos400 / hann400[400] (replace whisper_global_cache, wc:3018-3047).  Null pointers are skipped.  Returns 0. */
WHISPER_API int whisper_b200_test_host_tables(uint16_t * gelu_f16, float * sin400, float * cos400, float * hann400);
/* Host only (tests): the DTW alignment-head table of a whisper_alignment_heads_preset (replaces g_aheads, wc:417-443) as (text layer,
 * head) pairs; returns the number of heads, -1 for NONE / N_TOP_MOST / CUSTOM / unknown */
WHISPER_API int whisper_b200_test_aheads_preset(int preset, int * pairs, int cap);
/* Host only (tests): expand n_elements of a quantised ggml tensor (type ids of ggml.h: 2 q4_0, 3 q4_1, 6 q5_0, 7 q5_1, 8 q8_0, 10..14
 * q2_K..q6_K) to f32 exactly as the model loader does before rounding to f16 (replaces dequantize_row_*, ggml/src/ggml-quants.c:
 * 260-420, 706-1720).  Returns 0, or 1 for a type this loader does not expand / a partial block. */
WHISPER_API int whisper_b200_test_dequantize(int ggml_type, const void * blocks, int64_t n_elements, float * dst);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  The same GEMM as whisper_b200_test_gemm_wstream in its split-K form
 * (split <= 0: the planner's choice; the partial sums are added in split order, then the bias), all pointers DEVICE memory supplied by
 * the caller: out f32 [n][n_out] = x f16 [n][k] . w f16 [n_out][k]^T (+ bias f32 [n_out] or NULL), n in 1..256.  The split-K partial sums
 * live in a fixed 8 MiB array inside the library (2097152 floats; every GPU that loads the library holds it until the process exits), so
 * the call performs no cudaMalloc / cudaFree and uses no allocating library routine.  Returns 0 with `out` complete (the call
 * synchronises the default stream it launches on), 1 on a CUDA failure or bad argument, 2 (and logs why) when
 * splits * n * n_out exceeds 2097152.  All calls on a device share the array: calls that overlap on the GPU race and must be serialised
 * by the caller. */
WHISPER_API int whisper_b200_gemm_wstream_noalloc(const void * w_f16, const void * x_f16, float * out, int n, int n_out, int k,
                                                  const float * bias, int split);
/* the TMA-streamed decode-step cross attention: same contract as whisper_b200_test_cross_attn */
WHISPER_API int whisper_b200_test_cross_attn_step(const void * q_f16, const void * kv_f16, void * out_f16, int n_rows, int n_head, int n_ctx);
/* micro-benchmarks, CUDA-event timed on the launching stream, graph-replayed so that launch overhead is excluded:
 *   "cross_attn_step": a = heads, b = audio ctx, c = layers (n_rows sequences, each with its own cross cache)
 *   "gemm_wstream"   : a = n_out, b = k, c = number of distinct weight matrices streamed round-robin */
WHISPER_API int whisper_b200_bench_kernel(const char * kind, int n_rows, int a, int b, int c, int iters, double * us_per_launch,
                                          double * bytes_per_launch);

/* DTW token-timestamp kernels on HOST data (tests; replaces wc:8840-8890 + dtw_and_backtrace wc:8647-8731):
 * qk f32 [n_heads][n_tok][T] softmax rows -> optional cost_out f32 [n_tok][n_audio]; path over rows sot_len .. n_tok-2.
 * Returns the path length L (path_out[0..L) token indices, [L..2L) time indices) or < 0. */
WHISPER_API int whisper_b200_test_dtw(const float * qk, int n_heads, int n_tok, int n_ctx, int n_audio, int medfilt_width, int sot_len,
                                      float * cost_out, int32_t * path_out, int cap);
/* dtw_and_backtrace alone on a host cost matrix cost[N][M] (time index contiguous) */
WHISPER_API int whisper_b200_test_dtw_path(const float * cost, int N, int M, int32_t * path_out, int cap);

#ifdef __cplusplus
}
#endif
#endif
