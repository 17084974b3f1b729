// Device execution engine for the transcription path: context (weights + shared workspaces), per-stream
// state (mel, cross/self KV caches, results) and the encode / decode-step drivers.
// Reference counterparts: whisper_context wc:967-982, whisper_state wc:867-965, whisper_kv_cache wc:725-750,
// whisper_encode_internal wc:2376-2472, whisper_decode_internal wc:2864-2994.
#pragma once

#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <random>
#include <set>
#include <string>
#include <vector>

#include "gemm_tc.cuh"
#include "kernels.cuh"
#include "grammar.h"
#include "model.h"

namespace wb {

static constexpr int MAX_DECODERS = 8;  // WHISPER_MAX_DECODERS wc:140

// ---- unified self-attention KV cache bookkeeping (host metadata over device K/V arenas; wc:725-750, 998-1167)
struct KvCell {
    int pos = -1;
    uint32_t seq = 0;  // bit i set <=> sequence id i owns the cell (ids < 2*MAX_DECODERS)
    bool has(int id) const { return (seq >> id) & 1u; }
};
struct KvCache {
    int head = 0, size = 0, n = 0;
    std::vector<KvCell> cells;
    __half * k = nullptr;  // [L][size][S]
    __half * v = nullptr;
    void clear();
    bool find_slot(const int * pos, const int * seq, int n_tokens);  // wc:1049-1098
    int cell_max() const;                                            // wc:1101-1109
    void seq_rm(int seq_id, int p0, int p1);                         // wc:1121-1149
    void seq_cp(int src, int dst, int p0, int p1);                   // wc:1151-1167
};

struct Segment {
    int64_t t0, t1;
    std::string text;
    float no_speech_prob;
    std::vector<whisper_token_data> tokens;
    bool speaker_turn_next;
};

struct Sequence {
    std::vector<whisper_token_data> tokens;
    int result_len = 0;
    double sum_logprobs_all = 0, sum_logprobs = 0, avg_logprobs = 0, entropy = 0, score = 0;
};

struct Decoder {
    Sequence sequence;
    int i_batch = 0, seek_delta = 0;
    bool failed = false, completed = false, has_ts = false;
    // host copies of the filtered distributions: only materialised for temperature > 0 / beam search / callbacks
    std::vector<float> probs, logits, logprobs;
    SamplerOut picked{};  // device-side greedy pick for the current step
    bool want_dists = false;  // copy the distributions to the host after the next sampler launch
    std::mt19937 rng;
    Grammar grammar;                  // grammar-constrained sampling state (wc:802-808); inactive unless params.grammar_rules is set
    std::vector<uint8_t> grammar_reject;  // tokens the grammar rejects at the NEXT step (uploaded with the sampler row)
};

// Encoder activations for up to `windows` 30-second windows processed as one batch of rows.
struct EncWorkspace {
    int windows = 0, T = 0, Tpad = 0;
    __half * melT = nullptr;   // [W][2T+2][kc1]
    __half * h1   = nullptr;   // [W][2T+2][S]   (row 0 and 2T+1 are the conv padding)
    float * x     = nullptr;   // [W*T][S] residual stream
    __half * xa   = nullptr;   // [W*T][S]
    __half * qk   = nullptr;   // [2H][W*T][64]
    __half * vT   = nullptr;   // [H][W][64][Tpad]
    float * sc    = nullptr;   // [H*W][T][Tpad]
    __half * p    = nullptr;   // [H*W][T][Tpad]
    __half * att  = nullptr;   // [W*T][S]
    __half * hb   = nullptr;   // [W*T][4S]
    __half * enc16 = nullptr;  // [W*T][S]
    float * enc32 = nullptr;   // [W*T][S] (embd_enc, kept for the stage-level API / probes)
    uint8_t * base = nullptr;
    size_t bytes = 0;
};

struct DecWorkspace {
    int rows = 0, logit_rows = 0;
    float * x = nullptr;       // [R][S]
    __half * xa = nullptr;     // [R][S]
    __half * qkv = nullptr;    // [R][3S]
    __half * q = nullptr;      // [R][S]
    __half * att = nullptr;    // [R][S]
    __half * hb = nullptr;     // [R][4S]
    float * logits = nullptr;  // [RL][V]
    float * samp_work = nullptr;  // [RL][3][V]
    // staging: one pinned host block mirrored by one device block
    uint8_t * h_stage = nullptr;
    uint8_t * d_stage = nullptr;
    size_t stage_bytes = 0;
    SamplerOut * h_out = nullptr;  // pinned
    SamplerOut * d_out = nullptr;
    int * d_cells = nullptr;       // explicit cell lists [R][max_cells]
    int max_cells = 0;
    uint8_t * d_grammar = nullptr; // [grammar_rows][V] grammar reject masks, allocated on first use
    int grammar_rows = 0;
    uint8_t * base = nullptr;
};

struct Context;

// Static part of whisper_process_logits (wc:6195-6260) for one combination of (tdrz_enable, suppress_nst, suppress_regex): cached per
// context, shared by every job with the same rules; jobs with different rules may share a sampler launch (SamplerRow::suppress).
struct SuppressMask {
    int device = 0;
    uint8_t * d = nullptr;             // [2][V]: everything | the part that follows logits_filter_callback
    std::vector<uint8_t> pre_host;     // the part the reference applies before logits_filter_callback (wc:6197-6224)
    ~SuppressMask();
};

struct State {
    Context * ctx = nullptr;

    // timing / counters (wc:868-881)
    int64_t t_sample_us = 0, t_encode_us = 0, t_decode_us = 0, t_batchd_us = 0, t_prompt_us = 0, t_mel_us = 0;
    int32_t n_sample = 0, n_encode = 0, n_decode = 0, n_batchd = 0, n_prompt = 0, n_fail_p = 0, n_fail_h = 0;

    // mel (device resident; host copy only on request)
    float * d_mel = nullptr;
    size_t mel_cap = 0;
    int n_len = 0, n_len_org = 0, n_mel = 0;
    float * d_pcm = nullptr;
    size_t pcm_cap = 0;
    int * d_melmax = nullptr;

    int kv_self_n_dec = 0;
    KvCache kv_self;
    __half * kv_cross = nullptr;  // [L][2H][T][64]
    int cross_T = 0;              // audio context the cross cache currently holds

    Decoder decoders[MAX_DECODERS];

    std::vector<float> logits;  // host [n_tokens][V], rows flagged for output are filled (wc:2965-2971)
    std::vector<Segment> result_all;
    std::vector<whisper_token> prompt_past;
    int lang_id = 0;
    float no_speech_prob = 0.0f;
    int exp_n_audio_ctx = 0;

    // token-level timestamps (wc:936-943)
    int64_t t_beg = 0, t_last = 0;
    whisper_token tid_last = 0;
    std::vector<float> energy;

    // DTW capture
    float * d_qk = nullptr;  // [n_aheads][n_tok][T]
    size_t qk_cap = 0;

    std::vector<float> mel_host_cache;  // filled by whisper_b200_get_mel

    // VAD (wc:950-965): set by whisper_full / whisper_full_parallel when params.vad, read by the segment-time accessors
    struct VadMap { int64_t processed, original; };   // centiseconds in the filtered audio -> in the caller's audio
    bool has_vad_segments = false;
    std::vector<VadMap> vad_mapping;
    void * vad_context = nullptr;                      // whisper_vad_context, created on first use, freed with the state

    ~State();
};

struct Context {
    Model model;
    whisper_context_params params{};
    State * state = nullptr;  // default state of the *_with_state-less API
    std::string path_model;
    int64_t t_load_us = 0, t_start_us = 0;

    cudaStream_t stream = nullptr;
    std::mutex mu;  // guards the lazy creation of the scheduler
    EncWorkspace enc_ws;
    DecWorkspace dec_ws;
    void * step = nullptr;  // decode-step graphs + their workspace (step.cu)

    // alignment heads (wc:417-443, 8623-8642): per layer, slot index per head or -1
    int n_aheads = 0;
    std::vector<std::vector<int>> ahead_slot;  // [L][H]
    int * d_ahead_slot = nullptr;              // [L][H]

    std::map<std::string, std::shared_ptr<SuppressMask>> masks;  // only touched on the scheduler thread

    void * sched = nullptr;  // the batching scheduler (full.cu): one worker thread per context owns all device work

    ~Context();
};

int64_t time_us();

// --- lifecycle
Context * context_create(whisper_model_loader * loader, whisper_context_params params);
State * state_create(Context * ctx);
bool kv_self_alloc(State & st, int n_ctx);

// --- stages.  All return false on failure (logged).
bool stage_pcm_to_mel(State & st, const float * samples, int n_samples, bool sync = true);
bool stage_mel_sync(Context & ctx, State * const * states, int n);
bool stage_set_mel(State & st, const float * data, int n_len, int n_mel);
bool stage_get_mel(State & st, std::vector<float> & out);
// Encodes one window of each listed state (mel offsets per state) as one batch, then fills their cross caches.
bool stage_encode(Context & ctx, State * const * states, const int * mel_offsets, int n_states);
// VAD front end (vad_api.cu)
int alignment_heads_preset(int preset, int * pairs, int cap);   // engine.cu: the DTW preset tables (tests)
bool vad_filter(Context * ctx, State * st, const whisper_full_params & params, const float * samples, int n_samples, std::vector<float> & filtered);
int64_t vad_map_time(const State & st, int64_t t);
void vad_state_release(State & st);
bool stage_conv_probe(Context & ctx, State * st, int mel_offset, float * dst, int64_t cap);  // E1 alone, host dst [T][S]

struct BatchRow {
    State * st;
    whisper_token token;
    int pos;
    int seq_id;
    bool want_logits;
};
// Decoder forward over a batch of rows (rows of one state must be contiguous and share its KV cache).
// On return the logits of flagged rows sit in ctx.dec_ws.logits in flagged order; `save_qk` captures
// alignment-head cross-attention rows into st.d_qk (single-state batches only).
bool stage_decode(Context & ctx, const BatchRow * rows, int n_rows, bool save_qk);
// copy flagged logits rows to the states' host `logits` vectors ([n_tokens][V] per state)
bool stage_fetch_logits(Context & ctx, const BatchRow * rows, int n_rows);
// run the device sampler over flagged rows; results land in ctx.dec_ws.h_out[0..n_flagged)
// `prefilter` (logits_filter_callback): rows with srows[i].prefiltered get their raw logits row on the host, the callee filters it in
// place (temperature, pre-callback suppressions, the user's callback), and the result becomes the sampler's input for that row.
bool stage_sample(Context & ctx, const SamplerStatic & sst, const SamplerRow * srows, int n, bool copy_dists, Decoder ** decs,
                  const std::function<void(int, float *)> * prefilter = nullptr);

std::shared_ptr<SuppressMask> suppress_mask(Context & ctx, const whisper_full_params & params);

// ---- the batching scheduler (full.cu).  Every entry point that touches the device goes through the context's worker thread:
// whisper_full_with_state / whisper_b200_full_batch / whisper_full_parallel submit jobs and block (user callbacks are handed back
// to the calling thread); the worker advances all live jobs of the context in lock-step, admitting new ones at step boundaries.
// Returns 0, or the first non-zero per-state code.  `params_per_state`: params[i] belongs to states[i], else params[0] to all.
int full_with_states(Context * ctx, State * const * states, int n_states, const whisper_full_params * params, bool params_per_state,
                     const float * const * samples, const int * n_samples);
// runs `fn` on the context's worker thread between device iterations (stage API: mel / encode / decode / probes); false if refused
bool sched_run(Context * ctx, const std::function<void()> & fn);
void sched_shutdown(Context * ctx);
int lang_auto_detect(Context * ctx, State * st, int offset_ms, float * lang_probs);

}  // namespace wb

struct whisper_context : wb::Context {};
struct whisper_state : wb::State {};
