// extern "C" boundary: every symbol include/whisper.h and include/whisper_b200.h declare.
// Thin argument checking + locking; the work lives in engine.cu / full.cu.  No C++ exception crosses this file.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <thread>

#include "../../include/whisper_b200.h"
#include "engine.h"
#include "step.cuh"
#include "log.h"
#include "prof.h"
#include "vad.h"

using namespace wb;

static_assert(sizeof(whisper_full_params) == 296, "whisper_full_params must be 296 bytes (SURVEY.md 8b)");
static_assert(sizeof(whisper_context_params) == 48, "whisper_context_params must be 48 bytes");
static_assert(sizeof(whisper_token_data) == 56, "whisper_token_data must be 56 bytes");
static_assert(offsetof(whisper_full_params, language) == 96 && offsetof(whisper_full_params, greedy) == 136 &&
                  offsetof(whisper_full_params, new_segment_callback) == 152 && offsetof(whisper_full_params, grammar_rules) == 232 &&
                  offsetof(whisper_full_params, vad) == 260 && offsetof(whisper_full_params, vad_params) == 272,
              "whisper_full_params layout drifted from the reference header");

#define WB_TRY try {
#define WB_CATCH(retval)                                                        \
    }                                                                           \
    catch (const std::exception & e) {                                          \
        wb::log_error("%s: exception: %s\n", __func__, e.what());               \
        return retval;                                                          \
    }                                                                           \
    catch (...) {                                                               \
        wb::log_error("%s: unknown exception\n", __func__);                     \
        return retval;                                                          \
    }

extern "C" {

// ---------------------------------------------------------------------------------------------
// logging + ggml leakage
// ---------------------------------------------------------------------------------------------
void whisper_log_set(ggml_log_callback log_callback, void * user_data) { wb::log_set((wb::wb_log_callback) log_callback, user_data); }
void ggml_log_set(ggml_log_callback, void *) {}  // no ggml inside: accepted for the crate's install_logging_hooks (src/lib.rs:67-70)
int ggml_cpu_has_avx(void) { return __builtin_cpu_supports("avx") ? 1 : 0; }
int ggml_cpu_has_avx2(void) { return __builtin_cpu_supports("avx2") ? 1 : 0; }
int ggml_cpu_has_fma(void) { return __builtin_cpu_supports("fma") ? 1 : 0; }
int ggml_cpu_has_f16c(void) { return __builtin_cpu_supports("f16c") ? 1 : 0; }

const char * whisper_print_system_info(void) {
    static std::string s;
    s = "WHISPER : COREML = 0 | OPENVINO = 0 | B200 : ARCH = sm_100a | TCGEN05 = 1 | TMA = 1 | CPU_FALLBACK = 0 | ";
    return s.c_str();
}

// ---------------------------------------------------------------------------------------------
// defaults (wc:3622-3638, 5914-6019)
// ---------------------------------------------------------------------------------------------
struct whisper_context_params whisper_context_default_params(void) {
    struct whisper_context_params r;
    memset(&r, 0, sizeof(r));
    r.use_gpu = true;
    r.flash_attn = false;
    r.gpu_device = 0;
    r.dtw_token_timestamps = false;
    r.dtw_aheads_preset = WHISPER_AHEADS_NONE;
    r.dtw_n_top = -1;
    r.dtw_aheads.n_heads = 0;
    r.dtw_aheads.heads = NULL;
    r.dtw_mem_size = 1024 * 1024 * 128;
    return r;
}

struct whisper_vad_params whisper_vad_default_params(void) {  // wc:4384-4395
    struct whisper_vad_params r;
    r.threshold = 0.5f;
    r.min_speech_duration_ms = 250;
    r.min_silence_duration_ms = 100;
    r.max_speech_duration_s = 3.4028234663852886e38f;
    r.speech_pad_ms = 30;
    r.samples_overlap = 0.1f;
    return r;
}

struct whisper_full_params whisper_full_default_params(enum whisper_sampling_strategy strategy) {
    struct whisper_full_params r;
    memset(&r, 0, sizeof(r));
    r.strategy = strategy;
    r.n_threads = std::min(4, (int) std::thread::hardware_concurrency());
    r.n_max_text_ctx = 16384;
    r.no_context = true;
    r.print_progress = true;
    r.print_timestamps = true;
    r.thold_pt = 0.01f;
    r.thold_ptsum = 0.01f;
    r.language = "en";
    r.suppress_blank = true;
    r.temperature = 0.0f;
    r.max_initial_ts = 1.0f;
    r.length_penalty = -1.0f;
    r.temperature_inc = 0.2f;
    r.entropy_thold = 2.4f;
    r.logprob_thold = -1.0f;
    r.no_speech_thold = 0.6f;
    r.greedy.best_of = -1;
    r.beam_search.beam_size = -1;
    r.beam_search.patience = -1.0f;
    r.grammar_penalty = 100.0f;
    r.vad_params = whisper_vad_default_params();
    if (strategy == WHISPER_SAMPLING_GREEDY) r.greedy.best_of = 5;
    if (strategy == WHISPER_SAMPLING_BEAM_SEARCH) {
        r.beam_search.beam_size = 5;
        r.beam_search.patience = -1.0f;
    }
    return r;
}

struct whisper_context_params * whisper_context_default_params_by_ref(void) {
    struct whisper_context_params * p = new whisper_context_params;
    *p = whisper_context_default_params();
    return p;
}
struct whisper_full_params * whisper_full_default_params_by_ref(enum whisper_sampling_strategy strategy) {
    struct whisper_full_params * p = new whisper_full_params;
    *p = whisper_full_default_params(strategy);
    return p;
}
void whisper_free_params(struct whisper_full_params * params) { delete params; }
void whisper_free_context_params(struct whisper_context_params * params) { delete params; }

// ---------------------------------------------------------------------------------------------
// init / free (wc:3640-3889)
// ---------------------------------------------------------------------------------------------
struct whisper_context * whisper_init_with_params_no_state(struct whisper_model_loader * loader, struct whisper_context_params params) {
    WB_TRY
    if (!loader || !loader->read) return nullptr;
    if (!params.use_gpu) {
        log_error("%s: use_gpu = false, but this library has no CPU path (sm_100a kernels only)\n", __func__);
        if (loader->close) loader->close(loader->context);
        return nullptr;
    }
    if (params.flash_attn && params.dtw_token_timestamps) {
        // the reference drops DTW under flash attention (wc:3724-3727); this library's attention always exposes the
        // alignment-head softmax rows, so the request is honoured
        log_msg(GGML_LOG_LEVEL_INFO, "%s: flash_attn is a no-op here; dtw_token_timestamps stays enabled\n", __func__);
    }
    Context * c = context_create(loader, params);
    if (!c) {
        if (loader->close) loader->close(loader->context);
        return nullptr;
    }
    return static_cast<whisper_context *>(c);
    WB_CATCH(nullptr)
}

struct whisper_context * whisper_init_from_file_with_params_no_state(const char * path_model, struct whisper_context_params params) {
    if (!path_model) return nullptr;
    log_msg(GGML_LOG_LEVEL_INFO, "%s: loading model from '%s'\n", __func__, path_model);
    FILE * f = fopen(path_model, "rb");
    if (!f) {
        log_error("%s: failed to open '%s'\n", __func__, path_model);
        return nullptr;
    }
    whisper_model_loader loader = {};
    loader.context = f;
    loader.read = [](void * ctx, void * output, size_t read_size) { return fread(output, 1, read_size, (FILE *) ctx); };
    loader.eof = [](void * ctx) { return feof((FILE *) ctx) != 0; };
    loader.close = [](void * ctx) { (void) ctx; };
    whisper_context * c = whisper_init_with_params_no_state(&loader, params);
    fclose(f);
    if (c) c->path_model = path_model;
    return c;
}

struct whisper_context * whisper_init_from_buffer_with_params_no_state(void * buffer, size_t buffer_size, struct whisper_context_params params) {
    if (!buffer) return nullptr;
    struct buf_context {
        uint8_t * buffer;
        size_t size, off;
    } bc = {(uint8_t *) buffer, buffer_size, 0};
    log_msg(GGML_LOG_LEVEL_INFO, "%s: loading model from buffer\n", __func__);
    whisper_model_loader loader = {};
    loader.context = &bc;
    loader.read = [](void * ctx, void * output, size_t read_size) {
        buf_context * b = (buf_context *) ctx;
        const size_t n = b->off + read_size <= b->size ? read_size : b->size - b->off;
        memcpy(output, b->buffer + b->off, n);
        b->off += n;
        return n;
    };
    loader.eof = [](void * ctx) {
        buf_context * b = (buf_context *) ctx;
        return b->off >= b->size;
    };
    loader.close = [](void *) {};
    return whisper_init_with_params_no_state(&loader, params);
}

struct whisper_state * whisper_init_state(struct whisper_context * ctx) {
    WB_TRY
    if (!ctx) return nullptr;
    std::lock_guard<std::mutex> lk(ctx->mu);
    State * s = state_create(ctx);
    return static_cast<whisper_state *>(s);
    WB_CATCH(nullptr)
}

static whisper_context * with_default_state(whisper_context * ctx) {
    if (!ctx) return nullptr;
    ctx->state = whisper_init_state(ctx);
    if (!ctx->state) {
        whisper_free(ctx);
        return nullptr;
    }
    return ctx;
}
struct whisper_context * whisper_init_from_file_with_params(const char * path_model, struct whisper_context_params params) {
    return with_default_state(whisper_init_from_file_with_params_no_state(path_model, params));
}
struct whisper_context * whisper_init_from_buffer_with_params(void * buffer, size_t buffer_size, struct whisper_context_params params) {
    return with_default_state(whisper_init_from_buffer_with_params_no_state(buffer, buffer_size, params));
}
struct whisper_context * whisper_init_with_params(struct whisper_model_loader * loader, struct whisper_context_params params) {
    return with_default_state(whisper_init_with_params_no_state(loader, params));
}
struct whisper_context * whisper_init_from_file(const char * p) { return whisper_init_from_file_with_params(p, whisper_context_default_params()); }
struct whisper_context * whisper_init_from_buffer(void * b, size_t n) { return whisper_init_from_buffer_with_params(b, n, whisper_context_default_params()); }
struct whisper_context * whisper_init(struct whisper_model_loader * l) { return whisper_init_with_params(l, whisper_context_default_params()); }
struct whisper_context * whisper_init_from_file_no_state(const char * p) { return whisper_init_from_file_with_params_no_state(p, whisper_context_default_params()); }
struct whisper_context * whisper_init_from_buffer_no_state(void * b, size_t n) { return whisper_init_from_buffer_with_params_no_state(b, n, whisper_context_default_params()); }
struct whisper_context * whisper_init_no_state(struct whisper_model_loader * l) { return whisper_init_with_params_no_state(l, whisper_context_default_params()); }

void whisper_free_state(struct whisper_state * state) {
    if (!state) return;
    if (state->ctx) {
        std::lock_guard<std::mutex> lk(state->ctx->mu);
        cudaSetDevice(state->ctx->model.device);
        cudaStreamSynchronize(state->ctx->stream);
        delete state;
    } else {
        delete state;
    }
}
void whisper_free(struct whisper_context * ctx) {
    if (!ctx) return;
    if (ctx->state) whisper_free_state(static_cast<whisper_state *>(ctx->state));
    ctx->state = nullptr;
    delete ctx;
}

int whisper_ctx_init_openvino_encoder_with_state(struct whisper_context *, struct whisper_state *, const char *, const char *, const char *) { return 1; }
int whisper_ctx_init_openvino_encoder(struct whisper_context *, const char *, const char *, const char *) { return 1; }

// ---------------------------------------------------------------------------------------------
// stage entry points (wc:3891-3971)
// ---------------------------------------------------------------------------------------------
int whisper_pcm_to_mel_with_state(struct whisper_context * ctx, struct whisper_state * state, const float * samples, int n_samples, int) {
    WB_TRY
    if (!ctx || !state || !samples || n_samples < 0) return -1;
    bool ok = false;
    if (!sched_run(ctx, [&] { ok = stage_pcm_to_mel(*state, samples, n_samples); }) || !ok) {
        log_error("%s: failed to compute mel spectrogram\n", __func__);
        return -1;
    }
    return 0;
    WB_CATCH(-1)
}
int whisper_pcm_to_mel(struct whisper_context * ctx, const float * samples, int n_samples, int n_threads) {
    return ctx && ctx->state ? whisper_pcm_to_mel_with_state(ctx, static_cast<whisper_state *>(ctx->state), samples, n_samples, n_threads) : -1;
}

int whisper_set_mel_with_state(struct whisper_context * ctx, struct whisper_state * state, const float * data, int n_len, int n_mel) {
    WB_TRY
    if (!ctx || !state || !data) return -1;
    if (n_mel != ctx->model.filters_n_mel) {
        log_error("%s: invalid number of mel bands: %d (expected %d)\n", __func__, n_mel, ctx->model.filters_n_mel);
        return -1;
    }
    bool ok = false;
    return sched_run(ctx, [&] { ok = stage_set_mel(*state, data, n_len, n_mel); }) && ok ? 0 : -1;
    WB_CATCH(-1)
}
int whisper_set_mel(struct whisper_context * ctx, const float * data, int n_len, int n_mel) {
    return ctx && ctx->state ? whisper_set_mel_with_state(ctx, static_cast<whisper_state *>(ctx->state), data, n_len, n_mel) : -1;
}

int whisper_encode_with_state(struct whisper_context * ctx, struct whisper_state * state, int offset, int) {
    WB_TRY
    if (!ctx || !state) return -1;
    State * sp = state;
    bool ok = false;
    if (!sched_run(ctx, [&] { ok = stage_encode(*ctx, &sp, &offset, 1); }) || !ok) {
        log_error("%s: failed to eval\n", __func__);
        return -1;
    }
    return 0;
    WB_CATCH(-1)
}
int whisper_encode(struct whisper_context * ctx, int offset, int n_threads) {
    return ctx && ctx->state ? whisper_encode_with_state(ctx, static_cast<whisper_state *>(ctx->state), offset, n_threads) : -1;
}

int whisper_decode_with_state(struct whisper_context * ctx, struct whisper_state * state, const whisper_token * tokens, int n_tokens, int n_past,
                              int) {
    WB_TRY
    if (!ctx || !state || !tokens || n_tokens <= 0) return 1;
    std::vector<BatchRow> rows;
    for (int i = 0; i < n_tokens; ++i) {
        if (tokens[i] < 0 || tokens[i] >= ctx->model.hp.n_vocab || n_past + i >= ctx->model.hp.n_text_ctx) {
            log_error("%s: token %d / position %d out of range\n", __func__, tokens[i], n_past + i);
            return 1;
        }
        rows.push_back(BatchRow{state, tokens[i], n_past + i, 0, i == n_tokens - 1});  // whisper_batch_prep_legacy wc:544-556
    }
    bool ok = false;
    const bool ran = sched_run(ctx, [&] {
        state->kv_self.seq_rm(0, n_past, -1);
        ok = stage_decode(*ctx, rows.data(), n_tokens, false) && stage_fetch_logits(*ctx, rows.data(), n_tokens);
    });
    if (!ran || !ok) {
        log_error("%s: failed to eval\n", __func__);
        return 1;
    }
    return 0;
    WB_CATCH(1)
}
int whisper_decode(struct whisper_context * ctx, const whisper_token * tokens, int n_tokens, int n_past, int n_threads) {
    if (!ctx || ctx->state == nullptr) {
        log_error("%s: ERROR state was not loaded.\n", __func__);
        return -1;
    }
    return whisper_decode_with_state(ctx, static_cast<whisper_state *>(ctx->state), tokens, n_tokens, n_past, n_threads);
}

// ---------------------------------------------------------------------------------------------
// F6: tokenizer (wc:3288-3336, 3973-3988): GPT-2 style regex pre-split, then greedy longest match
// ---------------------------------------------------------------------------------------------
// Pre-split.  The reference runs std::regex over
//   's|'t|'re|'ve|'m|'ll|'d| ?[[:alpha:]]+| ?[[:digit:]]+| ?[^\s[:alpha:][:digit:]]+|\s+(?!\S)|\s+
// (ECMAScript: leftmost match, first alternative that succeeds, "C" locale classes on bytes).  Every byte belongs to exactly one of four
// classes, so a match always starts at the current position and the pattern is a small scanner: a contraction suffix; else an optional
// single ' ' glued to a run of one non-space class; else a whitespace run, which keeps its last character for the next piece unless it
// reaches the end of the text or is a single character.  Returns the length of the piece starting at s[p] (p < n).
enum : int { CH_SPACE = 0, CH_ALPHA = 1, CH_DIGIT = 2, CH_OTHER = 3 };
static inline int char_class(unsigned char c) {
    if (c == ' ' || (c >= 9 && c <= 13)) return CH_SPACE;
    if ((c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z')) return CH_ALPHA;
    if (c >= '0' && c <= '9') return CH_DIGIT;
    return CH_OTHER;   // punctuation, control characters, every byte of a multi-byte UTF-8 sequence
}
static size_t presplit_piece(const char * s, size_t p, size_t n) {
    if (s[p] == '\'' && p + 1 < n) {
        static const char * const suffix[] = {"s", "t", "re", "ve", "m", "ll", "d"};
        for (const char * suf : suffix) {
            const size_t len = strlen(suf);
            if (p + 1 + len <= n && memcmp(s + p + 1, suf, len) == 0) return 1 + len;
        }
    }
    size_t q = p;
    if (s[q] == ' ' && q + 1 < n && char_class((unsigned char) s[q + 1]) != CH_SPACE) ++q;   // " ?" binds only in front of a run
    const int cls = char_class((unsigned char) s[q]);
    if (cls != CH_SPACE) {
        while (q < n && char_class((unsigned char) s[q]) == cls) ++q;
        return q - p;
    }
    while (q < n && char_class((unsigned char) s[q]) == CH_SPACE) ++q;
    const size_t run = q - p;
    return (q == n || run == 1) ? run : run - 1;
}

static std::vector<int> tokenize_text(const Vocab & vocab, const std::string & text) {
    std::vector<int> tokens;
    const size_t n = text.size();
    for (size_t p = 0; p < n;) {
        const size_t end = p + presplit_piece(text.data(), p, n);
        // greedy longest match inside the piece (wc:3307-3331); a byte no token starts with is reported and skipped
        while (p < end) {
            size_t j = end;
            for (; j > p; --j) {
                auto it = vocab.token_to_id.find(text.substr(p, j - p));
                if (it != vocab.token_to_id.end()) {
                    tokens.push_back(it->second);
                    break;
                }
            }
            if (j == p) {
                log_error("unknown token\n");
                ++p;
            } else {
                p = j;
            }
        }
    }
    return tokens;
}

int whisper_tokenize(struct whisper_context * ctx, const char * text, whisper_token * tokens, int n_max_tokens) {
    WB_TRY
    if (!ctx || !text) return -1;
    const auto res = tokenize_text(ctx->model.vocab, text);
    if (n_max_tokens < (int) res.size()) {
        log_error("%s: too many resulting tokens: %d (max %d)\n", __func__, (int) res.size(), n_max_tokens);
        return -(int) res.size();
    }
    for (int i = 0; i < (int) res.size(); i++) tokens[i] = res[(size_t) i];
    return (int) res.size();
    WB_CATCH(-1)
}
// host only (tests): the tokenizer on a caller-supplied id -> text table (later ids win on duplicate texts, like the loader's map).
// texts[k] -> tokens[k * n_max_tokens ...), counts[k] = whisper_tokenize's return value for that text.
int whisper_b200_test_tokenize_vocab(const char * const * token_texts, int n_vocab, const char * const * texts, int n_texts, whisper_token * tokens,
                                     int n_max_tokens, int * counts) {
    WB_TRY
    if (!token_texts || n_vocab <= 0 || !texts || n_texts < 0 || !counts || n_max_tokens < 0) return -1;
    Vocab vocab;
    for (int i = 0; i < n_vocab; ++i)
        if (token_texts[i]) vocab.token_to_id[token_texts[i]] = i;
    for (int k = 0; k < n_texts; ++k) {
        const auto res = tokenize_text(vocab, texts[k] ? texts[k] : "");
        if (n_max_tokens < (int) res.size()) {
            counts[k] = -(int) res.size();
            continue;
        }
        for (int i = 0; i < (int) res.size(); i++) tokens[(size_t) k * n_max_tokens + i] = res[(size_t) i];
        counts[k] = (int) res.size();
    }
    return 0;
    WB_CATCH(-1)
}
int whisper_token_count(struct whisper_context * ctx, const char * text) { return -whisper_tokenize(ctx, text, NULL, 0); }

int whisper_lang_max_id(void) { return 99; }
int whisper_lang_id(const char * lang) {
    const int id = wb::lang_id(lang);
    if (id < 0) log_error("%s: unknown language '%s'\n", __func__, lang ? lang : "(null)");
    return id;
}
const char * whisper_lang_str(int id) {
    if (id < 0 || id >= 100) {
        log_error("%s: unknown language id %d\n", __func__, id);
        return nullptr;
    }
    return g_langs[id].code;
}
const char * whisper_lang_str_full(int id) {
    if (id < 0 || id >= 100) {
        log_error("%s: unknown language id %d\n", __func__, id);
        return nullptr;
    }
    return g_langs[id].name;
}

int whisper_lang_auto_detect_with_state(struct whisper_context * ctx, struct whisper_state * state, int offset_ms, int, float * lang_probs) {
    WB_TRY
    if (!ctx || !state) return -1;
    int r = -1;
    if (!sched_run(ctx, [&] { r = lang_auto_detect(ctx, state, offset_ms, lang_probs); })) return -1;
    return r;
    WB_CATCH(-1)
}
int whisper_lang_auto_detect(struct whisper_context * ctx, int offset_ms, int n_threads, float * lang_probs) {
    return ctx && ctx->state ? whisper_lang_auto_detect_with_state(ctx, static_cast<whisper_state *>(ctx->state), offset_ms, n_threads, lang_probs) : -1;
}

// ---------------------------------------------------------------------------------------------
// model info / tokens (wc:4185-4259)
// ---------------------------------------------------------------------------------------------
int whisper_n_len_from_state(struct whisper_state * state) { return state ? state->n_len_org : 0; }
int whisper_n_len(struct whisper_context * ctx) { return ctx && ctx->state ? ctx->state->n_len_org : 0; }
int whisper_n_vocab(struct whisper_context * ctx) { return ctx->model.vocab.n_vocab; }
int whisper_n_text_ctx(struct whisper_context * ctx) { return ctx->model.hp.n_text_ctx; }
int whisper_n_audio_ctx(struct whisper_context * ctx) { return ctx->model.hp.n_audio_ctx; }
int whisper_is_multilingual(struct whisper_context * ctx) { return ctx->model.vocab.is_multilingual() ? 1 : 0; }
int whisper_model_n_vocab(struct whisper_context * ctx) { return ctx->model.hp.n_vocab; }
int whisper_model_n_audio_ctx(struct whisper_context * ctx) { return ctx->model.hp.n_audio_ctx; }
int whisper_model_n_audio_state(struct whisper_context * ctx) { return ctx->model.hp.n_audio_state; }
int whisper_model_n_audio_head(struct whisper_context * ctx) { return ctx->model.hp.n_audio_head; }
int whisper_model_n_audio_layer(struct whisper_context * ctx) { return ctx->model.hp.n_audio_layer; }
int whisper_model_n_text_ctx(struct whisper_context * ctx) { return ctx->model.hp.n_text_ctx; }
int whisper_model_n_text_state(struct whisper_context * ctx) { return ctx->model.hp.n_text_state; }
int whisper_model_n_text_head(struct whisper_context * ctx) { return ctx->model.hp.n_text_head; }
int whisper_model_n_text_layer(struct whisper_context * ctx) { return ctx->model.hp.n_text_layer; }
int whisper_model_n_mels(struct whisper_context * ctx) { return ctx->model.hp.n_mels; }
int whisper_model_ftype(struct whisper_context * ctx) { return ctx->model.hp.ftype; }
int whisper_model_type(struct whisper_context * ctx) { return ctx->model.model_type; }
const char * whisper_model_type_readable(struct whisper_context * ctx) {
    static const char * names[] = {"unknown", "tiny", "base", "small", "medium", "large"};
    const int t = ctx->model.model_type;
    return names[t >= 0 && t <= 5 ? t : 0];
}
float * whisper_get_logits_from_state(struct whisper_state * state) { return state ? state->logits.data() : nullptr; }
float * whisper_get_logits(struct whisper_context * ctx) { return ctx && ctx->state ? ctx->state->logits.data() : nullptr; }
const char * whisper_token_to_str(struct whisper_context * ctx, whisper_token token) {
    const auto & v = ctx->model.vocab.id_to_token;
    if (token < 0 || token >= (int) v.size()) {
        log_error("%s: token %d out of range\n", __func__, token);  // the reference throws std::out_of_range here (wc:4218)
        return "";
    }
    return v[(size_t) token].c_str();
}
whisper_token whisper_token_eot(struct whisper_context * ctx) { return ctx->model.vocab.token_eot; }
whisper_token whisper_token_sot(struct whisper_context * ctx) { return ctx->model.vocab.token_sot; }
whisper_token whisper_token_solm(struct whisper_context * ctx) { return ctx->model.vocab.token_solm; }
whisper_token whisper_token_prev(struct whisper_context * ctx) { return ctx->model.vocab.token_prev; }
whisper_token whisper_token_nosp(struct whisper_context * ctx) { return ctx->model.vocab.token_nosp; }
whisper_token whisper_token_not(struct whisper_context * ctx) { return ctx->model.vocab.token_not; }
whisper_token whisper_token_beg(struct whisper_context * ctx) { return ctx->model.vocab.token_beg; }
whisper_token whisper_token_lang(struct whisper_context * ctx, int lang_id) { return ctx->model.vocab.token_sot + 1 + lang_id; }
whisper_token whisper_token_translate(struct whisper_context * ctx) { return ctx->model.vocab.token_translate; }
whisper_token whisper_token_transcribe(struct whisper_context * ctx) { return ctx->model.vocab.token_transcribe; }

// ---------------------------------------------------------------------------------------------
// timings (wc:4261-4312): like the reference these read ctx->state, which is NULL for *_no_state contexts
// ---------------------------------------------------------------------------------------------
struct whisper_timings * whisper_get_timings(struct whisper_context * ctx) {
    if (!ctx || ctx->state == nullptr) return nullptr;
    const State & s = *ctx->state;
    whisper_timings * t = new whisper_timings;
    t->sample_ms = 1e-3f * s.t_sample_us / std::max(1, s.n_sample);
    t->encode_ms = 1e-3f * s.t_encode_us / std::max(1, s.n_encode);
    t->decode_ms = 1e-3f * s.t_decode_us / std::max(1, s.n_decode);
    t->batchd_ms = 1e-3f * s.t_batchd_us / std::max(1, s.n_batchd);
    t->prompt_ms = 1e-3f * s.t_prompt_us / std::max(1, s.n_prompt);
    return t;
}
void whisper_print_timings(struct whisper_context * ctx) {
    if (!ctx) return;
    const int64_t t_end_us = time_us();
    log_msg(GGML_LOG_LEVEL_INFO, "\n");
    log_msg(GGML_LOG_LEVEL_INFO, "%s:     load time = %8.2f ms\n", __func__, ctx->t_load_us / 1000.0f);
    if (ctx->state != nullptr) {
        const State & s = *ctx->state;
        log_msg(GGML_LOG_LEVEL_INFO, "%s:     fallbacks = %3d p / %3d h\n", __func__, s.n_fail_p, s.n_fail_h);
        log_msg(GGML_LOG_LEVEL_INFO, "%s:      mel time = %8.2f ms\n", __func__, s.t_mel_us / 1000.0f);
        log_msg(GGML_LOG_LEVEL_INFO, "%s:   sample time = %8.2f ms / %5d runs\n", __func__, 1e-3f * s.t_sample_us, std::max(1, s.n_sample));
        log_msg(GGML_LOG_LEVEL_INFO, "%s:   encode time = %8.2f ms / %5d runs\n", __func__, 1e-3f * s.t_encode_us, std::max(1, s.n_encode));
        log_msg(GGML_LOG_LEVEL_INFO, "%s:   decode time = %8.2f ms / %5d runs\n", __func__, 1e-3f * s.t_decode_us, std::max(1, s.n_decode));
        log_msg(GGML_LOG_LEVEL_INFO, "%s:   batchd time = %8.2f ms / %5d runs\n", __func__, 1e-3f * s.t_batchd_us, std::max(1, s.n_batchd));
        log_msg(GGML_LOG_LEVEL_INFO, "%s:   prompt time = %8.2f ms / %5d runs\n", __func__, 1e-3f * s.t_prompt_us, std::max(1, s.n_prompt));
    }
    log_msg(GGML_LOG_LEVEL_INFO, "%s:    total time = %8.2f ms\n", __func__, (t_end_us - ctx->t_start_us) / 1000.0f);
}
void whisper_reset_timings(struct whisper_context * ctx) {
    if (!ctx) return;
    ctx->t_start_us = time_us();
    if (ctx->state != nullptr) {
        State & s = *ctx->state;
        s.t_mel_us = s.t_sample_us = s.t_encode_us = s.t_decode_us = s.t_batchd_us = s.t_prompt_us = 0;
        s.n_sample = s.n_encode = s.n_decode = s.n_batchd = s.n_prompt = 0;
    }
}

// ---------------------------------------------------------------------------------------------
// full transcription (wc:6795-7864) and result accessors (wc:7866-8033)
// ---------------------------------------------------------------------------------------------
// whisper_full_params options: everything the reference reads on this path is honoured (grammar: grammar.cpp + the sampler's penalty hook).
static int unsupported_params(const whisper_full_params & params, const char * fn) {
    (void) params;
    (void) fn;
    // params.vad is read by whisper_full / whisper_full_parallel only, exactly as in the reference (wc:7720, 7748): whisper_full_with_state
    // itself never looks at it (wc:6795-7711), so a vad = true arriving here transcribes the samples it is given
    return 0;
}

int whisper_full_with_state(struct whisper_context * ctx, struct whisper_state * state, struct whisper_full_params params, const float * samples,
                            int n_samples) {
    WB_TRY
    if (!ctx || !state) return -1;
    if (const int rc = unsupported_params(params, __func__)) return rc;
    State * sp = state;
    return full_with_states(ctx, &sp, 1, &params, false, &samples, &n_samples);
    WB_CATCH(-1)
}
// the VAD step shared by whisper_full and whisper_full_parallel (wc:7718-7733, 7746-7761): 1 = continue with the filtered samples,
// 0 = no speech (empty result, return 0), -1 = failure.  The PCM must be host memory on this path (the filter is a host-side gather).
static int vad_step(struct whisper_context * ctx, const whisper_full_params & params, const float *& samples, int & n_samples, std::vector<float> & keep,
                    bool clear_on_silence) {
    if (!params.vad) return 1;
    log_msg(GGML_LOG_LEVEL_INFO, "%s: VAD is enabled, processing speech segments only\n", __func__);
    bool ok = false;
    const bool ran = sched_run(ctx, [&] { ok = vad_filter(ctx, ctx->state, params, samples, n_samples, keep); });
    if (!ran || !ok) {
        log_error("%s: failed to compute VAD\n", __func__);
        return -1;
    }
    if (keep.empty()) {
        if (clear_on_silence) ctx->state->result_all.clear();   // whisper_full clears (wc:7727), whisper_full_parallel does not (wc:7753)
        return 0;
    }
    samples = keep.data();
    n_samples = (int) keep.size();
    return 1;
}

int whisper_full(struct whisper_context * ctx, struct whisper_full_params params, const float * samples, int n_samples) {
    WB_TRY
    if (!ctx || !ctx->state) return -1;
    std::vector<float> vad_samples;
    const int v = vad_step(ctx, params, samples, n_samples, vad_samples, true);
    if (v <= 0) return v;
    return whisper_full_with_state(ctx, static_cast<whisper_state *>(ctx->state), params, samples, n_samples);
    WB_CATCH(-1)
}

int whisper_b200_full_batch(struct whisper_context * ctx, struct whisper_state ** states, int n_states, struct whisper_full_params params,
                            const float * const * samples, const int * n_samples) {
    WB_TRY
    if (!ctx || !states || n_states <= 0 || !samples || !n_samples) return -1;
    std::vector<State *> sp((size_t) n_states);
    for (int i = 0; i < n_states; ++i) {
        if (!states[i] || states[i]->ctx != ctx) return -1;
        for (int j = 0; j < i; ++j)
            if (states[j] == states[i]) return -1;
        sp[(size_t) i] = states[i];
    }
    if (const int rc = unsupported_params(params, __func__)) return rc;
    return full_with_states(ctx, sp.data(), n_states, &params, false, samples, n_samples);
    WB_CATCH(-1)
}

// Host-side harness for the drop-in threading contract (rs/whisper_state.rs:19-21: states are Send+Sync): one std::thread per
// state, each making an ordinary whisper_full_with_state call on the shared context, exactly what a Rust caller's thread pool does.
// Nothing here is batched by the caller; the batching happens in the library's scheduler.
int whisper_b200_full_threads(struct whisper_context * ctx, struct whisper_state ** states, int n_states, struct whisper_full_params params,
                              const float * const * samples, const int * n_samples) {
    WB_TRY
    if (!ctx || !states || n_states <= 0 || !samples || !n_samples) return -1;
    std::vector<int> rc((size_t) n_states, 0);
    std::vector<std::thread> th;
    th.reserve((size_t) n_states);
    for (int i = 0; i < n_states; ++i)
        th.emplace_back([&, i] { rc[(size_t) i] = whisper_full_with_state(ctx, states[i], params, samples[i], n_samples[i]); });
    for (auto & t : th) t.join();
    for (int r : rc)
        if (r != 0) return r;
    return 0;
    WB_CATCH(-1)
}

// The same harness as a stream of work: n_states worker threads (one WhisperState each) drain a shared queue of rounds * n_states clips,
// samples[0 .. rounds * n_states), each thread taking the next clip as soon as its previous whisper_full_with_state call returns (a
// server's thread pool).  There is no barrier between rounds: the scheduler admits a thread's next clip while other clips are still
// decoding.  With read_back != 0 every call's result is read through the public accessors (segments, times, text, token data) like a
// caller would; *bytes_read sums them.  last_item[i] (optional) receives the queue index of the last clip state i transcribed.
int whisper_b200_full_stream(struct whisper_context * ctx, struct whisper_state ** states, int n_states, struct whisper_full_params params,
                             const float * const * samples, const int * n_samples, int rounds, int read_back, uint64_t * bytes_read,
                             int * last_item) {
    WB_TRY
    if (!ctx || !states || n_states <= 0 || !samples || !n_samples || rounds <= 0) return -1;
    const int n_items = rounds * n_states;
    std::atomic<int> next{0};
    std::atomic<int> first_error{0};
    std::vector<uint64_t> nb((size_t) n_states, 0);
    std::vector<std::thread> th;
    th.reserve((size_t) n_states);
    for (int i = 0; i < n_states; ++i)
        th.emplace_back([&, i] {
            whisper_state * st = states[i];
            if (last_item) last_item[i] = -1;
            while (first_error.load() == 0) {
                const int k = next.fetch_add(1);
                if (k >= n_items) break;
                const int rc = whisper_full_with_state(ctx, st, params, samples[k], n_samples[k]);
                if (rc != 0) {
                    int zero = 0;
                    first_error.compare_exchange_strong(zero, rc);
                    break;
                }
                if (last_item) last_item[i] = k;
                if (!read_back) continue;
                const int ns = whisper_full_n_segments_from_state(st);
                for (int sgi = 0; sgi < ns; ++sgi) {
                    const char * txt = whisper_full_get_segment_text_from_state(st, sgi);
                    nb[(size_t) i] += 16 + (txt ? strlen(txt) : 0);
                    (void) whisper_full_get_segment_t0_from_state(st, sgi);
                    (void) whisper_full_get_segment_t1_from_state(st, sgi);
                    const int nt = whisper_full_n_tokens_from_state(st, sgi);
                    for (int t = 0; t < nt; ++t) {
                        const whisper_token_data td = whisper_full_get_token_data_from_state(st, sgi, t);
                        nb[(size_t) i] += sizeof(td) + (td.id < 0 ? 1 : 0);
                    }
                }
            }
        });
    for (auto & t : th) t.join();
    if (bytes_read) {
        *bytes_read = 0;
        for (uint64_t b : nb) *bytes_read += b;
    }
    return first_error.load();
    WB_CATCH(-1)
}

// whisper_full_parallel (wc:7736-7864): equal sample spans, one state per span — here the spans advance as ONE batch on
// the device instead of one CPU thread each; the merge rule (time offsets, clamping, append order) is the reference's.
int whisper_full_parallel(struct whisper_context * ctx, struct whisper_full_params params, const float * samples, int n_samples, int n_processors) {
    WB_TRY
    if (!ctx || !ctx->state) return -1;
    if (n_processors <= 1) return whisper_full(ctx, params, samples, n_samples);
    std::vector<float> vad_samples;
    {
        const int v = vad_step(ctx, params, samples, n_samples, vad_samples, false);
        if (v <= 0) return v;
    }
    const int offset_samples = (WHISPER_SAMPLE_RATE * params.offset_ms) / 1000;
    const int n_per = (n_samples - offset_samples) / n_processors;
    std::vector<whisper_state *> extra;
    std::vector<State *> sp;
    std::vector<whisper_full_params> ps;
    std::vector<const float *> smp;
    std::vector<int> ns;
    {
        auto p0 = params;
        p0.print_realtime = false;
        sp.push_back(ctx->state);
        ps.push_back(p0);
        smp.push_back(samples);
        ns.push_back(offset_samples + n_per);
    }
    for (int i = 0; i < n_processors - 1; ++i) {
        whisper_state * s = whisper_init_state(ctx);
        if (!s) {
            for (auto * e : extra) whisper_free_state(e);
            return -1;
        }
        extra.push_back(s);
        const int start = offset_samples + (i + 1) * n_per;
        const int n_cur = (i == n_processors - 2) ? n_samples - start : n_per;
        auto pc = params;
        pc.offset_ms = 0;
        pc.print_progress = false;
        pc.print_realtime = false;
        pc.new_segment_callback = nullptr;
        pc.new_segment_callback_user_data = nullptr;
        pc.progress_callback = nullptr;
        pc.progress_callback_user_data = nullptr;
        sp.push_back(s);
        ps.push_back(pc);
        smp.push_back(samples + start);
        ns.push_back(n_cur);
    }
    const int ret = full_with_states(ctx, sp.data(), (int) sp.size(), ps.data(), true, smp.data(), ns.data());
    const int64_t offset_t = (int64_t) (params.offset_ms / 10.0);
    State & main = *ctx->state;
    for (int i = 0; i < n_processors - 1; ++i) {
        for (auto & r : extra[(size_t) i]->result_all) {
            r.t0 += 100 * ((i + 1) * n_per) / WHISPER_SAMPLE_RATE + offset_t;
            r.t1 += 100 * ((i + 1) * n_per) / WHISPER_SAMPLE_RATE + offset_t;
            if (!main.result_all.empty()) r.t0 = std::max(r.t0, main.result_all.back().t1);
            main.result_all.push_back(std::move(r));
            if (params.new_segment_callback)
                params.new_segment_callback(ctx, static_cast<whisper_state *>(ctx->state), 1, params.new_segment_callback_user_data);
        }
        main.n_encode += extra[(size_t) i]->n_encode;
        main.n_decode += extra[(size_t) i]->n_decode;
        main.n_batchd += extra[(size_t) i]->n_batchd;
        main.n_prompt += extra[(size_t) i]->n_prompt;
        whisper_free_state(extra[(size_t) i]);
    }
    return ret;
    WB_CATCH(-1)
}

#define SEG(state, i) ((state)->result_all[(size_t) (i)])
int whisper_full_n_segments_from_state(struct whisper_state * state) { return (int) state->result_all.size(); }
int whisper_full_n_segments(struct whisper_context * ctx) { return (int) ctx->state->result_all.size(); }
int whisper_full_lang_id_from_state(struct whisper_state * state) { return state->lang_id; }
int whisper_full_lang_id(struct whisper_context * ctx) { return ctx->state->lang_id; }
// segment times: mapped back to the caller's audio when the state's last whisper_full ran on VAD-filtered samples (wc:7927-7960)
int64_t whisper_full_get_segment_t0_from_state(struct whisper_state * state, int i) {
    const int64_t t0 = SEG(state, i).t0;
    if (!state->has_vad_segments || state->vad_mapping.empty()) return t0;
    return vad_map_time(*state, t0);
}
int64_t whisper_full_get_segment_t1_from_state(struct whisper_state * state, int i) {
    const int64_t t1 = SEG(state, i).t1;
    if (!state->has_vad_segments || state->vad_mapping.empty()) return t1;
    int64_t o1 = vad_map_time(*state, t1);
    const int64_t o0 = whisper_full_get_segment_t0_from_state(state, i);
    if (o1 - o0 < 10) o1 = o0 + 10;   // no zero-length segments
    return o1;
}
int64_t whisper_full_get_segment_t0(struct whisper_context * ctx, int i) { return whisper_full_get_segment_t0_from_state(static_cast<whisper_state *>(ctx->state), i); }
int64_t whisper_full_get_segment_t1(struct whisper_context * ctx, int i) { return whisper_full_get_segment_t1_from_state(static_cast<whisper_state *>(ctx->state), i); }
bool whisper_full_get_segment_speaker_turn_next_from_state(struct whisper_state * state, int i) { return SEG(state, i).speaker_turn_next; }
bool whisper_full_get_segment_speaker_turn_next(struct whisper_context * ctx, int i) { return SEG(ctx->state, i).speaker_turn_next; }
const char * whisper_full_get_segment_text_from_state(struct whisper_state * state, int i) { return SEG(state, i).text.c_str(); }
const char * whisper_full_get_segment_text(struct whisper_context * ctx, int i) { return SEG(ctx->state, i).text.c_str(); }
int whisper_full_n_tokens_from_state(struct whisper_state * state, int i) { return (int) SEG(state, i).tokens.size(); }
int whisper_full_n_tokens(struct whisper_context * ctx, int i) { return (int) SEG(ctx->state, i).tokens.size(); }
const char * whisper_full_get_token_text_from_state(struct whisper_context * ctx, struct whisper_state * state, int i, int j) {
    return whisper_token_to_str(ctx, SEG(state, i).tokens[(size_t) j].id);
}
const char * whisper_full_get_token_text(struct whisper_context * ctx, int i, int j) {
    return whisper_token_to_str(ctx, SEG(ctx->state, i).tokens[(size_t) j].id);
}
whisper_token whisper_full_get_token_id_from_state(struct whisper_state * state, int i, int j) { return SEG(state, i).tokens[(size_t) j].id; }
whisper_token whisper_full_get_token_id(struct whisper_context * ctx, int i, int j) { return SEG(ctx->state, i).tokens[(size_t) j].id; }
whisper_token_data whisper_full_get_token_data_from_state(struct whisper_state * state, int i, int j) { return SEG(state, i).tokens[(size_t) j]; }
whisper_token_data whisper_full_get_token_data(struct whisper_context * ctx, int i, int j) { return SEG(ctx->state, i).tokens[(size_t) j]; }
float whisper_full_get_token_p_from_state(struct whisper_state * state, int i, int j) { return SEG(state, i).tokens[(size_t) j].p; }
float whisper_full_get_token_p(struct whisper_context * ctx, int i, int j) { return SEG(ctx->state, i).tokens[(size_t) j].p; }
float whisper_full_get_segment_no_speech_prob_from_state(struct whisper_state * state, int i) { return SEG(state, i).no_speech_prob; }
float whisper_full_get_segment_no_speech_prob(struct whisper_context * ctx, int i) { return SEG(ctx->state, i).no_speech_prob; }
#undef SEG

// ---------------------------------------------------------------------------------------------
// VAD front end and ggml micro-benchmarks: outside the hot path; present for link completeness
// ---------------------------------------------------------------------------------------------
struct whisper_vad_context_params whisper_vad_default_context_params(void) {
    struct whisper_vad_context_params r;
    r.n_threads = 4;
    r.use_gpu = false;
    r.gpu_device = 0;
    return r;
}
// whisper_vad_* : vad_api.cu
int whisper_bench_memcpy(int) { return 0; }
const char * whisper_bench_memcpy_str(int) { return "not available: use bench.py\n"; }
int whisper_bench_ggml_mul_mat(int) { return 0; }
const char * whisper_bench_ggml_mul_mat_str(int) { return "not available: use bench.py\n"; }

// ---------------------------------------------------------------------------------------------
// whisper_b200.h extensions
// ---------------------------------------------------------------------------------------------
int whisper_b200_encode_batch(struct whisper_context * ctx, struct whisper_state ** states, const int * offsets, int n_states) {
    WB_TRY
    if (!ctx || !states || !offsets || n_states <= 0) return -1;
    std::vector<State *> sp(states, states + n_states);
    bool ok = false;
    return sched_run(ctx, [&] { ok = stage_encode(*ctx, sp.data(), offsets, n_states); }) && ok ? 0 : -1;
    WB_CATCH(-1)
}
int whisper_b200_decode_batch(struct whisper_context * ctx, struct whisper_state ** states, const whisper_token * tokens, const int * n_past,
                              int n_states) {
    WB_TRY
    if (!ctx || !states || !tokens || !n_past || n_states <= 0) return 1;
    std::vector<BatchRow> rows;
    for (int i = 0; i < n_states; ++i) {
        rows.push_back(BatchRow{states[i], tokens[i], n_past[i], 0, true});
        states[i]->kv_self.seq_rm(0, n_past[i], -1);
    }
    bool ok = false;
    if (!sched_run(ctx, [&] { ok = stage_decode(*ctx, rows.data(), n_states, false) && stage_fetch_logits(*ctx, rows.data(), n_states); }) || !ok) return 1;
    return 0;
    WB_CATCH(1)
}

int64_t whisper_b200_get_mel(struct whisper_state * state, float * dst, int64_t cap, int * n_len, int * n_len_org, int * n_mel) {
    if (!state || !state->ctx) return -1;
    if (n_len) *n_len = state->n_len;
    if (n_len_org) *n_len_org = state->n_len_org;
    if (n_mel) *n_mel = state->n_mel;
    const int64_t n = (int64_t) state->n_len * state->n_mel;
    if (dst && cap >= n && n > 0) {
        std::lock_guard<std::mutex> lk(state->ctx->mu);
        cudaSetDevice(state->ctx->model.device);
        if (cudaMemcpy(dst, state->d_mel, (size_t) n * 4, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    }
    return n;
}
int64_t whisper_b200_get_embd_enc(struct whisper_context * ctx, int window, float * dst, int64_t cap) {
    if (!ctx || !ctx->enc_ws.base || window < 0 || window >= ctx->enc_ws.windows) return -1;
    const int64_t n = (int64_t) ctx->enc_ws.T * ctx->model.hp.n_audio_state;
    if (dst && cap >= n) {
        std::lock_guard<std::mutex> lk(ctx->mu);
        cudaSetDevice(ctx->model.device);
        if (cudaMemcpy(dst, ctx->enc_ws.enc32 + (size_t) window * n, (size_t) n * 4, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    }
    return n;
}
int64_t whisper_b200_test_conv_stem(struct whisper_context * ctx, struct whisper_state * state, int mel_offset, float * dst, int64_t cap) {
    WB_TRY
    if (!ctx || !state) return -1;
    const int T = state->exp_n_audio_ctx > 0 ? state->exp_n_audio_ctx : ctx->model.hp.n_audio_ctx;
    const int64_t n = (int64_t) T * ctx->model.hp.n_audio_state;
    if (!dst || cap < n) return n;
    bool ok = false;
    return sched_run(ctx, [&] { ok = stage_conv_probe(*ctx, state, mel_offset, dst, cap); }) && ok ? n : -1;
    WB_CATCH(-1)
}
int64_t whisper_b200_get_kv_cross(struct whisper_state * state, uint16_t * dst, int64_t cap) {
    if (!state || !state->ctx || state->cross_T <= 0) return -1;
    const auto & hp = state->ctx->model.hp;
    const int64_t n = (int64_t) hp.n_text_layer * 2 * hp.n_text_head * state->cross_T * 64;
    if (dst && cap >= n) {
        std::lock_guard<std::mutex> lk(state->ctx->mu);
        cudaSetDevice(state->ctx->model.device);
        if (cudaMemcpy(dst, state->kv_cross, (size_t) n * 2, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    }
    return n;
}
int64_t whisper_b200_get_aheads_qk(struct whisper_state * state, float * dst, int64_t cap) {
    if (!state || !state->ctx || !state->d_qk) return -1;
    const int64_t n = (int64_t) state->qk_cap;
    if (dst && cap >= n) {
        std::lock_guard<std::mutex> lk(state->ctx->mu);
        cudaSetDevice(state->ctx->model.device);
        if (cudaMemcpy(dst, state->d_qk, (size_t) n * 4, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    }
    return n;
}
void whisper_b200_get_counters(struct whisper_state * s, int32_t * c, int64_t * us) {
    if (!s) return;
    if (c) {
        c[0] = s->n_sample; c[1] = s->n_encode; c[2] = s->n_decode; c[3] = s->n_batchd; c[4] = s->n_prompt; c[5] = s->n_fail_p; c[6] = s->n_fail_h;
    }
    if (us) {
        us[0] = s->t_sample_us; us[1] = s->t_encode_us; us[2] = s->t_decode_us; us[3] = s->t_batchd_us; us[4] = s->t_prompt_us; us[5] = s->t_mel_us;
    }
}
void whisper_b200_profile(int max_samples_per_category) {
    prof_enable(max_samples_per_category);
    prof_reset();
}
int whisper_b200_profile_read(int category, char * name, int name_cap, double * ms_sum, uint64_t * n_sampled, uint64_t * n_total,
                              double * bytes_sampled, double * flops_sampled) {
    if (category < 0 || category >= PC_COUNT) return -1;
    static double ms[PC_COUNT], by[PC_COUNT], fl[PC_COUNT];
    static uint64_t ns[PC_COUNT], nt[PC_COUNT];
    if (category == 0) prof_collect(ms, ns, nt, by, fl);  // reading category 0 snapshots all of them
    if (name && name_cap > 0) snprintf(name, (size_t) name_cap, "%s", prof_cat_name(category));
    if (ms_sum) *ms_sum = ms[category];
    if (n_sampled) *n_sampled = ns[category];
    if (n_total) *n_total = nt[category];
    if (bytes_sampled) *bytes_sampled = by[category];
    if (flops_sampled) *flops_sampled = fl[category];
    return PC_COUNT;
}
uint64_t whisper_b200_kernel_launches(void) { return gemm_launch_count() + kernel_launch_count() + step_launch_count() + enc_attn_launch_count() + vad_launch_count(); }
void whisper_b200_device_sync(struct whisper_context * ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->model.device);
    cudaStreamSynchronize(ctx->stream);
}

}  // extern "C" (closed around the device helper below)
// pseudo-random f16 fill in [-1, 1) for the micro-benchmarks (constant operands would under-state switching power)
__global__ void bench_fill_f16_kernel(__half * p, size_t n, uint32_t seed) {
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
        uint32_t x = (uint32_t) i * 2654435761u + seed;
        x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
        p[i] = __float2half_rn((float) (int32_t) x * (1.0f / 2147483648.0f));
    }
}
static void bench_fill_f16(void * p, size_t n, uint32_t seed) { bench_fill_f16_kernel<<<1184, 256>>>((__half *) p, n, seed); }
extern "C" {
int whisper_b200_test_gemm(const void * a, const void * b, void * c, int m, int n, int k, int lda, int ldb, int ldc, int c_is_f16,
                           const float * bias, int block_n) {
    WB_TRY
    GemmOperand A{}, B{};
    A.ptr = a; A.rows = m; A.row_stride = lda; A.batch = 1; A.phases = 1;
    B.ptr = b; B.rows = n; B.row_stride = ldb; B.batch = 1; B.phases = 1;
    GemmEpilogue e{};
    e.out = c; e.out_f16 = c_is_f16; e.out_mode = OUT_ROW; e.ldo = ldc; e.batch_div = 1; e.bias = bias;
    if (!gemm_f16_tc(A, B, m, n, k, 1, e, block_n, 0)) return 1;
    return cudaGetLastError() == cudaSuccess ? 0 : 2;
    WB_CATCH(3)
}
// flags: 1 = accumulate into c (f32 only: c += A.B^T + bias), 2 = per-head f16 panels c[n/64][m][64]
int whisper_b200_test_gemm_ex(const void * a, const void * b, void * c, int m, int n, int k, int lda, int ldb, int ldc, int c_is_f16,
                              const float * bias, int block_n, int flags) {
    WB_TRY
    GemmOperand A{}, B{};
    A.ptr = a; A.rows = m; A.row_stride = lda; A.batch = 1; A.phases = 1;
    B.ptr = b; B.rows = n; B.row_stride = ldb; B.batch = 1; B.phases = 1;
    GemmEpilogue e{};
    e.out = c; e.out_f16 = c_is_f16; e.out_mode = (flags & 2) ? OUT_HEAD : OUT_ROW; e.ldo = ldc; e.batch_div = 1; e.bias = bias;
    if (flags & 2) e.head_stride = (int64_t) m * 64;
    if (flags & 1) { e.res = (const float *) c; e.ldr = ldc; }
    if (!gemm_f16_tc(A, B, m, n, k, 1, e, block_n, 0)) return 1;
    return cudaDeviceSynchronize() == cudaSuccess ? 0 : 2;
    WB_CATCH(3)
}

// Encoder-GEMM micro-benchmark: C[m][n] = A[m][k] . B[n][k]^T on scratch device buffers, CUDA-event timed on the launching
// stream; n_weights distinct B matrices are used round-robin.  flags as whisper_b200_test_gemm_ex, plus 4 = bias.
int whisper_b200_bench_gemm(int m, int n, int k, int c_is_f16, int flags, int block_n, int n_weights, int iters, double * us_per_launch) {
    WB_TRY
    __half *a = nullptr, *b = nullptr;
    void * c = nullptr;
    float * bias = nullptr;
    const size_t wsz = (size_t) n * k;
    bool ok = cudaMalloc((void **) &a, (size_t) m * k * 2) == cudaSuccess && cudaMalloc((void **) &b, wsz * n_weights * 2) == cudaSuccess &&
              cudaMalloc(&c, (size_t) m * n * 4) == cudaSuccess && cudaMalloc((void **) &bias, (size_t) n * 4) == cudaSuccess;
    if (ok) {
        bench_fill_f16(a, (size_t) m * k, 1u);
        bench_fill_f16(b, wsz * n_weights, 2u);
        cudaMemset(c, 0, (size_t) m * n * 4);
        cudaMemset(bias, 0, (size_t) n * 4);
    }
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    ok = ok && cudaEventCreate(&e0) == cudaSuccess && cudaEventCreate(&e1) == cudaSuccess;
    auto run = [&](int i) {
        GemmOperand A{}, B{};
        A.ptr = a; A.rows = m; A.row_stride = k; A.batch = 1; A.phases = 1;
        B.ptr = b + wsz * (i % n_weights); B.rows = n; B.row_stride = k; B.batch = 1; B.phases = 1;
        GemmEpilogue e{};
        e.out = c; e.out_f16 = c_is_f16; e.out_mode = (flags & 2) ? OUT_HEAD : OUT_ROW; e.ldo = n; e.batch_div = 1;
        e.bias = (flags & 4) ? bias : nullptr;
        if (flags & 2) e.head_stride = (int64_t) m * 64;
        if (flags & 1) { e.res = (const float *) c; e.ldr = n; }
        return gemm_f16_tc(A, B, m, n, k, 1, e, block_n, 0);
    };
    for (int i = 0; i < 2 && ok; ++i) ok = run(i);
    if (ok) {
        cudaEventRecord(e0, 0);
        for (int i = 0; i < iters && ok; ++i) ok = run(i);
        cudaEventRecord(e1, 0);
        ok = ok && cudaEventSynchronize(e1) == cudaSuccess;
        float ms = 0.0f;
        if (ok) cudaEventElapsedTime(&ms, e0, e1);
        if (us_per_launch) *us_per_launch = 1e3 * ms / std::max(1, iters);
    }
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    cudaFree(a); cudaFree(b); cudaFree(c); cudaFree(bias);
    return ok ? 0 : 1;
    WB_CATCH(3)
}
// fused encoder attention on device buffers: q, k, v f16 [W][H][T][64] (plain layout; repacked here into the kernel's panels)
int whisper_b200_test_enc_attn(const void * q, const void * k, const void * v, void * out, int W, int H, int T, int v_transposed) {
    WB_TRY
    const int R = W * T, Tpad = (int) pad_to(T, 8), S = H * 64;
    __half *qk = nullptr, *vT = nullptr;
    if (cudaMalloc((void **) &qk, (size_t) 3 * H * R * 64 * 2) != cudaSuccess) return 1;
    if (cudaMalloc((void **) &vT, (size_t) H * W * 64 * Tpad * 2) != cudaSuccess) { cudaFree(qk); return 1; }
    cudaMemset(vT, 0, (size_t) H * W * 64 * Tpad * 2);
    std::vector<__half> hq((size_t) W * H * T * 64), hk(hq.size()), hv(hq.size()), pqk((size_t) 3 * H * R * 64), pv((size_t) H * W * 64 * Tpad);
    cudaMemcpy(hq.data(), q, hq.size() * 2, cudaMemcpyDeviceToHost);
    cudaMemcpy(hk.data(), k, hk.size() * 2, cudaMemcpyDeviceToHost);
    cudaMemcpy(hv.data(), v, hv.size() * 2, cudaMemcpyDeviceToHost);
    memset(pv.data(), 0, pv.size() * 2);
    for (int w = 0; w < W; ++w)
        for (int h = 0; h < H; ++h)
            for (int t = 0; t < T; ++t)
                for (int d = 0; d < 64; ++d) {
                    const size_t src = (((size_t) w * H + h) * T + t) * 64 + d;
                    pqk[((size_t) h * R + (size_t) w * T + t) * 64 + d] = hq[src];
                    pqk[((size_t) (H + h) * R + (size_t) w * T + t) * 64 + d] = hk[src];
                    pqk[((size_t) (2 * H + h) * R + (size_t) w * T + t) * 64 + d] = hv[src];
                    pv[(((size_t) h * W + w) * 64 + d) * Tpad + t] = hv[src];
                }
    cudaMemcpy(qk, pqk.data(), pqk.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(vT, pv.data(), pv.size() * 2, cudaMemcpyHostToDevice);
    const bool ok = enc_attention(qk, v_transposed ? vT : nullptr, (__half *) out, W, T, Tpad, H, S, 0.125f, 0) && cudaDeviceSynchronize() == cudaSuccess;
    cudaFree(qk); cudaFree(vT);
    return ok ? 0 : 2;
    WB_CATCH(3)
}
// micro-benchmark of the fused encoder attention: W windows x H heads x T frames; returns microseconds per launch
int whisper_b200_bench_enc_attn(int W, int H, int T, int iters, int v_transposed, double * us_per_launch) {
    WB_TRY
    const int R = W * T, Tpad = (int) pad_to(T, 8), S = H * 64;
    __half *qk = nullptr, *vT = nullptr, *out = nullptr;
    bool ok = cudaMalloc((void **) &qk, (size_t) 3 * H * R * 64 * 2) == cudaSuccess && cudaMalloc((void **) &vT, (size_t) H * W * 64 * Tpad * 2) == cudaSuccess &&
              cudaMalloc((void **) &out, (size_t) R * S * 2) == cudaSuccess;
    if (ok) { bench_fill_f16(qk, (size_t) 3 * H * R * 64, 3u); bench_fill_f16(vT, (size_t) H * W * 64 * Tpad, 4u); }
    const __half * vt_arg = v_transposed ? vT : nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    ok = ok && cudaEventCreate(&e0) == cudaSuccess && cudaEventCreate(&e1) == cudaSuccess;
    for (int i = 0; i < 2 && ok; ++i) ok = enc_attention(qk, vt_arg, out, W, T, Tpad, H, S, 0.125f, 0);
    if (ok) {
        cudaEventRecord(e0, 0);
        for (int i = 0; i < iters && ok; ++i) ok = enc_attention(qk, vt_arg, out, W, T, Tpad, H, S, 0.125f, 0);
        cudaEventRecord(e1, 0);
        ok = ok && cudaEventSynchronize(e1) == cudaSuccess;
        float ms = 0.0f;
        if (ok) cudaEventElapsedTime(&ms, e0, e1);
        if (us_per_launch) *us_per_launch = 1e3 * ms / std::max(1, iters);
    }
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    cudaFree(qk); cudaFree(vT); cudaFree(out);
    return ok ? 0 : 1;
    WB_CATCH(3)
}
int whisper_b200_test_mel(struct whisper_context * ctx, const float * d_pcm, int n_samples, float * d_mel) {
    WB_TRY
    if (!ctx) return 1;
    cudaSetDevice(ctx->model.device);
    static int * d_max = nullptr;
    if (!d_max && cudaMalloc((void **) &d_max, sizeof(int)) != cudaSuccess) return 1;
    MelTables tab{ctx->model.mel_sin, ctx->model.mel_cos, ctx->model.mel_hann, ctx->model.mel_filters_t, ctx->model.filters_n_mel};
    return mel_compute(tab, d_pcm, n_samples, mel_n_len(n_samples), d_mel, d_max, 0) ? 0 : 1;
    WB_CATCH(3)
}
int whisper_b200_test_cross_attn(const void * q, const void * kv, void * out, int n_rows, int n_head, int n_ctx) {
    WB_TRY
    static CrossKvRow * d_rows = nullptr;
    static int cap = 0;
    if (n_rows > cap) {
        cudaFree(d_rows);
        if (cudaMalloc((void **) &d_rows, (size_t) n_rows * sizeof(CrossKvRow)) != cudaSuccess) return 1;
        cap = n_rows;
    }
    {
        std::vector<CrossKvRow> h((size_t) n_rows);
        for (auto & r : h) r = CrossKvRow{(const __half *) kv, nullptr, 0, 0};
        if (cudaMemcpy(d_rows, h.data(), h.size() * sizeof(CrossKvRow), cudaMemcpyHostToDevice) != cudaSuccess) return 1;
    }
    const float scale = (float) pow(64.0, -0.25);
    return cross_attn_decode((const __half *) q, n_head * 64, d_rows, n_rows, n_head, n_ctx, 0, scale, nullptr, (__half *) out, n_head * 64, 0) ? 0 : 1;
    WB_CATCH(3)
}

int whisper_b200_test_gemm_wcluster(const void * w, const void * x, void * out, int n, int n_out, int k, const float * bias, int mode,
                                    float scale, int scale_f1) {
    WB_TRY
    return test_gemm_wcluster((const __half *) w, (const __half *) x, out, n, n_out, k, bias, mode, scale, scale_f1, nullptr) ? 0 : 1;
    WB_CATCH(3)
}
int whisper_b200_test_gemm_wstream(const void * w, const void * x, float * out, int n, int n_out, int k, const float * bias, int force_split) {
    WB_TRY
    return test_gemm_wstream((const __half *) w, (const __half *) x, out, n, n_out, k, bias, force_split) ? 0 : 1;
    WB_CATCH(3)
}
int whisper_b200_test_cross_attn_step(const void * q, const void * kv, void * out, int n_rows, int n_head, int n_ctx) {
    WB_TRY
    return test_cross_attn_step((const __half *) q, (const __half *) kv, (__half *) out, n_rows, n_head, n_ctx) ? 0 : 1;
    WB_CATCH(3)
}
// Host only: the loader's expansion of a quantised ggml tensor (model.cpp: dequantize_blocks) on caller data.  0 = done, 1 = type not
// expanded by this loader or n not a whole number of blocks.
int whisper_b200_test_dequantize(int ggml_type, const void * blocks, int64_t n_elements, float * dst) {
    WB_TRY
    return dequantize_tensor(ggml_type, blocks, n_elements, dst) ? 0 : 1;
    WB_CATCH(1)
}
// Host only (tests): the self-attention KV cache's cell bookkeeping (engine.cu: KvCache) driven by a script of operations; same
// encoding as oracle/ref_probe.cpp: ref_probe_kv_script.
int whisper_b200_test_kv_script(int size, const int * ops, int n_ops, int * per_op, int * cells_out) {
    WB_TRY
    if (size <= 0 || !ops || n_ops < 0 || !per_op || !cells_out) return -1;
    KvCache kv;
    kv.size = size;
    kv.cells.assign((size_t) size, KvCell{});
    for (int i = 0; i < n_ops; ++i) {
        const int * o = ops + 5 * i;
        int ret = 0;
        if (o[0] == 0) {
            kv.clear();
        } else if (o[0] == 1) {
            std::vector<int> pos((size_t) o[1]), seq((size_t) o[1], o[2]);
            for (int k = 0; k < o[1]; ++k) pos[(size_t) k] = o[3] + k;
            ret = kv.find_slot(pos.data(), seq.data(), o[1]) ? 1 : 0;
        } else if (o[0] == 2) {
            kv.seq_rm(o[1], o[2], o[3]);
        } else if (o[0] == 3) {
            kv.seq_cp(o[1], o[2], o[3], o[4]);
        } else if (o[0] == 4) {
            ret = kv.cell_max();
        }
        per_op[3 * i] = ret;
        per_op[3 * i + 1] = kv.head;
        per_op[3 * i + 2] = kv.cell_max();
    }
    for (int c = 0; c < size; ++c) {
        cells_out[2 * c] = kv.cells[(size_t) c].pos;
        cells_out[2 * c + 1] = (int) kv.cells[(size_t) c].seq;
    }
    return 0;
    WB_CATCH(-1)
}
int whisper_b200_test_host_tables(uint16_t * gelu_f16, float * sin400, float * cos400, float * hann400) {
    host_tables(gelu_f16, sin400, cos400, hann400);
    return 0;
}
int whisper_b200_test_aheads_preset(int preset, int * pairs, int cap) { return alignment_heads_preset(preset, pairs, cap); }
int whisper_b200_gemm_wstream_noalloc(const void * w, const void * x, float * out, int n, int n_out, int k, const float * bias, int split) {
    WB_TRY
    return gemm_wstream_noalloc((const __half *) w, (const __half *) x, out, n, n_out, k, bias, split);
    WB_CATCH(1)
}
int whisper_b200_bench_kernel(const char * kind, int n_rows, int a, int b, int c, int iters, double * us_per_launch, double * bytes_per_launch) {
    WB_TRY
    StepBenchResult r{};
    bool ok = false;
    if (strcmp(kind, "cross_attn_step") == 0) ok = bench_cross_attn_step(n_rows, a /*H*/, b /*T*/, c /*layers*/, iters, &r);
    else if (strcmp(kind, "gemm_wstream") == 0) ok = bench_gemm_wstream(n_rows, a /*N_out*/, b /*K*/, c /*n_weights*/, iters, &r);
    else if (strcmp(kind, "gemm_wcluster") == 0) ok = bench_gemm_wcluster(n_rows, a /*N_out*/, b /*K*/, c /*n_weights*/, iters, &r);
    else if (strcmp(kind, "overlap_gemm") == 0) {
        ok = bench_overlap_gemm(n_rows, a /*N_out*/, b /*K*/, c /*n_weights*/, iters, &r);
        if (ok && us_per_launch) us_per_launch[1] = 1e-3 * r.launches;  // background cross-attention us per launch
    }
    if (!ok) return 1;
    if (us_per_launch) *us_per_launch = r.us_per_launch;
    if (bytes_per_launch) *bytes_per_launch = r.bytes_per_launch;
    return 0;
    WB_CATCH(3)
}


// Grammar engine in isolation (tests): start a grammar on `rules`, accept the tokens accept[0..n_accept), report the reject mask of the
// next step (mask_out[n_vocab]).  Returns the number of live stacks.  Host logic only; the context supplies the vocabulary.
int whisper_b200_test_grammar(struct whisper_context * ctx, const whisper_grammar_element ** rules, size_t n_rules, size_t i_start_rule,
                              const whisper_token * accept, int n_accept, uint8_t * mask_out) {
    WB_TRY
    if (!ctx || !rules || !mask_out) return -1;
    const Vocab & vc = ctx->model.vocab;
    Grammar g = Grammar::init(rules, n_rules, i_start_rule);
    for (int i = 0; i < n_accept; ++i)
        if (accept[i] >= 0 && accept[i] < (int) vc.id_to_token.size()) g.accept_token(vc.id_to_token[(size_t) accept[i]]);
    std::vector<uint8_t> mask;
    g.reject_mask(vc.id_to_token, vc.token_eot, mask);
    memcpy(mask_out, mask.data(), std::min(mask.size(), (size_t) vc.n_vocab));
    return (int) g.stacks.size();
    WB_CATCH(-1)
}

// The same on a caller-supplied vocabulary (token_texts[n_vocab], NUL-terminated): no context, no device; the CPU suite checks the
// automaton against the reference with it.
int whisper_b200_test_grammar_vocab(const char * const * token_texts, int n_vocab, int token_eot, const whisper_grammar_element ** rules,
                                    size_t n_rules, size_t i_start_rule, const whisper_token * accept, int n_accept, uint8_t * mask_out) {
    WB_TRY
    if (!token_texts || n_vocab <= 0 || !rules || !mask_out) return -1;
    std::vector<std::string> id_to_token((size_t) n_vocab);
    for (int i = 0; i < n_vocab; ++i) id_to_token[(size_t) i] = token_texts[i] ? token_texts[i] : "";
    Grammar g = Grammar::init(rules, n_rules, i_start_rule);
    for (int i = 0; i < n_accept; ++i)
        if (accept[i] >= 0 && accept[i] < n_vocab) g.accept_token(id_to_token[(size_t) accept[i]]);
    std::vector<uint8_t> mask;
    g.reject_mask(id_to_token, token_eot, mask);
    memcpy(mask_out, mask.data(), std::min(mask.size(), (size_t) n_vocab));
    return (int) g.stacks.size();
    WB_CATCH(-1)
}

// X1/X2 on caller-supplied HOST data (tests): qk [n_heads][n_tok][T] softmax rows -> cost [n_tok][n_audio] (host copy optional) and, over the
// rows sot_len .. n_tok-2, the DTW path.  Returns the path length L (path_out[0..L) token indices, [L..2L) time indices), < 0 on failure.
int whisper_b200_test_dtw(const float * qk, int n_heads, int n_tok, int T, int n_audio, int medfilt_width, int sot_len, float * cost_out,
                          int32_t * path_out, int cap) {
    WB_TRY
    const int N = n_tok - sot_len - 1;
    if (!qk || N <= 0 || n_audio <= medfilt_width || n_audio > T) return -1;
    float *d_qk = nullptr, *d_nrm = nullptr, *d_cost = nullptr, *d_diag = nullptr;
    int8_t * d_trace = nullptr;
    int * d_path = nullptr;
    const size_t capp = (size_t) N + n_audio;
    std::vector<int> path(2 * capp + 1);
    bool ok = cudaMalloc((void **) &d_qk, (size_t) n_heads * n_tok * T * 4) == cudaSuccess &&
              cudaMalloc((void **) &d_nrm, (size_t) n_heads * n_tok * n_audio * 4) == cudaSuccess &&
              cudaMalloc((void **) &d_cost, (size_t) n_tok * n_audio * 4) == cudaSuccess &&
              cudaMalloc((void **) &d_diag, (size_t) 3 * (N + 1) * 4) == cudaSuccess &&
              cudaMalloc((void **) &d_trace, (size_t) (N + 1) * (n_audio + 1)) == cudaSuccess &&
              cudaMalloc((void **) &d_path, (2 * capp + 1) * 4) == cudaSuccess &&
              cudaMemcpy(d_qk, qk, (size_t) n_heads * n_tok * T * 4, cudaMemcpyHostToDevice) == cudaSuccess;
    if (ok) ok = dtw_cost(d_qk, n_heads, n_tok, T, n_audio, medfilt_width, d_nrm, d_cost, nullptr);
    if (ok) ok = dtw_path(d_cost + (size_t) sot_len * n_audio, N, n_audio, d_diag, d_trace, d_path, d_path + 2 * capp, nullptr);
    if (ok) ok = cudaMemcpy(path.data(), d_path, (2 * capp + 1) * 4, cudaMemcpyDeviceToHost) == cudaSuccess;
    if (ok && cost_out) ok = cudaMemcpy(cost_out, d_cost, (size_t) n_tok * n_audio * 4, cudaMemcpyDeviceToHost) == cudaSuccess;
    cudaFree(d_qk); cudaFree(d_nrm); cudaFree(d_cost); cudaFree(d_diag); cudaFree(d_trace); cudaFree(d_path);
    if (!ok) return -2;
    const int L = path[2 * capp];
    if (path_out && cap >= 2 * L) {
        memcpy(path_out, path.data() + (capp - (size_t) L), (size_t) L * 4);
        memcpy(path_out + L, path.data() + capp + (capp - (size_t) L), (size_t) L * 4);
    }
    return L;
    WB_CATCH(-3)
}
// dtw_and_backtrace alone (wc:8647-8731) on a HOST cost matrix cost[N][M] (time index contiguous)
int whisper_b200_test_dtw_path(const float * cost, int N, int M, int32_t * path_out, int cap) {
    WB_TRY
    if (!cost || N <= 0 || M <= 0) return -1;
    float *d_cost = nullptr, *d_diag = nullptr;
    int8_t * d_trace = nullptr;
    int * d_path = nullptr;
    const size_t capp = (size_t) N + M;
    std::vector<int> path(2 * capp + 1);
    bool ok = cudaMalloc((void **) &d_cost, (size_t) N * M * 4) == cudaSuccess && cudaMalloc((void **) &d_diag, (size_t) 3 * (N + 1) * 4) == cudaSuccess &&
              cudaMalloc((void **) &d_trace, (size_t) (N + 1) * (M + 1)) == cudaSuccess &&
              cudaMalloc((void **) &d_path, (2 * capp + 1) * 4) == cudaSuccess &&
              cudaMemcpy(d_cost, cost, (size_t) N * M * 4, cudaMemcpyHostToDevice) == cudaSuccess;
    if (ok) ok = dtw_path(d_cost, N, M, d_diag, d_trace, d_path, d_path + 2 * capp, nullptr);
    if (ok) ok = cudaMemcpy(path.data(), d_path, (2 * capp + 1) * 4, cudaMemcpyDeviceToHost) == cudaSuccess;
    cudaFree(d_cost); cudaFree(d_diag); cudaFree(d_trace); cudaFree(d_path);
    if (!ok) return -2;
    const int L = path[2 * capp];
    if (path_out && cap >= 2 * L) {
        memcpy(path_out, path.data() + (capp - (size_t) L), (size_t) L * 4);
        memcpy(path_out + L, path.data() + capp + (capp - (size_t) L), (size_t) L * 4);
    }
    return L;
    WB_CATCH(-3)
}

}  // extern "C"
