// F1: the transcription control loop (whisper_full_with_state, wc:6795-7711), restated as a resumable per-state
// job so that any number of states advance in lock-step: every iteration gathers the decoder rows of all jobs
// (prompt rows or one row per live decoder), runs ONE batched decoder forward + ONE sampler launch on the
// device, and lets each job do its host-side bookkeeping.  A single whisper_full_with_state call is a batch of 1.
//
// Host work here is integer/bookkeeping only: the state machine of wc:7297-7379, sequence ranking (wc:6567-6613),
// temperature fallback (wc:7484-7544), segment emission (wc:7569-7676), prompt carry-over.  Token picking at
// temperature 0 happens in the sampler kernel; temperature > 0 and beam search draw from the device-computed
// distributions with the same std::mt19937 / std::discrete_distribution sequence as the reference (wc:6476, 6549).
#include <math.h>
#include <string.h>

#include <algorithm>
#include <condition_variable>
#include <deque>
#include <functional>
#include <thread>
#include <map>
#include <regex>

#include "../../include/whisper_b200.h"
#include "engine.h"
#include "log.h"
#include "step.cuh"

namespace wb {

// --------------------------------------------------------------------------------------------
// suppress mask (the static part of whisper_process_logits, wc:6195-6260)
// --------------------------------------------------------------------------------------------
static const char * const k_non_speech_tokens[] = {
    "\"", "#", "(", ")", "*", "+", "/", ":", ";", "<", "=", ">", "@", "[", "\\", "]", "^", "_", "`", "{", "|", "}", "~",
    "\xe3\x80\x8c", "\xe3\x80\x8d", "\xe3\x80\x8e", "\xe3\x80\x8f", "<<", ">>", "<<<", ">>>", "--", "---", "-(", "-[", "('", "(\"", "((",
    "))", "(((", ")))", "[[", "]]", "{{", "}}", "\xe2\x99\xaa\xe2\x99\xaa", "\xe2\x99\xaa\xe2\x99\xaa\xe2\x99\xaa", "\xe2\x99\xa9",
    "\xe2\x99\xaa", "\xe2\x99\xab", "\xe2\x99\xac", "\xe2\x99\xad", "\xe2\x99\xae", "\xe2\x99\xaf"};

SuppressMask::~SuppressMask() {
    if (d) {
        cudaSetDevice(device);
        cudaFree(d);
    }
}

// Called on the scheduler thread only (the per-context cache is not locked).
std::shared_ptr<SuppressMask> suppress_mask(Context & ctx, const whisper_full_params & params) {
    const std::string key = std::string(params.tdrz_enable ? "1" : "0") + (params.suppress_nst ? "1" : "0") + (params.suppress_regex ? params.suppress_regex : "\x01");
    auto hit = ctx.masks.find(key);
    if (hit != ctx.masks.end()) return hit->second;
    const Vocab & vc = ctx.model.vocab;
    const int V = vc.n_vocab;
    std::vector<uint8_t> mk((size_t) V, 0);
    auto kill = [&](int id) {
        if (id >= 0 && id < V) mk[(size_t) id] = 1;
    };
    kill(vc.token_not);
    kill(vc.token_sot);
    kill(vc.token_nosp);
    if (!params.tdrz_enable) kill(vc.token_solm);
    kill(vc.token_translate);
    kill(vc.token_transcribe);
    kill(vc.token_prev);
    for (int i = 0; i < 100; ++i) kill(vc.token_sot + 1 + i);  // whisper_token_lang over the whole language table (wc:6219-6221)
    auto m = std::make_shared<SuppressMask>();
    m->device = ctx.model.device;
    m->pre_host = mk;  // everything above precedes logits_filter_callback in the reference (wc:6226-6228); what follows comes after it
    if (params.suppress_regex != nullptr) {
        try {
            std::regex re(params.suppress_regex);
            for (const auto & kv : vc.token_to_id)
                if (std::regex_match(kv.first, re)) kill(kv.second);
        } catch (const std::regex_error &) {
            log_error("%s: invalid suppress_regex\n", __func__);
        }
    }
    if (params.suppress_nst) {
        for (const char * t : k_non_speech_tokens) {
            const std::string a = t, b = " " + a;
            auto ia = vc.token_to_id.find(a);
            if (ia != vc.token_to_id.end()) kill(ia->second);
            auto ib = vc.token_to_id.find(b);
            if (ib != vc.token_to_id.end()) kill(ib->second);
        }
        auto i1 = vc.token_to_id.find(" -");
        if (i1 != vc.token_to_id.end()) kill(i1->second);
        auto i2 = vc.token_to_id.find(" '");
        if (i2 != vc.token_to_id.end()) kill(i2->second);
    }
    std::vector<uint8_t> post((size_t) V);
    for (int i = 0; i < V; ++i) post[(size_t) i] = mk[(size_t) i] && !m->pre_host[(size_t) i];
    if (cudaMalloc((void **) &m->d, 2 * (size_t) V) != cudaSuccess) return nullptr;
    if (cudaMemcpyAsync(m->d, mk.data(), (size_t) V, cudaMemcpyHostToDevice, ctx.stream) != cudaSuccess ||
        cudaMemcpyAsync(m->d + V, post.data(), (size_t) V, cudaMemcpyHostToDevice, ctx.stream) != cudaSuccess ||
        cudaStreamSynchronize(ctx.stream) != cudaSuccess)
        return nullptr;
    if (ctx.masks.size() >= 64) ctx.masks.clear();  // a caller cycling through regexes: jobs keep their masks alive through the shared_ptr
    ctx.masks[key] = m;
    return m;
}

static SamplerStatic sampler_static(Context & ctx) {
    const Vocab & vc = ctx.model.vocab;
    SamplerStatic s{};
    s.suppress = nullptr;       // every sampler row carries its job's masks (SamplerRow::suppress)
    s.suppress_post = nullptr;
    s.n_vocab = vc.n_vocab;
    s.token_eot = vc.token_eot;
    s.token_beg = vc.token_beg;
    auto it = vc.token_to_id.find(" ");
    s.token_space = it != vc.token_to_id.end() ? it->second : -1;
    s.token_nosp = vc.token_nosp;
    return s;
}

// --------------------------------------------------------------------------------------------
// host-side sampling from device-computed distributions (wc:6432-6564)
// --------------------------------------------------------------------------------------------
static void ts_stats(const Vocab & vc, const std::vector<float> & probs, whisper_token & tid, float & pt, float & ptsum) {
    double sum_ts = 0.0, max_ts = 0.0;
    for (int i = vc.token_beg; i < vc.n_vocab; i++) {
        if (probs[i] == -INFINITY) continue;
        sum_ts += probs[i];
        if (max_ts < probs[i]) {
            max_ts = probs[i];
            tid = i;
        }
    }
    pt = (float) (max_ts / (sum_ts + 1e-10));
    ptsum = (float) sum_ts;
}

static whisper_token_data sample_token_host(const Vocab & vc, Decoder & d) {
    whisper_token_data r = {0, 0, 0.0f, 0.0f, 0.0f, 0.0f, -1, -1, -1, 0.0f};
    ts_stats(vc, d.probs, r.tid, r.pt, r.ptsum);
    std::discrete_distribution<> dist(d.probs.begin(), d.probs.end());
    r.id = dist(d.rng);
    r.p = d.probs[r.id];
    r.plog = d.logprobs[r.id];
    if (r.id >= vc.token_beg) {
        r.tid = r.id;
        r.pt = r.p;
    }
    return r;
}

static std::vector<whisper_token_data> sample_token_topk_host(const Vocab & vc, Decoder & d, int k) {
    std::vector<whisper_token_data> result;
    result.reserve(k);
    whisper_token tid = vc.token_beg;
    float pt = 0.0f, ptsum = 0.0f;
    ts_stats(vc, d.probs, tid, pt, ptsum);
    std::discrete_distribution<> dist(d.probs.begin(), d.probs.end());
    for (int i = 0; i < k; ++i) {
        const int id = dist(d.rng);
        result.push_back({id, tid, d.probs[id], d.logprobs[id], pt, ptsum, -1, -1, -1, 0.0f});
        if (result[i].id >= vc.token_beg) {
            result[i].tid = result[i].id;
            result[i].pt = result[i].p;
        }
    }
    return result;
}

static whisper_token_data token_from_pick(const SamplerOut & o) {
    whisper_token_data r = {o.id, o.tid, o.p, o.plog, o.pt, o.ptsum, -1, -1, -1, 0.0f};
    return r;
}

// wc:6567-6613
static void sequence_score(const whisper_full_params & params, Sequence & sq) {
    if (sq.result_len == 0) return;
    double result = 0.0;
    for (int i = 0; i < sq.result_len; ++i) result += sq.tokens[i].plog;
    sq.sum_logprobs = result;
    sq.avg_logprobs = result / sq.result_len;
    double penalty = sq.result_len;
    if (params.length_penalty > 0.0f) penalty = pow((5.0 + penalty) / 6.0, params.length_penalty);
    sq.score = result / penalty;
    const int n = 32;
    int cnt = 0;
    double entropy = 0.0;
    std::map<whisper_token, int> counts;
    for (int i = std::max(0, sq.result_len - n); i < sq.result_len; ++i) {
        counts[sq.tokens[i].id]++;
        cnt++;
    }
    for (const auto & kv : counts) {
        const double p = kv.second / (double) cnt;
        entropy -= p * log(p);
    }
    sq.entropy = entropy;
}

static bool sequences_equal(const Sequence & a, const Sequence & b) {
    if (a.tokens.size() != b.tokens.size()) return false;
    for (int i = (int) a.tokens.size() - 1; i >= 0; i--)
        if (a.tokens[i].id != b.tokens[i].id) return false;
    return true;
}

static const char * token_str(const Vocab & vc, int id) {
    if (id < 0 || id >= (int) vc.id_to_token.size()) return "";
    return vc.id_to_token[(size_t) id].c_str();
}

// F5 (defined below): heuristic token times of one segment, wrapping of the newest segment
static void heuristic_token_timestamps(Context * ctx, State * st, int i_segment, float thold_pt, float thold_ptsum);
static int wrap_last_segment(Context * ctx, State * st, int max_len, bool split_on_word);

// --------------------------------------------------------------------------------------------
// the per-state job
// --------------------------------------------------------------------------------------------
struct BeamCandidate {
    int decoder_idx;
    int seek_delta;
    bool has_ts;
    Sequence sequence;
    Grammar grammar;   // wc:7201
};

struct Client;
struct FullJob;
// runs `fn` on the thread that submitted the job (blocked inside its whisper_full* call): user callbacks fire there (wh:461-480)
static void call_owner(FullJob * job, const std::function<void()> & fn, bool is_logits_filter = false);

struct FullJob {
    enum Phase { WINDOW_BEGIN, TEMP_BEGIN, AFTER_PROMPT, STEP, DONE };

    Context * ctx = nullptr;
    Client * client = nullptr;                 // who waits for this job
    std::shared_ptr<SuppressMask> mask;        // this job's static suppress rules
    int lane = -1;                             // decode lane the job currently belongs to
    bool admitted = false;
    int64_t t_submit = 0;
    State * st = nullptr;
    whisper_full_params params{};
    const float * samples = nullptr;
    int n_samples = 0;

    int ret = 0;
    Phase phase = WINDOW_BEGIN;

    int seek_start = 0, seek_end = 0, seek = 0;
    std::vector<float> temperatures;
    int n_decoders = 1;
    std::vector<whisper_token> prompt_init, prompt, prompt_tokens_buf;

    int it = 0, n_decoders_cur = 1, best_decoder_id = 0, i = 0, n_max = 0;
    float t_cur = 0.0f;
    bool host_sampling = false;

    std::vector<std::vector<BeamCandidate>> bc_per_dec;
    std::vector<BeamCandidate> beam_candidates;

    // contribution to the current device iteration
    std::vector<BatchRow> rows;
    std::vector<SamplerRow> srows;
    std::vector<Decoder *> sdecs;

    static constexpr int delta_min = 10;

    // grammar rows take the general decoder path: their reject mask is computed on the host every step and uploaded with the sampler row
    bool has_grammar() const { return params.grammar_rules != nullptr && params.n_grammar_rules > 0; }
    whisper_context * wctx() const { return static_cast<whisper_context *>(ctx); }
    whisper_state * wstate() const { return static_cast<whisper_state *>(st); }
    const Vocab & vocab() const { return ctx->model.vocab; }

    void finish(int r) {
        ret = r;
        phase = DONE;
        rows.clear();
        srows.clear();
        sdecs.clear();
    }

    bool init();
    bool window_begin();   // true: this job needs its window encoded at `seek`
    bool aborted();
    void after_encode();
    void temp_begin();
    void after_device();
    void after_prompt();
    void step();
    void finish_temperature();
    void emit_results();
    SamplerRow sampler_row(const Decoder & d, int logits_row, bool want_nosp) const;
    void refresh_grammar_mask(Decoder & d);
    void queue_step_rows();
};

// everything of wc:6801-6984 after the mel computation
bool FullJob::init() {
    const Vocab & vc = vocab();
    st->result_all.clear();

    // language detection (wc:6815-6830)
    if (params.language == nullptr || strlen(params.language) == 0 || strcmp(params.language, "auto") == 0 || params.detect_language) {
        std::vector<float> probs(100, 0.0f);
        const int lid = lang_auto_detect(ctx, st, 0, probs.data());
        if (lid < 0) {
            log_error("%s: failed to auto-detect language\n", __func__);
            finish(-3);
            return false;
        }
        st->lang_id = lid;
        params.language = g_langs[lid].code;
        log_msg(GGML_LOG_LEVEL_INFO, "%s: auto-detected language: %s (p = %f)\n", __func__, params.language, probs[lid]);
        if (params.detect_language) {
            finish(0);
            return false;
        }
    }
    if (params.token_timestamps) {  // wc:6832-6839
        st->t_beg = 0;
        st->t_last = 0;
        st->tid_last = 0;
        if (n_samples > 0) {
            // moving average of |pcm| over 65 samples, computed on the device copy of the PCM (the caller's pointer may be device
            // memory in the batch entry point) and brought to the host, where the heuristic's scalar walks run
            float * d_energy = nullptr;
            st->energy.assign((size_t) n_samples, 0.0f);
            const bool ok = cudaMalloc((void **) &d_energy, (size_t) n_samples * sizeof(float)) == cudaSuccess &&
                            signal_energy(st->d_pcm, n_samples, 32, d_energy, ctx->stream) &&
                            cudaMemcpyAsync(st->energy.data(), d_energy, (size_t) n_samples * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream) == cudaSuccess &&
                            cudaStreamSynchronize(ctx->stream) == cudaSuccess;
            cudaFree(d_energy);
            if (!ok) {
                log_error("%s: failed to compute the signal energy\n", __func__);
                finish(-2);
                return false;
            }
        }
    }
    seek_start = params.offset_ms / 10;
    seek_end = params.duration_ms == 0 ? st->n_len_org : seek_start + params.duration_ms / 10;
    if (seek_end < seek_start + delta_min) {
        log_msg(GGML_LOG_LEVEL_WARN, "%s: input is too short - %d ms < 100 ms. consider padding the input audio with silence\n", __func__,
                (seek_end - seek_start) * 10);
        finish(0);
        return false;
    }
    temperatures.clear();
    if (params.temperature_inc > 0.0f) {
        for (float t = params.temperature; t < 1.0f + 1e-6f; t += params.temperature_inc) temperatures.push_back(t);
    } else {
        temperatures.push_back(params.temperature);
    }
    n_decoders = 1;
    switch (params.strategy) {
        case WHISPER_SAMPLING_GREEDY: n_decoders = params.greedy.best_of; break;
        case WHISPER_SAMPLING_BEAM_SEARCH: n_decoders = std::max(params.greedy.best_of, params.beam_search.beam_size); break;
    }
    n_decoders = std::max(1, n_decoders);
    if (n_decoders > MAX_DECODERS) {
        log_error("%s: too many decoders requested (%d), max = %d\n", __func__, n_decoders, MAX_DECODERS);
        finish(-4);
        return false;
    }
    for (int j = 1; j < n_decoders; j++) st->decoders[j].rng = std::mt19937(j);  // wc:6897

    auto & prompt_past = st->prompt_past;
    if (params.no_context) prompt_past.clear();
    // initial prompt (wc:6907-6931)
    if (!params.prompt_tokens && params.initial_prompt) {
        prompt_tokens_buf.resize(1024);
        int n_needed = whisper_tokenize(wctx(), params.initial_prompt, prompt_tokens_buf.data(), (int) prompt_tokens_buf.size());
        if (n_needed < 0) {
            prompt_tokens_buf.resize((size_t) -n_needed);
            n_needed = whisper_tokenize(wctx(), params.initial_prompt, prompt_tokens_buf.data(), (int) prompt_tokens_buf.size());
        }
        prompt_tokens_buf.resize((size_t) std::max(0, n_needed));
        params.prompt_tokens = prompt_tokens_buf.data();
        params.prompt_n_tokens = (int) prompt_tokens_buf.size();
    }
    if (params.prompt_tokens && params.prompt_n_tokens > 0) {
        for (int k = 0; k < params.prompt_n_tokens; k++) prompt_past.push_back(params.prompt_tokens[k]);
        std::rotate(prompt_past.begin(), prompt_past.end() - params.prompt_n_tokens, prompt_past.end());
    }
    if (params.audio_ctx > ctx->model.hp.n_audio_ctx) {
        log_error("%s: audio_ctx is larger than the maximum allowed (%d > %d)\n", __func__, params.audio_ctx, ctx->model.hp.n_audio_ctx);
        finish(-5);
        return false;
    }
    st->exp_n_audio_ctx = params.audio_ctx;

    prompt_init = {vc.token_sot};
    if (vc.is_multilingual()) {
        const int lid = lang_id(params.language);
        if (lid < 0) log_error("%s: unknown language '%s'\n", "whisper_lang_id", params.language);
        st->lang_id = lid;
        prompt_init.push_back(vc.token_sot + 1 + lid);
        prompt_init.push_back(params.translate ? vc.token_translate : vc.token_transcribe);
    }
    {
        const bool is_distil = ctx->model.hp.n_text_layer == 2 && ctx->model.hp.n_vocab != 51866;
        if (is_distil && !params.no_timestamps) {
            log_msg(GGML_LOG_LEVEL_WARN, "%s: using first release distilled models - forcing no_timestamps\n", __func__);
            params.no_timestamps = true;
        }
    }
    if (params.no_timestamps) prompt_init.push_back(vc.token_not);

    seek = seek_start;
    prompt.reserve((size_t) ctx->model.hp.n_text_ctx);
    bc_per_dec.assign((size_t) n_decoders, {});
    n_max = ctx->model.hp.n_text_ctx / 2 - 4;
    phase = WINDOW_BEGIN;
    return true;
}

// head of the main loop (wc:6986-7004)
bool FullJob::window_begin() {
    if (params.progress_callback) {
        const int progress_cur = (100 * (seek - seek_start)) / (seek_end - seek_start);
        call_owner(this, [&] { params.progress_callback(wctx(), wstate(), progress_cur, params.progress_callback_user_data); });
    }
    if (seek + delta_min >= seek_end) {
        finish(0);
        return false;
    }
    if (params.encoder_begin_callback) {
        bool go = true;
        call_owner(this, [&] { go = params.encoder_begin_callback(wctx(), wstate(), params.encoder_begin_callback_user_data); });
        if (go == false) {
            log_error("%s: encoder_begin_callback returned false - aborting\n", __func__);
            finish(0);
            return false;
        }
    }
    return true;
}

bool FullJob::aborted() {
    if (!params.abort_callback) return false;
    bool stop = false;
    call_owner(this, [&] { stop = params.abort_callback(params.abort_callback_user_data); });
    return stop;
}

void FullJob::after_encode() {
    if (aborted()) {  // wc:2471
        log_error("%s: failed to encode\n", __func__);
        finish(-6);
        return;
    }
    if (seek > seek_start && seek + 500 >= seek_end) st->prompt_past.clear();  // wc:7014-7016
    best_decoder_id = 0;
    it = 0;
    phase = TEMP_BEGIN;
}

SamplerRow FullJob::sampler_row(const Decoder & d, int logits_row, bool want_nosp) const {
    const Vocab & vc = vocab();
    const auto & tk = d.sequence.tokens;
    SamplerRow r{};
    r.logits_row = logits_row;
    r.temperature = t_cur;
    r.is_initial = tk.empty() ? 1 : 0;
    r.suppress_blank = params.suppress_blank ? 1 : 0;
    r.no_timestamps = params.no_timestamps ? 1 : 0;
    r.last_was_ts = (!tk.empty() && tk.back().id >= vc.token_beg) ? 1 : 0;
    r.penult_was_ts = (tk.size() < 2 || tk[tk.size() - 2].id >= vc.token_beg) ? 1 : 0;
    r.max_initial_tid = vc.n_vocab;
    if (tk.empty() && params.max_initial_ts > 0.0f) {  // wc:6285-6292
        const float precision = float(WHISPER_CHUNK_SIZE) / ctx->model.hp.n_audio_ctx;
        const int tid0 = (int) std::round(params.max_initial_ts / precision);
        r.max_initial_tid = vc.token_beg + tid0 + 1;
    }
    r.has_ts = d.has_ts ? 1 : 0;
    r.tid0 = d.seek_delta / 2;
    r.want_nosp = want_nosp ? 1 : 0;
    r.work_slot = logits_row;
    r.suppress = mask->d;
    r.suppress_post = mask->d + vc.n_vocab;
    r.grammar_reject = nullptr;  // filled in by stage_sample from Decoder::grammar_reject
    r.grammar_penalty = params.grammar_penalty;
    return r;
}

// which tokens the decoder's grammar rejects at the step about to be sampled (whisper_suppress_invalid_grammar, wc:5824-5866);
// empty when the job has no grammar
void FullJob::refresh_grammar_mask(Decoder & d) {
    if (params.n_grammar_rules > 0 && d.grammar.active()) d.grammar.reject_mask(vocab().id_to_token, vocab().token_eot, d.grammar_reject);
    else d.grammar_reject.clear();
}

// start of one temperature pass (wc:7020-7155): decoder reset, prompt assembly, KV reset, prompt rows
void FullJob::temp_begin() {
    const Vocab & vc = vocab();
    const auto & hp = ctx->model.hp;
    t_cur = temperatures[(size_t) it];
    n_decoders_cur = 1;
    switch (params.strategy) {
        case WHISPER_SAMPLING_GREEDY:
            if (t_cur > 0.0f) n_decoders_cur = params.greedy.best_of;
            break;
        case WHISPER_SAMPLING_BEAM_SEARCH:
            n_decoders_cur = t_cur > 0.0f ? params.greedy.best_of : params.beam_search.beam_size;
            break;
    }
    n_decoders_cur = std::max(1, n_decoders_cur);
    host_sampling = params.strategy == WHISPER_SAMPLING_BEAM_SEARCH || !(t_cur < 1e-6f);

    for (int j = 0; j < n_decoders_cur; ++j) {
        Decoder & d = st->decoders[j];
        d.sequence.tokens.clear();
        d.sequence.result_len = 0;
        d.sequence.sum_logprobs_all = 0.0;
        d.sequence.sum_logprobs = -INFINITY;
        d.sequence.avg_logprobs = -INFINITY;
        d.sequence.entropy = 0.0;
        d.sequence.score = -INFINITY;
        d.seek_delta = 100 * WHISPER_CHUNK_SIZE;
        d.failed = false;
        d.completed = false;
        d.has_ts = false;
        d.want_dists = host_sampling;
        if (params.grammar_rules != nullptr && params.n_grammar_rules > 0)   // wc:7064-7068
            d.grammar = Grammar::init(params.grammar_rules, params.n_grammar_rules, params.i_start_rule);
        else
            d.grammar = Grammar();
        d.grammar_reject.clear();
    }
    prompt.clear();
    auto & prompt_past = st->prompt_past;
    if (!prompt_past.empty() && t_cur < 0.5f && params.n_max_text_ctx > 0) {
        const int n_take = std::min(std::min(params.n_max_text_ctx, hp.n_text_ctx / 2), int(prompt_past.size()));
        prompt = {vc.token_prev};
        prompt.insert(prompt.begin() + 1, prompt_past.end() - n_take, prompt_past.end());
    }
    prompt.insert(prompt.end(), prompt_init.begin(), prompt_init.end());

    if (st->kv_self_n_dec < n_decoders_cur) {  // wc:7095-7113
        const int factor = n_decoders_cur > 1 ? n_decoders_cur + 2 : 1;
        if (!kv_self_alloc(*st, (int) pad_to(hp.n_text_ctx, 256) * factor)) {
            log_error("%s: whisper_kv_cache_init() failed for self-attention cache\n", __func__);
            finish(-7);
            return;
        }
        st->kv_self_n_dec = n_decoders_cur;
    }
    st->kv_self.clear();

    rows.clear();
    srows.clear();
    sdecs.clear();
    for (int k = 0; k < (int) prompt.size(); ++k) rows.push_back(BatchRow{st, prompt[(size_t) k], k, 0, k == (int) prompt.size() - 1});
    st->decoders[0].i_batch = (int) prompt.size() - 1;
    refresh_grammar_mask(st->decoders[0]);
    srows.push_back(sampler_row(st->decoders[0], 0, true));
    sdecs.push_back(&st->decoders[0]);
    phase = AFTER_PROMPT;
}

// the device has run this job's rows: decoder forward + sampler
void FullJob::after_device() {
    if (aborted()) {  // wc:2993
        log_error("%s: failed to decode\n", __func__);
        finish(phase == AFTER_PROMPT ? -8 : -9);
        return;
    }
    if (phase == AFTER_PROMPT) {
        after_prompt();
    } else if (phase == STEP) {
        ++i;
        if (i < n_max) step();
        else finish_temperature();
    }
}

void FullJob::after_prompt() {
    st->no_speech_prob = st->decoders[0].picked.nosp_prob;  // wc:7126-7134
    for (int j = 1; j < n_decoders_cur; ++j) {              // wc:7143-7151
        Decoder & d = st->decoders[j];
        st->kv_self.seq_cp(0, j, -1, -1);
        d.picked = st->decoders[0].picked;
        if (host_sampling) {
            d.probs = st->decoders[0].probs;
            d.logits = st->decoders[0].logits;
            d.logprobs = st->decoders[0].logprobs;
        }
    }
    i = 0;
    phase = STEP;
    step();
}

// one iteration of the token loop, host part (wc:7157-7398): pick tokens, beam bookkeeping, decoder state machine
void FullJob::step() {
    const Vocab & vc = vocab();
    const int64_t t_start = time_us();
    const bool beam = params.strategy == WHISPER_SAMPLING_BEAM_SEARCH;
    if (beam)
        for (auto & bc : bc_per_dec) bc.clear();

    for (int j = 0; j < n_decoders_cur; ++j) {
        Decoder & d = st->decoders[j];
        if (d.completed || d.failed) continue;
        if (!beam) {
            if (t_cur < 1e-6f) d.sequence.tokens.push_back(token_from_pick(d.picked));
            else d.sequence.tokens.push_back(sample_token_host(vc, d));
            d.sequence.sum_logprobs_all += d.sequence.tokens.back().plog;
        } else {
            const auto tokens_new = sample_token_topk_host(vc, d, params.beam_search.beam_size);
            for (const auto & token : tokens_new) {
                bc_per_dec[(size_t) j].push_back({j, d.seek_delta, d.has_ts, d.sequence, d.grammar});
                bc_per_dec[(size_t) j].back().sequence.tokens.push_back(token);
                bc_per_dec[(size_t) j].back().sequence.sum_logprobs_all += token.plog;
            }
        }
    }
    beam_candidates.clear();
    for (const auto & bc : bc_per_dec) {
        beam_candidates.insert(beam_candidates.end(), bc.begin(), bc.end());
        if (!bc.empty()) st->n_sample += 1;
    }
    if (beam) {  // wc:7239-7291
        std::sort(beam_candidates.begin(), beam_candidates.end(), [](const BeamCandidate & a, const BeamCandidate & b) {
            if (a.sequence.sum_logprobs_all != b.sequence.sum_logprobs_all) return a.sequence.sum_logprobs_all > b.sequence.sum_logprobs_all;
            return a.decoder_idx < b.decoder_idx;
        });
        uint32_t cur_c = 0;
        for (int j = 0; j < n_decoders_cur; ++j) {
            Decoder & d = st->decoders[j];
            if (d.completed || d.failed) continue;
            if (cur_c >= beam_candidates.size()) cur_c = 0;
            auto & cur = beam_candidates[cur_c++];
            while (beam_candidates.size() > cur_c && sequences_equal(beam_candidates[cur_c].sequence, cur.sequence) && i > 0) ++cur_c;
            d.seek_delta = cur.seek_delta;
            d.has_ts = cur.has_ts;
            d.sequence = cur.sequence;
            d.grammar = cur.grammar;
            st->kv_self.seq_cp(cur.decoder_idx, MAX_DECODERS + j, -1, -1);
        }
        for (int j = 0; j < n_decoders_cur; ++j) {
            Decoder & d = st->decoders[j];
            if (d.completed || d.failed) continue;
            st->kv_self.seq_rm(j, -1, -1);
            st->kv_self.seq_cp(MAX_DECODERS + j, j, -1, -1);
            st->kv_self.seq_rm(MAX_DECODERS + j, -1, -1);
        }
    }

    // decoder state machine (wc:7297-7379)
    for (int j = 0; j < n_decoders_cur; ++j) {
        Decoder & d = st->decoders[j];
        if (d.completed || d.failed) continue;
        auto & has_ts = d.has_ts;
        auto & failed = d.failed;
        auto & completed = d.completed;
        auto & seek_delta = d.seek_delta;
        auto & result_len = d.sequence.result_len;
        {
            const auto & token = d.sequence.tokens.back();
            if (token.id > vc.token_beg) {
                const int seek_delta_new = 2 * (token.id - vc.token_beg);
                if (has_ts && seek_delta > seek_delta_new && result_len < i) {
                    failed = true;
                    continue;
                }
                seek_delta = seek_delta_new;
                result_len = i + 1;
                has_ts = true;
            }
            if (d.grammar.active()) d.grammar.accept_token(vc.id_to_token[(size_t) token.id]);  // wc:7329
            if (token.id == vc.token_eot || (params.max_tokens > 0 && i >= params.max_tokens) ||
                (has_ts && seek + seek_delta + delta_min >= seek_end)) {
                if (result_len == 0 && !params.no_timestamps) {
                    if (seek + seek_delta + delta_min >= seek_end) {
                        result_len = i + 1;
                    } else {
                        failed = true;
                        continue;
                    }
                }
                if (params.single_segment || params.no_timestamps) {
                    result_len = i + 1;
                    seek_delta = 100 * WHISPER_CHUNK_SIZE;
                }
                completed = true;
                continue;
            }
            if (ctx->model.n_loaded == 0) {  // header-only test models (wc:7364-7369)
                seek_delta = 100 * WHISPER_CHUNK_SIZE;
                completed = true;
                continue;
            }
        }
        if (i == n_max - 1 && (result_len == 0 || seek_delta < 100 * WHISPER_CHUNK_SIZE / 2)) {
            failed = true;
            continue;
        }
    }
    bool completed_all = true;
    for (int j = 0; j < n_decoders_cur; ++j) {
        const Decoder & d = st->decoders[j];
        if (d.completed || d.failed) continue;
        completed_all = false;
    }
    st->t_sample_us += time_us() - t_start;
    if (completed_all) {
        finish_temperature();
        return;
    }
    queue_step_rows();
}

// one row per live decoder for the next decoder forward (wc:7403-7427) + the matching sampler rows
void FullJob::queue_step_rows() {
    rows.clear();
    srows.clear();
    sdecs.clear();
    const int n_past = (int) prompt.size() + i;
    for (int j = 0; j < n_decoders_cur; ++j) {
        Decoder & d = st->decoders[j];
        if (d.failed || d.completed) continue;
        d.i_batch = (int) rows.size();
        rows.push_back(BatchRow{st, d.sequence.tokens.back().id, n_past, j, true});
        refresh_grammar_mask(d);
        srows.push_back(sampler_row(d, 0, false));
        sdecs.push_back(&d);
    }
    phase = STEP;
}

// rank, decide on fallback (wc:7483-7544), then either restart at the next temperature or emit
void FullJob::finish_temperature() {
    rows.clear();
    srows.clear();
    sdecs.clear();
    double best_score = -INFINITY;
    for (int j = 0; j < n_decoders_cur; ++j) {
        Decoder & d = st->decoders[j];
        if (d.failed) continue;
        d.sequence.tokens.resize((size_t) d.sequence.result_len);
        sequence_score(params, d.sequence);
        if (d.sequence.result_len > 32 && d.sequence.entropy < params.entropy_thold) {
            d.failed = true;
            st->n_fail_h++;
            continue;
        }
        if (best_score < d.sequence.score) {
            best_score = d.sequence.score;
            best_decoder_id = j;
        }
    }
    bool success = true;
    if (it != (int) temperatures.size() - 1) {
        const Decoder & d = st->decoders[best_decoder_id];
        if (d.failed || (d.sequence.avg_logprobs < params.logprob_thold && st->no_speech_prob < params.no_speech_thold)) {
            success = false;
            st->n_fail_p++;
        }
    }
    if (!success) {
        ++it;
        phase = TEMP_BEGIN;
        return;
    }
    emit_results();
}

// --------------------------------------------------------------------------------------------
// X1/X2: DTW token timestamps for the segments of one window (wc:8772-8933)
// --------------------------------------------------------------------------------------------
static bool dtw_token_timestamps(Context * ctx, State * st, const whisper_full_params & params, int i_segment, size_t n_segments, int seek,
                                 int n_frames, int medfilt_width) {
    const Vocab & vc = ctx->model.vocab;
    const int T = st->cross_T;
    std::vector<whisper_token> tokens = {vc.token_sot};
    if (vc.is_multilingual()) {
        const int lid = lang_id(params.language);
        st->lang_id = lid;
        tokens.push_back(vc.token_sot + 1 + lid);
    }
    const int sot_len = (int) tokens.size();
    tokens.push_back(vc.token_not);
    for (size_t s = (size_t) i_segment; s < (size_t) i_segment + n_segments; ++s)
        for (auto & t : st->result_all[s].tokens)
            if (t.id < vc.token_eot) tokens.push_back(t.id);
    tokens.push_back(vc.token_eot);

    st->kv_self.clear();
    const int n_tok = (int) tokens.size();
    std::vector<BatchRow> rows;
    for (int k = 0; k < n_tok; ++k) rows.push_back(BatchRow{st, tokens[(size_t) k], k, 0, false});
    st->kv_self.seq_rm(0, 0, -1);
    if (!stage_decode(*ctx, rows.data(), n_tok, true)) {
        log_error("%s: decoder failed while capturing alignment heads\n", __func__);
        return false;
    }
    const int n_audio = n_frames / 2;
    const int N = n_tok - sot_len - 1;  // rows kept: <|notimestamps|> + text tokens (wc:8882)
    if (n_audio <= medfilt_width || N <= 0 || ctx->n_aheads <= 0) {
        WB_CUDA_OK(cudaStreamSynchronize(ctx->stream));
        return true;  // the reference asserts here (wc:8742); nothing sensible to align
    }
    const int H = ctx->n_aheads;
    float *d_nrm = nullptr, *d_cost = nullptr, *d_diag = nullptr;
    int8_t * d_trace = nullptr;
    int * d_path = nullptr;
    const size_t cap = (size_t) N + n_audio;
    bool ok = cudaMalloc((void **) &d_nrm, (size_t) H * n_tok * n_audio * 4) == cudaSuccess &&
              cudaMalloc((void **) &d_cost, (size_t) n_tok * n_audio * 4) == cudaSuccess &&
              cudaMalloc((void **) &d_diag, (size_t) 3 * (N + 1) * 4) == cudaSuccess &&
              cudaMalloc((void **) &d_trace, (size_t) (N + 1) * (n_audio + 1)) == cudaSuccess &&
              cudaMalloc((void **) &d_path, (2 * cap + 1) * 4) == cudaSuccess;
    std::vector<int> path(2 * cap + 1);
    if (ok) ok = dtw_cost(st->d_qk, H, n_tok, T, n_audio, medfilt_width, d_nrm, d_cost, ctx->stream);
    if (ok) ok = dtw_path(d_cost + (size_t) sot_len * n_audio, N, n_audio, d_diag, d_trace, d_path, d_path + 2 * cap, ctx->stream);
    if (ok) ok = cudaMemcpyAsync(path.data(), d_path, (2 * cap + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream) == cudaSuccess;
    if (ok) ok = cudaStreamSynchronize(ctx->stream) == cudaSuccess;
    cudaFree(d_nrm);
    cudaFree(d_cost);
    cudaFree(d_diag);
    cudaFree(d_trace);
    cudaFree(d_path);
    if (!ok) {
        log_error("%s: DTW kernels failed\n", __func__);
        return false;
    }
    const int len = path[2 * cap];
    const int * pi = path.data() + (cap - (size_t) len);        // text index row
    const int * pj = path.data() + cap + (cap - (size_t) len);  // time index row
    // place timestamps (wc:8894-8920)
    int32_t last_v = 0;
    auto seg_i = st->result_all.begin() + i_segment;
    auto tok_i = seg_i->tokens.begin();
    const auto seg_end = st->result_all.begin() + i_segment + (long) n_segments;
    for (int k = 0; k < len; ++k) {
        const int32_t v = pi[k];
        if (v != last_v) {
            const int64_t timestamp = (int64_t) pj[k] * 2 + seek;
            last_v = v;
            while (seg_i != seg_end && !(tok_i->id < vc.token_eot)) {
                ++tok_i;
                if (tok_i == seg_i->tokens.end()) {
                    ++seg_i;
                    if (seg_i == seg_end) break;
                    tok_i = seg_i->tokens.begin();
                }
            }
            if (seg_i == seg_end) break;
            tok_i->t_dtw = timestamp;
            ++tok_i;
            if (tok_i == seg_i->tokens.end()) {
                ++seg_i;
                if (seg_i == seg_end) break;
                tok_i = seg_i->tokens.begin();
            }
        }
    }
    return true;
}

// segment emission and window advance (wc:7546-7707)
void FullJob::emit_results() {
    const Vocab & vc = vocab();
    auto & result_all = st->result_all;
    const Decoder & best = st->decoders[best_decoder_id];
    int seek_delta = best.seek_delta;
    const int result_len = best.sequence.result_len;
    const auto & tokens_cur = best.sequence.tokens;
    const size_t n_segments_before = result_all.size();
    const bool is_no_speech = (st->no_speech_prob > params.no_speech_thold && best.sequence.avg_logprobs < params.logprob_thold);

    auto & prompt_past = st->prompt_past;
    prompt_past.clear();
    if (prompt.front() == vc.token_prev) prompt_past.insert(prompt_past.end(), prompt.begin() + 1, prompt.end() - (long) prompt_init.size());
    for (int k = 0; k < result_len && !is_no_speech; ++k) prompt_past.push_back(tokens_cur[(size_t) k].id);

    // wc:7619-7626 / 7664-7671: heuristic token times for the segment just pushed, then optional wrapping; returns n_new
    auto token_times = [&]() {
        if (!params.token_timestamps) return 1;
        heuristic_token_timestamps(ctx, st, (int) result_all.size() - 1, params.thold_pt, params.thold_ptsum);
        return params.max_len > 0 ? wrap_last_segment(ctx, st, params.max_len, params.split_on_word) : 1;
    };
    auto notify = [&](int n_new) {
        if (params.new_segment_callback && !ctx->params.dtw_token_timestamps)
            call_owner(this, [&] { params.new_segment_callback(wctx(), wstate(), n_new, params.new_segment_callback_user_data); });
    };

    if (!tokens_cur.empty() && ctx->model.n_loaded > 0 && !is_no_speech) {
        int i0 = 0;
        int64_t t0 = seek + 2 * (tokens_cur.front().tid - vc.token_beg);
        std::string text;
        bool speaker_turn_next = false;
        for (int k = 0; k < (int) tokens_cur.size(); k++) {
            if (params.print_special || tokens_cur[(size_t) k].id < vc.token_eot) text += token_str(vc, tokens_cur[(size_t) k].id);
            if (params.tdrz_enable && tokens_cur[(size_t) k].id == vc.token_solm) speaker_turn_next = true;
            if (tokens_cur[(size_t) k].id > vc.token_beg && !params.single_segment) {
                const int64_t t1 = seek + 2 * (tokens_cur[(size_t) k].tid - vc.token_beg);
                if (!text.empty()) {
                    if (params.print_realtime) {
                        printf("%s", text.c_str());
                        fflush(stdout);
                    }
                    result_all.push_back({t0, t1, text, st->no_speech_prob, {}, speaker_turn_next});
                    for (int j = i0; j <= k; j++) result_all.back().tokens.push_back(tokens_cur[(size_t) j]);
                    notify(token_times());
                }
                text = "";
                while (k < (int) tokens_cur.size() && tokens_cur[(size_t) k].id > vc.token_beg) k++;
                k--;
                t0 = t1;
                i0 = k + 1;
                speaker_turn_next = false;
            }
        }
        if (!text.empty()) {
            const int64_t t1 = seek + seek_delta;
            if (params.print_realtime) {
                printf("%s", text.c_str());
                fflush(stdout);
            }
            result_all.push_back({t0, t1, text, st->no_speech_prob, {}, speaker_turn_next});
            for (int j = i0; j < (int) tokens_cur.size(); j++) result_all.back().tokens.push_back(tokens_cur[(size_t) j]);
            notify(token_times());
        }
    }
    {   // DTW (wc:7680-7692)
        const int n_segments = (int) (result_all.size() - n_segments_before);
        if (ctx->params.dtw_token_timestamps && n_segments) {
            const int n_frames = std::min(std::min(WHISPER_CHUNK_SIZE * 100, seek_delta), seek_end - seek);
            dtw_token_timestamps(ctx, st, params, (int) result_all.size() - n_segments, (size_t) n_segments, seek, n_frames, 7);
            if (params.new_segment_callback) {
                for (int seg = (int) result_all.size() - n_segments; seg < n_segments; seg++)
                    call_owner(this, [&] { params.new_segment_callback(wctx(), wstate(), seg, params.new_segment_callback_user_data); });
            }
        }
    }
    const bool single_timestamp_ending = tokens_cur.size() > 1 && tokens_cur[tokens_cur.size() - 2].id < vc.token_beg &&
                                         tokens_cur[tokens_cur.size() - 1].id > vc.token_beg;
    if (single_timestamp_ending) seek_delta = std::min(seek_end - seek, WHISPER_CHUNK_SIZE * 100);
    seek += seek_delta;
    phase = WINDOW_BEGIN;
}

// --------------------------------------------------------------------------------------------
// F5: heuristic token timestamps (`token_timestamps = true`; wc:8335-8615) and segment wrapping (`max_len`; wc:6049-6100)
// --------------------------------------------------------------------------------------------
// Anchors: a token whose timestamp-token statistics are confident (pt, ptsum above the thresholds, tid advancing) pins the
// boundary between itself and its predecessor at t_beg + 2 (tid - beg).  Runs of tokens between anchors share their span in
// proportion to a "voice length" of their text; finally every text token's edges are pulled to where the moving average of
// |signal| crosses half of its local mean.  Host arithmetic in the reference's types (int64 times, double split, f32 energy).
static float text_voice_length(const char * text) {  // wc:8335-8357
    float len = 0.0f;
    for (const char * c = text; *c; ++c) {
        if (*c == ' ') len += 0.01f;
        else if (*c == ',') len += 2.00f;
        else if (*c == '.' || *c == '!' || *c == '?' || (*c >= '0' && *c <= '9')) len += 3.00f;
        else len += 1.00f;
    }
    return len;
}

// The computation on explicit inputs (the state's t_beg / t_last / tid_last are carried from segment to segment by the caller)
static void heuristic_token_timestamps_core(const Vocab & vc, Segment & seg, const std::vector<float> & energy, int64_t & t_beg,
                                            int64_t & t_last, whisper_token & tid_last, float thold_pt, float thold_ptsum) {
    auto & tk = seg.tokens;
    const int n_samples = (int) energy.size(), n = (int) tk.size();
    if (n_samples == 0) {
        log_error("%s: no signal data available\n", __func__);
        return;
    }
    if (n == 0) return;
    const int64_t seg_t0 = seg.t0, seg_t1 = seg.t1;
    if (n == 1) {
        tk[0].t0 = seg_t0;
        tk[0].t1 = seg_t1;
        return;
    }
    // ---- anchors (wc:8432-8468); t_beg / t_last / tid_last persist across the segments of a call
    for (int j = 0; j < n; ++j) {
        whisper_token_data & t = tk[(size_t) j];
        if (j == 0) {
            if (t.id == vc.token_beg) {
                t.t0 = seg_t0;
                t.t1 = seg_t0;
                tk[1].t0 = seg_t0;
                t_beg = seg_t0;
                t_last = seg_t0;
                tid_last = vc.token_beg;
            } else {
                t.t0 = t_last;
            }
        }
        const int64_t tt = t_beg + 2 * (int64_t) (t.tid - vc.token_beg);
        t.vlen = text_voice_length(token_str(vc, t.id));
        if (t.pt > thold_pt && t.ptsum > thold_ptsum && t.tid > tid_last && tt <= seg_t1) {
            if (j > 0) tk[(size_t) j - 1].t1 = tt;
            t.t0 = tt;
            tid_last = t.tid;
        }
    }
    tk[(size_t) n - 2].t1 = seg_t1;
    tk[(size_t) n - 1].t0 = seg_t1;
    tk[(size_t) n - 1].t1 = seg_t1;
    t_last = seg_t1;

    // ---- runs without a known end share [first.t0, last.t1] in proportion to the voice lengths (wc:8472-8510)
    for (int first = 0; first < n;) {
        int last = first;
        while (last < n && tk[(size_t) last].t1 < 0) ++last;
        if (last >= n) last = n - 1;
        if (last > first) {
            double total = 0.0;
            for (int j = first; j <= last; ++j) total += tk[(size_t) j].vlen;
            const double span = (double) (tk[(size_t) last].t1 - tk[(size_t) first].t0);
            for (int j = first + 1; j <= last; ++j) {
                const double cut = tk[(size_t) j - 1].t0 + span * tk[(size_t) j - 1].vlen / total;
                tk[(size_t) j - 1].t1 = (int64_t) cut;
                tk[(size_t) j].t0 = (int64_t) cut;
            }
        }
        first = last + 1;
    }
    // ---- consistency pass (wc:8513-8525)
    for (int j = 0; j < n - 1; ++j) {
        if (tk[(size_t) j].t1 < 0) tk[(size_t) j + 1].t0 = tk[(size_t) j].t1;
        if (j > 0 && tk[(size_t) j - 1].t1 > tk[(size_t) j].t0) {
            tk[(size_t) j].t0 = tk[(size_t) j - 1].t1;
            tk[(size_t) j].t1 = std::max(tk[(size_t) j].t0, tk[(size_t) j].t1);
        }
    }
    // ---- energy edges (wc:8529-8593).  Times map to samples RELATIVE to the segment start, as in the reference
    const std::vector<float> & en = energy;
    auto to_sample = [&](int64_t t) {
        const int sidx = (int) (((t - seg_t0) * WHISPER_SAMPLE_RATE) / 100);
        return std::max(0, std::min(n_samples - 1, sidx));
    };
    auto to_time = [&](int sidx) { return (100ll * sidx) / WHISPER_SAMPLE_RATE + seg_t0; };
    const int reach = WHISPER_SAMPLE_RATE / 8;
    for (int j = 0; j < n; ++j) {
        whisper_token_data & t = tk[(size_t) j];
        if (t.id >= vc.token_eot) continue;
        int s_lo = to_sample(t.t0), s_hi = to_sample(t.t1);
        const int w_lo = std::max(s_lo - reach, 0), w_hi = std::min(s_hi + reach, n_samples);
        float acc = 0.0f;
        for (int k = w_lo; k < w_hi; ++k) acc += en[(size_t) k];
        const float level = (float) (0.5 * acc / (w_hi - w_lo));
        {   // leading edge: walk out while loud (but not into the previous token), else walk in to the first loud sample
            int k = s_lo;
            if (en[(size_t) k] > level && j > 0) {
                while (k > 0 && en[(size_t) k] > level) --k;
                t.t0 = to_time(k);
                if (t.t0 < tk[(size_t) j - 1].t1) t.t0 = tk[(size_t) j - 1].t1;
                else s_lo = k;
            } else {
                while (en[(size_t) k] < level && k < s_hi) ++k;
                s_lo = k;
                t.t0 = to_time(k);
            }
        }
        {   // trailing edge, mirrored
            int k = s_hi;
            if (en[(size_t) k] > level) {
                while (k < n_samples - 1 && en[(size_t) k] > level) ++k;
                t.t1 = to_time(k);
                if (j < n - 1 && t.t1 > tk[(size_t) j + 1].t0) t.t1 = tk[(size_t) j + 1].t0;
                else s_hi = k;
            } else {
                while (en[(size_t) k] < level && k > s_lo) --k;
                s_hi = k;
                t.t1 = to_time(k);
            }
        }
    }
}

static void heuristic_token_timestamps(Context * ctx, State * st, int i_segment, float thold_pt, float thold_ptsum) {
    heuristic_token_timestamps_core(ctx->model.vocab, st->result_all[(size_t) i_segment], st->energy, st->t_beg, st->t_last, st->tid_last,
                                    thold_pt, thold_ptsum);
}

// Splits the newest segment so that no piece's text exceeds max_len bytes (optionally only at tokens that start a word);
// pieces take over the token times computed above.  Returns the number of segments the newest one became.
static int wrap_last_segment_core(const Vocab & vc, std::vector<Segment> & all, int max_len, bool split_on_word) {
    Segment cur = all.back();  // the piece being scanned (a copy: `all` grows underneath)
    int pieces = 1, used = 0;
    std::string text;
    for (int i = 0; i < (int) cur.tokens.size(); ++i) {
        const whisper_token_data & t = cur.tokens[(size_t) i];
        if (t.id >= vc.token_eot) continue;
        const char * txt = token_str(vc, t.id);
        const int len = (int) strlen(txt);
        const bool may_split = !split_on_word || txt[0] == ' ';
        if (used + len > max_len && i > 0 && may_split) {
            Segment & head = all.back();
            head.text = std::move(text);
            head.t1 = t.t0;
            head.tokens.resize((size_t) i);
            head.speaker_turn_next = false;
            Segment tail{};
            tail.t0 = t.t0;
            tail.t1 = cur.t1;
            tail.tokens.assign(cur.tokens.begin() + i, cur.tokens.end());
            tail.speaker_turn_next = cur.speaker_turn_next;
            all.push_back(tail);
            used = 0;
            text.clear();
            cur = all.back();
            i = -1;
            ++pieces;
        } else {
            used += len;
            text += txt;
        }
    }
    all.back().text = std::move(text);
    return pieces;
}
static int wrap_last_segment(Context * ctx, State * st, int max_len, bool split_on_word) {
    return wrap_last_segment_core(ctx->model.vocab, st->result_all, max_len, split_on_word);
}

}  // namespace wb

// Host logic without a device (tests): heuristic token timestamps and segment wrapping on caller-supplied segments, exactly as
// whisper_full applies them to each new segment (timestamps, then the wrap when max_len > 0; t_beg / t_last / tid_last carried from
// segment to segment).  Input: n_segments segments, segment s owning seg_n_tokens[s] consecutive entries of tokens[].  Output: the
// resulting segments flattened the same way.  Returns the number of output segments, or -1 (capacity / arguments).
extern "C" int whisper_b200_test_token_timestamps(const char * const * token_texts, int n_vocab, int token_eot, int token_beg, const float * energy,
                                                  int n_energy, const whisper_token_data * tokens, const int * seg_n_tokens, const int64_t * seg_t0,
                                                  const int64_t * seg_t1, int n_segments, float thold_pt, float thold_ptsum, int max_len,
                                                  int split_on_word, whisper_token_data * out_tokens, int out_cap_tokens, int64_t * out_t0,
                                                  int64_t * out_t1, int * out_n_tokens, int out_cap_segments) {
    try {
        if (!token_texts || n_vocab <= 0 || !tokens || !seg_n_tokens || !seg_t0 || !seg_t1 || n_segments < 0 || n_energy < 0) return -1;
        wb::Vocab vc;
        vc.n_vocab = n_vocab;
        vc.token_eot = token_eot;
        vc.token_beg = token_beg;
        vc.id_to_token.resize((size_t) n_vocab);
        for (int i = 0; i < n_vocab; ++i) vc.id_to_token[(size_t) i] = token_texts[i] ? token_texts[i] : "";
        const std::vector<float> en(energy, energy + (energy ? n_energy : 0));
        std::vector<wb::Segment> all;
        int64_t t_beg = 0, t_last = 0;
        whisper_token tid_last = 0;
        for (int sgi = 0, at = 0; sgi < n_segments; at += seg_n_tokens[sgi], ++sgi) {
            wb::Segment sg{};
            sg.t0 = seg_t0[sgi];
            sg.t1 = seg_t1[sgi];
            sg.tokens.assign(tokens + at, tokens + at + seg_n_tokens[sgi]);
            all.push_back(std::move(sg));
            wb::heuristic_token_timestamps_core(vc, all.back(), en, t_beg, t_last, tid_last, thold_pt, thold_ptsum);
            if (max_len > 0) wb::wrap_last_segment_core(vc, all, max_len, split_on_word != 0);
        }
        if ((int) all.size() > out_cap_segments) return -1;
        int at = 0;
        for (size_t k = 0; k < all.size(); ++k) {
            const int n = (int) all[k].tokens.size();
            if (at + n > out_cap_tokens) return -1;
            out_t0[k] = all[k].t0;
            out_t1[k] = all[k].t1;
            out_n_tokens[k] = n;
            for (int j = 0; j < n; ++j) out_tokens[at + j] = all[k].tokens[(size_t) j];
            at += n;
        }
        return (int) all.size();
    } catch (...) {
        return -1;
    }
}

namespace wb {

// --------------------------------------------------------------------------------------------
// F4: language auto-detection (wc:4037-4110): encode at the offset, decode [sot], softmax over the language logits
// --------------------------------------------------------------------------------------------
int lang_auto_detect(Context * ctx, State * st, int offset_ms, float * lang_probs) {
    const Vocab & vc = ctx->model.vocab;
    const int seek = offset_ms / 10;
    if (seek < 0) {
        log_error("%s: offset %dms is before the start of the audio\n", __func__, offset_ms);
        return -1;
    }
    if (seek >= st->n_len_org) {
        log_error("%s: offset %dms is past the end of the audio (%dms)\n", __func__, offset_ms, st->n_len_org * 10);
        return -2;
    }
    State * sp = st;
    if (!stage_encode(*ctx, &sp, &seek, 1)) {
        log_error("%s: failed to encode\n", __func__);
        return -6;
    }
    BatchRow row{st, vc.token_sot, 0, 0, true};
    st->kv_self.seq_rm(0, 0, -1);
    if (!stage_decode(*ctx, &row, 1, false) || !stage_fetch_logits(*ctx, &row, 1)) {
        log_error("%s: failed to decode\n", __func__);
        return -7;
    }
    // the reference walks its language map in key (code) order before sorting by logit
    std::vector<int> order(100);
    for (int i = 0; i < 100; ++i) order[(size_t) i] = i;
    std::sort(order.begin(), order.end(), [](int a, int b) { return strcmp(g_langs[a].code, g_langs[b].code) < 0; });
    std::vector<std::pair<double, int>> li;
    for (int id : order) {
        const int tok = vc.token_sot + 1 + id;
        if (tok >= vc.n_vocab) continue;
        li.emplace_back((double) st->logits[(size_t) tok], id);
    }
    std::sort(li.begin(), li.end(), [](const std::pair<double, int> & a, const std::pair<double, int> & b) { return a.first > b.first; });
    const double mx = li[0].first;
    double sum = 0.0;
    for (auto & kv : li) {
        kv.first = exp(kv.first - mx);
        sum += kv.first;
    }
    for (auto & kv : li) kv.first /= sum;
    if (lang_probs)
        for (const auto & kv : li) lang_probs[kv.second] = (float) kv.first;
    return li[0].second;
}

// --------------------------------------------------------------------------------------------
// the batching scheduler: continuous batching behind the unchanged whisper_full_with_state
// --------------------------------------------------------------------------------------------
// The reference lets many whisper_states of one context run concurrently from different threads (rs/whisper_state.rs:19-21,
// wh:45-46), each call being its own CPU computation.  Here one worker thread per context owns the device: calls enqueue
// their FullJob and block; the worker advances every live job of the context in lock-step (one batched decoder forward per
// token), admits newly submitted jobs — and later windows of running ones — at step boundaries into the decode lanes, and
// wakes each caller when its job is done.  whisper_b200_full_batch and whisper_full_parallel are clients of the same queue.
// User callbacks are handed back to the thread that made the call and run there while the worker waits.
// WB200_SCHED_LOG=<file>: one line per scheduler event (microseconds since the first event): where the wall time of a call goes
static void sched_log(const char * what, long a = 0, long b = 0) {
    static FILE * f = [] {
        const char * p = getenv("WB200_SCHED_LOG");
        return p ? fopen(p, "a") : (FILE *) nullptr;
    }();
    if (!f) return;
    static const int64_t t0 = time_us();
    fprintf(f, "%lld %s %ld %ld\n", (long long) (time_us() - t0), what, a, b);
    fflush(f);
}

struct Client {
    std::condition_variable cv;
    const std::function<void()> * pending_fn = nullptr;  // a callback the worker wants run on the owner's thread
    bool fn_done = false;
    int n_jobs = 0, n_done = 0;
};

struct Command {
    const std::function<void()> * fn;
    bool done = false;
};

struct Lane {
    std::vector<FullJob *> jobs, enc_jobs, dev_jobs;
    std::vector<State *> enc_states;
    std::vector<int> enc_offsets;
    std::vector<BatchRow> rows;
    std::vector<SamplerRow> srows;
    std::vector<Decoder *> sdecs;
    std::vector<FullJob *> sjobs;  // the job behind every sampler row
    bool pending = false;          // a fused step is in flight on this lane's stream
    int last_logged_rows = -1;     // WB200_SCHED_LOG: row count of the last logged fused step
    int n_lanes_at_launch = 1;
};

struct Scheduler {
    Context * ctx;
    std::mutex mu;
    std::condition_variable cv_worker;
    std::deque<FullJob *> submitted;    // jobs waiting for admission, in arrival order
    std::deque<Command *> commands;     // stage-API closures to run on the worker
    int64_t last_submit_us = 0;
    bool stop = false;
    bool in_logits_filter = false;      // the worker is inside a logits_filter_callback hand-off: shared workspaces are live
    std::thread worker;
    std::thread::id worker_id;

    // worker-only state
    std::vector<Lane> lanes;
    int n_live = 0;
    int max_lanes = 4, lane_rows = 256;

    explicit Scheduler(Context * c) : ctx(c) {}
    void main();
    void run_commands(std::unique_lock<std::mutex> & lk);
    template <typename Pred> void wait_serving(std::unique_lock<std::mutex> & lk, Pred pred);
    void admit(std::vector<FullJob *> & incoming);
    void encode_jobs(Lane & ln, const std::vector<FullJob *> & cand);
    void lane_advance(int lane_id);
    void lane_complete(int lane_id);
    void reap();
    void job_done(FullJob * j);
    int active_lanes() const;
};

static Scheduler * scheduler_of(Context * ctx) {
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (!ctx->sched) {
        Scheduler * s = new Scheduler(ctx);
        s->worker = std::thread([s] { s->main(); });
        s->worker_id = s->worker.get_id();
        ctx->sched = s;
    }
    return (Scheduler *) ctx->sched;
}

void sched_shutdown(Context * ctx) {
    Scheduler * s;
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        s = (Scheduler *) ctx->sched;
        ctx->sched = nullptr;
    }
    if (!s) return;
    {
        std::lock_guard<std::mutex> lk(s->mu);
        s->stop = true;
    }
    s->cv_worker.notify_all();
    if (s->worker.joinable()) s->worker.join();
    delete s;
}

// worker side: run queued stage-API closures (lock held on entry and exit)
void Scheduler::run_commands(std::unique_lock<std::mutex> & lk) {
    while (!commands.empty()) {
        Command * c = commands.front();
        commands.pop_front();
        lk.unlock();
        try {
            (*c->fn)();
        } catch (...) {
            log_error("%s: exception in a stage call\n", __func__);
        }
        lk.lock();
        c->done = true;
        cv_worker.notify_all();  // command issuers wait on cv_worker too
    }
}

// worker side: wait for pred() while staying responsive to stage-API closures (a user callback may call back into the library)
template <typename Pred> void Scheduler::wait_serving(std::unique_lock<std::mutex> & lk, Pred pred) {
    while (true) {
        if (!in_logits_filter) run_commands(lk);
        if (pred()) return;
        cv_worker.wait(lk);
    }
}

static void call_owner(FullJob * job, const std::function<void()> & fn, bool is_logits_filter) {
    Scheduler * s = (Scheduler *) job->ctx->sched;
    Client * c = job->client;
    if (!s || !c || std::this_thread::get_id() != s->worker_id) {  // not under the scheduler (never the case today): run in place
        fn();
        return;
    }
    std::unique_lock<std::mutex> lk(s->mu);
    c->pending_fn = &fn;
    c->fn_done = false;
    s->in_logits_filter = is_logits_filter;
    c->cv.notify_all();
    s->wait_serving(lk, [&] { return c->fn_done; });
    s->in_logits_filter = false;
}

bool sched_run(Context * ctx, const std::function<void()> & fn) {
    Scheduler * s = scheduler_of(ctx);
    if (std::this_thread::get_id() == s->worker_id) {  // a stage call made from code already running on the worker
        fn();
        return true;
    }
    Command cmd{&fn};
    std::unique_lock<std::mutex> lk(s->mu);
    if (s->in_logits_filter) {
        log_error("%s: stage calls on the same context from inside logits_filter_callback are not supported\n", __func__);
        return false;
    }
    if (s->stop) return false;
    s->commands.push_back(&cmd);
    s->cv_worker.notify_all();
    s->cv_worker.wait(lk, [&] { return cmd.done; });
    return true;
}

void Scheduler::job_done(FullJob * j) {
    std::lock_guard<std::mutex> lk(mu);
    Client * c = j->client;
    c->n_done++;
    c->cv.notify_all();
}

// client side: submit the jobs, then serve callback hand-offs until every job has finished
int full_with_states(Context * ctx, State * const * states, int n_states, const whisper_full_params * params_arr, bool params_per_state,
                     const float * const * samples, const int * n_samples) {
    if (n_states <= 0) return 0;
    Scheduler * s = scheduler_of(ctx);
    if (std::this_thread::get_id() == s->worker_id) {
        log_error("%s: whisper_full* called from inside a callback of the same context\n", __func__);
        return -1;
    }
    Client client;
    std::vector<FullJob> jobs((size_t) n_states);
    for (int k = 0; k < n_states; ++k) {
        FullJob & j = jobs[(size_t) k];
        j.ctx = ctx;
        j.st = states[k];
        j.client = &client;
        j.params = params_per_state ? params_arr[k] : params_arr[0];
        j.samples = samples ? samples[k] : nullptr;
        j.n_samples = n_samples ? n_samples[k] : 0;
        j.t_submit = time_us();
    }
    client.n_jobs = n_states;
    std::unique_lock<std::mutex> lk(s->mu);
    if (s->stop) return -1;
    for (auto & j : jobs) s->submitted.push_back(&j);
    s->last_submit_us = time_us();
    s->cv_worker.notify_all();
    while (client.n_done < client.n_jobs) {
        if (client.pending_fn) {
            const std::function<void()> * fn = client.pending_fn;
            client.pending_fn = nullptr;
            lk.unlock();
            try {
                (*fn)();
            } catch (...) {
                log_error("%s: a user callback threw\n", __func__);
            }
            lk.lock();
            client.fn_done = true;
            s->cv_worker.notify_all();
            continue;
        }
        client.cv.wait(lk);
    }
    lk.unlock();
    int ret = 0;
    for (auto & j : jobs)
        if (j.ret != 0 && ret == 0) ret = j.ret;
    return ret;
}

// ---- worker side ------------------------------------------------------------------------------------------------------
int Scheduler::active_lanes() const {
    int n = 0;
    for (const Lane & ln : lanes) n += ln.jobs.empty() ? 0 : 1;
    return std::max(1, n);
}

// window heads (callbacks, end-of-audio test) and ONE batched encoder pass for every candidate that starts a window
void Scheduler::encode_jobs(Lane & ln, const std::vector<FullJob *> & cand) {
    ln.enc_states.clear();
    ln.enc_offsets.clear();
    ln.enc_jobs.clear();
    for (FullJob * j : cand) {
        if (j->phase != FullJob::WINDOW_BEGIN) continue;
        if (j->window_begin()) {
            ln.enc_states.push_back(j->st);
            ln.enc_offsets.push_back(j->seek);
            ln.enc_jobs.push_back(j);
        }
    }
    if (ln.enc_jobs.empty()) return;
    sched_log("encode_begin", (long) ln.enc_jobs.size(), n_live);
    if (!stage_encode(*ctx, ln.enc_states.data(), ln.enc_offsets.data(), (int) ln.enc_jobs.size())) {
        log_error("%s: failed to encode\n", __func__);
        for (auto * j : ln.enc_jobs) j->finish(-6);
    } else {
        for (auto * j : ln.enc_jobs) j->after_encode();
    }
    ln.enc_jobs.clear();
}

// mel, call set-up (wc:6801-6984), lane assignment and the first encoder pass of newly submitted jobs
void Scheduler::admit(std::vector<FullJob *> & incoming) {
    std::vector<FullJob *> fresh;
    // the log-mel front end of every arrival is launched back to back and waited for ONCE (per-state synchronisation made the
    // front end launch-bound: 0.19 ms per 30 s window for 3.5 MB of traffic)
    std::vector<State *> mel_states;
    for (FullJob * j : incoming) {
        j->admitted = true;
        j->st->result_all.clear();
        j->mask = suppress_mask(*ctx, j->params);
        if (!j->mask) {
            j->finish(-6);
        } else if (j->n_samples > 0) {
            if (!stage_pcm_to_mel(*j->st, j->samples, j->n_samples, false)) {  // wc:6806-6812
                log_error("%s: failed to compute log mel spectrogram\n", "whisper_full_with_state");
                j->finish(-2);
            } else {
                mel_states.push_back(j->st);
            }
        }
    }
    sched_log("admit_begin", (long) incoming.size(), n_live);
    if (!mel_states.empty() && !stage_mel_sync(*ctx, mel_states.data(), (int) mel_states.size())) {
        for (FullJob * j : incoming)
            if (j->phase != FullJob::DONE) j->finish(-2);
    }
    for (FullJob * j : incoming) {
        if (j->phase != FullJob::DONE) j->init();
        if (j->phase == FullJob::DONE) job_done(j);
        else fresh.push_back(j);
    }
    if (fresh.empty()) return;
    // lanes (step.cuh): groups of jobs that advance independently.  Measured on B200: 4 lanes of 64-96 rows beat one
    // of 256 by ~5 %; below 128 sequences one lane streams the decoder weights once for all of them
    const int total = n_live + (int) fresh.size();
    int target;
    const char * e = getenv("WB200_LANES");
    if (e && atoi(e) > 0) target = std::min(atoi(e), total);
    else target = std::max(std::min(max_lanes, total / 64), (total + lane_rows - 1) / lane_rows);
    target = std::max(1, std::min(target, (int) lanes.size()));
    for (size_t k = 0; k < fresh.size(); ++k) {
        int best = 0;
        for (int l = 1; l < target; ++l)
            if (lanes[(size_t) l].jobs.size() < lanes[(size_t) best].jobs.size()) best = l;
        fresh[k]->lane = best;
        lanes[(size_t) best].jobs.push_back(fresh[k]);
    }
    n_live += (int) fresh.size();
    sched_log("mel_done", (long) fresh.size(), n_live);
    Lane scratch;
    encode_jobs(scratch, fresh);  // the first windows of every lane go through the encoder as one batch
    sched_log("admit_encode_done", (long) fresh.size(), n_live);
}

void Scheduler::reap() {
    for (Lane & ln : lanes) {
        bool any = false;
        for (FullJob * j : ln.jobs) any = any || j->phase == FullJob::DONE;
        if (!any) continue;
        std::vector<FullJob *> keep, done;
        for (FullJob * j : ln.jobs) (j->phase == FullJob::DONE ? done : keep).push_back(j);
        ln.jobs.swap(keep);
        if (!ln.pending) {
            ln.dev_jobs.clear();
            ln.sjobs.clear();
            ln.sdecs.clear();
        }
        n_live -= (int) done.size();
        sched_log("jobs_done", (long) done.size(), n_live);
        for (FullJob * j : done) job_done(j);  // the job's memory belongs to its caller from here on
    }
}

// advances a lane until it has a fused step in flight (pending) or nothing left to do
void Scheduler::lane_advance(int lane_id) {
    Lane & ln = lanes[(size_t) lane_id];
    const SamplerStatic sst = sampler_static(*ctx);
    while (true) {
        // Jobs that come back for a later window of their audio have fallen out of step with their lane; with several lanes
        // active they are handed to lane 0, which runs them as one batch instead of each decoding alone on its own lane.
        if (lane_id != 0 && active_lanes() > 1) {
            for (size_t q = 0; q < ln.jobs.size();) {
                FullJob * j = ln.jobs[q];
                if (j->phase == FullJob::WINDOW_BEGIN && j->seek > j->seek_start && j->seek + FullJob::delta_min < j->seek_end &&
                    (int) lanes[0].jobs.size() < lane_rows) {
                    j->lane = 0;
                    lanes[0].jobs.push_back(j);
                    ln.jobs.erase(ln.jobs.begin() + (long) q);
                } else {
                    ++q;
                }
            }
        }
        // ---- windows that need their encoder pass
        encode_jobs(ln, ln.jobs);
        // ---- gather the decoder rows of every live job
        ln.rows.clear();
        ln.srows.clear();
        ln.sdecs.clear();
        ln.sjobs.clear();
        ln.dev_jobs.clear();
        bool any_live = false, any_dists = false;
        // Jobs that contribute several rows (prompt passes, beams) take the general decoder path, jobs with one row per
        // state the fused step.  When both kinds are waiting, the multi-row jobs go first on their own: which path a
        // token is computed on then depends on the job alone, never on what else happens to share the iteration.
        bool has_multi = false, has_single = false;
        for (FullJob * jp : ln.jobs) {
            FullJob & j = *jp;
            if (j.phase == FullJob::DONE) continue;
            if (j.phase == FullJob::TEMP_BEGIN) j.temp_begin();
            if ((j.phase != FullJob::AFTER_PROMPT && j.phase != FullJob::STEP) || j.rows.empty()) continue;
            if (j.rows.size() == 1 && j.srows.size() == 1 && !j.params.logits_filter_callback && !j.has_grammar()) has_single = true;
            else has_multi = true;
        }
        for (FullJob * jp : ln.jobs) {
            FullJob & j = *jp;
            if (j.phase == FullJob::DONE) continue;
            any_live = true;
            if (j.phase != FullJob::AFTER_PROMPT && j.phase != FullJob::STEP) continue;
            if (j.rows.empty()) continue;
            const bool single = j.rows.size() == 1 && j.srows.size() == 1 && !j.params.logits_filter_callback && !j.has_grammar();
            if (has_multi && has_single && single) continue;  // waits for the next iteration
            if (single && (int) ln.rows.size() >= lane_rows) continue;
            // flagged rows of this job map to consecutive logits rows
            int flagged = 0;
            for (const auto & r : j.rows) flagged += r.want_logits ? 1 : 0;
            const int base = (int) ln.srows.size();
            if (flagged != (int) j.srows.size()) {
                log_error("%s: internal error: %d flagged rows vs %zu sampler rows\n", __func__, flagged, j.srows.size());
                j.finish(-9);
                continue;
            }
            ln.rows.insert(ln.rows.end(), j.rows.begin(), j.rows.end());
            for (size_t q = 0; q < j.srows.size(); ++q) {
                SamplerRow sr = j.srows[q];
                sr.logits_row = base + (int) q;
                sr.work_slot = base + (int) q;
                sr.prefiltered = j.params.logits_filter_callback ? 1 : 0;
                ln.srows.push_back(sr);
                ln.sdecs.push_back(j.sdecs[q]);
                ln.sjobs.push_back(&j);
                any_dists = any_dists || j.sdecs[q]->want_dists;
            }
            ln.dev_jobs.push_back(&j);
        }
        if (!any_live) return;
        if (ln.dev_jobs.empty()) {
            bool progress = false;  // a job may have just finished or moved to another window: look again, else stop
            for (FullJob * jp : ln.jobs) progress = progress || jp->phase == FullJob::WINDOW_BEGIN || jp->phase == FullJob::TEMP_BEGIN;
            if (!progress) return;
            continue;
        }

        // logits_filter_callback (wh:474-480, wc:6226-6228) mutates HOST logits between the static suppressions and the
        // timestamp rules: those rows take the general path, the host applies everything the reference applies before the
        // callback, calls it on the caller's thread, and the device sampler continues from there (SamplerRow::prefiltered)
        bool any_callback = false;
        for (const SamplerRow & sr : ln.srows) any_callback = any_callback || sr.prefiltered;
        const std::function<void(int, float *)> prefilter = [&](int q, float * lg) {
            FullJob & j = *ln.sjobs[(size_t) q];
            const Decoder & d = *ln.sdecs[(size_t) q];
            const SamplerRow & r = ln.srows[(size_t) q];
            const Vocab & vc = ctx->model.vocab;
            const int V = vc.n_vocab;
            if (r.temperature > 0.0f)
                for (int k = 0; k < V; ++k) lg[k] /= r.temperature;
            if (r.suppress_blank && r.is_initial) {
                lg[vc.token_eot] = -INFINITY;
                if (sst.token_space >= 0) lg[sst.token_space] = -INFINITY;
            }
            for (int k = 0; k < V; ++k)
                if (j.mask->pre_host[(size_t) k]) lg[k] = -INFINITY;
            if (r.no_timestamps)
                for (int k = vc.token_beg; k < V; ++k) lg[k] = -INFINITY;
            call_owner(&j, [&] {
                j.params.logits_filter_callback(j.wctx(), j.wstate(), d.sequence.tokens.data(), (int) d.sequence.tokens.size(), lg,
                                                j.params.logits_filter_callback_user_data);
            }, true);
        };
        const int n_rows = (int) ln.rows.size();
        bool any_grammar = false;  // grammar rows carry a host-computed reject mask that only stage_sample uploads
        for (FullJob * gj : ln.dev_jobs) any_grammar = any_grammar || gj->has_grammar();
        const bool fused_step = !any_callback && !any_grammar && step_eligible(*ctx, ln.rows.data(), n_rows, any_dists) &&
                                ln.rows.size() == ln.srows.size();
        bool ok;
        if (fused_step) {
            if (n_rows != ln.last_logged_rows) {
                sched_log("lane_step_rows", lane_id, n_rows);
                ln.last_logged_rows = n_rows;
            }
            // one token per sequence: the whole decoder forward + sampler is one CUDA-graph launch (step.cu)
            ok = stage_step_launch(*ctx, lane_id, active_lanes(), ln.rows.data(), n_rows, sst, ln.srows.data());
            if (ok) {
                ln.pending = true;
                return;
            }
        } else {
            sched_log("general_decode_begin", lane_id, n_rows);
            ok = stage_decode(*ctx, ln.rows.data(), n_rows, false) &&
                 stage_sample(*ctx, sst, ln.srows.data(), (int) ln.srows.size(), any_dists, ln.sdecs.data(), any_callback ? &prefilter : nullptr);
            sched_log("general_decode_end", lane_id, n_rows);
            ln.last_logged_rows = -1;
        }
        if (!ok) {
            log_error("%s: failed to decode\n", __func__);
            for (auto * j : ln.dev_jobs) j->finish(j->phase == FullJob::AFTER_PROMPT ? -8 : -9);
            continue;
        }
        for (auto * j : ln.dev_jobs) j->after_device();
    }
}

void Scheduler::lane_complete(int lane_id) {
    Lane & ln = lanes[(size_t) lane_id];
    ln.pending = false;
    if (!stage_step_wait(*ctx, lane_id, ln.rows.data(), (int) ln.rows.size(), ln.sdecs.data())) {
        log_error("%s: failed to decode\n", __func__);
        for (auto * j : ln.dev_jobs) j->finish(j->phase == FullJob::AFTER_PROMPT ? -8 : -9);
        return;
    }
    for (auto * j : ln.dev_jobs) j->after_device();
}

void Scheduler::main() {
    {
        std::lock_guard<std::mutex> lk(mu);
        worker_id = std::this_thread::get_id();
    }
    cudaSetDevice(ctx->model.device);
    lanes.assign((size_t) step_max_lanes(), Lane());
    int trace_steps = getenv("WB200_STEP_TRACE") ? 0 : -1;
    std::unique_lock<std::mutex> lk(mu);
    while (true) {
        wait_serving(lk, [&] { return stop || !submitted.empty() || n_live > 0; });
        if (stop && n_live == 0) {
            // jobs still queued at shutdown are failed, their callers released
            while (!submitted.empty()) {
                FullJob * j = submitted.front();
                submitted.pop_front();
                j->ret = -1;
                j->client->n_done++;
                j->client->cv.notify_all();
            }
            break;
        }
        // ---- admission: wait until arrivals have paused (callers arriving in a burst become ONE encoder batch and one decode
        // cohort), bounded by how long the oldest caller has waited; longer pauses are accepted while other jobs keep the device busy
        std::vector<FullJob *> incoming;
        if (!submitted.empty()) {
            const int64_t now = time_us();
            const bool idle = n_live == 0;
            const int64_t settle = idle ? 150 : 1000, max_wait = idle ? 2000 : 20000;
            const bool go = now - last_submit_us >= settle || now - submitted.front()->t_submit >= max_wait || stop;
            if (go) {
                const int room = (int) lanes.size() * lane_rows - n_live;
                while (!submitted.empty() && (int) incoming.size() < room) {
                    incoming.push_back(submitted.front());
                    submitted.pop_front();
                }
            } else if (idle) {
                cv_worker.wait_for(lk, std::chrono::microseconds(settle));
                continue;
            }
        }
        lk.unlock();
        try {
            if (!incoming.empty()) admit(incoming);
            int n_pending = 0;
            for (size_t l = 0; l < lanes.size(); ++l) {
                if (!lanes[l].pending && !lanes[l].jobs.empty()) lane_advance((int) l);
                n_pending += lanes[l].pending ? 1 : 0;
            }
            bool served = false;
            for (size_t l = 0; l < lanes.size(); ++l) {
                Lane & ln = lanes[l];
                if (!ln.pending) continue;
                // serve whichever lane finishes first (a fixed round-robin would hold every lane to the slowest one's pace); a lane
                // that is alone is simply waited for
                if (n_pending > 1 && !stage_step_ready(*ctx, (int) l)) continue;
                served = true;
                lane_complete((int) l);
                lane_advance((int) l);
                if (trace_steps >= 0 && ++trace_steps == 100) {
                    // in-situ timeline (step.cuh): drain every lane, dump the stamps of their last (concurrent) steps, resume
                    for (size_t q = 0; q < lanes.size(); ++q)
                        if (lanes[q].pending) lane_complete((int) q);
                    step_trace_dump(*ctx);
                }
            }
            reap();
            if (!served && n_pending > 0) std::this_thread::yield();
        } catch (const std::exception & ex) {
            log_error("%s: exception on the scheduler thread: %s\n", __func__, ex.what());
            for (Lane & ln : lanes) {
                if (ln.pending) cudaStreamSynchronize(nullptr);
                ln.pending = false;
                for (FullJob * j : ln.jobs)
                    if (j->phase != FullJob::DONE) j->finish(-1);
            }
            reap();
        }
        lk.lock();
    }
}

}  // namespace wb
