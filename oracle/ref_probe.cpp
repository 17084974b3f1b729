// Oracle probe translation unit (TEST INFRASTRUCTURE — never linked into the product).
//
// Compiles the reference engine *in place* by including its source file from
// /root/reference (nothing is copied), and adds a few read-only accessors for
// internal stage outputs that the public whisper.h ABI does not expose, so the
// parity tests can compare stage by stage:
//   mel           -> whisper_state::mel            (sys/whisper.cpp/src/whisper.cpp:896)
//   conv output   -> whisper_state::embd_conv      (whisper.cpp:911)
//   encoder out   -> whisper_state::embd_enc       (whisper.cpp:912)
//   cross K/V     -> whisper_state::kv_cross       (whisper.cpp:891)
//   self K/V      -> whisper_state::kv_self        (whisper.cpp:887)
#include "whisper.cpp"

extern "C" {

const float * ref_probe_mel(struct whisper_state * st, int * n_len, int * n_len_org, int * n_mel) {
    *n_len = st->mel.n_len; *n_len_org = st->mel.n_len_org; *n_mel = st->mel.n_mel;
    return st->mel.data.data();
}

static int64_t probe_copy(struct ggml_tensor * t, void * dst, int64_t cap_bytes) {
    if (!t) return -1;
    const int64_t nb = (int64_t) ggml_nbytes(t);
    if (dst && cap_bytes >= nb) ggml_backend_tensor_get(t, dst, 0, nb);
    return nb;
}

// each returns the tensor's byte size (or -1); copies when dst has room. ne[4] receives the ggml shape.
int64_t ref_probe_embd_enc(struct whisper_state * st, void * dst, int64_t cap, int64_t * ne) {
    if (st->embd_enc && ne) for (int i = 0; i < 4; ++i) ne[i] = st->embd_enc->ne[i];
    return probe_copy(st->embd_enc, dst, cap);
}
int64_t ref_probe_embd_conv(struct whisper_state * st, void * dst, int64_t cap, int64_t * ne) {
    if (st->embd_conv && ne) for (int i = 0; i < 4; ++i) ne[i] = st->embd_conv->ne[i];
    return probe_copy(st->embd_conv, dst, cap);
}
int64_t ref_probe_kv_cross_k(struct whisper_state * st, void * dst, int64_t cap) { return probe_copy(st->kv_cross.k, dst, cap); }
int64_t ref_probe_kv_cross_v(struct whisper_state * st, void * dst, int64_t cap) { return probe_copy(st->kv_cross.v, dst, cap); }
int64_t ref_probe_kv_self_k (struct whisper_state * st, void * dst, int64_t cap) { return probe_copy(st->kv_self.k,  dst, cap); }
int64_t ref_probe_kv_self_v (struct whisper_state * st, void * dst, int64_t cap) { return probe_copy(st->kv_self.v,  dst, cap); }
int     ref_probe_kv_self_size(struct whisper_state * st) { return (int) st->kv_self.size; }

// counters the reference keeps per state (whisper.cpp:868-881)
void ref_probe_counters(struct whisper_state * st, int32_t * out7, int64_t * us6) {
    out7[0] = st->n_sample; out7[1] = st->n_encode; out7[2] = st->n_decode; out7[3] = st->n_batchd;
    out7[4] = st->n_prompt; out7[5] = st->n_fail_p; out7[6] = st->n_fail_h;
    us6[0] = st->t_sample_us; us6[1] = st->t_encode_us; us6[2] = st->t_decode_us;
    us6[3] = st->t_batchd_us; us6[4] = st->t_prompt_us; us6[5] = st->t_mel_us;
}

// last decode's alignment-head cross-attention weights (whisper.cpp:947-948)
const float * ref_probe_aheads_qks(struct whisper_state * st, int64_t * n) {
    *n = (int64_t) st->aheads_cross_QKs_data.size();
    return st->aheads_cross_QKs_data.data();
}


// X1/X2 in isolation: the reference's own alignment post-processing (whisper.cpp:8840-8890: token-axis ggml_norm eps 1e-9, median
// filter, mean over heads, negate, drop the sot sequence and eot) and dtw_and_backtrace (whisper.cpp:8647-8731) on a caller-supplied
// alignment-head tensor in the layout of whisper_state::aheads_cross_QKs_data: data[head][audio_ctx][token].
// cost_out (optional): [n_audio_tokens][N] floats, N = n_tokens - sot_len - 1 (token index contiguous, the ggml order of `w`).
// Returns the path length L; path_out[0..L) = token indices, path_out[L..2L) = time indices (when cap >= 2L).
int ref_probe_dtw(const float * data, int n_tokens, int n_audio_ctx, int n_heads, int n_audio_tokens, int sot_len, int medfilt_width,
                  float * cost_out, int32_t * path_out, int cap) {
    const size_t mem = (size_t) 64 * 1024 * 1024 + (size_t) n_tokens * n_audio_ctx * n_heads * 4 * 8;
    struct ggml_init_params gparams = { mem, nullptr, false };
    struct ggml_context * gctx = ggml_init(gparams);
    ggml_tensor * w = ggml_new_tensor_3d(gctx, GGML_TYPE_F32, n_tokens, n_audio_tokens, n_heads);
    for (int k = 0; k < n_heads; ++k)
        for (int j = 0; j < n_audio_tokens; ++j)
            memcpy((char *) w->data + j * w->nb[1] + k * w->nb[2], data + (size_t) j * n_tokens + (size_t) k * n_tokens * n_audio_ctx,
                   n_tokens * sizeof(float));
    w = ggml_norm(gctx, w, 1e-9f);
    w = ggml_permute(gctx, ggml_permute(gctx, w, 2, 1, 0, 3), 0, 2, 1, 3);
    median_filter_user_data mf_user_data = {medfilt_width};
    w = ggml_map_custom1(gctx, w, median_filter, 1, &mf_user_data);
    w = ggml_mean(gctx, w);
    w = ggml_scale(gctx, w, -1.0);
    w = ggml_reshape_2d(gctx, w, w->ne[1], w->ne[2]);
    w = ggml_view_2d(gctx, w, w->ne[0] - sot_len - 1, w->ne[1], w->nb[1], sot_len * w->nb[0]);
    struct ggml_cgraph * gf = ggml_new_graph(gctx);
    ggml_build_forward_expand(gf, w);
    ggml_backend_ptr backend { ggml_backend_init_by_type(GGML_BACKEND_DEVICE_TYPE_CPU, nullptr) };
    ggml_backend_graph_compute(backend.get(), gf);
    const int64_t N = w->ne[0], M = w->ne[1];
    if (cost_out)
        for (int64_t j = 0; j < M; ++j)
            for (int64_t i = 0; i < N; ++i) cost_out[j * N + i] = whisper_get_f32_nd(w, i, j, 0, 0);
    ggml_tensor * al = dtw_and_backtrace(gctx, w);
    const int L = (int) al->ne[1];
    if (path_out && cap >= 2 * L)
        for (int i = 0; i < L; ++i) {
            path_out[i] = whisper_get_i32_nd(al, 0, i, 0, 0);
            path_out[L + i] = whisper_get_i32_nd(al, 1, i, 0, 0);
        }
    ggml_free(gctx);
    return L;
}

// dtw_and_backtrace alone on a caller-supplied cost matrix cost[M][N] (token index contiguous)
int ref_probe_dtw_path(const float * cost, int N, int M, int32_t * path_out, int cap) {
    struct ggml_init_params gparams = { (size_t) 64 * 1024 * 1024 + (size_t) (N + 1) * (M + 1) * 16, nullptr, false };
    struct ggml_context * gctx = ggml_init(gparams);
    ggml_tensor * x = ggml_new_tensor_2d(gctx, GGML_TYPE_F32, N, M);
    memcpy(x->data, cost, (size_t) N * M * sizeof(float));
    ggml_tensor * al = dtw_and_backtrace(gctx, x);
    const int L = (int) al->ne[1];
    if (path_out && cap >= 2 * L)
        for (int i = 0; i < L; ++i) {
            path_out[i] = whisper_get_i32_nd(al, 0, i, 0, 0);
            path_out[L + i] = whisper_get_i32_nd(al, 1, i, 0, 0);
        }
    ggml_free(gctx);
    return L;
}

}

// Grammar (whisper.cpp:5784-5890) in isolation: start the reference's grammar on `rules`, feed it the tokens `accept[0..n_accept)`
// through whisper_grammar_accept_token, then ask whisper_suppress_invalid_grammar which tokens it penalises next.
// mask_out[n_vocab] receives 1 for the penalised ids.  Returns the number of grammar stacks alive (0 = the grammar is dead / complete).
extern "C" int ref_probe_grammar(struct whisper_context * ctx, const whisper_grammar_element ** rules, size_t n_rules, size_t i_start_rule,
                                 const whisper_token * accept, int n_accept, uint8_t * mask_out) {
    whisper_grammar g = whisper_grammar_init(rules, n_rules, i_start_rule);
    for (int i = 0; i < n_accept; ++i) whisper_grammar_accept_token(*ctx, g, accept[i]);
    const int n_vocab = whisper_n_vocab(ctx);
    std::vector<float> logits((size_t) n_vocab, 0.0f);
    whisper_full_params p = whisper_full_default_params(WHISPER_SAMPLING_GREEDY);
    p.grammar_penalty = 1.0f;
    whisper_suppress_invalid_grammar(*ctx, p, logits, g);
    for (int i = 0; i < n_vocab; ++i) mask_out[i] = logits[(size_t) i] != 0.0f ? 1 : 0;
    return (int) g.stacks.size();
}

// The reference's VAD pre-pass (static whisper_vad, whisper.cpp:6615-6793) on the context's default state: runs its Silero graph on
// `samples`, filters them, builds the state's time-mapping table.  out[cap] receives the filtered samples, table[2 * table_cap] the
// (processed, original) pairs, map_in[n_map] -> map_out through map_processed_to_original_time (whisper.cpp:7882-7925).
// Returns the number of filtered samples (0 when it found no speech), -1 on failure.
extern "C" int ref_probe_vad_filter(struct whisper_context * ctx, struct whisper_full_params params, const float * samples, int n_samples,
                                    float * out, int cap, int64_t * table, int table_cap, int * n_table, const int64_t * map_in,
                                    int64_t * map_out, int n_map) {
    whisper_state * st = ctx->state;
    if (st == nullptr) return -1;
    std::vector<float> filtered;
    if (!whisper_vad(ctx, st, params, samples, n_samples, filtered)) return -1;
    const auto & tab = st->vad_mapping_table;
    *n_table = (int) tab.size();
    for (int i = 0; i < (int) tab.size() && i < table_cap; ++i) {
        table[2 * i] = tab[(size_t) i].processed_time;
        table[2 * i + 1] = tab[(size_t) i].original_time;
    }
    if (out) memcpy(out, filtered.data(), std::min(filtered.size(), (size_t) std::max(cap, 0)) * sizeof(float));
    for (int i = 0; i < n_map; ++i)
        map_out[i] = (st->has_vad_segments && !tab.empty()) ? map_processed_to_original_time(map_in[i], tab) : map_in[i];
    return st->has_vad_segments ? (int) filtered.size() : 0;
}

// Heuristic token timestamps + segment wrapping (static whisper_exp_compute_token_level_timestamps, whisper.cpp:8390-8615, and
// whisper_wrap_segment, whisper.cpp:6049-6100) on caller-supplied segments, applied the way whisper_full_with_state applies them to
// each new segment (whisper.cpp:7615-7630).  `state` is a scratch state of `ctx`: its result_all / energy / t_beg.. are overwritten.
// Same flattened in/out layout as whisper_b200_test_token_timestamps.  Returns the number of output segments or -1.
extern "C" int ref_probe_token_timestamps(struct whisper_context * ctx, struct whisper_state * state, const float * energy, int n_energy,
                                          const whisper_token_data * tokens, const int * seg_n_tokens, const int64_t * seg_t0,
                                          const int64_t * seg_t1, int n_segments, float thold_pt, float thold_ptsum, int max_len,
                                          int split_on_word, whisper_token_data * out_tokens, int out_cap_tokens, int64_t * out_t0,
                                          int64_t * out_t1, int * out_n_tokens, int out_cap_segments) {
    state->energy.assign(energy, energy + n_energy);
    state->result_all.clear();
    state->t_beg = 0;
    state->t_last = 0;
    state->tid_last = 0;
    for (int s = 0, at = 0; s < n_segments; at += seg_n_tokens[s], ++s) {
        state->result_all.push_back({});
        auto & seg = state->result_all.back();
        seg.t0 = seg_t0[s];
        seg.t1 = seg_t1[s];
        seg.tokens.assign(tokens + at, tokens + at + seg_n_tokens[s]);
        seg.speaker_turn_next = false;
        whisper_exp_compute_token_level_timestamps(*ctx, *state, (int) state->result_all.size() - 1, thold_pt, thold_ptsum);
        if (max_len > 0) whisper_wrap_segment(*ctx, *state, max_len, split_on_word != 0);
    }
    const auto & all = state->result_all;
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
}

// the reference's alignment-head preset table (g_aheads, whisper.cpp:417-443) as (text layer, head) pairs; -1 when the preset has no table
extern "C" int ref_probe_aheads_preset(int preset, int * pairs, int cap) {
    auto it = g_aheads.find((whisper_alignment_heads_preset) preset);
    if (it == g_aheads.end()) return -1;
    for (size_t i = 0; i < it->second.n_heads && (int) i < cap; ++i) {
        pairs[2 * i] = it->second.heads[i].n_text_layer;
        pairs[2 * i + 1] = it->second.heads[i].n_head;
    }
    return (int) it->second.n_heads;
}

// The reference's unified self-attention KV cache bookkeeping (static whisper_kv_cache_{find_slot,cell_max,seq_rm,seq_cp},
// whisper.cpp:1049-1167) driven by a script of operations on a cache of `size` cells with no tensors behind it.
// ops[5 * i ..]: {0 clear} {1 find_slot: n_tokens, seq, first pos} {2 seq_rm: seq, p0, p1} {3 seq_cp: src, dst, p0, p1} {4 cell_max}
// per_op[3 * i ..] = {return value (bool / cell_max / 0), head, cell_max after the op}; cells_out[2 * c ..] = {pos, seq-id bit mask}.
extern "C" int ref_probe_kv_script(int size, const int * ops, int n_ops, int * per_op, int * cells_out) {
    whisper_kv_cache kv;
    kv.size = (uint32_t) size;
    kv.cells.clear();
    kv.cells.resize((size_t) size);
    for (int i = 0; i < n_ops; ++i) {
        const int * o = ops + 5 * i;
        int ret = 0;
        if (o[0] == 0) {   // whisper_kv_cache_clear without its ggml_backend_buffer_clear (there is no buffer here)
            for (auto & c : kv.cells) { c.pos = -1; c.seq_id.clear(); }
            kv.head = 0;
        } else if (o[0] == 1) {
            const int n = o[1];
            std::vector<whisper_token> tok((size_t) n, 0);
            std::vector<whisper_pos> pos((size_t) n);
            std::vector<int32_t> nseq((size_t) n, 1);
            std::vector<whisper_seq_id> sid((size_t) n, (whisper_seq_id) o[2]);
            std::vector<whisper_seq_id *> sidp((size_t) n);
            std::vector<int8_t> lg((size_t) n, 0);
            for (int k = 0; k < n; ++k) { pos[(size_t) k] = o[3] + k; sidp[(size_t) k] = &sid[(size_t) k]; }
            whisper_batch b{};
            b.n_tokens = n; b.token = tok.data(); b.pos = pos.data(); b.n_seq_id = nseq.data(); b.seq_id = sidp.data(); b.logits = lg.data();
            ret = whisper_kv_cache_find_slot(kv, b) ? 1 : 0;
        } else if (o[0] == 2) {
            whisper_kv_cache_seq_rm(kv, o[1], o[2], o[3]);
        } else if (o[0] == 3) {
            whisper_kv_cache_seq_cp(kv, o[1], o[2], o[3], o[4]);
        } else if (o[0] == 4) {
            ret = whisper_kv_cache_cell_max(kv);
        }
        per_op[3 * i] = ret;
        per_op[3 * i + 1] = (int) kv.head;
        per_op[3 * i + 2] = whisper_kv_cache_cell_max(kv);
    }
    for (int c = 0; c < size; ++c) {
        cells_out[2 * c] = kv.cells[(size_t) c].pos;
        int mask = 0;
        for (auto s : kv.cells[(size_t) c].seq_id) mask |= 1 << s;
        cells_out[2 * c + 1] = mask;
    }
    return 0;
}

// the reference's FFT tables of the mel front end (whisper_global_cache, whisper.cpp:3010-3047): 400 entries each
extern "C" void ref_probe_mel_tables(float * sn, float * cs, float * hn) {
    memcpy(sn, global_cache.sin_vals, sizeof(global_cache.sin_vals));
    memcpy(cs, global_cache.cos_vals, sizeof(global_cache.cos_vals));
    memcpy(hn, global_cache.hann_window, sizeof(global_cache.hann_window));
}
