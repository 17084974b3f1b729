// Device execution engine: lifecycle, mel / encoder / decoder-step drivers.  See engine.h.
#include "engine.h"

#include <math.h>
#include <string.h>

#include <algorithm>
#include <chrono>

#include "log.h"
#include "prof.h"
#include "step.cuh"

namespace wb {

int64_t time_us() {
    return std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// =============================================================================================
// KV cache bookkeeping — same cell model as the reference (wc:1049-1167) so that slot reuse, beam
// sequence copies and the attended-cell sets are identical; the device side only sees cell lists.
// =============================================================================================
void KvCache::clear() {
    for (auto & c : cells) {
        c.pos = -1;
        c.seq = 0;
    }
    head = 0;
}

bool KvCache::find_slot(const int * pos, const int * seq, int n_tokens) {
    const int n_ctx = size;
    if (n_tokens > n_ctx) {
        log_error("%s: n_tokens=%d > n_ctx=%d\n", __func__, n_tokens, n_ctx);
        return false;
    }
    int n_tested = 0;
    while (true) {
        if (head + n_tokens > n_ctx) {
            n_tested += n_ctx - head;
            head = 0;
            continue;
        }
        bool found = true;
        for (int i = 0; i < n_tokens; i++) {
            if (cells[head + i].pos >= 0) {
                found = false;
                head += i + 1;
                n_tested += i + 1;
                break;
            }
        }
        if (found) break;
        if (n_tested >= n_ctx) return false;
    }
    for (int i = 0; i < n_tokens; i++) {
        cells[head + i].pos = pos[i];
        cells[head + i].seq |= 1u << seq[i];
    }
    return true;
}

int KvCache::cell_max() const {
    for (int i = size - 1; i > 0; --i)
        if (cells[i].pos >= 0 && cells[i].seq != 0) return i + 1;
    return 1;
}

void KvCache::seq_rm(int seq_id, int p0, int p1) {
    int new_head = size;
    if (p0 < 0) p0 = 0;
    if (p1 < 0) p1 = 0x7fffffff;
    for (int i = 0; i < size; ++i) {
        if (cells[i].pos >= p0 && cells[i].pos < p1) {
            if (seq_id < 0) cells[i].seq = 0;
            else if (cells[i].has(seq_id)) cells[i].seq &= ~(1u << seq_id);
            else continue;
            if (cells[i].seq == 0) {
                cells[i].pos = -1;
                if (new_head == size) new_head = i;
            }
        }
    }
    if (new_head != size) head = new_head;
}

void KvCache::seq_cp(int src, int dst, int p0, int p1) {
    if (p0 < 0) p0 = 0;
    if (p1 < 0) p1 = 0x7fffffff;
    head = 0;
    for (int i = 0; i < size; ++i)
        if (cells[i].has(src) && cells[i].pos >= p0 && cells[i].pos < p1) cells[i].seq |= 1u << dst;
}

// =============================================================================================
// lifecycle
// =============================================================================================
State::~State() {
    if (ctx) cudaSetDevice(ctx->model.device);
    cudaFree(d_mel);
    cudaFree(d_pcm);
    cudaFree(d_melmax);
    cudaFree(kv_self.k);
    cudaFree(kv_self.v);
    cudaFree(kv_cross);
    cudaFree(d_qk);
    vad_state_release(*this);
}

Context::~Context() {
    sched_shutdown(this);
    cudaSetDevice(model.device);
    if (stream) cudaStreamSynchronize(stream);
    step_release(*this);
    cudaFree(enc_ws.base);
    cudaFree(dec_ws.base);
    if (dec_ws.h_stage) cudaFreeHost(dec_ws.h_stage);
    cudaFree(dec_ws.d_stage);
    if (dec_ws.h_out) cudaFreeHost(dec_ws.h_out);
    cudaFree(dec_ws.d_out);
    cudaFree(dec_ws.d_grammar);
    cudaFree(d_ahead_slot);
    masks.clear();
    if (stream) cudaStreamDestroy(stream);
}

// alignment-head presets (wc:417-443): {text layer, head}
namespace {
struct AH { int l, h; };
static const AH ah_tiny_en[]   = {{1, 0}, {2, 0}, {2, 5}, {3, 0}, {3, 1}, {3, 2}, {3, 3}, {3, 4}};
static const AH ah_tiny[]      = {{2, 2}, {3, 0}, {3, 2}, {3, 3}, {3, 4}, {3, 5}};
static const AH ah_base_en[]   = {{3, 3}, {4, 7}, {5, 1}, {5, 5}, {5, 7}};
static const AH ah_base[]      = {{3, 1}, {4, 2}, {4, 3}, {4, 7}, {5, 1}, {5, 2}, {5, 4}, {5, 6}};
static const AH ah_small_en[]  = {{6, 6}, {7, 0}, {7, 3}, {7, 8}, {8, 2}, {8, 5}, {8, 7}, {9, 0}, {9, 4}, {9, 8}, {9, 10}, {10, 0},
                                  {10, 1}, {10, 2}, {10, 3}, {10, 6}, {10, 11}, {11, 2}, {11, 4}};
static const AH ah_small[]     = {{5, 3}, {5, 9}, {8, 0}, {8, 4}, {8, 7}, {8, 8}, {9, 0}, {9, 7}, {9, 9}, {10, 5}};
static const AH ah_medium_en[] = {{11, 4}, {14, 1}, {14, 12}, {14, 14}, {15, 4}, {16, 0}, {16, 4}, {16, 9}, {17, 12}, {17, 14},
                                  {18, 7}, {18, 10}, {18, 15}, {20, 0}, {20, 3}, {20, 9}, {20, 14}, {21, 12}};
static const AH ah_medium[]    = {{13, 15}, {15, 4}, {15, 15}, {16, 1}, {20, 0}, {23, 4}};
static const AH ah_large_v1[]  = {{9, 19}, {11, 2}, {11, 4}, {11, 17}, {22, 7}, {22, 11}, {22, 17}, {23, 2}, {23, 15}};
static const AH ah_large_v2[]  = {{10, 12}, {13, 17}, {16, 11}, {16, 12}, {16, 13}, {17, 15}, {17, 16}, {18, 4}, {18, 11}, {18, 19},
                                  {19, 11}, {21, 2}, {21, 3}, {22, 3}, {22, 9}, {22, 12}, {23, 5}, {23, 7}, {23, 13}, {25, 5},
                                  {26, 1}, {26, 12}, {27, 15}};
static const AH ah_large_v3[]  = {{7, 0}, {10, 17}, {12, 18}, {13, 12}, {16, 1}, {17, 14}, {19, 11}, {21, 4}, {24, 1}, {25, 6}};
static const AH ah_large_v3_turbo[] = {{2, 4}, {2, 11}, {3, 3}, {3, 6}, {3, 11}, {3, 14}};
#define AHP(a) {a, (int) (sizeof(a) / sizeof(a[0]))}
static const struct { const AH * p; int n; } ah_presets[] = {
    {nullptr, 0}, {nullptr, 0}, {nullptr, 0}, AHP(ah_tiny_en), AHP(ah_tiny), AHP(ah_base_en), AHP(ah_base), AHP(ah_small_en),
    AHP(ah_small), AHP(ah_medium_en), AHP(ah_medium), AHP(ah_large_v1), AHP(ah_large_v2), AHP(ah_large_v3), AHP(ah_large_v3_turbo)};
#undef AHP
}  // namespace

// the preset table alone (tests): (text layer, head) pairs of `preset` in table order; returns their number, -1 for a non-table preset
int alignment_heads_preset(int preset, int * pairs, int cap) {
    if (preset < 0 || preset >= (int) (sizeof(ah_presets) / sizeof(ah_presets[0])) || ah_presets[preset].p == nullptr) return -1;
    for (int i = 0; i < ah_presets[preset].n && i < cap && pairs; ++i) {
        pairs[2 * i] = ah_presets[preset].p[i].l;
        pairs[2 * i + 1] = ah_presets[preset].p[i].h;
    }
    return ah_presets[preset].n;
}

// Validates and tabulates the alignment heads (aheads_masks_init wc:1190-1303; order: get_alignment_heads_by_layer wc:8623-8642).
// Slot numbering = concatenation over layers in layer order, heads in preset order (ggml_concat order, wc:2746-2750).
static bool setup_alignment_heads(Context & ctx) {
    const auto & cp = ctx.params;
    const int L = ctx.model.hp.n_text_layer, H = ctx.model.hp.n_text_head;
    ctx.ahead_slot.assign(L, std::vector<int>(H, -1));
    ctx.n_aheads = 0;
    if (!cp.dtw_token_timestamps) return true;
    if (cp.dtw_aheads_preset == WHISPER_AHEADS_NONE) {
        log_error("%s: dtw_aheads_preset should be != DTW_AHEADS_NONE\n", __func__);
        return false;
    }
    std::vector<AH> heads;
    if (cp.dtw_aheads_preset == WHISPER_AHEADS_N_TOP_MOST) {
        if (cp.dtw_n_top > L || cp.dtw_n_top <= 0) {
            log_error("%s: dtw_n_top must be between %d and %d for this model.", __func__, 1, L);
            return false;
        }
        for (int l = L - cp.dtw_n_top; l < L; ++l)
            for (int h = 0; h < H; ++h) heads.push_back({l, h});
    } else if (cp.dtw_aheads_preset == WHISPER_AHEADS_CUSTOM) {
        if (cp.dtw_aheads.n_heads == 0) {
            log_error("%s: dtw_aheads.n_heads should be > 0", __func__);
            return false;
        }
        if (cp.dtw_aheads.heads == nullptr) {
            log_error("%s: dtw_aheads.heads unset", __func__);
            return false;
        }
        for (size_t i = 0; i < cp.dtw_aheads.n_heads; ++i) heads.push_back({cp.dtw_aheads.heads[i].n_text_layer, cp.dtw_aheads.heads[i].n_head});
    } else {
        const int pi = (int) cp.dtw_aheads_preset;
        if (pi < 0 || pi >= (int) (sizeof(ah_presets) / sizeof(ah_presets[0]))) {
            log_error("%s: unknown alignment heads preset %d\n", __func__, pi);
            return false;
        }
        heads.assign(ah_presets[pi].p, ah_presets[pi].p + ah_presets[pi].n);
    }
    for (const auto & a : heads) {
        if (a.l >= L || a.l < 0 || a.h >= H || a.h < 0) {
            log_error("%s: alignment head (%d, %d) is outside this model (%d text layers, %d heads)\n", __func__, a.l, a.h, L, H);
            return false;
        }
    }
    int slot = 0;
    for (int l = 0; l < L; ++l)
        for (const auto & a : heads)
            if (a.l == l) ctx.ahead_slot[l][a.h] = slot++;
    ctx.n_aheads = slot;
    std::vector<int> flat((size_t) L * H);
    for (int l = 0; l < L; ++l)
        for (int h = 0; h < H; ++h) flat[(size_t) l * H + h] = ctx.ahead_slot[l][h];
    WB_CUDA_OK(cudaMalloc((void **) &ctx.d_ahead_slot, flat.size() * sizeof(int)));
    WB_CUDA_OK(cudaMemcpy(ctx.d_ahead_slot, flat.data(), flat.size() * sizeof(int), cudaMemcpyHostToDevice));
    return true;
}

Context * context_create(whisper_model_loader * loader, whisper_context_params params) {
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) {
        log_error("%s: no CUDA device is visible: this library has no CPU path and cannot run\n", __func__);
        return nullptr;
    }
    if (params.gpu_device < 0 || params.gpu_device >= n_dev) {
        log_error("%s: gpu_device %d out of range (%d devices)\n", __func__, params.gpu_device, n_dev);
        return nullptr;
    }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, params.gpu_device) != cudaSuccess || prop.major != 10) {
        log_error("%s: device %d is sm_%d%d; this library is built for sm_100a (B200) only\n", __func__, params.gpu_device, prop.major,
                  prop.minor);
        return nullptr;
    }
    whisper_context * ctx = new whisper_context;
    ctx->params = params;
    ctx->t_start_us = time_us();
    if (!load_model(loader, params.gpu_device, ctx->model)) {
        delete ctx;
        return nullptr;
    }
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
        log_error("%s: failed to create stream\n", __func__);
        delete ctx;
        return nullptr;
    }
    ctx->t_load_us = time_us() - ctx->t_start_us;
    return ctx;
}

bool kv_self_alloc(State & st, int n_ctx) {
    const auto & hp = st.ctx->model.hp;
    WB_CUDA_OK(cudaSetDevice(st.ctx->model.device));
    cudaFree(st.kv_self.k);
    cudaFree(st.kv_self.v);
    st.kv_self.k = st.kv_self.v = nullptr;
    const size_t n = (size_t) hp.n_text_layer * n_ctx * hp.n_text_state;
    WB_CUDA_OK(cudaMalloc((void **) &st.kv_self.k, n * sizeof(__half)));
    WB_CUDA_OK(cudaMalloc((void **) &st.kv_self.v, n * sizeof(__half)));
    st.kv_self.size = n_ctx;
    st.kv_self.cells.assign((size_t) n_ctx, KvCell());
    st.kv_self.head = 0;
    return true;
}

State * state_create(Context * ctx) {
    const auto & hp = ctx->model.hp;
    if (cudaSetDevice(ctx->model.device) != cudaSuccess) return nullptr;
    if (ctx->params.dtw_token_timestamps && ctx->d_ahead_slot == nullptr) {
        if (!setup_alignment_heads(*ctx)) {
            log_error("%s: aheads_masks_init() failed for alignment heads masks\n", __func__);
            return nullptr;
        }
    }
    whisper_state * st = new whisper_state;
    st->ctx = ctx;
    st->kv_self_n_dec = 1;
    bool ok = kv_self_alloc(*st, (int) pad_to(hp.n_text_ctx, 256));  // wc:3403-3406
    const size_t cross = (size_t) hp.n_text_layer * 2 * hp.n_text_head * hp.n_audio_ctx * 64;
    ok = ok && cudaMalloc((void **) &st->kv_cross, cross * sizeof(__half)) == cudaSuccess;
    ok = ok && cudaMalloc((void **) &st->d_melmax, sizeof(int)) == cudaSuccess;
    if (!ok) {
        log_error("%s: failed to allocate the KV caches\n", __func__);
        delete st;
        return nullptr;
    }
    st->decoders[0].rng = std::mt19937(0);  // wc:3486
    return st;
}

// =============================================================================================
// workspaces
// =============================================================================================
// WB200_ATTN=legacy: materialised scores + softmax kernel + P.V GEMM (the round-1 v2 path; kept for A/B measurements)
static bool legacy_attention() {
    static int v = -1;
    if (v < 0) {
        const char * e = getenv("WB200_ATTN");
        v = (e && strcmp(e, "legacy") == 0) ? 1 : 0;
    }
    return v == 1;
}

static bool ensure_enc_ws(Context & ctx, int windows, int T) {
    EncWorkspace & ws = ctx.enc_ws;
    if (ws.base && ws.windows >= windows && ws.T == T) return true;
    const auto & m = ctx.model;
    const int S = m.hp.n_audio_state, H = m.hp.n_audio_head, W = std::max(windows, ws.T == T ? ws.windows : 0);
    const int Tpad = (int) pad_to(T, 8);
    WB_CUDA_OK(cudaStreamSynchronize(ctx.stream));
    cudaFree(ws.base);
    ws = EncWorkspace();
    size_t off = 0;
    auto take = [&](size_t bytes) {
        const size_t o = off;
        off += (bytes + 1023) & ~(size_t) 1023;
        return o;
    };
    const size_t R = (size_t) W * T;
    const size_t o_melT = take((size_t) W * (2 * T + 2) * m.kc1 * 2), o_h1 = take((size_t) W * (2 * T + 2) * S * 2);
    const size_t o_x = take(R * S * 4), o_xa = take(R * S * 2), o_qk = take(3 * R * S * 2);  // Q, K (, V) per-head panels
    const bool legacy = legacy_attention();
    const size_t o_vT = take((size_t) H * W * 64 * Tpad * 2), o_sc = take(legacy ? (size_t) H * W * T * Tpad * 4 : 16);
    const size_t o_p = take(legacy ? (size_t) H * W * T * Tpad * 2 : 16), o_att = take(R * S * 2), o_hb = take(R * 4 * S * 2);
    const size_t o_e16 = take(R * S * 2), o_e32 = take(R * S * 4);
    WB_CUDA_OK(cudaMalloc((void **) &ws.base, off));
    WB_CUDA_OK(cudaMemset(ws.base, 0, off));  // conv padding rows and the V^T / P tails must start as zeros
    ws.bytes = off;
    ws.melT = (__half *) (ws.base + o_melT); ws.h1 = (__half *) (ws.base + o_h1); ws.x = (float *) (ws.base + o_x);
    ws.xa = (__half *) (ws.base + o_xa); ws.qk = (__half *) (ws.base + o_qk); ws.vT = (__half *) (ws.base + o_vT);
    ws.sc = (float *) (ws.base + o_sc); ws.p = (__half *) (ws.base + o_p); ws.att = (__half *) (ws.base + o_att);
    ws.hb = (__half *) (ws.base + o_hb); ws.enc16 = (__half *) (ws.base + o_e16); ws.enc32 = (float *) (ws.base + o_e32);
    ws.windows = W;
    ws.T = T;
    ws.Tpad = Tpad;
    return true;
}

static bool ensure_dec_ws(Context & ctx, int rows, int logit_rows, int max_cells) {
    DecWorkspace & ws = ctx.dec_ws;
    if (ws.base && ws.rows >= rows && ws.logit_rows >= logit_rows && ws.max_cells >= max_cells) return true;
    const auto & hp = ctx.model.hp;
    const int S = hp.n_text_state, V = hp.n_vocab;
    // Growing means cudaFree / cudaFreeHost / cudaHostAlloc on a process that holds ~150 GB of device memory and hundreds of pinned
    // regions: measured 0.85 s for ONE regrow in the middle of a 448-state run (a second-window prompt with 101 carried-over tokens
    // raised max_cells from 3 to 105).  So: start at the sizes the greedy path can reach (512 cells = one decoder's whole cache,
    // wc:7101-7106) and grow geometrically.
    auto grow = [](int need, int have, int floor_) { return need <= have ? have : std::max(std::max(need, floor_), have + have / 2); };
    const int R = grow(rows, ws.rows, 64), RL = grow(logit_rows, ws.logit_rows, MAX_DECODERS), MC = grow(max_cells, ws.max_cells, 512);
    WB_CUDA_OK(cudaStreamSynchronize(ctx.stream));
    cudaFree(ws.base);
    if (ws.h_stage) cudaFreeHost(ws.h_stage);
    cudaFree(ws.d_stage);
    if (ws.h_out) cudaFreeHost(ws.h_out);
    cudaFree(ws.d_out);
    cudaFree(ws.d_grammar);
    ws = DecWorkspace();
    size_t off = 0;
    auto take = [&](size_t bytes) {
        const size_t o = off;
        off += (bytes + 1023) & ~(size_t) 1023;
        return o;
    };
    const size_t o_x = take((size_t) R * S * 4), o_xa = take((size_t) R * S * 2), o_qkv = take((size_t) R * 3 * S * 2);
    const size_t o_q = take((size_t) R * S * 2), o_att = take((size_t) R * S * 2), o_hb = take((size_t) R * 4 * S * 2);
    const size_t o_lg = take((size_t) RL * V * 4), o_sw = take((size_t) RL * 3 * V * 4), o_cells = take((size_t) R * MC * 4);
    WB_CUDA_OK(cudaMalloc((void **) &ws.base, off));
    ws.x = (float *) (ws.base + o_x); ws.xa = (__half *) (ws.base + o_xa); ws.qkv = (__half *) (ws.base + o_qkv);
    ws.q = (__half *) (ws.base + o_q); ws.att = (__half *) (ws.base + o_att); ws.hb = (__half *) (ws.base + o_hb);
    ws.logits = (float *) (ws.base + o_lg); ws.samp_work = (float *) (ws.base + o_sw); ws.d_cells = (int *) (ws.base + o_cells);
    // a multiple of 16: the sampler rows sit at the END of the block (stage_sample) and hold pointers (8-byte aligned loads)
    ws.stage_bytes = (((size_t) R * (8 + sizeof(SelfKvRow) + sizeof(CrossKvRow) + sizeof(SamplerRow) + (size_t) MC * 4) + 4096) + 15) & ~(size_t) 15;
    WB_CUDA_OK(cudaMallocHost((void **) &ws.h_stage, ws.stage_bytes));
    WB_CUDA_OK(cudaMalloc((void **) &ws.d_stage, ws.stage_bytes));
    WB_CUDA_OK(cudaMallocHost((void **) &ws.h_out, (size_t) RL * sizeof(SamplerOut)));
    WB_CUDA_OK(cudaMalloc((void **) &ws.d_out, (size_t) RL * sizeof(SamplerOut)));
    ws.rows = R;
    ws.logit_rows = RL;
    ws.max_cells = MC;
    return true;
}

// =============================================================================================
// M1: mel
// =============================================================================================
bool stage_pcm_to_mel(State & st, const float * samples, int n_samples, bool sync) {
    Context & ctx = *st.ctx;
    const int64_t t0 = time_us();
    WB_CUDA_OK(cudaSetDevice(ctx.model.device));
    const int n_mel = ctx.model.filters_n_mel;
    const int n_len = mel_n_len(n_samples);
    if ((size_t) n_samples > st.pcm_cap) {
        cudaFree(st.d_pcm);
        st.d_pcm = nullptr;
        WB_CUDA_OK(cudaMalloc((void **) &st.d_pcm, (size_t) std::max(n_samples, 16) * sizeof(float)));
        st.pcm_cap = (size_t) std::max(n_samples, 16);
    }
    const size_t need = (size_t) n_mel * n_len;
    if (need > st.mel_cap) {
        cudaFree(st.d_mel);
        st.d_mel = nullptr;
        WB_CUDA_OK(cudaMalloc((void **) &st.d_mel, need * sizeof(float)));
        st.mel_cap = need;
    }
    WB_CUDA_OK(cudaMemcpyAsync(st.d_pcm, samples, (size_t) n_samples * sizeof(float), cudaMemcpyDefault /* host or device PCM */, ctx.stream));
    MelTables tab{ctx.model.mel_sin, ctx.model.mel_cos, ctx.model.mel_hann, ctx.model.mel_filters_t, n_mel};
    if (!mel_compute(tab, st.d_pcm, n_samples, n_len, st.d_mel, st.d_melmax, ctx.stream)) return false;
    // sync == false: the caller batches the front end of many states and synchronises once (stage_mel_sync); nothing on the host
    // depends on the result, and the encoder consumes it on the same stream
    if (sync) WB_CUDA_OK(cudaStreamSynchronize(ctx.stream));
    st.n_mel = n_mel;
    st.n_len = n_len;
    st.n_len_org = mel_n_len_org(n_samples);
    st.t_mel_us += time_us() - t0;
    return true;
}

// one synchronisation for a batch of stage_pcm_to_mel(..., sync = false); the wait is shared out to the states' mel timers
bool stage_mel_sync(Context & ctx, State * const * states, int n) {
    const int64_t t0 = time_us();
    WB_CUDA_OK(cudaStreamSynchronize(ctx.stream));
    const int64_t dt = time_us() - t0;
    for (int i = 0; i < n; ++i) states[i]->t_mel_us += dt / std::max(1, n);
    return true;
}

bool stage_set_mel(State & st, const float * data, int n_len, int n_mel) {
    Context & ctx = *st.ctx;
    WB_CUDA_OK(cudaSetDevice(ctx.model.device));
    const size_t need = (size_t) n_mel * n_len;
    if (need > st.mel_cap) {
        cudaFree(st.d_mel);
        st.d_mel = nullptr;
        WB_CUDA_OK(cudaMalloc((void **) &st.d_mel, std::max<size_t>(need, 16) * sizeof(float)));
        st.mel_cap = std::max<size_t>(need, 16);
    }
    WB_CUDA_OK(cudaMemcpyAsync(st.d_mel, data, need * sizeof(float), cudaMemcpyHostToDevice, ctx.stream));
    WB_CUDA_OK(cudaStreamSynchronize(ctx.stream));
    st.n_len = n_len;
    st.n_len_org = n_len;
    st.n_mel = n_mel;
    return true;
}

bool stage_get_mel(State & st, std::vector<float> & out) {
    out.resize((size_t) st.n_mel * st.n_len);
    if (out.empty()) return true;
    WB_CUDA_OK(cudaSetDevice(st.ctx->model.device));
    WB_CUDA_OK(cudaMemcpy(out.data(), st.d_mel, out.size() * sizeof(float), cudaMemcpyDeviceToHost));
    return true;
}

// =============================================================================================
// E0-E3: encoder
// =============================================================================================
static inline GemmOperand opnd(const void * p, int64_t rows, int64_t row_stride, int64_t batch = 1, int64_t batch_stride = 0) {
    GemmOperand o{};
    o.ptr = p; o.rows = rows; o.row_stride = row_stride; o.batch = batch; o.batch_stride = batch_stride; o.phases = 1; o.phase_stride = 0;
    return o;
}
static inline GemmEpilogue epi_row(void * out, bool f16, int64_t ldo) {
    GemmEpilogue e{};
    e.out = out; e.out_f16 = f16 ? 1 : 0; e.out_mode = OUT_ROW; e.ldo = ldo; e.batch_div = 1;
    return e;
}

// E1 (wc:1994-2054, 2106-2107): the mel slice of every window through the two convolutions into the residual stream ws.x
static bool conv_stem(Context & ctx, State * const * states, const int * mel_offsets, int W, int T, bool add_pos_emb) {
    const Model & m = ctx.model;
    const auto & hp = m.hp;
    const int S = hp.n_audio_state;
    EncWorkspace & ws = ctx.enc_ws;
    const int F = 2 * T;  // F mel frames per window
    g_prof_gemm_cat = PC_GEMM_ENC;
    cudaStream_t s = ctx.stream;

    for (int w = 0; w < W; ++w) {
        State & st = *states[w];
        WB_CHECK(st.n_mel == hp.n_mels, "mel bands %d != model n_mels %d\n", st.n_mel, hp.n_mels);
        if (!mel_window_to_f16(st.d_mel, st.n_mel, st.n_len, mel_offsets[w], F, ws.melT + (size_t) w * (F + 2) * m.kc1, m.kc1, s))
            return false;
    }
    // E1: conv1d(k3,s1,p1)+bias+GELU -> f16 ; conv1d(k3,s2,p1)+bias+GELU + positional embedding -> f32   (wc:2024-2032, 2106-2107)
    {
        GemmOperand A = opnd(ws.melT, F + 2, m.kc1, W, (int64_t) (F + 2) * m.kc1);
        GemmOperand B = opnd(m.conv1_w, S, 3 * m.kc1);
        GemmTaps taps{3, {0, 0, 0}, {0, 1, 2}, m.kc1};
        GemmEpilogue e = epi_row(ws.h1 + S, true, S);
        e.out_batch_stride = (int64_t) (F + 2) * S;
        e.bias = m.conv1_b;
        e.gelu_table = m.gelu_table;
        if (!gemm_f16_tc(A, B, F, S, m.kc1, W, e, 0, s, &taps)) return false;
    }
    {
        GemmOperand A = opnd(ws.h1, T + 1, 2 * S, W, (int64_t) (F + 2) * S);
        A.phases = 2;
        A.phase_stride = S;
        GemmOperand B = opnd(m.conv2_w, S, 3 * S);
        GemmTaps taps{3, {0, 1, 0}, {0, 0, 1}, S};
        GemmEpilogue e = epi_row(ws.x, false, S);
        e.out_batch_stride = (int64_t) T * S;
        e.bias = m.conv2_b;
        e.gelu_table = m.gelu_table;
        if (add_pos_emb) {
            e.res = m.e_pe;
            e.ldr = S;
        }
        if (!gemm_f16_tc(A, B, T, S, S, W, e, 0, s, &taps)) return false;
    }
    return true;
}

// test probe (whisper_b200_test_conv_stem): the conv stem alone, WITHOUT the positional embedding = the reference's embd_conv
bool stage_conv_probe(Context & ctx, State * st, int mel_offset, float * dst, int64_t cap) {
    WB_CUDA_OK(cudaSetDevice(ctx.model.device));
    const int T = st->exp_n_audio_ctx > 0 ? st->exp_n_audio_ctx : ctx.model.hp.n_audio_ctx;
    const int64_t n = (int64_t) T * ctx.model.hp.n_audio_state;
    if (cap < n) return false;
    if (!ensure_enc_ws(ctx, 1, T) || !conv_stem(ctx, &st, &mel_offset, 1, T, false)) return false;
    WB_CUDA_OK(cudaMemcpyAsync(dst, ctx.enc_ws.x, (size_t) n * 4, cudaMemcpyDeviceToHost, ctx.stream));
    WB_CUDA_OK(cudaStreamSynchronize(ctx.stream));
    return true;
}

static bool encode_chunk(Context & ctx, State * const * states, const int * mel_offsets, int W, int T) {
    const Model & m = ctx.model;
    const auto & hp = m.hp;
    const int S = hp.n_audio_state, H = hp.n_audio_head, L = hp.n_audio_layer, Lt = hp.n_text_layer;
    if (!ensure_enc_ws(ctx, W, T)) return false;
    EncWorkspace & ws = ctx.enc_ws;
    const int Tpad = ws.Tpad, R = W * T;
    g_prof_gemm_cat = PC_GEMM_ENC;
    cudaStream_t s = ctx.stream;
    if (!conv_stem(ctx, states, mel_offsets, W, T, true)) return false;
    // E2: encoder blocks (wc:2116-2255)
    const float kq_scale = 1.0f / sqrtf((float) 64);
    for (int il = 0; il < L; ++il) {
        const EncLayer & l = m.enc[il];
        if (!layernorm_rows(ws.x, R, S, l.ln1.g, l.ln1.b, hp.eps, ws.xa, nullptr, s)) return false;
        if (!legacy_attention()) {
            // Q (bias), K (no bias), V (bias) in ONE GEMM -> f16 per-head [t][64] panels [3H][R][64]  (wc:2131-2148; the f16 casts of
            // wc:2183-2200 are the epilogue's rounding)
            GemmEpilogue e{};
            e.out = ws.qk; e.out_f16 = 1; e.out_mode = OUT_HEAD; e.head_stride = (int64_t) R * 64; e.batch_div = 1;
            e.bias = l.qk.b;
            if (!gemm_f16_tc(opnd(ws.xa, R, S), opnd(l.qk.w, 3 * S, S), R, 3 * S, S, 1, e, 0, s)) return false;
        } else {
        {   // Q (with bias) and K (no bias) -> per-head [t][64] panels, f16 (the casts of wc:2183-2186 and mul_mat's f32->f16)
            GemmEpilogue e{};
            e.out = ws.qk; e.out_f16 = 1; e.out_mode = OUT_HEAD; e.head_stride = (int64_t) R * 64; e.batch_div = 1;
            e.bias = l.qk.b;
            if (!gemm_f16_tc(opnd(ws.xa, R, S), opnd(l.qk.w, 2 * S, S), R, 2 * S, S, 1, e, 0, s)) return false;
        }
        {   // V (with bias) -> per-head transposed [64][t] panels, f16 (wc:2193-2200)
            GemmEpilogue e{};
            e.out = ws.vT; e.out_f16 = 1; e.out_mode = OUT_HEAD_COL; e.ldo = Tpad; e.head_stride = (int64_t) W * 64 * Tpad;
            e.rows_per_panel = T; e.panel_stride = (int64_t) 64 * Tpad; e.batch_div = 1;
            e.bias = l.v.b;
            if (!gemm_f16_tc(opnd(ws.xa, R, S), opnd(l.v.w, S, S), R, S, S, 1, e, 0, s)) return false;
        }
        }
        if (!legacy_attention()) {
            // fused: softmax(Q K^T * D^-1/2) V with the scores in TMEM (wc:2183-2206)
            if (!enc_attention(ws.qk, nullptr, ws.att, W, T, Tpad, H, S, kq_scale, s)) return false;
        } else {
        {   // scores = Q K^T per (head, window)  (wc:2189)
            GemmEpilogue e = epi_row(ws.sc, false, Tpad);
            e.out_batch_stride = (int64_t) T * Tpad;
            if (!gemm_f16_tc(opnd(ws.qk, T, 64, (int64_t) H * W, (int64_t) T * 64),
                             opnd(ws.qk + (size_t) H * R * 64, T, 64, (int64_t) H * W, (int64_t) T * 64), T, T, 64, H * W, e, 0, s))
                return false;
        }
        if (!softmax_rows_f16(ws.sc, (int64_t) H * W * T, T, Tpad, kq_scale, ws.p, Tpad, s)) return false;  // wc:2191
        {   // P V -> merged heads [t][S] f16 (wc:2202-2206)
            GemmEpilogue e = epi_row(ws.att, true, S);
            e.batch_div = W;
            e.out_batch_stride = 64;                  // head index
            e.out_batch_stride2 = (int64_t) T * S;    // window index
            if (!gemm_f16_tc(opnd(ws.p, T, Tpad, (int64_t) H * W, (int64_t) T * Tpad),
                             opnd(ws.vT, 64, Tpad, (int64_t) H * W, (int64_t) 64 * Tpad), T, 64, T, H * W, e, 64, s))
                return false;
        }
        }
        {   // output projection + bias + residual (wc:2212-2220)
            GemmEpilogue e = epi_row(ws.x, false, S);
            e.bias = l.o.b; e.res = ws.x; e.ldr = S;
            if (!gemm_f16_tc(opnd(ws.att, R, S), opnd(l.o.w, S, S), R, S, S, 1, e, 0, s)) return false;
        }
        if (!layernorm_rows(ws.x, R, S, l.ln2.g, l.ln2.b, hp.eps, ws.xa, nullptr, s)) return false;
        {   // MLP (wc:2237-2254)
            // fc1 stores h = f16(W1.x + b1), the index of the reference's GELU table lookup; the lookup itself is a separate
            // shared-memory-table pass (a gather from the 128 KB table inside the GEMM epilogue costs more than the GEMM)
            GemmEpilogue e = epi_row(ws.hb, true, 4 * S);
            e.bias = l.fc1.b;
            if (!gemm_f16_tc(opnd(ws.xa, R, S), opnd(l.fc1.w, 4 * S, S), R, 4 * S, S, 1, e, 0, s)) return false;
            if (!gelu_table_inplace(ws.hb, (int64_t) R * 4 * S, m.gelu_table, s)) return false;
            GemmEpilogue e2 = epi_row(ws.x, false, S);
            e2.bias = l.fc2.b; e2.res = ws.x; e2.ldr = S;
            if (!gemm_f16_tc(opnd(ws.hb, R, 4 * S), opnd(l.fc2.w, S, 4 * S), R, S, 4 * S, 1, e2, 0, s)) return false;
        }
    }
    if (!layernorm_rows(ws.x, R, S, m.e_ln.g, m.e_ln.b, hp.eps, ws.enc16, ws.enc32, s)) return false;  // wc:2259-2267

    // E3: cross-attention K (pre-scaled by D^-1/4) and V for every text layer, straight into each state's cache (wc:2318-2357)
    const float k_scale = (float) pow((double) 64.0f, -0.25);
    for (int w = 0; w < W; ++w) {
        State & st = *states[w];
        GemmEpilogue e{};
        e.out = st.kv_cross; e.out_f16 = 1; e.out_mode = OUT_HEAD; e.head_stride = (int64_t) T * 64; e.batch_div = 1;
        e.out_batch_stride = (int64_t) 2 * H * T * 64;
        e.bias = m.ckv_b; e.bias_batch_stride = 2 * S;
        e.scale = k_scale; e.scale_n0 = 0; e.scale_n1 = S;
        if (!gemm_f16_tc(opnd(ws.enc16 + (size_t) w * T * S, T, S), opnd(m.ckv_w, 2 * S, S, Lt, (int64_t) 2 * S * S), T, 2 * S, S, Lt, e, 0, s))
            return false;
        st.cross_T = T;
    }
    return true;
}

bool stage_encode(Context & ctx, State * const * states, const int * mel_offsets, int n_states) {
    WB_CUDA_OK(cudaSetDevice(ctx.model.device));
    const int64_t t0 = time_us();
    const auto & hp = ctx.model.hp;
    // windows per encoder pass: the legacy attention path materialises [H][W][T][T] scores; the fused path only needs
    // activations (64 windows of large-v3: ~2.6 GB), and more rows per GEMM means fuller waves on the 148 SMs
    const int kMaxWindows = legacy_attention() ? 8 : 64;
    int i = 0;
    while (i < n_states) {
        const int T = states[i]->exp_n_audio_ctx > 0 ? states[i]->exp_n_audio_ctx : hp.n_audio_ctx;
        int W = 1;
        while (i + W < n_states && W < kMaxWindows) {
            const int T2 = states[i + W]->exp_n_audio_ctx > 0 ? states[i + W]->exp_n_audio_ctx : hp.n_audio_ctx;
            if (T2 != T) break;
            ++W;
        }
        if (!encode_chunk(ctx, states + i, mel_offsets + i, W, T)) return false;
        i += W;
    }
    WB_CUDA_OK(cudaStreamSynchronize(ctx.stream));
    const int64_t dt = time_us() - t0;
    for (int k = 0; k < n_states; ++k) {
        states[k]->t_encode_us += dt / n_states;
        states[k]->n_encode++;
    }
    return true;
}

// =============================================================================================
// D0/D1: decoder forward over a batch of rows
// =============================================================================================
bool stage_decode(Context & ctx, const BatchRow * rows, int n, bool save_qk) {
    WB_CUDA_OK(cudaSetDevice(ctx.model.device));
    const int64_t t0 = time_us();
    static const bool dlog = getenv("WB200_DECODE_LOG") != nullptr;  // debugging aid: host-clock phases of the general path (adds syncs)
    auto mark = [&](const char * what) {
        if (!dlog) return;
        cudaStreamSynchronize(ctx.stream);
        fprintf(stderr, "stage_decode n=%d %s +%lld us\n", n, what, (long long) (time_us() - t0));
    };
    const Model & m = ctx.model;
    const auto & hp = m.hp;
    const int S = hp.n_text_state, H = hp.n_text_head, L = hp.n_text_layer, V = hp.n_vocab;
    cudaStream_t s = ctx.stream;

    // ---- KV slots (wc:2885-2897) and attended-cell lists (mask construction wc:2924-2956)
    int n_flag = 0, max_cells = 1;
    for (int i = 0; i < n;) {
        State * st = rows[i].st;
        int cnt = 1;
        while (i + cnt < n && rows[i + cnt].st == st) ++cnt;
        std::vector<int> pos(cnt), seq(cnt);
        for (int k = 0; k < cnt; ++k) {
            pos[k] = rows[i + k].pos;
            seq[k] = rows[i + k].seq_id;
        }
        if (!st->kv_self.find_slot(pos.data(), seq.data(), cnt)) {
            log_error("%s: failed to find a KV slot for %d tokens\n", __func__, cnt);
            return false;
        }
        st->kv_self.n = std::min(st->kv_self.size, std::max(1, st->kv_self.cell_max()));
        max_cells = std::max(max_cells, st->kv_self.n);
        i += cnt;
    }
    for (int i = 0; i < n; ++i) n_flag += rows[i].want_logits ? 1 : 0;
    mark("slots");
    if (!ensure_dec_ws(ctx, std::max(n, 8), std::max(n_flag, MAX_DECODERS), max_cells)) return false;
    DecWorkspace & ws = ctx.dec_ws;
    mark("workspace");

    // staging layout: tokens | pos | SelfKvRow[] | CrossKvRow[] | cell lists
    uint8_t * hs = ws.h_stage;
    size_t so = 0;
    auto sub = [&](size_t bytes) {
        const size_t o = so;
        so += (bytes + 15) & ~(size_t) 15;
        return o;
    };
    const size_t o_tok = sub((size_t) n * 4), o_pos = sub((size_t) n * 4), o_self = sub((size_t) n * sizeof(SelfKvRow));
    const size_t o_cross = sub((size_t) n * sizeof(CrossKvRow)), o_cells = sub((size_t) n * max_cells * 4);
    WB_CHECK(so <= ws.stage_bytes, "decode staging overflow\n");
    int * h_tok = (int *) (hs + o_tok);
    int * h_pos = (int *) (hs + o_pos);
    SelfKvRow * h_self = (SelfKvRow *) (hs + o_self);
    CrossKvRow * h_cross = (CrossKvRow *) (hs + o_cross);
    int * h_cells = (int *) (hs + o_cells);
    int * d_cells = (int *) (ws.d_stage + o_cells);

    int n_tok_total = 0;
    for (int i = 0; i < n;) {
        State * st = rows[i].st;
        int cnt = 1;
        while (i + cnt < n && rows[i + cnt].st == st) ++cnt;
        KvCache & kv = st->kv_self;
        const int T = st->cross_T;
        WB_CHECK(T > 0, "decode called before encode\n");
        if (save_qk) {
            WB_CHECK(cnt == n, "alignment capture needs a single-state batch\n");
            const size_t need = (size_t) ctx.n_aheads * cnt * T;
            if (need > st->qk_cap) {
                cudaFree(st->d_qk);
                st->d_qk = nullptr;
                WB_CUDA_OK(cudaMalloc((void **) &st->d_qk, need * sizeof(float)));
                st->qk_cap = need;
            }
            n_tok_total = cnt;
        }
        for (int k = 0; k < cnt; ++k) {
            const BatchRow & r = rows[i + k];
            h_tok[i + k] = r.token;
            h_pos[i + k] = r.pos;
            SelfKvRow & sr = h_self[i + k];
            sr.k_base = kv.k;
            sr.v_base = kv.v;
            sr.n_ctx = kv.size;
            sr.store_cell = kv.head + k;
            int * list = h_cells + (size_t) (i + k) * max_cells;
            int nc = 0;
            bool identity = true;
            for (int c = 0; c < kv.n; ++c) {
                if (kv.cells[c].has(r.seq_id) && kv.cells[c].pos <= r.pos) {
                    if (c != nc) identity = false;
                    list[nc++] = c;
                }
            }
            sr.n_cells = nc;
            sr.cells = identity ? nullptr : d_cells + (size_t) (i + k) * max_cells;
            CrossKvRow & cr = h_cross[i + k];
            cr.kv_base = st->kv_cross;
            cr.qk_out = save_qk ? st->d_qk : nullptr;
            cr.tok_index = k;
            cr.n_tok_total = n_tok_total;
        }
        i += cnt;
    }
    WB_CUDA_OK(cudaMemcpyAsync(ws.d_stage, hs, so, cudaMemcpyHostToDevice, s));
    const int * d_tok = (const int *) (ws.d_stage + o_tok);
    const int * d_pos = (const int *) (ws.d_stage + o_pos);
    const SelfKvRow * d_self = (const SelfKvRow *) (ws.d_stage + o_self);
    const CrossKvRow * d_cross = (const CrossKvRow *) (ws.d_stage + o_cross);

    const int T = rows[0].st->cross_T;
    for (int i = 1; i < n; ++i) WB_CHECK(rows[i].st->cross_T == T, "mixed audio contexts in one decode batch\n");

    mark("staged");
    g_prof_gemm_cat = PC_GEMM_DEC;
    if (!embed_tokens(m.d_te, m.d_pe, d_tok, d_pos, n, S, ws.x, s)) return false;
    const float kq_scale = (float) pow((double) 64.0f, -0.25);  // wc:2522
    for (int il = 0; il < L; ++il) {
        const DecLayer & l = m.dec[il];
        if (!layernorm_rows(ws.x, n, S, l.ln1.g, l.ln1.b, hp.eps, ws.xa, nullptr, s)) return false;
        {   // fused Q|K|V: (q + bq) * D^-1/4, k * D^-1/4, v + bv  -> f16 (wc:2558-2583)
            GemmEpilogue e = epi_row(ws.qkv, true, 3 * S);
            e.bias = l.qkv.b; e.scale = kq_scale; e.scale_n0 = 0; e.scale_n1 = 2 * S;
            if (!gemm_f16_tc(opnd(ws.xa, n, S), opnd(l.qkv.w, 3 * S, S), n, 3 * S, S, 1, e, 0, s)) return false;
        }
        if (!self_kv_store(ws.qkv, d_self, n, S, il, s)) return false;
        if (!self_attn_decode(ws.qkv, d_self, n, S, H, il, ws.att, s)) return false;
        {
            GemmEpilogue e = epi_row(ws.x, false, S);
            e.bias = l.o.b; e.res = ws.x; e.ldr = S;
            if (!gemm_f16_tc(opnd(ws.att, n, S), opnd(l.o.w, S, S), n, S, S, 1, e, 0, s)) return false;
        }
        if (!layernorm_rows(ws.x, n, S, l.lnc.g, l.lnc.b, hp.eps, ws.xa, nullptr, s)) return false;
        {
            GemmEpilogue e = epi_row(ws.q, true, S);
            e.bias = l.cq.b;
            if (!gemm_f16_tc(opnd(ws.xa, n, S), opnd(l.cq.w, S, S), n, S, S, 1, e, 0, s)) return false;
        }
        const int * slots = (save_qk && ctx.d_ahead_slot) ? ctx.d_ahead_slot + (size_t) il * H : nullptr;
        if (!cross_attn_decode(ws.q, S, d_cross, n, H, T, il, kq_scale, slots, ws.att, S, s)) return false;
        {
            GemmEpilogue e = epi_row(ws.x, false, S);
            e.bias = l.co.b; e.res = ws.x; e.ldr = S;
            if (!gemm_f16_tc(opnd(ws.att, n, S), opnd(l.co.w, S, S), n, S, S, 1, e, 0, s)) return false;
        }
        if (!layernorm_rows(ws.x, n, S, l.ln2.g, l.ln2.b, hp.eps, ws.xa, nullptr, s)) return false;
        {
            GemmEpilogue e = epi_row(ws.hb, true, 4 * S);
            e.bias = l.fc1.b; e.gelu_table = m.gelu_table;
            if (!gemm_f16_tc(opnd(ws.xa, n, S), opnd(l.fc1.w, 4 * S, S), n, 4 * S, S, 1, e, 0, s)) return false;
            GemmEpilogue e2 = epi_row(ws.x, false, S);
            e2.bias = l.fc2.b; e2.res = ws.x; e2.ldr = S;
            if (!gemm_f16_tc(opnd(ws.hb, n, 4 * S), opnd(l.fc2.w, S, 4 * S), n, S, 4 * S, 1, e2, 0, s)) return false;
        }
    }
    mark("layers");
    // final norm + tied-embedding logits, only for the rows whose logits are requested (wc:2821-2835, 2965-2971)
    if (n_flag > 0) {
        if (!layernorm_rows(ws.x, n, S, m.d_ln.g, m.d_ln.b, hp.eps, ws.xa, nullptr, s)) return false;
        const __half * a_rows = ws.xa;
        if (n_flag != n) {
            // gather the flagged rows (contiguous runs) into the attention buffer
            int dst = 0;
            for (int i = 0; i < n; ++i) {
                if (!rows[i].want_logits) continue;
                WB_CUDA_OK(cudaMemcpyAsync(ws.att + (size_t) dst * S, ws.xa + (size_t) i * S, (size_t) S * 2, cudaMemcpyDeviceToDevice, s));
                ++dst;
            }
            a_rows = ws.att;
        }
        g_prof_gemm_cat = PC_GEMM_LOGITS;
        GemmEpilogue e = epi_row(ws.logits, false, V);
        if (!gemm_f16_tc(opnd(a_rows, n_flag, S), opnd(m.d_te, V, S), n_flag, V, S, 1, e, 0, s)) return false;
    }
    mark("logits");
    // timing buckets of the reference (wc:2982-2991) per state
    int n_states = 0;
    for (int i = 0; i < n; ++i) n_states += (i == 0 || rows[i].st != rows[i - 1].st) ? 1 : 0;
    const int64_t dt = (time_us() - t0) / std::max(1, n_states);  // a batched call is shared by its states
    for (int i = 0; i < n;) {
        State * st = rows[i].st;
        int cnt = 1;
        while (i + cnt < n && rows[i + cnt].st == st) ++cnt;
        if (cnt == 1) {
            st->t_decode_us += dt;
            st->n_decode++;
        } else if (cnt < 16) {
            st->t_batchd_us += dt;
            st->n_batchd += cnt;
        } else {
            st->t_prompt_us += dt;
            st->n_prompt += cnt;
        }
        i += cnt;
    }
    return true;
}

bool stage_fetch_logits(Context & ctx, const BatchRow * rows, int n) {
    const int V = ctx.model.hp.n_vocab;
    cudaStream_t s = ctx.stream;
    int flag = 0;
    for (int i = 0; i < n;) {
        State * st = rows[i].st;
        int cnt = 1;
        while (i + cnt < n && rows[i + cnt].st == st) ++cnt;
        st->logits.resize((size_t) cnt * V);
        for (int k = 0; k < cnt; ++k) {
            if (!rows[i + k].want_logits) continue;
            WB_CUDA_OK(cudaMemcpyAsync(st->logits.data() + (size_t) k * V, ctx.dec_ws.logits + (size_t) flag * V, (size_t) V * 4,
                                       cudaMemcpyDeviceToHost, s));
            ++flag;
        }
        i += cnt;
    }
    WB_CUDA_OK(cudaStreamSynchronize(s));
    return true;
}

bool stage_sample(Context & ctx, const SamplerStatic & sst, const SamplerRow * srows, int n, bool copy_dists, Decoder ** decs,
                  const std::function<void(int, float *)> * prefilter) {
    if (n <= 0) return true;
    DecWorkspace & ws = ctx.dec_ws;
    const int V = ctx.model.hp.n_vocab;
    cudaStream_t s = ctx.stream;
    WB_CHECK(n <= ws.logit_rows, "sampler rows %d exceed the workspace (%d)\n", n, ws.logit_rows);
    // sampler rows live at the tail of the staging block (the head is still in flight for the decode kernels)
    const size_t o = ws.stage_bytes - (((size_t) n * sizeof(SamplerRow) + 15) & ~(size_t) 15);
    memcpy(ws.h_stage + o, srows, (size_t) n * sizeof(SamplerRow));
    // grammar-constrained rows: the host computed which tokens the row's grammar rejects (decs[i]->grammar_reject); the mask goes to the
    // row's slot of a device buffer and the staged sampler row points at it
    {
        SamplerRow * hr = (SamplerRow *) (ws.h_stage + o);
        bool any = false;
        for (int i = 0; i < n; ++i) any = any || (decs && decs[i] && !decs[i]->grammar_reject.empty());
        if (any) {
            if (ws.grammar_rows < ws.logit_rows) {
                cudaFree(ws.d_grammar);
                ws.d_grammar = nullptr;
                WB_CUDA_OK(cudaMalloc((void **) &ws.d_grammar, (size_t) ws.logit_rows * V));
                ws.grammar_rows = ws.logit_rows;
            }
            for (int i = 0; i < n; ++i) {
                hr[i].grammar_reject = nullptr;
                if (!decs[i] || decs[i]->grammar_reject.empty()) continue;
                WB_CHECK((int) decs[i]->grammar_reject.size() == V && srows[i].work_slot < ws.grammar_rows, "grammar mask size\n");
                uint8_t * dst = ws.d_grammar + (size_t) srows[i].work_slot * V;
                WB_CUDA_OK(cudaMemcpyAsync(dst, decs[i]->grammar_reject.data(), (size_t) V, cudaMemcpyHostToDevice, s));
                hr[i].grammar_reject = dst;
            }
        } else {
            for (int i = 0; i < n; ++i) hr[i].grammar_reject = nullptr;
        }
    }
    WB_CUDA_OK(cudaMemcpyAsync(ws.d_stage + o, ws.h_stage + o, (size_t) n * sizeof(SamplerRow), cudaMemcpyHostToDevice, s));
    if (prefilter) {
        // the callback mutates HOST logits (wh:474-480): one D2H / H2D round trip per flagged row, only when a callback is set
        std::vector<float> host((size_t) V);
        for (int i = 0; i < n; ++i) {
            if (!srows[i].prefiltered) continue;
            WB_CUDA_OK(cudaMemcpyAsync(host.data(), ws.logits + (size_t) srows[i].logits_row * V, (size_t) V * 4, cudaMemcpyDeviceToHost, s));
            WB_CUDA_OK(cudaStreamSynchronize(s));
            (*prefilter)(i, host.data());
            WB_CUDA_OK(cudaMemcpyAsync(ws.samp_work + (size_t) srows[i].work_slot * 3 * V, host.data(), (size_t) V * 4, cudaMemcpyHostToDevice, s));
            WB_CUDA_OK(cudaStreamSynchronize(s));
        }
    }
    if (!sample_rows(sst, ws.logits, V, (const SamplerRow *) (ws.d_stage + o), n, ws.samp_work, ws.d_out, s)) return false;
    WB_CUDA_OK(cudaMemcpyAsync(ws.h_out, ws.d_out, (size_t) n * sizeof(SamplerOut), cudaMemcpyDeviceToHost, s));
    if (copy_dists) {
        for (int i = 0; i < n; ++i) {
            if (!decs[i] || !decs[i]->want_dists) continue;
            Decoder & d = *decs[i];
            d.logits.resize(V);
            d.logprobs.resize(V);
            d.probs.resize(V);
            const float * w = ws.samp_work + (size_t) srows[i].work_slot * 3 * V;
            WB_CUDA_OK(cudaMemcpyAsync(d.logits.data(), w, (size_t) V * 4, cudaMemcpyDeviceToHost, s));
            WB_CUDA_OK(cudaMemcpyAsync(d.logprobs.data(), w + V, (size_t) V * 4, cudaMemcpyDeviceToHost, s));
            WB_CUDA_OK(cudaMemcpyAsync(d.probs.data(), w + 2 * (size_t) V, (size_t) V * 4, cudaMemcpyDeviceToHost, s));
        }
    }
    WB_CUDA_OK(cudaStreamSynchronize(s));
    for (int i = 0; i < n; ++i)
        if (decs && decs[i]) decs[i]->picked = ws.h_out[i];
    return true;
}

}  // namespace wb
