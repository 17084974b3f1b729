// whisper_vad_* entry points (wh VAD section; reference definitions whisper.cpp:4445-5480) and the VAD step of whisper_full /
// whisper_full_parallel (whisper.cpp:6615-6793, 7720-7735, 7748-7763).  Device work: vad.cu.  Everything below is host logic in the
// reference's integer / float types: the speech-segment rules, the sample filter and the processed-time -> original-time table.
#include <limits.h>
#include <string.h>

#include <algorithm>

#include "../../include/whisper_b200.h"
#include "engine.h"
#include "log.h"
#include "vad.h"

using namespace wb;

struct whisper_vad_context {
    VadModel model;
    VadDevice dev;
    std::vector<float> probs;
    int n_threads = 4;
    int64_t t_vad_us = 0;
};
struct whisper_vad_segments {
    struct Seg { int64_t start, end; };   // centiseconds
    std::vector<Seg> data;
};

static int cs_to_samples(int64_t cs) { return (int) ((cs / 100.0) * WHISPER_SAMPLE_RATE + 0.5); }                    // wc:4467-4469
static int64_t samples_to_cs(int samples) { return (int64_t) ((samples / (double) WHISPER_SAMPLE_RATE) * 100.0 + 0.5); }  // wc:4471-4473

extern "C" {

struct whisper_vad_context * whisper_vad_init_with_params(struct whisper_model_loader * loader, struct whisper_vad_context_params params) {
    try {
        if (!loader) return nullptr;
        whisper_vad_context * v = new whisper_vad_context;
        v->n_threads = params.n_threads;
        v->dev.device = params.gpu_device;   // params.use_gpu is not consulted: there is no CPU path (the reference forces CPU here, wc:4674)
        if (!vad_load(loader, v->model)) {
            delete v;
            return nullptr;
        }
        log_msg(GGML_LOG_LEVEL_INFO, "%s: model type: %s, version %s, window %d samples\n", __func__, v->model.type.c_str(), v->model.version.c_str(),
                v->model.n_window);
        return v;
    } catch (...) {
        return nullptr;
    }
}

struct whisper_vad_context * whisper_vad_init_from_file_with_params(const char * path_model, struct whisper_vad_context_params params) {
    if (!path_model) return nullptr;
    log_msg(GGML_LOG_LEVEL_INFO, "%s: loading VAD model from '%s'\n", __func__, path_model);
    FILE * f = fopen(path_model, "rb");
    if (!f) {
        log_error("%s: failed to open VAD model '%s'\n", __func__, path_model);
        return nullptr;
    }
    whisper_model_loader loader = {};
    loader.context = f;
    loader.read = [](void * ctx, void * output, size_t read_size) { return fread(output, 1, read_size, (FILE *) ctx); };
    loader.eof = [](void * ctx) { return feof((FILE *) ctx) != 0; };
    loader.close = [](void *) {};
    whisper_vad_context * v = whisper_vad_init_with_params(&loader, params);
    fclose(f);
    return v;
}

bool whisper_vad_detect_speech(struct whisper_vad_context * vctx, const float * samples, int n_samples) {
    try {
        if (!vctx || !samples || n_samples < 0) return false;
        const int64_t t0 = time_us();
        const bool ok = vad_detect(vctx->model, vctx->dev, samples, n_samples, vctx->probs);
        vctx->t_vad_us += time_us() - t0;
        if (!ok) log_error("%s: failed to compute VAD graph\n", __func__);
        return ok;
    } catch (...) {
        return false;
    }
}

int whisper_vad_n_probs(struct whisper_vad_context * vctx) { return vctx ? (int) vctx->probs.size() : 0; }
float * whisper_vad_probs(struct whisper_vad_context * vctx) { return vctx ? vctx->probs.data() : nullptr; }

// test hook (host logic without a device): replace the context's probabilities
int whisper_b200_test_vad_set_probs(struct whisper_vad_context * vctx, const float * probs, int n) {
    if (!vctx || !probs || n < 0) return -1;
    vctx->probs.assign(probs, probs + n);
    return n;
}

// whisper_vad_segments_from_probs (wc:5202-5436): hysteresis (threshold / threshold - 0.15), minimum silence and speech lengths, forced
// splits at max_speech_duration, merging of gaps under 200 ms, padding, samples -> centiseconds
struct whisper_vad_segments * whisper_vad_segments_from_probs(struct whisper_vad_context * vctx, struct whisper_vad_params params) {
    try {
        if (!vctx) return nullptr;
        const int n_probs = (int) vctx->probs.size();
        const float * probs = vctx->probs.data();
        const int rate = WHISPER_SAMPLE_RATE, n_window = vctx->model.n_window;
        const float thr_on = params.threshold;
        float thr_off = thr_on - 0.15f;
        if (thr_off < 0.01f) thr_off = 0.01f;
        const int min_silence = rate * params.min_silence_duration_ms / 1000;
        const int min_speech = rate * params.min_speech_duration_ms / 1000;
        const int pad = rate * params.speech_pad_ms / 1000;
        const int audio_len = n_probs * n_window;
        int max_speech;
        if (params.max_speech_duration_s > 100000.0f) {
            max_speech = INT_MAX / 2;
        } else {
            const int64_t t = (int64_t) rate * (int64_t) (params.max_speech_duration_s) - n_window - 2 * pad;
            max_speech = t > INT_MAX ? INT_MAX / 2 : (int) t;
            if (max_speech < 0) max_speech = INT_MAX / 2;
        }
        const int split_silence = rate * 98 / 1000;   // a silence this long is remembered as a place to split an over-long segment

        struct Span { int start, end; };
        std::vector<Span> spans;
        bool in_speech = false, open = false;
        int tentative_end = 0, split_end = 0, resume_at = 0, start = 0;
        for (int i = 0; i < n_probs; ++i) {
            const float p = probs[i];
            const int at = n_window * i;
            if (p >= thr_on && tentative_end) {          // speech again: the tentative end is withdrawn
                tentative_end = 0;
                if (resume_at < split_end) resume_at = at;
            }
            if (p >= thr_on && !in_speech) {             // a new segment opens
                in_speech = true;
                open = true;
                start = at;
                continue;
            }
            if (in_speech && at - start > max_speech) {  // too long: split at the remembered silence, or right here
                if (split_end) {
                    spans.push_back({start, split_end});
                    open = true;
                    if (resume_at < split_end) {
                        in_speech = false;
                        open = false;
                    } else {
                        start = resume_at;
                    }
                    split_end = resume_at = tentative_end = 0;
                } else {
                    spans.push_back({start, at});
                    split_end = resume_at = tentative_end = 0;
                    in_speech = false;
                    open = false;
                    continue;
                }
            }
            if (p < thr_off && in_speech) {              // silence inside a segment
                if (!tentative_end) tentative_end = at;
                if (at - tentative_end > split_silence) split_end = tentative_end;
                if (at - tentative_end < min_silence) continue;
                if (tentative_end - start > min_speech) spans.push_back({start, tentative_end});
                split_end = resume_at = tentative_end = 0;
                in_speech = false;
                open = false;
                continue;
            }
        }
        if (open && audio_len - start > min_speech) spans.push_back({start, audio_len});

        if (spans.size() > 1) {                          // merge neighbours closer than 200 ms
            const int merge_gap = rate * 200 / 1000;
            for (int i = 0; i < (int) spans.size() - 1; i++) {
                if (spans[i + 1].start - spans[i].end < merge_gap) {
                    spans[i].end = spans[i + 1].end;
                    spans.erase(spans.begin() + i + 1);
                    i--;
                }
            }
        }
        for (int i = 0; i < (int) spans.size(); i++) {
            if (spans[i].end - spans[i].start < min_speech) {
                spans.erase(spans.begin() + i);
                i--;
            }
        }
        whisper_vad_segments * out = new whisper_vad_segments;
        out->data.resize(spans.size());
        for (int i = 0; i < (int) spans.size(); i++) {   // padding: full pad where there is room, else the gap is shared
            if (i == 0) spans[i].start = spans[i].start > pad ? spans[i].start - pad : 0;
            if (i < (int) spans.size() - 1) {
                const int gap = spans[i + 1].start - spans[i].end;
                if (gap < 2 * pad) {
                    spans[i].end += gap / 2;
                    spans[i + 1].start = spans[i + 1].start > gap / 2 ? spans[i + 1].start - gap / 2 : 0;
                } else {
                    spans[i].end = spans[i].end + pad < audio_len ? spans[i].end + pad : audio_len;
                    spans[i + 1].start = spans[i + 1].start > pad ? spans[i + 1].start - pad : 0;
                }
            } else {
                spans[i].end = spans[i].end + pad < audio_len ? spans[i].end + pad : audio_len;
            }
            out->data[(size_t) i].start = samples_to_cs(spans[i].start);
            out->data[(size_t) i].end = samples_to_cs(spans[i].end);
        }
        return out;
    } catch (...) {
        return nullptr;
    }
}

struct whisper_vad_segments * whisper_vad_segments_from_samples(struct whisper_vad_context * vctx, struct whisper_vad_params params,
                                                                const float * samples, int n_samples) {
    if (!whisper_vad_detect_speech(vctx, samples, n_samples)) {
        log_error("%s: failed to detect speech\n", __func__);
        return nullptr;
    }
    return whisper_vad_segments_from_probs(vctx, params);
}

int whisper_vad_segments_n_segments(struct whisper_vad_segments * segments) { return segments ? (int) segments->data.size() : 0; }
float whisper_vad_segments_get_segment_t0(struct whisper_vad_segments * segments, int i) { return (float) segments->data[(size_t) i].start; }
float whisper_vad_segments_get_segment_t1(struct whisper_vad_segments * segments, int i) { return (float) segments->data[(size_t) i].end; }
void whisper_vad_free_segments(struct whisper_vad_segments * segments) { delete segments; }
void whisper_vad_free(struct whisper_vad_context * ctx) { delete ctx; }

}  // extern "C"

namespace wb {

bool vad_filter_segments(const whisper_vad_segments & seg_list, float samples_overlap, const float * samples, int n_samples,
                         std::vector<float> & filtered, std::vector<State::VadMap> & mapping);
int64_t vad_map_time_table(const std::vector<State::VadMap> & tab, int64_t t);

void vad_state_release(State & st) {
    whisper_vad_free((whisper_vad_context *) st.vad_context);
    st.vad_context = nullptr;
}

// whisper_vad (wc:6615-6793): speech segments of the clip -> the filtered sample buffer whisper_full transcribes (segments joined by
// 0.1 s of silence, the non-final ones extended by samples_overlap) and the state's processed-time -> original-time table
bool vad_filter(Context * ctx, State * st, const whisper_full_params & params, const float * samples, int n_samples, std::vector<float> & filtered) {
    st->vad_mapping.clear();
    st->has_vad_segments = false;
    if (!st->vad_context) {
        whisper_vad_context_params vp = whisper_vad_default_context_params();
        vp.gpu_device = ctx->model.device;
        whisper_vad_context * v = params.vad_model_path ? whisper_vad_init_from_file_with_params(params.vad_model_path, vp) : nullptr;
        if (!v) {
            log_error("%s: failed to initialize VAD context\n", __func__);
            return false;
        }
        st->vad_context = v;
    }
    whisper_vad_context * vctx = (whisper_vad_context *) st->vad_context;
    const whisper_vad_params & vp = params.vad_params;
    whisper_vad_segments * segs = whisper_vad_segments_from_samples(vctx, vp, samples, n_samples);
    if (!segs) return false;
    st->has_vad_segments = vad_filter_segments(*segs, vp.samples_overlap, samples, n_samples, filtered, st->vad_mapping);
    whisper_vad_free_segments(segs);
    return true;
}

// The host half of whisper_vad: segments (centiseconds) -> filtered samples + mapping table.  Returns whether there were segments
// (no segments: `filtered` and `mapping` are left untouched, wc:6643-6645).
bool vad_filter_segments(const whisper_vad_segments & seg_list, float samples_overlap, const float * samples, int n_samples,
                         std::vector<float> & filtered, std::vector<State::VadMap> & mapping) {
    const whisper_vad_segments * segs = &seg_list;
    const int n_seg = (int) segs->data.size();
    if (n_seg > 0) {
        const int overlap = (int) (samples_overlap * WHISPER_SAMPLE_RATE);
        const int silence = (int) (0.1 * WHISPER_SAMPLE_RATE);
        int total = 0;
        for (int i = 0; i < n_seg; i++) {
            const int a = cs_to_samples(segs->data[(size_t) i].start);
            int b = cs_to_samples(segs->data[(size_t) i].end);
            if (i < n_seg - 1) b += overlap;
            b = std::min(b, n_samples - 1);
            total += b - a;
        }
        filtered.resize((size_t) std::max(0, total + (n_seg > 1 ? (n_seg - 1) * silence : 0)));
        auto add = [&](int64_t processed, int64_t original) { mapping.push_back({processed, original}); };
        int off = 0;
        for (int i = 0; i < n_seg; i++) {
            int a = cs_to_samples(segs->data[(size_t) i].start);
            int b = cs_to_samples(segs->data[(size_t) i].end);
            if (i < n_seg - 1) b += overlap;
            a = std::min(a, n_samples - 1);
            b = std::min(b, n_samples);
            const int len = b - a;
            if (len <= 0) continue;
            const int64_t o0 = segs->data[(size_t) i].start, o1 = segs->data[(size_t) i].end;
            const int64_t v0 = samples_to_cs(off), v1 = samples_to_cs(off + len);
            add(v0, o0);
            add(v1, o1);
            if (v1 - v0 > 100) {   // interior points every 200 ms of segments longer than 1 s
                const int64_t span = v1 - v0;
                const int n_pts = (int) (span / 20) - 1;
                for (int j = 1; j <= n_pts; j++) {
                    const int64_t vt = v0 + j * 20;
                    if (vt >= v1) continue;
                    add(vt, o0 + ((vt - v0) * (o1 - o0)) / span);
                }
            }
            // The reference sizes the buffer with min(end, n - 1) above and copies with min(end, n) here, i.e. one float past the end for
            // a segment that reaches the end of the clip; the buffer KEEPS its first size (whisper_full transcribes vad_samples.size()
            // samples, wc:7732).  Same observable behaviour without the overflow: copy what fits.
            const size_t room = filtered.size() > (size_t) off ? filtered.size() - (size_t) off : 0;
            memcpy(filtered.data() + off, samples + a, std::min((size_t) len, room) * sizeof(float));
            off += len;
            if (i < n_seg - 1) {
                add(samples_to_cs(off), o1);
                add(samples_to_cs(off + silence), segs->data[(size_t) i + 1].start);
                const size_t room2 = filtered.size() > (size_t) off ? filtered.size() - (size_t) off : 0;
                memset(filtered.data() + off, 0, std::min((size_t) silence, room2) * sizeof(float));
                off += silence;
            }
        }
        std::sort(mapping.begin(), mapping.end(), [](const State::VadMap & x, const State::VadMap & y) { return x.processed < y.processed; });
        mapping.erase(std::unique(mapping.begin(), mapping.end(),
                                  [](const State::VadMap & x, const State::VadMap & y) { return x.processed == y.processed; }),
                      mapping.end());
    }
    return n_seg > 0;
}

// map_processed_to_original_time (wc:7882-7925): clamp outside the table, exact hits, integer linear interpolation between points
int64_t vad_map_time(const State & st, int64_t t) { return vad_map_time_table(st.vad_mapping, t); }
int64_t vad_map_time_table(const std::vector<State::VadMap> & tab, int64_t t) {
    if (tab.empty()) return t;
    if (t <= tab.front().processed) return tab.front().original;
    if (t >= tab.back().processed) return tab.back().original;
    auto hi = std::lower_bound(tab.begin(), tab.end(), t, [](const State::VadMap & e, int64_t v) { return e.processed < v; });
    if (hi->processed == t) return hi->original;
    auto lo = hi - 1;
    const int64_t dp = hi->processed - lo->processed;
    if (dp == 0) return lo->original;
    return lo->original + ((t - lo->processed) * (hi->original - lo->original)) / dp;
}

}  // namespace wb

// Host logic of whisper_vad without a device (tests): speech segments from the context's current probabilities (see
// whisper_b200_test_vad_set_probs), then the sample filter and the time-mapping table.  out[cap] receives the filtered samples,
// table[2 * table_cap] the (processed, original) pairs; map_in[n_map] -> map_out[n_map] through the table.  Returns the number of
// filtered samples (0 with *n_table = 0 when there is no speech segment), or -1.
extern "C" int whisper_b200_test_vad_filter(struct whisper_vad_context * vctx, struct whisper_vad_params params, const float * samples, int n_samples,
                                            float * out, int cap, int64_t * table, int table_cap, int * n_table, const int64_t * map_in,
                                            int64_t * map_out, int n_map) {
    try {
        if (!vctx || !samples || n_samples < 0 || !n_table) return -1;
        whisper_vad_segments * segs = whisper_vad_segments_from_probs(vctx, params);
        if (!segs) return -1;
        std::vector<float> filtered;
        std::vector<wb::State::VadMap> mapping;
        const bool any = wb::vad_filter_segments(*segs, params.samples_overlap, samples, n_samples, filtered, mapping);
        whisper_vad_free_segments(segs);
        *n_table = (int) mapping.size();
        for (int i = 0; i < (int) mapping.size() && i < table_cap && table; ++i) {
            table[2 * i] = mapping[(size_t) i].processed;
            table[2 * i + 1] = mapping[(size_t) i].original;
        }
        if (out) memcpy(out, filtered.data(), std::min(filtered.size(), (size_t) std::max(cap, 0)) * sizeof(float));
        for (int i = 0; i < n_map && map_in && map_out; ++i) map_out[i] = any ? wb::vad_map_time_table(mapping, map_in[i]) : map_in[i];
        return (int) filtered.size();
    } catch (...) {
        return -1;
    }
}
