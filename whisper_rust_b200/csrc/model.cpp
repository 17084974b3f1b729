// Model file parser + device upload.  Restates the reference loader's checks (wc:1513-1965) over its
// whisper_model_loader callback interface; the file format is documented in SURVEY.md Appendix B.
#include "model.h"

#include <math.h>
#include <string.h>

#include <functional>

#include "common.cuh"
#include "log.h"

namespace wb {

// language table: id order of the reference (wc:313-415)
const LangEntry g_langs[100] = {
    {"en", "english"}, {"zh", "chinese"}, {"de", "german"}, {"es", "spanish"}, {"ru", "russian"}, {"ko", "korean"},
    {"fr", "french"}, {"ja", "japanese"}, {"pt", "portuguese"}, {"tr", "turkish"}, {"pl", "polish"}, {"ca", "catalan"},
    {"nl", "dutch"}, {"ar", "arabic"}, {"sv", "swedish"}, {"it", "italian"}, {"id", "indonesian"}, {"hi", "hindi"},
    {"fi", "finnish"}, {"vi", "vietnamese"}, {"he", "hebrew"}, {"uk", "ukrainian"}, {"el", "greek"}, {"ms", "malay"},
    {"cs", "czech"}, {"ro", "romanian"}, {"da", "danish"}, {"hu", "hungarian"}, {"ta", "tamil"}, {"no", "norwegian"},
    {"th", "thai"}, {"ur", "urdu"}, {"hr", "croatian"}, {"bg", "bulgarian"}, {"lt", "lithuanian"}, {"la", "latin"},
    {"mi", "maori"}, {"ml", "malayalam"}, {"cy", "welsh"}, {"sk", "slovak"}, {"te", "telugu"}, {"fa", "persian"},
    {"lv", "latvian"}, {"bn", "bengali"}, {"sr", "serbian"}, {"az", "azerbaijani"}, {"sl", "slovenian"}, {"kn", "kannada"},
    {"et", "estonian"}, {"mk", "macedonian"}, {"br", "breton"}, {"eu", "basque"}, {"is", "icelandic"}, {"hy", "armenian"},
    {"ne", "nepali"}, {"mn", "mongolian"}, {"bs", "bosnian"}, {"kk", "kazakh"}, {"sq", "albanian"}, {"sw", "swahili"},
    {"gl", "galician"}, {"mr", "marathi"}, {"pa", "punjabi"}, {"si", "sinhala"}, {"km", "khmer"}, {"sn", "shona"},
    {"yo", "yoruba"}, {"so", "somali"}, {"af", "afrikaans"}, {"oc", "occitan"}, {"ka", "georgian"}, {"be", "belarusian"},
    {"tg", "tajik"}, {"sd", "sindhi"}, {"gu", "gujarati"}, {"am", "amharic"}, {"yi", "yiddish"}, {"lo", "lao"},
    {"uz", "uzbek"}, {"fo", "faroese"}, {"ht", "haitian creole"}, {"ps", "pashto"}, {"tk", "turkmen"}, {"nn", "nynorsk"},
    {"mt", "maltese"}, {"sa", "sanskrit"}, {"lb", "luxembourgish"}, {"my", "myanmar"}, {"bo", "tibetan"}, {"tl", "tagalog"},
    {"mg", "malagasy"}, {"as", "assamese"}, {"tt", "tatar"}, {"haw", "hawaiian"}, {"ln", "lingala"}, {"ha", "hausa"},
    {"ba", "bashkir"}, {"jw", "javanese"}, {"su", "sundanese"}, {"yue", "cantonese"},
};

int lang_id(const char * lang) {
    if (!lang) return -1;
    for (int i = 0; i < 100; ++i)
        if (strcmp(g_langs[i].code, lang) == 0) return i;
    for (int i = 0; i < 100; ++i)
        if (strcmp(g_langs[i].name, lang) == 0) return i;
    return -1;
}

Model::~Model() {
    if (arena) {
        cudaSetDevice(device);
        cudaFree(arena);
    }
}

namespace {

struct Reader {
    whisper_model_loader * l;
    bool ok = true;
    bool read(void * dst, size_t n) {
        if (n == 0) return true;
        const size_t got = l->read(l->context, dst, n);
        if (got != n) ok = false;
        return got == n;
    }
    template <typename T> bool pod(T & v) { return read(&v, sizeof(T)); }
};

// host f16 conversions (IEEE round-to-nearest-even, the F16C behaviour the reference relies on)
static inline float h2f(uint16_t h) {
    __half_raw r;
    r.x = h;
    return __half2float(__half(r));
}
static inline uint16_t f2h(float f) {
    __half_raw r = __float2half_rn(f);
    return r.x;
}

enum class Kind { F32_VEC, F16_MAT, CONV1, CONV2 };

// ---- quantised ggml tensor types (ftype -> wtype: wc:1567-1577; block layouts: ggml/src/ggml-common.h, dequantize_row_*:
// ggml/src/ggml-quants.c).  The B200 path computes in f16 x f16 -> f32: a quantised matrix is expanded ONCE at load,
// x = q * d (+ m) evaluated in f32 like dequantize_row_*, then rounded to f16.  (The reference instead keeps the blocks and
// quantises the ACTIVATIONS to q8_0/q8_1 inside vec_dot, so agreement with it on such checkpoints is tolerance-level.)
enum : int { GT_F32 = 0, GT_F16 = 1, GT_Q4_0 = 2, GT_Q4_1 = 3, GT_Q5_0 = 6, GT_Q5_1 = 7, GT_Q8_0 = 8,
             GT_Q2_K = 10, GT_Q3_K = 11, GT_Q4_K = 12, GT_Q5_K = 13, GT_Q6_K = 14 };
static constexpr int QK = 32;     // elements per block, the five legacy types
static constexpr int QK_K = 256;  // elements per super-block, the five K types

static int ftype_to_wtype(int ftype) {  // ggml_ftype_to_ggml_type (ggml.c); -1: not supported here
    switch (ftype) {
        case 0: return GT_F32;
        case 1: return GT_F16;
        case 2: return GT_Q4_0;
        case 3: return GT_Q4_1;
        case 7: return GT_Q8_0;
        case 8: return GT_Q5_0;
        case 9: return GT_Q5_1;
        case 10: return GT_Q2_K;
        case 11: return GT_Q3_K;
        case 12: return GT_Q4_K;
        case 13: return GT_Q5_K;
        case 14: return GT_Q6_K;
        default: return -1;   // 4 (Q4_1_SOME_F16) is rejected by ggml itself; 15+ are the codebook (IQ) types and bf16
    }
}
static size_t quant_block_bytes(int t) {
    switch (t) {
        case GT_Q4_0: return 2 + 16;
        case GT_Q4_1: return 4 + 16;
        case GT_Q5_0: return 2 + 4 + 16;
        case GT_Q5_1: return 4 + 4 + 16;
        case GT_Q8_0: return 2 + 32;
        case GT_Q2_K: return 16 + 64 + 4;            // scales | qs | d dmin
        case GT_Q3_K: return 32 + 64 + 12 + 2;       // hmask | qs | scales | d
        case GT_Q4_K: return 4 + 12 + 128;           // d dmin | scales | qs
        case GT_Q5_K: return 4 + 12 + 32 + 128;      // d dmin | scales | qh | qs
        case GT_Q6_K: return 128 + 64 + 16 + 2;      // ql | qh | scales | d
        default: return 0;
    }
}
static int quant_block_elems(int t) { return t >= GT_Q2_K && t <= GT_Q6_K ? QK_K : QK; }

// One 256-element super-block of a K type (block_q{2,3,4,5,6}_K, ggml-common.h:252-325; dequantize_row_q*_K, ggml-quants.c:706-1720).
// Every K type is "sixteen groups of 16 elements, each with a scale (and, for q2/q4/q5, a minimum)": the block is first unpacked into
// integer levels and per-group f32 scale / minimum, then y = scale * level (- minimum) in the reference's order of operations.
static inline uint16_t rd16(const uint8_t * p) {
    uint16_t v;
    memcpy(&v, p, 2);
    return v;
}
// the 6-bit (scale, minimum) pair k of the 12-byte table q4_K / q5_K share (get_scale_min_k4, ggml-quants.c:625-632)
static inline void scale_min_6bit(const uint8_t * tab, int k, int & sc, int & mn) {
    if (k < 4) {
        sc = tab[k] & 63;
        mn = tab[k + 4] & 63;
    } else {
        sc = (tab[k + 4] & 0x0F) | ((tab[k - 4] >> 6) << 4);
        mn = (tab[k + 4] >> 4) | ((tab[k] >> 6) << 4);
    }
}
static void expand_k_block(int t, const uint8_t * p, float * y) {
    int level[QK_K];
    float scale[16], minimum[16];
    bool has_min = false;
    if (t == GT_Q2_K || t == GT_Q3_K) {
        // 2-bit planes: element e of half hf (128 elements) and plane pl (32 elements) sits in byte qs[hf * 32 + e % 32], bits 2 pl
        const uint8_t * qs = p + (t == GT_Q2_K ? 16 : 32);
        for (int e = 0; e < QK_K; ++e) level[e] = (qs[(e >> 7) * 32 + (e & 31)] >> (2 * ((e >> 5) & 3))) & 3;
        if (t == GT_Q2_K) {
            const float d = h2f(rd16(p + 80)), dmin = h2f(rd16(p + 82));
            for (int g = 0; g < 16; ++g) {
                scale[g] = d * (float) (p[g] & 0x0F);
                minimum[g] = dmin * (float) (p[g] >> 4);
            }
            has_min = true;
        } else {
            // third bit: hmask[e % 32] bit e / 32, CLEAR means "subtract 4"; 6-bit scales: low nibbles in bytes 0..7, top two bits in 8..11
            for (int e = 0; e < QK_K; ++e)
                if (!((p[e & 31] >> (e >> 5)) & 1)) level[e] -= 4;
            const uint8_t * tab = p + 96;
            const float d = h2f(rd16(p + 108));
            for (int g = 0; g < 16; ++g) {
                const int lo = g < 8 ? (tab[g] & 0x0F) : (tab[g - 8] >> 4);
                const int hi = (tab[8 + (g & 3)] >> (2 * (g >> 2))) & 3;
                scale[g] = d * (float) ((lo | (hi << 4)) - 32);
            }
        }
    } else if (t == GT_Q4_K || t == GT_Q5_K) {
        // nibbles: chunk c (64 elements) uses bytes qs[c * 32 ..): low nibbles first, then high nibbles; q5_K adds bit e / 32 of qh[e % 32]
        const uint8_t * qs = p + (t == GT_Q4_K ? 16 : 48);
        for (int e = 0; e < QK_K; ++e) {
            const uint8_t b = qs[(e >> 6) * 32 + (e & 31)];
            level[e] = (e & 32) ? (b >> 4) : (b & 0x0F);
        }
        if (t == GT_Q5_K)
            for (int e = 0; e < QK_K; ++e)
                if ((p[16 + (e & 31)] >> (e >> 5)) & 1) level[e] += 16;
        const float d = h2f(rd16(p)), dmin = h2f(rd16(p + 2));
        for (int k = 0; k < 8; ++k) {   // eight groups of 32 = pairs of 16-element groups
            int sc, mn;
            scale_min_6bit(p + 4, k, sc, mn);
            scale[2 * k] = scale[2 * k + 1] = d * (float) sc;
            minimum[2 * k] = minimum[2 * k + 1] = dmin * (float) mn;
        }
        has_min = true;
    } else {  // GT_Q6_K
        // per half (128 elements, 64 bytes of ql, 32 of qh): quarter qt of the half takes nibble qt / 2 of ql[e % 32 + 32 (qt & 1)] and
        // bits 2 qt of qh[e % 32]; levels are offset by 32; int8 scales
        const uint8_t * ql = p, * qh = p + 128;
        const int8_t * sc = (const int8_t *) (p + 192);
        const float d = h2f(rd16(p + 208));
        for (int e = 0; e < QK_K; ++e) {
            const int hf = e >> 7, qt = (e >> 5) & 3, l = e & 31;
            const uint8_t b = ql[hf * 64 + l + 32 * (qt & 1)];
            const int lo = (qt & 2) ? (b >> 4) : (b & 0x0F);
            const int hi = (qh[hf * 32 + l] >> (2 * qt)) & 3;
            level[e] = (lo | (hi << 4)) - 32;
        }
        for (int g = 0; g < 16; ++g) scale[g] = d * (float) sc[g];
    }
    if (has_min) {
        for (int e = 0; e < QK_K; ++e) y[e] = scale[e >> 4] * (float) level[e] - minimum[e >> 4];
    } else {
        for (int e = 0; e < QK_K; ++e) y[e] = scale[e >> 4] * (float) level[e];
    }
}

static void dequantize_blocks(int t, const uint8_t * src, int64_t n, float * dst) {
    if (quant_block_elems(t) == QK_K) {
        const size_t kbs = quant_block_bytes(t);
        for (int64_t b = 0; b < n / QK_K; ++b) expand_k_block(t, src + (size_t) b * kbs, dst + b * QK_K);
        return;
    }
    const int64_t nb = n / QK;
    const size_t bs = quant_block_bytes(t);
    for (int64_t b = 0; b < nb; ++b) {
        const uint8_t * p = src + (size_t) b * bs;
        float * y = dst + b * QK;
        uint16_t dh, mh = 0;
        memcpy(&dh, p, 2);
        const float d = h2f(dh);
        if (t == GT_Q8_0) {
            const int8_t * q = (const int8_t *) (p + 2);
            for (int j = 0; j < QK; ++j) y[j] = q[j] * d;
        } else if (t == GT_Q4_0) {
            const uint8_t * q = p + 2;
            for (int j = 0; j < QK / 2; ++j) {
                y[j] = ((int) (q[j] & 0x0F) - 8) * d;
                y[j + QK / 2] = ((int) (q[j] >> 4) - 8) * d;
            }
        } else if (t == GT_Q4_1) {
            memcpy(&mh, p + 2, 2);
            const float mn = h2f(mh);
            const uint8_t * q = p + 4;
            for (int j = 0; j < QK / 2; ++j) {
                y[j] = (q[j] & 0x0F) * d + mn;
                y[j + QK / 2] = (q[j] >> 4) * d + mn;
            }
        } else if (t == GT_Q5_0) {
            uint32_t qh;
            memcpy(&qh, p + 2, 4);
            const uint8_t * q = p + 6;
            for (int j = 0; j < QK / 2; ++j) {
                const uint8_t xh0 = ((qh >> (j + 0)) << 4) & 0x10, xh1 = (qh >> (j + 12)) & 0x10;
                y[j] = ((int) ((q[j] & 0x0F) | xh0) - 16) * d;
                y[j + QK / 2] = ((int) ((q[j] >> 4) | xh1) - 16) * d;
            }
        } else {  // GT_Q5_1
            memcpy(&mh, p + 2, 2);
            const float mn = h2f(mh);
            uint32_t qh;
            memcpy(&qh, p + 4, 4);
            const uint8_t * q = p + 8;
            for (int j = 0; j < QK / 2; ++j) {
                const uint8_t xh0 = ((qh >> (j + 0)) << 4) & 0x10, xh1 = (qh >> (j + 12)) & 0x10;
                y[j] = ((q[j] & 0x0F) | xh0) * d + mn;
                y[j + QK / 2] = ((q[j] >> 4) | xh1) * d + mn;
            }
        }
    }
}

struct Dest {
    Kind kind;
    void * ptr;          // device destination
    int64_t n;           // element count expected
    int n_dims;
    int ne[3];           // expected ggml shape (reversed numpy shape)
};

}  // namespace

// The constant tables the kernels read, computed on the host exactly as the reference computes its own: the 65,536-entry f16 GELU table
// (ggml_table_gelu_f16, ggml-cpu.c:3515 over ggml_gelu_f32, ggml-cpu/vec.h:552-554) and the FFT sin / cos / Hann tables of the mel front
// end (whisper_global_cache, wc:3018-3047).  gelu[65536], sn / cs / hn[400]; any pointer may be null.
void host_tables(uint16_t * gelu, float * sn, float * cs, float * hn) {
    if (gelu) {
        const float GELU_COEF_A = 0.044715f, SQRT_2_OVER_PI = 0.79788456080286535587989211986876f;
        for (int i = 0; i < 65536; ++i) {
            const float x = h2f((uint16_t) i);
            // this translation unit is compiled with -ffp-contract=off like the ISO-C build of the reference's table initialiser
            const float g = 0.5f * x * (1.0f + tanhf(SQRT_2_OVER_PI * x * (1.0f + GELU_COEF_A * x * x)));
            gelu[i] = f2h(g);
        }
    }
    for (int i = 0; i < 400; ++i) {
        const double theta = (2 * M_PI * i) / 400;
        if (sn) sn[i] = sinf((float) theta);
        if (cs) cs[i] = cosf((float) theta);
        if (hn) hn[i] = (float) (0.5 * (1.0 - cosf((float) ((2.0 * M_PI * i) / 400))));
    }
}

// the loader's block expansion alone (tests): n elements of ggml type `ggml_type` -> f32; false for a type the loader does not expand
bool dequantize_tensor(int ggml_type, const void * blocks, int64_t n, float * dst) {
    if (quant_block_bytes(ggml_type) == 0 || n < 0 || (n % quant_block_elems(ggml_type)) != 0 || !blocks || !dst) return false;
    dequantize_blocks(ggml_type, (const uint8_t *) blocks, n, dst);
    return true;
}

bool load_model(whisper_model_loader * loader, int device, Model & m) {
    Reader rd{loader};
    m.device = device;
    WB_CUDA_OK(cudaSetDevice(device));

    uint32_t magic = 0;
    rd.pod(magic);
    if (!rd.ok || magic != 0x67676d6c) {
        log_error("%s: invalid model data (bad magic)\n", __func__);
        return false;
    }
    HParams & hp = m.hp;
    int32_t h[11];
    if (!rd.read(h, sizeof(h))) {
        log_error("%s: truncated header\n", __func__);
        return false;
    }
    hp.n_vocab = h[0]; hp.n_audio_ctx = h[1]; hp.n_audio_state = h[2]; hp.n_audio_head = h[3]; hp.n_audio_layer = h[4];
    hp.n_text_ctx = h[5]; hp.n_text_state = h[6]; hp.n_text_head = h[7]; hp.n_text_layer = h[8]; hp.n_mels = h[9]; hp.ftype = h[10];

    switch (hp.n_audio_layer) {  // wc:1543-1565
        case 4: m.model_type = 1; break;
        case 6: m.model_type = 2; break;
        case 12: m.model_type = 3; break;
        case 24: m.model_type = 4; break;
        case 32: m.model_type = 5; break;
        default: m.model_type = 0; break;
    }
    const int qntvr = hp.ftype / 1000;
    hp.ftype %= 1000;
    (void) qntvr;
    const int wtype = ftype_to_wtype(hp.ftype);  // wc:1567-1577
    if (wtype < 0) {
        log_error("%s: invalid model (bad ftype value %d): supported are f32, f16, q4_0, q4_1, q5_0, q5_1, q8_0, q2_K, q3_K, q4_K, q5_K, "
                  "q6_K (the codebook IQ types and bf16 are not expanded by this loader)\n", __func__, hp.ftype);
        return false;
    }
    const int S = hp.n_audio_state, H = hp.n_audio_head, T = hp.n_audio_ctx, M = hp.n_mels, V = hp.n_vocab;
    const int La = hp.n_audio_layer, Lt = hp.n_text_layer, St = hp.n_text_state, Tt = hp.n_text_ctx;
    if (S <= 0 || H <= 0 || S % H != 0 || S / H != 64 || St != S || hp.n_text_head != H || M <= 0 || M > 128 || V <= 0 || T <= 0 ||
        T > 1500 || Tt <= 0 || La <= 0 || Lt <= 0 || (S % 64) != 0) {
        log_error("%s: unsupported architecture (state %d/%d heads %d/%d mels %d ctx %d/%d): head width must be 64\n", __func__, S, St,
                  H, hp.n_text_head, M, T, Tt);
        return false;
    }

    // mel filters (wc:1595-1604)
    int32_t n_mel = 0, n_fft = 0;
    rd.pod(n_mel);
    rd.pod(n_fft);
    if (!rd.ok || n_mel <= 0 || n_mel > 128 || n_fft != 201) {
        log_error("%s: bad mel filter header (%d x %d)\n", __func__, n_mel, n_fft);
        return false;
    }
    m.filters_n_mel = n_mel;
    m.filters.resize((size_t) n_mel * n_fft);
    if (!rd.read(m.filters.data(), m.filters.size() * sizeof(float))) {
        log_error("%s: truncated mel filters\n", __func__);
        return false;
    }

    // vocab (wc:1607-1693)
    Vocab & vc = m.vocab;
    int32_t n_file_vocab = 0;
    rd.pod(n_file_vocab);
    if (!rd.ok || n_file_vocab < 0 || n_file_vocab > (1 << 20)) {
        log_error("%s: bad vocab size\n", __func__);
        return false;
    }
    vc.id_to_token.assign((size_t) std::max(n_file_vocab, V), std::string());
    std::vector<char> tmp;
    for (int i = 0; i < n_file_vocab; ++i) {
        uint32_t len = 0;
        rd.pod(len);
        if (!rd.ok || len > (1u << 16)) {
            log_error("%s: bad token length\n", __func__);
            return false;
        }
        std::string word;
        if (len > 0) {
            tmp.resize(len);
            if (!rd.read(tmp.data(), len)) {
                log_error("%s: truncated vocab\n", __func__);
                return false;
            }
            word.assign(tmp.data(), len);
        }
        vc.token_to_id[word] = i;
        vc.id_to_token[i] = word;
    }
    vc.n_vocab = V;
    if (vc.is_multilingual()) {
        vc.token_eot++;
        vc.token_sot++;
        const int dt = vc.num_languages() - 98;
        vc.token_translate += dt; vc.token_transcribe += dt; vc.token_solm += dt; vc.token_prev += dt;
        vc.token_nosp += dt; vc.token_not += dt; vc.token_beg += dt;
    }
    if (n_file_vocab < V) {
        for (int i = n_file_vocab; i < V; ++i) {
            std::string word;
            if (i > vc.token_beg) word = "[_TT_" + std::to_string(i - vc.token_beg) + "]";
            else if (i == vc.token_eot) word = "[_EOT_]";
            else if (i == vc.token_sot) word = "[_SOT_]";
            else if (i == vc.token_translate) word = "[_TRANSLATE_]";
            else if (i == vc.token_transcribe) word = "[_TRANSCRIBE_]";
            else if (i == vc.token_solm) word = "[_SOLM_]";
            else if (i == vc.token_prev) word = "[_PREV_]";
            else if (i == vc.token_nosp) word = "[_NOSP_]";
            else if (i == vc.token_not) word = "[_NOT_]";
            else if (i == vc.token_beg) word = "[_BEG_]";
            else if (i > vc.token_sot && i <= vc.token_sot + vc.num_languages()) {
                const int li = i - vc.token_sot - 1;
                word = "[_LANG_" + std::string(li >= 0 && li < 100 ? g_langs[li].code : "??") + "]";
            } else word = "[_extra_token_" + std::to_string(i) + "]";
            vc.token_to_id[word] = i;
            vc.id_to_token[i] = word;
        }
    }

    // ---------------- device arena ----------------
    m.kc1 = (int) pad_to(M, 64);
    size_t off = 0;
    auto take = [&](size_t bytes) {
        const size_t o = off;
        off += (bytes + 255) & ~(size_t) 255;
        return o;
    };
    struct Slot { void ** pp; size_t off; };
    std::vector<Slot> slots;
    auto want = [&](void ** pp, size_t bytes) { slots.push_back({pp, take(bytes)}); };
#define WANT(p, count) want((void **) &(p), sizeof(*(p)) * (size_t) (count))
    m.enc.resize(La);
    m.dec.resize(Lt);
    WANT(m.e_pe, (size_t) T * S);
    WANT(m.conv1_w, (size_t) S * 3 * m.kc1);
    WANT(m.conv1_b, S);
    WANT(m.conv2_w, (size_t) S * 3 * S);
    WANT(m.conv2_b, S);
    WANT(m.e_ln.g, S); WANT(m.e_ln.b, S);
    for (auto & l : m.enc) {
        WANT(l.ln1.g, S); WANT(l.ln1.b, S); WANT(l.ln2.g, S); WANT(l.ln2.b, S);
        // query, key, value rows contiguous ([3S][S], bias (bq, 0, bv)): one GEMM launch produces all three
        // (2*S*S*2 and 2*S*4 bytes are multiples of the 256-byte slot size for every Whisper width; checked after allocation)
        WANT(l.qk.w, (size_t) 2 * S * S); WANT(l.v.w, (size_t) S * S);
        WANT(l.qk.b, 2 * S); WANT(l.v.b, S);
        WANT(l.o.w, (size_t) S * S); WANT(l.o.b, S);
        WANT(l.fc1.w, (size_t) 4 * S * S); WANT(l.fc1.b, 4 * S);
        WANT(l.fc2.w, (size_t) 4 * S * S); WANT(l.fc2.b, S);
    }
    WANT(m.d_pe, (size_t) Tt * S);
    WANT(m.d_te, (size_t) V * S);
    WANT(m.d_ln.g, S); WANT(m.d_ln.b, S);
    for (auto & l : m.dec) {
        WANT(l.ln1.g, S); WANT(l.ln1.b, S); WANT(l.lnc.g, S); WANT(l.lnc.b, S); WANT(l.ln2.g, S); WANT(l.ln2.b, S);
        WANT(l.qkv.w, (size_t) 3 * S * S); WANT(l.qkv.b, 3 * S);
        WANT(l.o.w, (size_t) S * S); WANT(l.o.b, S);
        WANT(l.cq.w, (size_t) S * S); WANT(l.cq.b, S);
        WANT(l.co.w, (size_t) S * S); WANT(l.co.b, S);
        WANT(l.fc1.w, (size_t) 4 * S * S); WANT(l.fc1.b, 4 * S);
        WANT(l.fc2.w, (size_t) 4 * S * S); WANT(l.fc2.b, S);
    }
    WANT(m.ckv_w, (size_t) Lt * 2 * S * S);
    WANT(m.ckv_b, (size_t) Lt * 2 * S);
    WANT(m.gelu_table, 65536);
    WANT(m.mel_sin, 400); WANT(m.mel_cos, 400); WANT(m.mel_hann, 400);
    WANT(m.mel_filters_t, (size_t) 201 * n_mel);
#undef WANT
    m.arena_size = off;
    WB_CUDA_OK(cudaMalloc((void **) &m.arena, m.arena_size));
    WB_CUDA_OK(cudaMemset(m.arena, 0, m.arena_size));  // fused-bias gaps (key projections have no bias) stay 0
    for (auto & s : slots) *s.pp = m.arena + s.off;
    for (auto & l : m.enc) {
        WB_CHECK(l.v.w == l.qk.w + (size_t) 2 * S * S && l.v.b == l.qk.b + 2 * S,
                 "encoder Q/K/V weights are not contiguous in the arena (n_audio_state %d)\n", S);
    }

    // ---------------- tables ----------------
    {
        std::vector<uint16_t> tab(65536);
        float sn[400], cs[400], hn[400];
        host_tables(tab.data(), sn, cs, hn);
        WB_CUDA_OK(cudaMemcpy(m.gelu_table, tab.data(), 65536 * 2, cudaMemcpyHostToDevice));
        WB_CUDA_OK(cudaMemcpy(m.mel_sin, sn, sizeof(sn), cudaMemcpyHostToDevice));
        WB_CUDA_OK(cudaMemcpy(m.mel_cos, cs, sizeof(cs), cudaMemcpyHostToDevice));
        WB_CUDA_OK(cudaMemcpy(m.mel_hann, hn, sizeof(hn), cudaMemcpyHostToDevice));
        std::vector<float> ft((size_t) 201 * n_mel);
        for (int j = 0; j < n_mel; ++j)
            for (int k = 0; k < 201; ++k) ft[(size_t) k * n_mel + j] = m.filters[(size_t) j * 201 + k];
        WB_CUDA_OK(cudaMemcpy(m.mel_filters_t, ft.data(), ft.size() * sizeof(float), cudaMemcpyHostToDevice));
    }

    // ---------------- tensor directory (names: whisper-arch.h:42-109) ----------------
    std::map<std::string, Dest> dir;
    auto vec = [&](const std::string & name, float * p, int n) { dir[name] = Dest{Kind::F32_VEC, p, n, 1, {n, 1, 1}}; };
    auto mat = [&](const std::string & name, __half * p, int n_out, int n_in) {
        dir[name] = Dest{Kind::F16_MAT, p, (int64_t) n_out * n_in, 2, {n_in, n_out, 1}};
    };
    dir["encoder.positional_embedding"] = Dest{Kind::F32_VEC, m.e_pe, (int64_t) T * S, 2, {S, T, 1}};
    dir["encoder.conv1.weight"] = Dest{Kind::CONV1, m.conv1_w, (int64_t) S * M * 3, 3, {3, M, S}};
    dir["encoder.conv1.bias"]   = Dest{Kind::F32_VEC, m.conv1_b, S, 2, {1, S, 1}};
    dir["encoder.conv2.weight"] = Dest{Kind::CONV2, m.conv2_w, (int64_t) S * S * 3, 3, {3, S, S}};
    dir["encoder.conv2.bias"]   = Dest{Kind::F32_VEC, m.conv2_b, S, 2, {1, S, 1}};
    vec("encoder.ln_post.weight", m.e_ln.g, S);
    vec("encoder.ln_post.bias", m.e_ln.b, S);
    for (int i = 0; i < La; ++i) {
        auto & l = m.enc[i];
        const std::string p = "encoder.blocks." + std::to_string(i) + ".";
        vec(p + "attn_ln.weight", l.ln1.g, S); vec(p + "attn_ln.bias", l.ln1.b, S);
        vec(p + "mlp_ln.weight", l.ln2.g, S); vec(p + "mlp_ln.bias", l.ln2.b, S);
        mat(p + "attn.query.weight", l.qk.w, S, S); vec(p + "attn.query.bias", l.qk.b, S);
        mat(p + "attn.key.weight", l.qk.w + (size_t) S * S, S, S);
        mat(p + "attn.value.weight", l.v.w, S, S); vec(p + "attn.value.bias", l.v.b, S);
        mat(p + "attn.out.weight", l.o.w, S, S); vec(p + "attn.out.bias", l.o.b, S);
        mat(p + "mlp.0.weight", l.fc1.w, 4 * S, S); vec(p + "mlp.0.bias", l.fc1.b, 4 * S);
        mat(p + "mlp.2.weight", l.fc2.w, S, 4 * S); vec(p + "mlp.2.bias", l.fc2.b, S);
    }
    dir["decoder.positional_embedding"] = Dest{Kind::F32_VEC, m.d_pe, (int64_t) Tt * S, 2, {S, Tt, 1}};
    mat("decoder.token_embedding.weight", m.d_te, V, S);
    vec("decoder.ln.weight", m.d_ln.g, S);
    vec("decoder.ln.bias", m.d_ln.b, S);
    for (int i = 0; i < Lt; ++i) {
        auto & l = m.dec[i];
        const std::string p = "decoder.blocks." + std::to_string(i) + ".";
        vec(p + "attn_ln.weight", l.ln1.g, S); vec(p + "attn_ln.bias", l.ln1.b, S);
        vec(p + "cross_attn_ln.weight", l.lnc.g, S); vec(p + "cross_attn_ln.bias", l.lnc.b, S);
        vec(p + "mlp_ln.weight", l.ln2.g, S); vec(p + "mlp_ln.bias", l.ln2.b, S);
        mat(p + "attn.query.weight", l.qkv.w, S, S); vec(p + "attn.query.bias", l.qkv.b, S);
        mat(p + "attn.key.weight", l.qkv.w + (size_t) S * S, S, S);
        mat(p + "attn.value.weight", l.qkv.w + (size_t) 2 * S * S, S, S); vec(p + "attn.value.bias", l.qkv.b + 2 * S, S);
        mat(p + "attn.out.weight", l.o.w, S, S); vec(p + "attn.out.bias", l.o.b, S);
        mat(p + "cross_attn.query.weight", l.cq.w, S, S); vec(p + "cross_attn.query.bias", l.cq.b, S);
        mat(p + "cross_attn.key.weight", m.ckv_w + (size_t) i * 2 * S * S, S, S);
        mat(p + "cross_attn.value.weight", m.ckv_w + (size_t) i * 2 * S * S + (size_t) S * S, S, S);
        vec(p + "cross_attn.value.bias", m.ckv_b + (size_t) i * 2 * S + S, S);
        mat(p + "cross_attn.out.weight", l.co.w, S, S); vec(p + "cross_attn.out.bias", l.co.b, S);
        mat(p + "mlp.0.weight", l.fc1.w, 4 * S, S); vec(p + "mlp.0.bias", l.fc1.b, 4 * S);
        mat(p + "mlp.2.weight", l.fc2.w, S, 4 * S); vec(p + "mlp.2.bias", l.fc2.b, S);
    }

    // ---------------- stream the tensor records (wc:1887-1955) ----------------
    std::vector<uint8_t> buf;
    std::vector<uint16_t> re;
    std::vector<float> deq;
    size_t total = 0;
    m.n_loaded = 0;
    while (true) {
        int32_t n_dims = 0, name_len = 0, ttype = 0;
        rd.ok = true;
        // a clean end of file lands on the first field of a record (the reference tests eof() after these reads, wc:1892-1898)
        if (!rd.pod(n_dims) || !rd.pod(name_len) || !rd.pod(ttype)) break;
        if (n_dims < 1 || n_dims > 3 || name_len <= 0 || name_len > 256) {
            log_error("%s: corrupt tensor record (n_dims %d, name_len %d)\n", __func__, n_dims, name_len);
            return false;
        }
        int32_t ne[3] = {1, 1, 1};
        int64_t nelements = 1;
        for (int i = 0; i < n_dims; ++i) {
            if (!rd.pod(ne[i])) return false;
            nelements *= ne[i];
        }
        std::string name((size_t) name_len, '\0');
        if (!rd.read(&name[0], (size_t) name_len)) return false;
        auto it = dir.find(name);
        if (it == dir.end()) {
            log_error("%s: unknown tensor '%s' in model file\n", __func__, name.c_str());
            return false;
        }
        const Dest & d = it->second;
        if (nelements != d.n) {
            log_error("%s: tensor '%s' has wrong size in model file\n", __func__, name.c_str());
            return false;
        }
        if (ne[0] != d.ne[0] || ne[1] != d.ne[1] || ne[2] != d.ne[2]) {
            log_error("%s: tensor '%s' has wrong shape in model file: got [%d, %d, %d], expected [%d, %d, %d]\n", __func__, name.c_str(),
                      ne[0], ne[1], ne[2], d.ne[0], d.ne[1], d.ne[2]);
            return false;
        }
        const size_t qbs = quant_block_bytes(ttype);
        if (ttype != GT_F32 && ttype != GT_F16 && qbs == 0) {
            log_error("%s: tensor '%s' has unsupported type %d\n", __func__, name.c_str(), ttype);
            return false;
        }
        const bool want_f16 = d.kind != Kind::F32_VEC;
        if (!want_f16 && ttype != GT_F32) {
            log_error("%s: tensor '%s' has wrong size in model file: expected f32 data\n", __func__, name.c_str());
            return false;
        }
        if (qbs != 0 && (d.kind != Kind::F16_MAT || ttype != wtype || (ne[0] % quant_block_elems(ttype)) != 0)) {
            // the reference sizes every matrix by the file's wtype and the conv kernels by f16 (wc:1696): anything else is
            // "wrong size in model file" there too
            log_error("%s: tensor '%s' has wrong size in model file (type %d, model wtype %d)\n", __func__, name.c_str(), ttype, wtype);
            return false;
        }
        const size_t file_bytes = qbs != 0 ? (size_t) (nelements / quant_block_elems(ttype)) * qbs : (size_t) nelements * (ttype == GT_F16 ? 2 : 4);
        buf.resize(file_bytes);
        if (!rd.read(buf.data(), buf.size())) {
            log_error("%s: tensor '%s' is truncated\n", __func__, name.c_str());
            return false;
        }
        total += file_bytes;
        if (qbs != 0) {  // expand the blocks to f32 (dequantize_row_*), then treat the tensor like an f32 checkpoint's
            deq.resize((size_t) nelements);
            dequantize_blocks(ttype, buf.data(), nelements, deq.data());
            buf.resize((size_t) nelements * 4);
            memcpy(buf.data(), deq.data(), buf.size());
            ttype = GT_F32;
        }
        const void * src = buf.data();
        size_t nbytes = buf.size();
        if (want_f16) {
            if (ttype == 0) {  // f32 checkpoint: round the matrices to f16 once at load
                re.resize((size_t) nelements);
                const float * f = (const float *) buf.data();
                for (int64_t i = 0; i < nelements; ++i) re[(size_t) i] = f2h(f[i]);
                src = re.data();
                nbytes = (size_t) nelements * 2;
            }
            if (d.kind == Kind::CONV1 || d.kind == Kind::CONV2) {
                // [oc][ic][kw] -> tap-major [oc][kw][ic_pad]: the implicit-GEMM conv reads one tap per k-slab
                const int IC = d.ne[1], OC = d.ne[2];
                const int ICP = d.kind == Kind::CONV1 ? m.kc1 : IC;
                std::vector<uint16_t> rr((size_t) OC * 3 * ICP, 0);
                const uint16_t * s16 = (const uint16_t *) src;
                for (int oc = 0; oc < OC; ++oc)
                    for (int ic = 0; ic < IC; ++ic)
                        for (int kw = 0; kw < 3; ++kw) rr[((size_t) oc * 3 + kw) * ICP + ic] = s16[((size_t) oc * IC + ic) * 3 + kw];
                WB_CUDA_OK(cudaMemcpy(d.ptr, rr.data(), rr.size() * 2, cudaMemcpyHostToDevice));
            } else {
                WB_CUDA_OK(cudaMemcpy(d.ptr, src, nbytes, cudaMemcpyHostToDevice));
            }
        } else {
            WB_CUDA_OK(cudaMemcpy(d.ptr, src, nbytes, cudaMemcpyHostToDevice));
        }
        m.n_loaded++;
    }
    log_msg(GGML_LOG_LEVEL_INFO, "%s: model size    = %7.2f MB (%d tensors, device arena %.2f MB)\n", __func__, total / 1e6, m.n_loaded,
            m.arena_size / 1e6);
    if (m.n_loaded == 0) {
        log_msg(GGML_LOG_LEVEL_WARN, "%s: WARN no tensors loaded from model file - assuming empty model for testing\n", __func__);
    } else if (m.n_loaded != (int) dir.size()) {
        log_error("%s: ERROR not all tensors loaded from model file - expected %zu, got %d\n", __func__, dir.size(), m.n_loaded);
        return false;
    }
    return true;
}

}  // namespace wb
