// Model description + device weight arena (reference types: whisper_hparams wc:623-636, whisper_model
// wc:752-795, whisper_vocab wc:462-491, loader wc:1503-1974).  "wc" = sys/whisper.cpp/src/whisper.cpp.
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/whisper.h"

namespace wb {

struct HParams {
    int32_t n_vocab = 51864, n_audio_ctx = 1500, n_audio_state = 384, n_audio_head = 6, n_audio_layer = 4;
    int32_t n_text_ctx = 448, n_text_state = 384, n_text_head = 6, n_text_layer = 4, n_mels = 80, ftype = 1;
    float eps = 1e-5f;
};

struct Vocab {
    int n_vocab = 51864;
    std::map<std::string, int> token_to_id;
    std::vector<std::string> id_to_token;
    // defaults are the English-only ids; fixed up for multilingual vocabularies (wc:472-482, 1642-1657)
    int token_eot = 50256, token_sot = 50257, token_translate = 50357, token_transcribe = 50358, token_solm = 50359;
    int token_prev = 50360, token_nosp = 50361, token_not = 50362, token_beg = 50363;
    bool is_multilingual() const { return n_vocab >= 51865; }
    int num_languages() const { return n_vocab - 51765 - (is_multilingual() ? 1 : 0); }
};

struct LangEntry {
    const char * code;
    const char * name;
};
extern const LangEntry g_langs[100];
int lang_id(const char * lang);  // -1 unknown

struct Linear {
    __half * w = nullptr;  // [n_out][n_in] f16
    float * b  = nullptr;  // [n_out] f32 (may be null)
};
struct Norm {
    float * g = nullptr;
    float * b = nullptr;
};

struct EncLayer {
    Norm ln1, ln2;
    Linear qk;  // fused [2S][S]: query rows then key rows; bias = (bq, 0)
    Linear v;
    Linear o, fc1, fc2;
};
struct DecLayer {
    Norm ln1, lnc, ln2;
    Linear qkv;  // fused [3S][S]; bias = (bq, 0, bv)
    Linear o;
    Linear cq, co;
    Linear fc1, fc2;
};

struct Model {
    HParams hp;
    Vocab vocab;
    int model_type = 0;  // e_model numbering of the reference: 0 unknown, 1 tiny, 2 base, 3 small, 4 medium, 5 large
    int n_loaded   = 0;  // tensor records found in the file (0 = header-only test model, wc:1959-1960)
    int device     = 0;

    std::vector<float> filters;  // host copy [n_mel][201]
    int filters_n_mel = 0;

    // device arena
    uint8_t * arena = nullptr;
    size_t arena_size = 0;

    // encoder
    float * e_pe = nullptr;                            // [n_audio_ctx][S]
    __half * conv1_w = nullptr;                        // rearranged [S][3][KC1] (tap-major, zero padded to KC1)
    __half * conv2_w = nullptr;                        // rearranged [S][3][S]
    float *conv1_b = nullptr, *conv2_b = nullptr;      // [S]
    int kc1 = 128;                                     // padded input-channel count of conv1
    Norm e_ln;
    std::vector<EncLayer> enc;
    // decoder
    float * d_pe = nullptr;   // [n_text_ctx][S]
    __half * d_te = nullptr;  // [V][S]
    Norm d_ln;
    std::vector<DecLayer> dec;
    __half * ckv_w = nullptr;  // [L][2S][S]  cross-attention key rows then value rows, all layers contiguous
    float * ckv_b  = nullptr;  // [L][2S]     (0, bv)

    // tables
    __half * gelu_table = nullptr;  // [65536] f16 -> f16 (ggml_table_gelu_f16, ggml-cpu.c:3515)
    float *mel_sin = nullptr, *mel_cos = nullptr, *mel_hann = nullptr, *mel_filters_t = nullptr;

    ~Model();
};

// Streaming loader over the ABI's whisper_model_loader (wh:153-158).  Returns false (with a logged reason) on any
// malformed input, exactly where the reference's loader would (wc:1513-1965).
bool load_model(whisper_model_loader * loader, int device, Model & model);
// host only: the GELU f16 table and the mel front end's sin / cos / Hann tables as load_model uploads them (null pointers are skipped)
void host_tables(uint16_t * gelu, float * sn, float * cs, float * hn);
// host only: expand n elements of a quantised ggml type (q4_0 ... q8_0, q2_K ... q6_K) to f32 the way load_model does
bool dequantize_tensor(int ggml_type, const void * blocks, int64_t n, float * dst);

}  // namespace wb
