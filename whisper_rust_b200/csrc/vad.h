// VAD front end: model, device state and the entry points abi.cu builds the whisper_vad_* API on.  See vad.cu.
#pragma once

#include <atomic>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/whisper.h"

struct __half;

namespace wb {

struct VadModel {   // host copy of the file (whisper_vad_model / whisper_vad_hparams, whisper.cpp:4361-4414)
    std::string type, version;
    int32_t n_window = 0, n_context = 0, n_encoder_layers = 0;
    int32_t enc_in[4] = {0, 0, 0, 0}, enc_out[4] = {0, 0, 0, 0}, enc_k[4] = {0, 0, 0, 0};
    int32_t lstm_input = 0, lstm_hidden = 0, final_in = 0, final_out = 0;
    std::vector<uint8_t> stft_basis, enc_w[4], enc_b[4], w_ih, w_hh, b_ih, b_hh, final_w, final_b;   // raw tensor bytes as in the file
};

struct VadDevice {  // weights and work buffers on the GPU, created on first use
    int device = 0;
    uint8_t * arena = nullptr;
    const __half * basis = nullptr;
    const __half * w[4] = {nullptr, nullptr, nullptr, nullptr};
    const float * b[4] = {nullptr, nullptr, nullptr, nullptr};
    const float *w_ih_t = nullptr, *w_hh_t = nullptr, *b_ih = nullptr, *b_hh = nullptr, *final_b = nullptr;
    const __half * final_w = nullptr;
    float * d_pcm = nullptr;
    size_t pcm_cap = 0;
    float * d_work = nullptr;
    size_t work_cap = 0;
    ~VadDevice();
};

bool vad_load(whisper_model_loader * loader, VadModel & m);
bool vad_detect(const VadModel & m, VadDevice & d, const float * samples, int n_samples, std::vector<float> & probs);
uint64_t vad_launch_count();

}  // namespace wb
