// Voice-activity-detection front end: the Silero-VAD graph of the reference (whisper.cpp:4361-5480) as three CUDA
// kernels, the speech-segment logic (whisper.cpp:5202-5436) and the sample filter + time-mapping table of whisper_full (whisper.cpp:
// 6615-6793, 7882-7960) on the host.
//
// Per 512-sample window (whisper_vad_build_graph, wc:4627-4668):
//   reflect-pad 64 | conv1d(basis [258 x 256], stride = lstm_input_size = 128) -> 4 frames x 258 | magnitude of (real, imag) halves
//   -> [129 x 4] | four conv1d(k = 3, pad 1, strides 1, 2, 2, 1) + bias + ReLU -> [128 x 1] | one LSTM cell step (state carried from
//   window to window) | ReLU | dot with the final conv kernel + bias | sigmoid = speech probability.
// Rounding points mirror ggml: conv1d = im2col to F16 (ggml.c:3929-3950) x F16 kernel with f32 accumulation, so every conv INPUT is
// rounded to f16; the LSTM matrices are f32 x f32; sigmoid / tanh in f32.
//   vad_features_kernel : one CTA per window, everything up to the 128-feature vector (windows are independent)
//   vad_gates_kernel    : input-to-hidden gate pre-activations W_ih x + b_ih for every window (independent)
//   vad_lstm_kernel     : ONE CTA walks the windows in time order (the recurrence), 512 threads = one per gate row; W_hh is read
//                         transposed ([128][512], coalesced, L2-resident: 256 KB) every step
#include "vad.h"

#include <cuda_fp16.h>
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <climits>
#include <map>
#include <string>

#include "common.cuh"
#include "log.h"

namespace wb {

namespace {

// fixed structure of the graph the reference builds (it hard-codes them too: wc:4537-4579, 4655)
constexpr int VAD_PAD = 64, VAD_TAPS = 256, VAD_FRAMES = 4, VAD_BINS = 129, VAD_HID = 128;

__device__ __forceinline__ float f16r(float x) { return __half2float(__float2half_rn(x)); }  // the im2col rounding

// grid = n_windows, 256 threads.  out: x [n_windows][128]
__global__ void __launch_bounds__(256)
vad_features_kernel(const float * __restrict__ pcm, int n_samples, int n_window, const __half * __restrict__ basis /*[258][256]*/,
                    const __half * __restrict__ w0, const float * __restrict__ b0,   // [128][129][3]
                    const __half * __restrict__ w1, const float * __restrict__ b1,   // [64][128][3]
                    const __half * __restrict__ w2, const float * __restrict__ b2,   // [64][64][3]
                    const __half * __restrict__ w3, const float * __restrict__ b3,   // [128][64][3]
                    float * __restrict__ x_out) {
    __shared__ float s_pad[512 + 2 * VAD_PAD];          // reflect-padded window, rounded to f16 (conv input)
    __shared__ float s_stft[2 * VAD_BINS][VAD_FRAMES];  // conv output [258][4]
    __shared__ float s_mag[VAD_BINS][VAD_FRAMES];       // f16-rounded magnitudes (input of encoder 0)
    __shared__ float s_a[128][VAD_FRAMES];              // encoder 0 out [128][4]
    __shared__ float s_b[64][2];                        // encoder 1 out [64][2]
    __shared__ float s_c[64];                           // encoder 2 out [64][1]
    const int w = blockIdx.x, tid = threadIdx.x;
    const int64_t base = (int64_t) w * n_window;
    // window samples (zero beyond the clip, wc:5139-5152), then ggml_pad_reflect_1d(64, 64)
    for (int i = tid; i < n_window + 2 * VAD_PAD; i += 256) {
        int j = i - VAD_PAD;
        if (j < 0) j = -j;
        else if (j >= n_window) j = 2 * (n_window - 1) - j;
        const int64_t g = base + j;
        s_pad[i] = f16r(g < n_samples ? pcm[g] : 0.0f);
    }
    __syncthreads();
    // STFT as a strided conv: s[c][f] = sum_t basis[c][t] * padded[f * 128 + t]
    for (int o = tid; o < 2 * VAD_BINS * VAD_FRAMES; o += 256) {
        const int c = o / VAD_FRAMES, f = o % VAD_FRAMES;
        const __half * bw = basis + (size_t) c * VAD_TAPS;
        const float * px = s_pad + f * 128;
        float acc = 0.0f;
        for (int t = 0; t < VAD_TAPS; ++t) acc = fmaf(__half2float(bw[t]), px[t], acc);
        s_stft[c][f] = acc;
    }
    __syncthreads();
    for (int o = tid; o < VAD_BINS * VAD_FRAMES; o += 256) {
        const int c = o / VAD_FRAMES, f = o % VAD_FRAMES;
        const float re = s_stft[c][f], im = s_stft[c + VAD_BINS][f];
        s_mag[c][f] = f16r(sqrtf(__fadd_rn(__fmul_rn(re, re), __fmul_rn(im, im))));
    }
    __syncthreads();
    // encoder 0: [129][4] -> [128][4], stride 1, pad 1
    for (int o = tid; o < 128 * VAD_FRAMES; o += 256) {
        const int oc = o / VAD_FRAMES, l = o % VAD_FRAMES;
        const __half * ww = w0 + (size_t) oc * VAD_BINS * 3;
        float acc = 0.0f;
        for (int ic = 0; ic < VAD_BINS; ++ic)
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int p = l + k - 1;
                if (p >= 0 && p < VAD_FRAMES) acc = fmaf(__half2float(ww[ic * 3 + k]), s_mag[ic][p], acc);
            }
        s_a[oc][l] = f16r(fmaxf(__fadd_rn(acc, b0[oc]), 0.0f));
    }
    __syncthreads();
    // encoder 1: [128][4] -> [64][2], stride 2, pad 1
    for (int o = tid; o < 64 * 2; o += 256) {
        const int oc = o / 2, l = o % 2;
        const __half * ww = w1 + (size_t) oc * 128 * 3;
        float acc = 0.0f;
        for (int ic = 0; ic < 128; ++ic)
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int p = 2 * l + k - 1;
                if (p >= 0 && p < VAD_FRAMES) acc = fmaf(__half2float(ww[ic * 3 + k]), s_a[ic][p], acc);
            }
        s_b[oc][l] = f16r(fmaxf(__fadd_rn(acc, b1[oc]), 0.0f));
    }
    __syncthreads();
    // encoder 2: [64][2] -> [64][1], stride 2, pad 1
    for (int oc = tid; oc < 64; oc += 256) {
        const __half * ww = w2 + (size_t) oc * 64 * 3;
        float acc = 0.0f;
        for (int ic = 0; ic < 64; ++ic)
#pragma unroll
            for (int k = 1; k < 3; ++k) acc = fmaf(__half2float(ww[ic * 3 + k]), s_b[ic][k - 1], acc);  // tap 0 falls on the left padding
        s_c[oc] = f16r(fmaxf(__fadd_rn(acc, b2[oc]), 0.0f));
    }
    __syncthreads();
    // encoder 3: [64][1] -> [128][1], stride 1, pad 1 (only the centre tap sees data); the graph keeps time index 0 (wc:4655)
    for (int oc = tid; oc < 128; oc += 256) {
        const __half * ww = w3 + (size_t) oc * 64 * 3;
        float acc = 0.0f;
        for (int ic = 0; ic < 64; ++ic) acc = fmaf(__half2float(ww[ic * 3 + 1]), s_c[ic], acc);
        x_out[(size_t) w * VAD_HID + oc] = fmaxf(__fadd_rn(acc, b3[oc]), 0.0f);
    }
}

// grid = n_windows, 512 threads: gx[w][r] = W_ih[r] . x_w + b_ih[r]   (wc:4590-4592)
__global__ void __launch_bounds__(512)
vad_gates_kernel(const float * __restrict__ x, const float * __restrict__ w_ih_t /*[128][512]*/, const float * __restrict__ b_ih,
                 float * __restrict__ gx) {
    __shared__ float s_x[VAD_HID];
    const int w = blockIdx.x, r = threadIdx.x;
    if (r < VAD_HID) s_x[r] = x[(size_t) w * VAD_HID + r];
    __syncthreads();
    float acc = 0.0f;
#pragma unroll 8
    for (int j = 0; j < VAD_HID; ++j) acc = fmaf(w_ih_t[j * 512 + r], s_x[j], acc);
    gx[(size_t) w * 512 + r] = __fadd_rn(acc, b_ih[r]);
}

// one CTA, 512 threads: the recurrence over the windows of one clip (wc:4594-4624, 4657-4662)
__global__ void __launch_bounds__(512)
vad_lstm_kernel(const float * __restrict__ gx, int n_windows, const float * __restrict__ w_hh_t /*[128][512]*/, const float * __restrict__ b_hh,
                const __half * __restrict__ w_final /*[128]*/, const float * __restrict__ b_final, float * __restrict__ probs) {
    __shared__ float s_h[VAD_HID], s_c[VAD_HID], s_g[512], s_red[4];
    const int r = threadIdx.x;
    if (r < VAD_HID) {
        s_h[r] = 0.0f;  // ggml_backend_buffer_clear at the start of every clip (wc:5115)
        s_c[r] = 0.0f;
    }
    const float bh = b_hh[r];
    const float wf = r < VAD_HID ? __half2float(w_final[r]) : 0.0f;
    __syncthreads();
    for (int w = 0; w < n_windows; ++w) {
        float acc = 0.0f;
#pragma unroll 8
        for (int j = 0; j < VAD_HID; ++j) acc = fmaf(w_hh_t[j * 512 + r], s_h[j], acc);
        const float pre = __fadd_rn(gx[(size_t) w * 512 + r], __fadd_rn(acc, bh));  // (W_ih x + b_ih) + (W_hh h + b_hh)
        const int gate = r >> 7;                                                       // i, f, g, o
        s_g[r] = gate == 2 ? tanhf(pre) : 1.0f / (1.0f + expf(-pre));
        __syncthreads();
        float contrib = 0.0f;
        if (r < VAD_HID) {
            const float c = __fadd_rn(__fmul_rn(s_g[VAD_HID + r], s_c[r]), __fmul_rn(s_g[r], s_g[2 * VAD_HID + r]));
            const float h = __fmul_rn(s_g[3 * VAD_HID + r], tanhf(c));
            s_c[r] = c;
            s_h[r] = h;
            contrib = __fmul_rn(wf, f16r(fmaxf(h, 0.0f)));  // final conv: F16 kernel x F16(im2col(relu(h)))
        }
        // sum of the 128 products: warp shuffles, then the four warp sums
        if (r < VAD_HID) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, o);
            if ((r & 31) == 0) s_red[r >> 5] = contrib;
        }
        __syncthreads();
        if (r == 0) {
            const float z = __fadd_rn((s_red[0] + s_red[1]) + (s_red[2] + s_red[3]), b_final[0]);
            probs[w] = 1.0f / (1.0f + expf(-z));
        }
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------------------
// model file (wc:4775-5098): magic | type string | version (3 x i32) | n_window | n_context | encoder table | lstm / final dims | tensors
// ---------------------------------------------------------------------------------------------------------------------
bool vad_load(whisper_model_loader * loader, VadModel & m) {
    auto rd = [&](void * dst, size_t n) { return loader->read(loader->context, dst, n) == n; };
    uint32_t magic = 0;
    if (!rd(&magic, 4) || magic != 0x67676d6c) {
        log_error("%s: invalid model data (bad magic)\n", __func__);
        return false;
    }
    int32_t len = 0;
    if (!rd(&len, 4) || len < 0 || len > 256) return false;
    m.type.resize((size_t) len);
    if (len && !rd(&m.type[0], (size_t) len)) return false;
    int32_t ver[3];
    if (!rd(ver, sizeof(ver))) return false;
    m.version = std::to_string(ver[0]) + "." + std::to_string(ver[1]) + "." + std::to_string(ver[2]);
    if (!rd(&m.n_window, 4) || !rd(&m.n_context, 4) || !rd(&m.n_encoder_layers, 4)) return false;
    if (m.n_encoder_layers != 4) {
        log_error("%s: %d encoder layers: the graph has four (wc:4557-4579)\n", __func__, m.n_encoder_layers);
        return false;
    }
    for (int i = 0; i < 4; ++i)
        if (!rd(&m.enc_in[i], 4) || !rd(&m.enc_out[i], 4) || !rd(&m.enc_k[i], 4)) return false;
    if (!rd(&m.lstm_input, 4) || !rd(&m.lstm_hidden, 4) || !rd(&m.final_in, 4) || !rd(&m.final_out, 4)) return false;
    const int exp_in[4] = {VAD_BINS, 128, 64, 64}, exp_out[4] = {128, 64, 64, 128};
    bool shape_ok = m.n_window == 512 && m.lstm_input == 128 && m.lstm_hidden == VAD_HID && m.final_in == VAD_HID;
    for (int i = 0; i < 4; ++i) shape_ok = shape_ok && m.enc_in[i] == exp_in[i] && m.enc_out[i] == exp_out[i] && m.enc_k[i] == 3;
    if (!shape_ok) {
        log_error("%s: unsupported VAD architecture (only the Silero v5 16 kHz shape the reference's graph hard-codes: 512-sample windows, "
                  "129-128-64-64-128 encoder, 128-unit LSTM)\n", __func__);
        return false;
    }
    struct Want { std::vector<uint8_t> * dst; int64_t n; int ttype; int ne[3]; };
    std::map<std::string, Want> dir;
    auto want = [&](const char * name, std::vector<uint8_t> & dst, int ttype, int n0, int n1, int n2) {
        dir[name] = Want{&dst, (int64_t) n0 * n1 * n2, ttype, {n0, n1, n2}};
    };
    want("_model.stft.forward_basis_buffer", m.stft_basis, 1, VAD_TAPS, 1, 2 * VAD_BINS);
    const char * wn[4] = {"_model.encoder.0.reparam_conv.weight", "_model.encoder.1.reparam_conv.weight", "_model.encoder.2.reparam_conv.weight",
                          "_model.encoder.3.reparam_conv.weight"};
    const char * bn[4] = {"_model.encoder.0.reparam_conv.bias", "_model.encoder.1.reparam_conv.bias", "_model.encoder.2.reparam_conv.bias",
                          "_model.encoder.3.reparam_conv.bias"};
    for (int i = 0; i < 4; ++i) {
        want(wn[i], m.enc_w[i], 1, 3, m.enc_in[i], m.enc_out[i]);
        want(bn[i], m.enc_b[i], 0, m.enc_out[i], 1, 1);
    }
    want("_model.decoder.rnn.weight_ih", m.w_ih, 0, VAD_HID, 4 * VAD_HID, 1);
    want("_model.decoder.rnn.weight_hh", m.w_hh, 0, VAD_HID, 4 * VAD_HID, 1);
    want("_model.decoder.rnn.bias_ih", m.b_ih, 0, 4 * VAD_HID, 1, 1);
    want("_model.decoder.rnn.bias_hh", m.b_hh, 0, 4 * VAD_HID, 1, 1);
    want("_model.decoder.decoder.2.weight", m.final_w, 1, m.final_in, 1, 1);
    want("_model.decoder.decoder.2.bias", m.final_b, 0, 1, 1, 1);
    int n_loaded = 0;
    while (true) {
        int32_t n_dims = 0, name_len = 0, ttype = 0;
        if (!rd(&n_dims, 4) || !rd(&name_len, 4) || !rd(&ttype, 4)) break;  // clean end of file (wc:5019-5025)
        if (n_dims < 1 || n_dims > 3 || name_len <= 0 || name_len > 256) {
            log_error("%s: corrupt tensor record\n", __func__);
            return false;
        }
        int32_t ne[3] = {1, 1, 1};
        int64_t nel = 1;
        for (int i = 0; i < n_dims; ++i) {
            if (!rd(&ne[i], 4)) return false;
            nel *= ne[i];
        }
        std::string name((size_t) name_len, '\0');
        if (!rd(&name[0], (size_t) name_len)) return false;
        auto it = dir.find(name);
        if (it == dir.end()) {
            log_error("%s: unknown tensor '%s' in model file\n", __func__, name.c_str());
            return false;
        }
        const Want & wv = it->second;
        if (nel != wv.n || ne[0] != wv.ne[0] || ne[1] != wv.ne[1] || ne[2] != wv.ne[2]) {
            log_error("%s: tensor '%s' has wrong shape in model file: got [%d, %d, %d], expected [%d, %d, %d]\n", __func__, name.c_str(), ne[0],
                      ne[1], ne[2], wv.ne[0], wv.ne[1], wv.ne[2]);
            return false;
        }
        if (ttype != wv.ttype) {
            log_error("%s: tensor '%s' has wrong size in model file (type %d, expected %d)\n", __func__, name.c_str(), ttype, wv.ttype);
            return false;
        }
        wv.dst->resize((size_t) nel * (ttype == 1 ? 2 : 4));
        if (!rd(wv.dst->data(), wv.dst->size())) {
            log_error("%s: tensor '%s' is truncated\n", __func__, name.c_str());
            return false;
        }
        ++n_loaded;
    }
    if (n_loaded != (int) dir.size()) {
        log_error("%s: ERROR not all tensors loaded from model file - expected %zu, got %d\n", __func__, dir.size(), n_loaded);
        return false;
    }
    return true;
}

VadDevice::~VadDevice() {
    if (arena) {
        cudaSetDevice(device);
        cudaFree(arena);
    }
    if (d_pcm) cudaFree(d_pcm);
    if (d_work) cudaFree(d_work);
}

// weights to the device on first use (so that the host-side API - loading, segments_from_probs - works without a GPU)
static bool vad_upload(const VadModel & m, VadDevice & d) {
    if (d.arena) return true;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) {
        log_error("%s: no CUDA device: this library has no CPU path\n", __func__);
        return false;
    }
    WB_CUDA_OK(cudaSetDevice(d.device));
    size_t off = 0;
    auto take = [&](size_t b) {
        const size_t o = off;
        off += (b + 255) & ~(size_t) 255;
        return o;
    };
    const size_t o_basis = take(m.stft_basis.size());
    size_t o_w[4], o_b[4];
    for (int i = 0; i < 4; ++i) {
        o_w[i] = take(m.enc_w[i].size());
        o_b[i] = take(m.enc_b[i].size());
    }
    const size_t o_ih = take(m.w_ih.size()), o_hh = take(m.w_hh.size()), o_bih = take(m.b_ih.size()), o_bhh = take(m.b_hh.size());
    const size_t o_fw = take(m.final_w.size()), o_fb = take(m.final_b.size());
    WB_CUDA_OK(cudaMalloc((void **) &d.arena, off));
    auto up = [&](size_t o, const std::vector<uint8_t> & v) { return cudaMemcpy(d.arena + o, v.data(), v.size(), cudaMemcpyHostToDevice) == cudaSuccess; };
    // the recurrent matrices transposed to [128][512]: thread r of the gate kernels reads column r, coalesced
    auto transpose = [](const std::vector<uint8_t> & v) {
        std::vector<uint8_t> t(v.size());
        const float * s = (const float *) v.data();
        float * o = (float *) t.data();
        for (int r = 0; r < 4 * VAD_HID; ++r)
            for (int j = 0; j < VAD_HID; ++j) o[(size_t) j * 4 * VAD_HID + r] = s[(size_t) r * VAD_HID + j];
        return t;
    };
    bool ok = up(o_basis, m.stft_basis) && up(o_ih, transpose(m.w_ih)) && up(o_hh, transpose(m.w_hh)) && up(o_bih, m.b_ih) && up(o_bhh, m.b_hh) &&
              up(o_fw, m.final_w) && up(o_fb, m.final_b);
    for (int i = 0; i < 4; ++i) ok = ok && up(o_w[i], m.enc_w[i]) && up(o_b[i], m.enc_b[i]);
    WB_CHECK(ok, "VAD weight upload failed\n");
    d.basis = (const __half *) (d.arena + o_basis);
    for (int i = 0; i < 4; ++i) {
        d.w[i] = (const __half *) (d.arena + o_w[i]);
        d.b[i] = (const float *) (d.arena + o_b[i]);
    }
    d.w_ih_t = (const float *) (d.arena + o_ih); d.w_hh_t = (const float *) (d.arena + o_hh);
    d.b_ih = (const float *) (d.arena + o_bih); d.b_hh = (const float *) (d.arena + o_bhh);
    d.final_w = (const __half *) (d.arena + o_fw); d.final_b = (const float *) (d.arena + o_fb);
    return true;
}

static std::atomic<uint64_t> g_vad_launches{0};
uint64_t vad_launch_count() { return g_vad_launches.load(); }

// whisper_vad_detect_speech (wc:5100-5180): probabilities of every 512-sample window of the clip; samples may be host or device memory
bool vad_detect(const VadModel & m, VadDevice & d, const float * samples, int n_samples, std::vector<float> & probs) {
    const int n_win = n_samples / m.n_window + (n_samples % m.n_window != 0 ? 1 : 0);
    probs.assign((size_t) std::max(n_win, 0), 0.0f);
    if (n_win <= 0) return true;
    if (!vad_upload(m, d)) return false;
    WB_CUDA_OK(cudaSetDevice(d.device));
    if ((size_t) n_samples > d.pcm_cap) {
        cudaFree(d.d_pcm);
        d.d_pcm = nullptr;
        WB_CUDA_OK(cudaMalloc((void **) &d.d_pcm, (size_t) n_samples * 4));
        d.pcm_cap = (size_t) n_samples;
    }
    const size_t work = (size_t) n_win * (VAD_HID + 512 + 1);
    if (work > d.work_cap) {
        cudaFree(d.d_work);
        d.d_work = nullptr;
        WB_CUDA_OK(cudaMalloc((void **) &d.d_work, work * 4));
        d.work_cap = work;
    }
    float * d_x = d.d_work, * d_gx = d_x + (size_t) n_win * VAD_HID, * d_probs = d_gx + (size_t) n_win * 512;
    WB_CUDA_OK(cudaMemcpy(d.d_pcm, samples, (size_t) n_samples * 4, cudaMemcpyDefault));
    vad_features_kernel<<<n_win, 256>>>(d.d_pcm, n_samples, m.n_window, d.basis, d.w[0], d.b[0], d.w[1], d.b[1], d.w[2], d.b[2], d.w[3], d.b[3], d_x);
    WB_CUDA_OK(cudaGetLastError());
    vad_gates_kernel<<<n_win, 512>>>(d_x, d.w_ih_t, d.b_ih, d_gx);
    WB_CUDA_OK(cudaGetLastError());
    vad_lstm_kernel<<<1, 512>>>(d_gx, n_win, d.w_hh_t, d.b_hh, d.final_w, d.final_b, d_probs);
    WB_CUDA_OK(cudaGetLastError());
    g_vad_launches.fetch_add(3);
    WB_CUDA_OK(cudaMemcpy(probs.data(), d_probs, (size_t) n_win * 4, cudaMemcpyDeviceToHost));
    return true;
}

}  // namespace wb
