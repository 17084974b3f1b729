#include "prof.h"

#include <mutex>
#include <vector>

namespace wb {

static const char * k_names[PC_COUNT] = {"mel", "gemm_encoder", "encoder_softmax", "layernorm", "gemm_decoder", "gemm_logits",
                                         "self_attn_decode", "cross_attn_decode", "sampler", "embed", "dtw", "other", "layernorm_decode",
                                         "event_bracket_null_kernel", "cross_attn_step", "gemm_step_qkv", "gemm_step_attn_proj",
                                         "gemm_step_fc1", "gemm_step_fc2"};
const char * prof_cat_name(int cat) { return cat >= 0 && cat < PC_COUNT ? k_names[cat] : "?"; }

int g_prof_gemm_cat = PC_OTHER;

namespace {
struct Cat {
    std::vector<cudaEvent_t> ev;  // pairs
    uint64_t total = 0;
    double bytes = 0.0, flops = 0.0;
    int used = 0;  // pairs in use
};
std::mutex g_mu;
int g_max = 0;
int g_suspended = 0;
uint64_t * g_capture_counts = nullptr;
Cat g_cat[PC_COUNT];
}  // namespace

bool prof_enabled() { return g_max > 0; }

void prof_capture_counts(uint64_t * counts) {
    std::lock_guard<std::mutex> lk(g_mu);
    g_capture_counts = counts;
}

void prof_add_totals(const uint64_t * counts) {
    if (g_max <= 0) return;
    std::lock_guard<std::mutex> lk(g_mu);
    for (int k = 0; k < PC_COUNT; ++k) g_cat[k].total += counts[k];
}

void prof_enable(int max_samples_per_cat) {
    std::lock_guard<std::mutex> lk(g_mu);
    g_max = max_samples_per_cat;
}

void prof_suspend(bool on) {
    std::lock_guard<std::mutex> lk(g_mu);
    g_suspended += on ? 1 : -1;
}

void prof_reset() {
    std::lock_guard<std::mutex> lk(g_mu);
    for (auto & c : g_cat) {
        c.total = 0;
        c.used = 0;
        c.bytes = c.flops = 0.0;
    }
}

ProfScope::ProfScope(int cat_, cudaStream_t s_, double bytes, double flops) : cat(cat_), s(s_), slot(-1) {
    if (g_capture_counts != nullptr) {
        std::lock_guard<std::mutex> lk(g_mu);
        if (g_capture_counts != nullptr && g_suspended > 0) g_capture_counts[cat]++;
    }
    if (g_max <= 0) return;
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_suspended > 0) return;
    Cat & c = g_cat[cat];
    c.total++;
    // sample every launch until the cap, then every 16th launch into a rotating window is not needed: a prefix of
    // g_max launches of each category is representative for steady-state loops
    if (c.used >= g_max) return;
    if ((int) c.ev.size() < 2 * (c.used + 1)) {
        cudaEvent_t a, b;
        if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) return;
        c.ev.push_back(a);
        c.ev.push_back(b);
    }
    slot = c.used++;
    c.bytes += bytes;
    c.flops += flops;
    cudaEventRecord(c.ev[2 * slot], s);
}

ProfScope::~ProfScope() {
    if (slot < 0) return;
    std::lock_guard<std::mutex> lk(g_mu);
    cudaEventRecord(g_cat[cat].ev[2 * slot + 1], s);
}

void prof_collect(double * ms_sum, uint64_t * n_sampled, uint64_t * n_total, double * bytes_sampled, double * flops_sampled) {
    cudaDeviceSynchronize();
    std::lock_guard<std::mutex> lk(g_mu);
    for (int k = 0; k < PC_COUNT; ++k) {
        Cat & c = g_cat[k];
        double sum = 0.0;
        int n = 0;
        for (int i = 0; i < c.used; ++i) {
            float ms = 0.0f;
            if (cudaEventElapsedTime(&ms, c.ev[2 * i], c.ev[2 * i + 1]) == cudaSuccess) {
                sum += ms;
                ++n;
            }
        }
        ms_sum[k] = sum;
        n_sampled[k] = (uint64_t) n;
        n_total[k] = c.total;
        bytes_sampled[k] = c.bytes;
        flops_sampled[k] = c.flops;
    }
}

}  // namespace wb
