// In-situ kernel timing: CUDA events recorded on the launching stream around sampled launches, per category.
// Off by default (zero overhead beyond one branch); bench.py switches it on to report the dominant kernel's
// average launch duration inside the timed region.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace wb {

enum ProfCat : int {
    PC_MEL = 0, PC_GEMM_ENC, PC_ENC_SOFTMAX, PC_LN, PC_GEMM_DEC, PC_GEMM_LOGITS, PC_SELF_ATTN, PC_CROSS_ATTN, PC_SAMPLER, PC_EMBED,
    PC_DTW, PC_OTHER, PC_LN_DEC, PC_CALIB, PC_CROSS_STEP, PC_GEMM_STEP_QKV, PC_GEMM_STEP_SQ,
    PC_GEMM_STEP_FC1, PC_GEMM_STEP_FC2, PC_COUNT
};
const char * prof_cat_name(int cat);

void prof_enable(int max_samples_per_cat);  // 0 disables
void prof_reset();
void prof_suspend(bool on);  // no event records while a stream capture is in progress
bool prof_enabled();
// Graph replays: while `counts` (PC_COUNT entries) is set, scopes opened under prof_suspend only count their category there;
// prof_add_totals adds such a tally to the launch totals each time the captured graph is replayed.
void prof_capture_counts(uint64_t * counts);
void prof_add_totals(const uint64_t * counts);
// returns sampled-milliseconds sum, sampled count and total launch count per category (arrays of PC_COUNT)
void prof_collect(double * ms_sum, uint64_t * n_sampled, uint64_t * n_total, double * bytes_sampled, double * flops_sampled);

extern int g_prof_gemm_cat;  // category the next gemm_f16_tc launches are accounted to

struct ProfScope {
    int cat;
    cudaStream_t s;
    int slot;
    // bytes / flops: ALGORITHMIC work of the launches inside this scope (accumulated only when the scope is sampled)
    ProfScope(int cat, cudaStream_t s, double bytes = 0.0, double flops = 0.0);
    ~ProfScope();
};

}  // namespace wb
