// tcgen05/TMEM GEMM for the transcription path:  C[b][m][n] = sum_k A[b][m][k] * B[b][n][k]
// A, B: f16, K-major, fetched by TMA (128-byte swizzle); accumulate f32 in TMEM; fused epilogue.
#pragma once

#include "common.cuh"

namespace wb {

enum GemmOutMode : int {
    OUT_ROW       = 0,  // out[b*bs + m*ldo + n]
    OUT_COL       = 1,  // out[b*bs + n*ldo + m]                               (transposed)
    OUT_HEAD      = 2,  // out[b*bs + (n/64)*head_stride + m*64 + n%64]        (per-head [t][64] panels)
    OUT_HEAD_COL  = 3,  // out[b*bs + (n/64)*head_stride + (m/rpp)*panel_stride + (n%64)*ldo + m%rpp]  (per-head [64][t] panels)
};

// One operand: a 4-D view {k (contiguous), phase, rows, batch} over f16 global memory.  `phase` is 1 for
// plain matrices; the strided conv1d view uses 2 phases (even/odd input rows), see GemmTaps.
struct GemmOperand {
    const void * ptr;
    int64_t      rows;          // extent of the row dimension (TMA zero-fills beyond it)
    int64_t      row_stride;    // elements between rows   (bytes must be a multiple of 16)
    int64_t      batch;         // extent of the batch dimension (>= 1); 1 + any launch batch = broadcast
    int64_t      batch_stride;  // elements between batches
    int64_t      phases;        // 0/1: none
    int64_t      phase_stride;  // elements between phases
};

// Implicit-GEMM conv1d(k=3): the reduction runs over `taps` shifted views of A.  Tap i reads A rows
// (phase p[i], row m + r[i]) and B columns [i*k_tap, i*k_tap + k_tap).  taps = 1 is a plain GEMM.
struct GemmTaps {
    int taps;
    int p[3];
    int r[3];
    int k_tap;  // multiple of 64
};

struct GemmEpilogue {
    void *         out;
    int            out_f16;      // 1: store f16, 0: store f32
    int            out_mode;     // GemmOutMode
    int64_t        ldo;
    int64_t        head_stride;
    int64_t        out_batch_stride;   // out offset = (b / batch_div) * out_batch_stride + (b % batch_div) * out_batch_stride2
    int64_t        out_batch_stride2;
    int            batch_div;          // >= 1
    int            rows_per_panel;     // OUT_HEAD_COL: rows m are split as (m / rows_per_panel, m % rows_per_panel); 0 = no split
    int64_t        panel_stride;       //   and the panel index advances the offset by panel_stride
    const float *  bias;         // [N] or null:  v += bias[b*bias_batch_stride + n]
    int64_t        bias_batch_stride;
    float          scale;        // columns in [scale_n0, scale_n1): v *= scale   (after bias)
    int            scale_n0, scale_n1;
    const __half * gelu_table;   // non-null: v = gelu(v) with the reference's f16-table semantics
    const float *  res;          // non-null: v += res[b*res_batch_stride + (m % res_row_mod)*ldr + n]
    int64_t        ldr;
    int64_t        res_batch_stride;
    int            res_row_mod;  // 0: no modulo
};

struct GemmStats {
    uint64_t launches;
};

// Launches the kernel on `stream`. block_n in {0 (auto), 32, 64, 128, 256}.  K is the logical
// reduction length (TMA zero-fills the tail of the last 64-wide k-block).
bool gemm_f16_tc(const GemmOperand & A, const GemmOperand & B, int M, int N, int K, int batch, const GemmEpilogue & epi,
                 int block_n, cudaStream_t stream, const GemmTaps * taps = nullptr);

uint64_t gemm_launch_count();
void gemm_launch_count_add(uint64_t n);  // launches issued through CUDA graphs

bool tmap_encode_2d(CUtensorMap * out, const void * ptr, uint64_t inner, uint64_t rows, uint64_t row_stride, uint32_t box_inner,
                    uint32_t box_rows);

}  // namespace wb
