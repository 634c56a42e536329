// gemm_sm100.cuh -- internal interface of the tcgen05 GEMM (gemm_sm100.cu).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda.h>

struct GemmEpilogue {
  void *out;                        // [rows, N] bf16 / f32 / f16 (row remapped when group_rows > 0)
  const float *bias;                // [N] or null
  const void *residual;             // [rows, N] of res_type or null; indexed by the OUTPUT row; may alias out (each
                                    // element is read, then written, by one thread / one warp's TMA box)
  const float *pos;                 // patch-embed: [group_rows + 1, N] f32 position table (bias folded in)
  int group_rows;                   // patch-embed: patches per crop (196); 0 = identity row map
  bool gelu;
  int out_type;                     // 0 bf16, 1 f32, 2 f16
  int res_type;                     // 0 bf16, 1 f32 (LSU epilogue), 2 f16 (fused TMA residual: needs out_type 2, no GELU)
  // LayerNorm folded into this Linear (EPI_TMA_LN): A is the raw fp16 stream, W = fp16(g (.) W_orig),
  // D = rstd_m (acc - mu_m col_s[n]) + bias[n] with bias = b . W_orig^T + bias_orig, col_s[n] = sum_k W[n, k],
  // row_stats[m] = (mu_m, rstd_m) from yv_row_stats_launch.  null = off.
  const float *row_stats = nullptr;
  const float *col_s = nullptr;
  // Fused TMA residual epilogue only: partial LayerNorm statistics of the UPDATED rows, one (sum, M2) pair per row and
  // 64-column group, stats_part[(n / 64) * M + m] as float2 (M2 = sum of squares about the group's own mean); N % 64 == 0.
  // yv_stats_finalize_launch combines the N / 64 groups of a row into (mean, rstd).  null = off.
  float *stats_part = nullptr;
  int dbg = 0;                      // measurement only (YV_GEMM_DBG): 1 = no global stores, 2 = no epilogue work, 8 = no MMAs (operand feed only)
};

// D = A[M,K] * W[N,K]^T with the epilogue above.  K % 64 == 0, N % 8 == 0.
int yv_gemm_launch(const void *A, const void *W, int M, int N, int K, const GemmEpilogue &ep, cudaStream_t s);

// TMA descriptor of a row-major bf16 [rows, cols] matrix: box = box_rows x 64 columns (128 B), 128-byte
// swizzle, zero fill out of bounds -- the K-major (or, read the other way, MN-major) UMMA operand layout.
int yv_tmap_bf16_2d(CUtensorMap *m, const void *ptr, uint64_t rows, uint64_t cols, uint32_t box_rows);
