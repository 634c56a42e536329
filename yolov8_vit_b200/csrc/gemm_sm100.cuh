// gemm_sm100.cuh -- internal interface of the tcgen05 GEMM (gemm_sm100.cu).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda.h>

// width of one row-statistics slot of the fused residual / folded LayerNorm modes (= the column group of one
// epilogue warp)
constexpr int GEMM_STAT_COLS = 64;

struct GemmEpilogue {
  void *out;                        // [rows, N] bf16 or f32 (row remapped when group_rows > 0)
  const float *bias;                // [N] or null
  const void *residual;             // [rows, N] bf16 (or f32 when res_f32) or null; indexed by the OUTPUT
                                    // row; may alias out (each element is read then written by one thread)
  const float *pos;                 // patch-embed: [group_rows + 1, N] f32 position table (bias folded in)
  int group_rows;                   // patch-embed: patches per crop (196); 0 = identity row map
  bool gelu;
  bool out_f32;
  bool res_f32;
  // LayerNorm folded into the GEMM (mode 2): A holds the RAW bf16 residual stream x, W = g (.) W_orig,
  // D = rstd_m * (acc - mu_m * col_s[n]) + bias[n]  with  bias = b . W_orig^T + bias_orig,  col_s[n] = sum_k W[n,k],
  // (mu_m, rstd_m) from row_stats[m][0..ln_slots) = partial (sum_k x, sum_k x^2) over ln_dim features.
  const float *row_stats = nullptr;
  const float *col_s = nullptr;
  int ln_dim = 0;
  int ln_slots = 0;
  float ln_eps = 0.f;
  // Residual update fused into the producer GEMM (mode 3): x32[m, n] += acc + bias (fp32, in place),
  // out = bf16(x32) and stats_out[m][n / GEMM_STAT_COLS] = (sum, sum of squares) of the new entries of that column
  // block (one slot per block: deterministic, nothing to zero).  null = off.
  float *x32 = nullptr;
  float *stats_out = nullptr;
  int dbg = 0;                      // measurement only (YV_GEMM_DBG): 1 = no global stores, 2 = no epilogue work
};

// D = A[M,K] * W[N,K]^T with the epilogue above.  K % 64 == 0, N % 8 == 0.
int yv_gemm_launch(const void *A, const void *W, int M, int N, int K, const GemmEpilogue &ep, cudaStream_t s);

// TMA descriptor of a row-major bf16 [rows, cols] matrix: box = box_rows x 64 columns (128 B), 128-byte
// swizzle, zero fill out of bounds -- the K-major (or, read the other way, MN-major) UMMA operand layout.
int yv_tmap_bf16_2d(CUtensorMap *m, const void *ptr, uint64_t rows, uint64_t cols, uint32_t box_rows);
