// gemm_sm100.cuh -- internal interface of the tcgen05 GEMM (gemm_sm100.cu).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda.h>

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
};

// D = A[M,K] * W[N,K]^T with the epilogue above.  K % 64 == 0, N % 8 == 0.
int yv_gemm_launch(const void *A, const void *W, int M, int N, int K, const GemmEpilogue &ep, cudaStream_t s);

// TMA descriptor of a row-major bf16 [rows, cols] matrix: box = box_rows x 64 columns (128 B), 128-byte
// swizzle, zero fill out of bounds -- the K-major (or, read the other way, MN-major) UMMA operand layout.
int yv_tmap_bf16_2d(CUtensorMap *m, const void *ptr, uint64_t rows, uint64_t cols, uint32_t box_rows);
