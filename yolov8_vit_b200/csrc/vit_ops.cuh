// vit_ops.cuh -- internal interface of the non-GEMM ViT kernels (vit_ops.cu).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

// y[r, :] = LayerNorm(x[r * in_stride : +dim] (+ delta[r * delta_stride : +dim])) * gamma + beta; fp32
// statistics, bf16 output.  x_type: 0 bf16, 1 f32 (the fp32 residual stream), 2 f16 (the fp16 residual stream).
// With `delta` (fp32 stream only) the residual add of the preceding GEMM is fused here (x += delta, written back
// when `writeback`), so that GEMM's epilogue is a plain bf16 store.  dim % 256 == 0, dim <= 1024.
int yv_ln_launch(const void *x, int x_type, size_t in_stride, const __nv_bfloat16 *delta, size_t delta_stride,
                 bool writeback, const float *gamma, const float *beta, __nv_bfloat16 *y, int rows, int dim, float eps,
                 cudaStream_t s);

// Same with a second pending delta: normalises (x + delta) + delta2.  Lets the block keep the fp32 stream
// untouched at norm2 (x + d_attn is formed on the fly, no write-back) and fold both deltas in at the next
// norm1: 16.9 KB instead of 18.4 KB of LayerNorm traffic per token per block, bit-identical sums.
int yv_ln2_launch(const void *x, int x_type, size_t in_stride, const __nv_bfloat16 *delta, const __nv_bfloat16 *delta2,
                  size_t delta_stride, bool writeback, const float *gamma, const float *beta, __nv_bfloat16 *y,
                  int rows, int dim, float eps, cudaStream_t s);

// softmax(Q K^T / sqrt(dh)) V; qkv [n*T, 3*heads*dh] bf16 (timm packing: q | k | v, head-major),
// out [n*T, heads*dh] bf16.  dh == 64.  Both kernels are tcgen05 / TMEM / TMA: attention_sm100.cu for sequences whose
// score row fits twice in tensor memory (<= 224 keys: ViT-B/16), attention_fa.cu (flash-style, key blocks of <= 128,
// online softmax) for every other length (ViT-B/8: 785 tokens).
int yv_attention_launch(const __nv_bfloat16 *qkv, __nv_bfloat16 *out, int n, int T, int heads, int dh,
                        cudaStream_t s);
int yv_attention_fa_launch(const __nv_bfloat16 *qkv, __nv_bfloat16 *out, int n, int T, int heads, int dh,
                           cudaStream_t s);

// LayerNorm folded into the next Linear: per-row (mean, rstd) of the fp16 stream -> stats [rows, 2] f32
// (the stream is only read), and the one-off weight fold wf16 = fp16(g (.) w), colsum, cf = bias + beta . w^T.
int yv_row_stats_launch(const void *x16, size_t in_stride, float *stats, int rows, int dim, float eps, cudaStream_t s);
// (mean, rstd) per row from the partial statistics of the GEMM's fused residual epilogue (GemmEpilogue::stats_part,
// [dim / 64][rows] float2): the statistics pass over the stream disappears.
int yv_stats_finalize_launch(const float *part, int rows, int dim, float eps, float *stats, cudaStream_t s);
int yv_fold_ln_launch(const __nv_bfloat16 *w, const float *bias, const float *g, const float *beta, int N, int K,
                      void *wf16, float *colsum, float *cf, cudaStream_t s);

// x[c*T, :] = cls + pos[0]  (the class-token row of every crop); x is the residual stream, f32 or f16
int yv_cls_rows_launch(void *x, bool x_f16, const float *cls_pos0, int n, int T, int dim, cudaStream_t s);

// Network_Wrapper.fc on the backbone output: logits = W3 relu(W1 relu(h) + b1) + b3; top1 = first argmax
int yv_wrapper_head_launch(const float *hidden, int ld_hidden, const float *w1, const float *b1, const float *w3,
                           const float *b3, int n, int head_out, int hid, int ncls, float *logits, int32_t *top1,
                           cudaStream_t s);
