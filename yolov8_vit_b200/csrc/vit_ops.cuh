// vit_ops.cuh -- internal interface of the non-GEMM ViT kernels (vit_ops.cu).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

// y[r, :] = LayerNorm(x[r * in_stride : +dim] (+ delta[r * delta_stride : +dim])) * gamma + beta; fp32
// statistics, bf16 output.  x is f32 (x_f32 = true, the residual stream) or bf16.  With `delta` the
// residual add of the preceding GEMM is fused here (x += delta, written back when `writeback`), so the
// GEMM epilogue is a plain bf16 store.  dim % 256 == 0, dim <= 1024.
int yv_ln_launch(const void *x, bool x_f32, size_t in_stride, const __nv_bfloat16 *delta, size_t delta_stride,
                 bool writeback, const float *gamma, const float *beta, __nv_bfloat16 *y, int rows, int dim, float eps,
                 cudaStream_t s);

// Same with a second pending delta: normalises (x + delta) + delta2.  Lets the block keep the fp32 stream
// untouched at norm2 (x + d_attn is formed on the fly, no write-back) and fold both deltas in at the next
// norm1: 16.9 KB instead of 18.4 KB of LayerNorm traffic per token per block, bit-identical sums.
int yv_ln2_launch(const void *x, bool x_f32, size_t in_stride, const __nv_bfloat16 *delta, const __nv_bfloat16 *delta2,
                  size_t delta_stride, bool writeback, const float *gamma, const float *beta, __nv_bfloat16 *y,
                  int rows, int dim, float eps, cudaStream_t s);

// softmax(Q K^T / sqrt(dh)) V; qkv [n*T, 3*heads*dh] bf16 (timm packing: q | k | v, head-major),
// out [n*T, heads*dh] bf16.  dh == 64.  T <= 256: tcgen05 kernels (attention_sm100.cu); longer sequences
// (ViT-B/8) are forwarded to the mma.sync multi-block kernel of attention.cu.
int yv_attention_launch(const __nv_bfloat16 *qkv, __nv_bfloat16 *out, int n, int T, int heads, int dh,
                        cudaStream_t s);
int yv_attention_long_launch(const __nv_bfloat16 *qkv, __nv_bfloat16 *out, int n, int T, int heads, int dh,
                             cudaStream_t s);

// Folded-LayerNorm pipeline: xb = bf16(x32) + one-pass row statistics in slot 0 of
// stats[rows][slots][2]; and the one-off weight fold wf = g (.) w, colsum, cf = bias + beta . w^T.
int yv_row_stats_launch(const float *x32, __nv_bfloat16 *xb, float *stats, int slots, int rows, int dim, cudaStream_t s);
int yv_fold_ln_launch(const __nv_bfloat16 *w, const float *bias, const float *g, const float *beta, int N, int K,
                      __nv_bfloat16 *wf, float *colsum, float *cf, cudaStream_t s);

// x32[c*T, :] = cls + pos[0]  (the class-token row of every crop)
int yv_cls_rows_launch(float *x32, const float *cls_pos0, int n, int T, int dim, cudaStream_t s);

// Network_Wrapper.fc on the backbone output: logits = W3 relu(W1 relu(h) + b1) + b3; top1 = first argmax
int yv_wrapper_head_launch(const float *hidden, int ld_hidden, const float *w1, const float *b1, const float *w3,
                           const float *b3, int n, int head_out, int hid, int ncls, float *logits, int32_t *top1,
                           cudaStream_t s);
