// vit_ops.cu -- the non-GEMM kernels of the ViT forward (architecture = timm
// vit_base_patch{16,8}_224 as wrapped by utils/utils.py:59-72):
//   K6  LayerNorm (fp32 statistics, one warp per token row, 16-byte vector I/O)
//   K7  softmax attention per (crop, head) on tensor cores (mma.sync m16n8k16 bf16 -> f32,
//       ldmatrix-fed, online softmax in registers)
//   K8  class-token rows, and the Network_Wrapper head (ReLU -> 1000x128 -> ReLU -> 128x5) + argmax
#include <cstdlib>
#include <cuda_fp16.h>
#include "common.cuh"
#include "vit_ops.cuh"

namespace {

// ------------------------------------------------------------------------------------ LayerNorm
constexpr int LN_WARPS = 8;
constexpr int LN_MAXCH = 4;    // dim <= 1024 (chunks of 256 = 32 lanes x 8 elements)

// CS: streaming cache operators on the operands nobody re-reads soon (the fp32 stream and the consumed deltas:
// ld.global.cs / st.global.cs = evict-first), so that y -- the A operand of the GEMM launched next -- is what
// stays in the 126 MB L2.
// XT: element type of x -- 0 bf16, 1 f32 (the fp32 residual stream, with the pending deltas), 2 f16 (the fp16
// residual stream: 2 B in + 2 B out per element, nothing else)
template <int XT, bool CS = false>
__global__ void __launch_bounds__(LN_WARPS * 32)
layernorm_kernel(const void *xv, size_t in_stride, const __nv_bfloat16 *__restrict__ delta,
                 const __nv_bfloat16 *__restrict__ delta2, size_t delta_stride, int writeback, const float *__restrict__ gamma, const float *__restrict__ beta,
                 __nv_bfloat16 *__restrict__ y, int rows, int dim, float eps, int reverse) {
  const int lane = threadIdx.x & 31;
  const int nblk = (rows + LN_WARPS - 1) / LN_WARPS;
  // grid == nblk: one 8-row block per CTA; a smaller (persistent) grid strides over the blocks.
  // reverse: walk the rows from the end -- the rows the preceding GEMM wrote last are still in L2, and the rows
  // written last here (the low ones) are the first the next GEMM reads
  for (int bi = blockIdx.x; bi < nblk; bi += gridDim.x) {
  const int blk = reverse ? nblk - 1 - bi : bi;
  const int row = blk * LN_WARPS + (threadIdx.x >> 5);
  if (row >= rows) continue;
  const int nch = dim >> 8;
  float v[LN_MAXCH][8];
#pragma unroll
  for (int c = 0; c < LN_MAXCH; ++c) {
    if (c < nch) {
      const int e = c * 256 + lane * 8;
      if (XT == 1) {
        const float4 *p = reinterpret_cast<const float4 *>((const float *)xv + (size_t)row * in_stride + e);
        const float4 a = CS ? __ldcs(p) : p[0], b = CS ? __ldcs(p + 1) : p[1];
        v[c][0] = a.x; v[c][1] = a.y; v[c][2] = a.z; v[c][3] = a.w;
        v[c][4] = b.x; v[c][5] = b.y; v[c][6] = b.z; v[c][7] = b.w;
      } else if (XT == 2) {
        const uint4 u = *reinterpret_cast<const uint4 *>((const __half *)xv + (size_t)row * in_stride + e);
        const __half2 *h = reinterpret_cast<const __half2 *>(&u);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float2 f = __half22float2(h[k]);
          v[c][2 * k] = f.x;
          v[c][2 * k + 1] = f.y;
        }
      } else {
        const uint4 u = *reinterpret_cast<const uint4 *>((const __nv_bfloat16 *)xv + (size_t)row * in_stride + e);
        const __nv_bfloat162 *h = reinterpret_cast<const __nv_bfloat162 *>(&u);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float2 f = __bfloat1622float2(h[k]);
          v[c][2 * k] = f.x;
          v[c][2 * k + 1] = f.y;
        }
      }
      if (delta) {     // fused residual add: x += delta (the bf16 output of the preceding proj / fc2 GEMM)
        const uint4 *dp = reinterpret_cast<const uint4 *>(delta + (size_t)row * delta_stride + e);
        const uint4 u = CS ? __ldcs(dp) : *dp;
        const __nv_bfloat162 *h = reinterpret_cast<const __nv_bfloat162 *>(&u);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float2 f = __bfloat1622float2(h[k]);
          v[c][2 * k] += f.x;
          v[c][2 * k + 1] += f.y;
        }
        if (delta2) {    // second pending delta ((x + delta) + delta2, the order of two separate adds)
          const uint4 *dp2 = reinterpret_cast<const uint4 *>(delta2 + (size_t)row * delta_stride + e);
          const uint4 u2 = CS ? __ldcs(dp2) : *dp2;
          const __nv_bfloat162 *h2 = reinterpret_cast<const __nv_bfloat162 *>(&u2);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const float2 f = __bfloat1622float2(h2[k]);
            v[c][2 * k] += f.x;
            v[c][2 * k + 1] += f.y;
          }
        }
        if (XT == 1 && writeback) {
          float4 *p = reinterpret_cast<float4 *>((float *)const_cast<void *>(xv) + (size_t)row * in_stride + e);
          if (CS) {
            __stcs(p, make_float4(v[c][0], v[c][1], v[c][2], v[c][3]));
            __stcs(p + 1, make_float4(v[c][4], v[c][5], v[c][6], v[c][7]));
          } else {
            p[0] = make_float4(v[c][0], v[c][1], v[c][2], v[c][3]);
            p[1] = make_float4(v[c][4], v[c][5], v[c][6], v[c][7]);
          }
        }
      }
    }
  }
  float sum = 0.f;
#pragma unroll
  for (int c = 0; c < LN_MAXCH; ++c)
    if (c < nch)
#pragma unroll
      for (int k = 0; k < 8; ++k) sum += v[c][k];
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
  const float mean = sum / (float)dim;
  float sq = 0.f;
#pragma unroll
  for (int c = 0; c < LN_MAXCH; ++c)
    if (c < nch)
#pragma unroll
      for (int k = 0; k < 8; ++k) { const float d = v[c][k] - mean; sq += d * d; }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, d);
  const float rstd = rsqrtf(sq / (float)dim + eps);
#pragma unroll
  for (int c = 0; c < LN_MAXCH; ++c) {
    if (c < nch) {
      const int e = c * 256 + lane * 8;
      const float4 g0 = *reinterpret_cast<const float4 *>(gamma + e), g1 = *reinterpret_cast<const float4 *>(gamma + e + 4);
      const float4 b0 = *reinterpret_cast<const float4 *>(beta + e), b1 = *reinterpret_cast<const float4 *>(beta + e + 4);
      const float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
      uint4 u;
      __nv_bfloat162 h[4];
#pragma unroll
      for (int k = 0; k < 4; ++k)
        h[k] = __floats2bfloat162_rn((v[c][2 * k] - mean) * rstd * g[2 * k] + bb[2 * k],
                                     (v[c][2 * k + 1] - mean) * rstd * g[2 * k + 1] + bb[2 * k + 1]);
      u.x = *reinterpret_cast<unsigned *>(&h[0]);
      u.y = *reinterpret_cast<unsigned *>(&h[1]);
      u.z = *reinterpret_cast<unsigned *>(&h[2]);
      u.w = *reinterpret_cast<unsigned *>(&h[3]);
      *reinterpret_cast<uint4 *>(y + (size_t)row * dim + e) = u;
    }
  }
  }  // block loop
}

// fp16 residual stream -> bf16 normalised rows, 4 B of traffic per element.  With so few bytes per row (1.5 KB at
// dim 768) one row per warp leaves too little in flight to cover the HBM latency (measured 3.2 TB/s with 48 KB per
// SM, 3.6 TB/s with 72 KB); here a warp owns LN_R rows per iteration, issues all their loads before the first
// reduction and keeps them as RAW fp16 in registers (3 x 16 B per row and lane), converting on the fly in each of the
// three passes (sum, centred sum of squares, normalise), so that 2 CTAs x 8 warps x 6 rows = 144 KB are in flight per SM
// (128 registers; capping the kernel at 80 registers for a third CTA spills).
constexpr int ln_rows_per_warp(int nch) { return nch <= 3 ? 6 : 4; }
template <int NCH>                                  // dim = 256 * NCH (compile-time: the row lives in registers)
__global__ void __launch_bounds__(LN_WARPS * 32, 2)
layernorm_f16_kernel(const __half *__restrict__ x, size_t in_stride, const float *__restrict__ gamma,
                     const float *__restrict__ beta, __nv_bfloat16 *__restrict__ y, int rows, float eps) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int dim = 256 * NCH, LN_R = ln_rows_per_warp(NCH);
  const int ngrp = (rows + LN_WARPS * LN_R - 1) / (LN_WARPS * LN_R);          // groups of LN_WARPS * LN_R rows
  for (int gi = blockIdx.x; gi < ngrp; gi += gridDim.x) {
    const int row0 = (gi * LN_WARPS + warp) * LN_R;
    uint4 raw[LN_R][NCH];
#pragma unroll
    for (int r = 0; r < LN_R; ++r)
#pragma unroll
      for (int c = 0; c < NCH; ++c)
        if (row0 + r < rows)
          raw[r][c] = __ldcs(reinterpret_cast<const uint4 *>(x + (size_t)(row0 + r) * in_stride + c * 256 + lane * 8));
#pragma unroll
    for (int r = 0; r < LN_R; ++r) {
      const int row = row0 + r;
      if (row >= rows) continue;                                               // warp-uniform
      float sum = 0.f;
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        const __half2 *h = reinterpret_cast<const __half2 *>(&raw[r][c]);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float2 f = __half22float2(h[k]);
          sum += f.x + f.y;
        }
      }
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
      const float mean = sum / (float)dim;
      float sq = 0.f;
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        const __half2 *h = reinterpret_cast<const __half2 *>(&raw[r][c]);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float2 f = __half22float2(h[k]);
          const float d0 = f.x - mean, d1 = f.y - mean;
          sq += d0 * d0 + d1 * d1;
        }
      }
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, d);
      const float rstd = rsqrtf(sq / (float)dim + eps);
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        const int e = c * 256 + lane * 8;
        const float4 g0 = __ldg(reinterpret_cast<const float4 *>(gamma + e)), g1 = __ldg(reinterpret_cast<const float4 *>(gamma + e + 4));
        const float4 b0 = __ldg(reinterpret_cast<const float4 *>(beta + e)), b1 = __ldg(reinterpret_cast<const float4 *>(beta + e + 4));
        const float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
        const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
        const __half2 *h = reinterpret_cast<const __half2 *>(&raw[r][c]);
        __nv_bfloat162 hh[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float2 f = __half22float2(h[k]);
          hh[k] = __floats2bfloat162_rn((f.x - mean) * rstd * g[2 * k] + bb[2 * k],
                                        (f.y - mean) * rstd * g[2 * k + 1] + bb[2 * k + 1]);
        }
        *reinterpret_cast<uint4 *>(y + (size_t)row * dim + e) = *reinterpret_cast<const uint4 *>(hh);
      }
    }
  }
}

// Row statistics of the fp16 residual stream for the LayerNorm that is folded into the next Linear (gemm_sm100.cu,
// EPI_TMA_LN): stats[row] = (mean, 1 / sqrt(var + eps)), two-pass on registers like the LayerNorm kernels.  The
// stream is only READ (2 B per element); nothing of its size is written.  Software-pipelined: a warp holds two
// groups of RS_ROWS rows and loads one while it reduces the other (the reduction -- conversions + 10 shuffles per
// row -- is ~40 % of the load time and would otherwise sit between the loads).
constexpr int RS_ROWS = 3;

template <int NCH>
__device__ __forceinline__ void rs_load(uint4 (&raw)[RS_ROWS][NCH], const __half *__restrict__ x, size_t in_stride,
                                        int row0, int rows, int lane) {
#pragma unroll
  for (int r = 0; r < RS_ROWS; ++r)
#pragma unroll
    for (int c = 0; c < NCH; ++c)
      if (row0 + r < rows)
        raw[r][c] = __ldcs(reinterpret_cast<const uint4 *>(x + (size_t)(row0 + r) * in_stride + c * 256 + lane * 8));
}

template <int NCH>
__device__ __forceinline__ void rs_reduce(const uint4 (&raw)[RS_ROWS][NCH], float2 *__restrict__ stats, int row0,
                                          int rows, int lane, float eps) {
  constexpr int dim = 256 * NCH;
  float sum[RS_ROWS], sq[RS_ROWS];
#pragma unroll
  for (int r = 0; r < RS_ROWS; ++r) {
    sum[r] = 0.f;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const __half2 *h = reinterpret_cast<const __half2 *>(&raw[r][c]);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 f = __half22float2(h[k]);
        sum[r] += f.x + f.y;
      }
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1)
#pragma unroll
    for (int r = 0; r < RS_ROWS; ++r) sum[r] += __shfl_xor_sync(0xffffffffu, sum[r], d);
#pragma unroll
  for (int r = 0; r < RS_ROWS; ++r) {
    const float mean = sum[r] * (1.f / (float)dim);
    sum[r] = mean;
    sq[r] = 0.f;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const __half2 *h = reinterpret_cast<const __half2 *>(&raw[r][c]);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 f = __half22float2(h[k]);
        const float d0 = f.x - mean, d1 = f.y - mean;
        sq[r] = fmaf(d0, d0, fmaf(d1, d1, sq[r]));
      }
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1)
#pragma unroll
    for (int r = 0; r < RS_ROWS; ++r) sq[r] += __shfl_xor_sync(0xffffffffu, sq[r], d);
  if (lane < RS_ROWS && row0 + lane < rows) {
    float m = sum[0], q = sq[0];
#pragma unroll
    for (int r = 1; r < RS_ROWS; ++r)
      if (lane == r) { m = sum[r]; q = sq[r]; }
    stats[row0 + lane] = make_float2(m, rsqrtf(q * (1.f / (float)dim) + eps));
  }
}

template <int NCH>
__global__ void __launch_bounds__(LN_WARPS * 32, NCH <= 3 ? 2 : 1)
row_stats_f16_kernel(const __half *__restrict__ x, size_t in_stride, float2 *__restrict__ stats, int rows, float eps) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int ngrp = (rows + LN_WARPS * RS_ROWS - 1) / (LN_WARPS * RS_ROWS);
  const int step = gridDim.x;
  uint4 a[RS_ROWS][NCH], b[RS_ROWS][NCH];
  int gi = blockIdx.x;
  if (gi < ngrp) rs_load<NCH>(a, x, in_stride, (gi * LN_WARPS + warp) * RS_ROWS, rows, lane);
  while (gi < ngrp) {
    const int g1 = gi + step, g2 = g1 + step;
    if (g1 < ngrp) rs_load<NCH>(b, x, in_stride, (g1 * LN_WARPS + warp) * RS_ROWS, rows, lane);
    rs_reduce<NCH>(a, stats, (gi * LN_WARPS + warp) * RS_ROWS, rows, lane, eps);
    if (g1 >= ngrp) break;
    if (g2 < ngrp) rs_load<NCH>(a, x, in_stride, (g2 * LN_WARPS + warp) * RS_ROWS, rows, lane);
    rs_reduce<NCH>(b, stats, (g1 * LN_WARPS + warp) * RS_ROWS, rows, lane, eps);
    gi = g2;
  }
}

// Combines the per-(row, 64-column group) partial statistics written by the GEMM's fused residual epilogue
// (gemm_sm100.cu, stats_part: (sum, M2 about the group's own mean), [groups][rows] float2) into the (mean, rstd) pair
// per row the folded-LayerNorm epilogue reads: mean = sum_g sum_g / dim, M2 = sum_g (M2_g + 64 (mean_g - mean)^2)
// (Chan's parallel update).  One thread per row, coalesced over rows: 8 B x groups read, 8 B written per row.
__global__ void __launch_bounds__(256)
stats_finalize_kernel(const float2 *__restrict__ part, int groups, int rows, float eps, float2 *__restrict__ stats) {
  const int row = blockIdx.x * 256 + threadIdx.x;
  if (row >= rows) return;
  float2 p[16];
  float sum = 0.f;
#pragma unroll
  for (int g = 0; g < 16; ++g)
    if (g < groups) {
      p[g] = __ldcs(part + (size_t)g * rows + row);
      sum += p[g].x;
    }
  const float inv_dim = 1.f / (64.f * (float)groups);
  const float mean = sum * inv_dim;
  float m2 = 0.f;
#pragma unroll
  for (int g = 0; g < 16; ++g)
    if (g < groups) {
      const float d = fmaf(p[g].x, 1.f / 64.f, -mean);
      m2 += fmaf(64.f * d, d, p[g].y);
    }
  stats[row] = make_float2(mean, rsqrtf(m2 * inv_dim + eps));
}

// One-off weight fold of a LayerNorm (g, beta) into the Linear (w, bias) that follows it, one block per output feature:
//   wf[n, k] = fp16(g[k] w[n, k]),  colsum[n] = sum_k wf[n, k] (of the ROUNDED values the tensor cores will see),
//   cf[n] = bias[n] + sum_k beta[k] w[n, k]     so that   Linear(LN(x)) = rstd (x . wf^T - mu colsum) + cf.
__global__ void __launch_bounds__(256)
fold_ln_f16_kernel(const __nv_bfloat16 *__restrict__ w, const float *__restrict__ bias, const float *__restrict__ g,
                   const float *__restrict__ beta, int K, __half *__restrict__ wf, float *__restrict__ colsum,
                   float *__restrict__ cf) {
  const int n = blockIdx.x;
  float s = 0.f, c = 0.f;
  for (int k = threadIdx.x; k < K; k += 256) {
    const float wv = __bfloat162float(w[(size_t)n * K + k]);
    const __half f = __float2half_rn(g[k] * wv);
    wf[(size_t)n * K + k] = f;
    s += __half2float(f);
    c += beta[k] * wv;
  }
  __shared__ float sh[2][8];
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, d);
    c += __shfl_xor_sync(0xffffffffu, c, d);
  }
  if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = s; sh[1][threadIdx.x >> 5] = c; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float ts = 0.f, tc = 0.f;
    for (int i = 0; i < 8; ++i) { ts += sh[0][i]; tc += sh[1][i]; }
    colsum[n] = ts;
    cf[n] = bias[n] + tc;
  }
}

// ------------------------------------------------------------------------------------ cls rows
template <typename XT>
__global__ void cls_rows_kernel(XT *__restrict__ x, const float *__restrict__ cls_pos0, int n, int T, int dim) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * dim) return;
  const int c = i / dim, e = i % dim;
  x[(size_t)c * T * dim + e] = (XT)cls_pos0[e];
}

// ------------------------------------------------------------------------------------ wrapper head
// One CTA per WH_CP crops: relu(h) [head_out] -> hid (<= 256) -> ncls (<= 32); fp32 throughout.
// Each w1 row is read once per CTA (16-byte loads when VEC) and applied to all WH_CP crops, two rows
// in flight per warp; the per-crop summation order does not depend on how crops are grouped.
constexpr int WH_T = 512;
constexpr int WH_CP = 2;
template <bool VEC>
__global__ void __launch_bounds__(WH_T)
wrapper_head_kernel(const float *__restrict__ hidden, int ld, const float *__restrict__ w1,
                    const float *__restrict__ b1, const float *__restrict__ w3, const float *__restrict__ b3,
                    int n, int head_out, int hid, int ncls, float *__restrict__ logits, int32_t *__restrict__ top1) {
  extern __shared__ __align__(16) float sh[];
  const int in_ld = (head_out + 3) & ~3;
  float *s_in = sh;                          // [WH_CP][in_ld]
  float *s_mid = s_in + WH_CP * in_ld;       // [WH_CP][hid]
  float *s_out = s_mid + WH_CP * hid;        // [WH_CP][ncls]
  const int c0 = blockIdx.x * WH_CP, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < WH_CP * in_ld; i += WH_T) {
    const int cc = i / in_ld, e = i - cc * in_ld;
    s_in[i] = (c0 + cc < n && e < head_out) ? fmaxf(hidden[(size_t)(c0 + cc) * ld + e], 0.f) : 0.f;
  }
  __syncthreads();
  for (int j0 = warp * 2; j0 < hid; j0 += (WH_T / 32) * 2) {
    const int j1 = min(j0 + 1, hid - 1);
    float acc[2][WH_CP];
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
      for (int cc = 0; cc < WH_CP; ++cc) acc[r][cc] = 0.f;
    if (VEC) {
      const float4 *wa = reinterpret_cast<const float4 *>(w1 + (size_t)j0 * head_out);
      const float4 *wb = reinterpret_cast<const float4 *>(w1 + (size_t)j1 * head_out);
#pragma unroll 4
      for (int i = lane; i < head_out / 4; i += 32) {
        const float4 a = wa[i], b = wb[i];
#pragma unroll
        for (int cc = 0; cc < WH_CP; ++cc) {
          const float4 x = *reinterpret_cast<const float4 *>(s_in + cc * in_ld + 4 * i);
          acc[0][cc] += a.x * x.x; acc[0][cc] += a.y * x.y; acc[0][cc] += a.z * x.z; acc[0][cc] += a.w * x.w;
          acc[1][cc] += b.x * x.x; acc[1][cc] += b.y * x.y; acc[1][cc] += b.z * x.z; acc[1][cc] += b.w * x.w;
        }
      }
    } else {
      const float *wa = w1 + (size_t)j0 * head_out, *wb = w1 + (size_t)j1 * head_out;
      for (int i = lane; i < head_out; i += 32) {
        const float a = wa[i], b = wb[i];
#pragma unroll
        for (int cc = 0; cc < WH_CP; ++cc) {
          acc[0][cc] += a * s_in[cc * in_ld + i];
          acc[1][cc] += b * s_in[cc * in_ld + i];
        }
      }
    }
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
      for (int cc = 0; cc < WH_CP; ++cc) {
        float v = acc[r][cc];
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
        const int j = r ? j0 + 1 : j0;
        if (lane == 0 && j < hid) s_mid[cc * hid + j] = fmaxf(v + b1[j], 0.f);
      }
  }
  __syncthreads();
  for (int q = warp; q < WH_CP * ncls; q += WH_T / 32) {
    const int cc = q / ncls, j = q - cc * ncls;
    const float *w = w3 + (size_t)j * hid;
    float acc = 0.f;
    for (int i = lane; i < hid; i += 32) acc += w[i] * s_mid[cc * hid + i];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if (lane == 0) {
      acc += b3[j];
      s_out[cc * ncls + j] = acc;
      if (c0 + cc < n) logits[(size_t)(c0 + cc) * ncls + j] = acc;
    }
  }
  __syncthreads();
  if (tid < WH_CP && c0 + tid < n) {
    const float *o = s_out + tid * ncls;
    int best = 0;
    for (int j = 1; j < ncls; ++j)
      if (o[j] > o[best]) best = j;
    top1[c0 + tid] = best;
  }
}

}  // namespace

int yv_ln_launch(const void *x, int x_type, size_t in_stride, const __nv_bfloat16 *delta, size_t delta_stride,
                 bool writeback, const float *gamma, const float *beta, __nv_bfloat16 *y, int rows, int dim, float eps,
                 cudaStream_t s) {
  return yv_ln2_launch(x, x_type, in_stride, delta, nullptr, delta_stride, writeback, gamma, beta, y, rows, dim, eps, s);
}

int yv_ln2_launch(const void *x, int x_type, size_t in_stride, const __nv_bfloat16 *delta, const __nv_bfloat16 *delta2,
                  size_t delta_stride,
                 bool writeback, const float *gamma, const float *beta, __nv_bfloat16 *y, int rows, int dim, float eps,
                 cudaStream_t s) {
  YV_REQUIRE(rows >= 1 && dim >= 256 && dim % 256 == 0 && dim <= 256 * LN_MAXCH, "layernorm: dim=%d must be a multiple of 256, <= %d", dim, 256 * LN_MAXCH);
  YV_REQUIRE(x && gamma && beta && y, "layernorm: null pointer");
  YV_REQUIRE(x_type >= 0 && x_type <= 2, "layernorm: x_type %d (0 bf16, 1 f32, 2 f16)", x_type);
  YV_REQUIRE(in_stride % 8 == 0, "layernorm: row stride must be a multiple of 8 elements");
  YV_REQUIRE(!delta2 || delta, "layernorm: delta2 needs delta");
  YV_REQUIRE(!delta || x_type == 1, "layernorm: deltas belong to the fp32 residual stream");
  const bool x_f32 = x_type == 1;
  YvProf prof(YV_K_LN, s, (double)rows * dim * ((x_f32 ? 6.0 : 4.0) + (delta ? 2.0 : 0.0) + (delta2 ? 2.0 : 0.0) + (delta && writeback ? 4.0 : 0.0)));
  static const bool cs = [] { const char *e = getenv("YV_LN_STREAM"); return e && e[0] == '1'; }();   // A/B switches
  static const int rev = [] { const char *e = getenv("YV_LN_REVERSE"); return (e && e[0] == '1') ? 1 : 0; }();
  // persistent grid: 4 CTAs per SM (= the occupancy at 60 registers) stride over the 8-row blocks; measured 7.93 -> 7.45 ms
  // per step against one CTA per block (26 398 CTAs per launch), profiles/r01_ln_ab.json.  YV_LN_PERSIST=0 restores that.
  static const int persist = [] { const char *e = getenv("YV_LN_PERSIST"); return e ? atoi(e) : 4; }();   // CTAs per SM
  int grid = (rows + LN_WARPS - 1) / LN_WARPS;
  if (persist > 0 && grid > yv_num_sms() * persist) grid = yv_num_sms() * persist;
  if (x_f32 && cs) layernorm_kernel<1, true><<<grid, LN_WARPS * 32, 0, s>>>(x, in_stride, delta, delta2, delta_stride, writeback ? 1 : 0, gamma, beta, y, rows, dim, eps, rev);
  else if (x_f32) layernorm_kernel<1><<<grid, LN_WARPS * 32, 0, s>>>(x, in_stride, delta, delta2, delta_stride, writeback ? 1 : 0, gamma, beta, y, rows, dim, eps, rev);
  else if (x_type == 2) {
    static const int f16_ctas = [] { const char *e = getenv("YV_LN_F16_CTAS"); return e ? atoi(e) : 2; }();    // per SM (128 registers)
    const int rpw = ln_rows_per_warp(dim >> 8);
    int g16 = (rows + LN_WARPS * rpw - 1) / (LN_WARPS * rpw);
    if (f16_ctas > 0 && g16 > yv_num_sms() * f16_ctas) g16 = yv_num_sms() * f16_ctas;
    if (f16_ctas == 0) layernorm_kernel<2><<<grid, LN_WARPS * 32, 0, s>>>(x, in_stride, nullptr, nullptr, 0, 0, gamma, beta, y, rows, dim, eps, rev);
    else if (dim == 768) layernorm_f16_kernel<3><<<g16, LN_WARPS * 32, 0, s>>>((const __half *)x, in_stride, gamma, beta, y, rows, eps);
    else if (dim == 256) layernorm_f16_kernel<1><<<g16, LN_WARPS * 32, 0, s>>>((const __half *)x, in_stride, gamma, beta, y, rows, eps);
    else if (dim == 512) layernorm_f16_kernel<2><<<g16, LN_WARPS * 32, 0, s>>>((const __half *)x, in_stride, gamma, beta, y, rows, eps);
    else layernorm_f16_kernel<4><<<g16, LN_WARPS * 32, 0, s>>>((const __half *)x, in_stride, gamma, beta, y, rows, eps);
  }
  else layernorm_kernel<0><<<grid, LN_WARPS * 32, 0, s>>>(x, in_stride, nullptr, nullptr, 0, 0, gamma, beta, y, rows, dim, eps, rev);
  YV_LAUNCH_CHECK();
  return YV_OK;
}

int yv_row_stats_launch(const void *x16, size_t in_stride, float *stats, int rows, int dim, float eps, cudaStream_t s) {
  YV_REQUIRE(rows >= 1 && dim % 256 == 0 && dim >= 256 && dim <= 1024 && in_stride % 8 == 0, "row_stats: dim=%d unsupported", dim);
  YvProf prof(YV_K_LN, s, (double)rows * dim * 2.0);
  int grid = (rows + LN_WARPS * RS_ROWS - 1) / (LN_WARPS * RS_ROWS);
  static const int per_sm = []() { const char *e = getenv("YV_STATS_CTAS"); int v = e ? atoi(e) : 0; return v > 0 ? v : 2; }();
  if (grid > yv_num_sms() * per_sm) grid = yv_num_sms() * per_sm;
  float2 *st = reinterpret_cast<float2 *>(stats);
  const __half *x = (const __half *)x16;
  if (dim == 768) row_stats_f16_kernel<3><<<grid, LN_WARPS * 32, 0, s>>>(x, in_stride, st, rows, eps);
  else if (dim == 256) row_stats_f16_kernel<1><<<grid, LN_WARPS * 32, 0, s>>>(x, in_stride, st, rows, eps);
  else if (dim == 512) row_stats_f16_kernel<2><<<grid, LN_WARPS * 32, 0, s>>>(x, in_stride, st, rows, eps);
  else row_stats_f16_kernel<4><<<grid, LN_WARPS * 32, 0, s>>>(x, in_stride, st, rows, eps);
  YV_LAUNCH_CHECK();
  return YV_OK;
}

int yv_stats_finalize_launch(const float *part, int rows, int dim, float eps, float *stats, cudaStream_t s) {
  YV_REQUIRE(rows >= 1 && dim % 64 == 0 && dim >= 64 && dim <= 1024, "stats_finalize: dim=%d unsupported", dim);
  YvProf prof(YV_K_LN, s, (double)rows * (dim / 64 + 1) * 8.0);
  stats_finalize_kernel<<<(rows + 255) / 256, 256, 0, s>>>(reinterpret_cast<const float2 *>(part), dim / 64, rows, eps,
                                                           reinterpret_cast<float2 *>(stats));
  YV_LAUNCH_CHECK();
  return YV_OK;
}

int yv_fold_ln_launch(const __nv_bfloat16 *w, const float *bias, const float *g, const float *beta, int N, int K,
                      void *wf16, float *colsum, float *cf, cudaStream_t s) {
  YvProf prof(YV_K_MISC, s, 0.0);
  fold_ln_f16_kernel<<<N, 256, 0, s>>>(w, bias, g, beta, K, (__half *)wf16, colsum, cf);
  YV_LAUNCH_CHECK();
  return YV_OK;
}

int yv_cls_rows_launch(void *x, bool x_f16, const float *cls_pos0, int n, int T, int dim, cudaStream_t s) {
  const int total = n * dim;
  YvProf prof(YV_K_MISC, s, 0.0);
  if (x_f16) cls_rows_kernel<__half><<<(total + 255) / 256, 256, 0, s>>>((__half *)x, cls_pos0, n, T, dim);
  else cls_rows_kernel<float><<<(total + 255) / 256, 256, 0, s>>>((float *)x, cls_pos0, n, T, dim);
  YV_LAUNCH_CHECK();
  return YV_OK;
}

int yv_wrapper_head_launch(const float *hidden, int ld_hidden, const float *w1, const float *b1, const float *w3,
                           const float *b3, int n, int head_out, int hid, int ncls, float *logits, int32_t *top1,
                           cudaStream_t s) {
  YV_REQUIRE(n >= 1 && head_out >= 1 && hid >= 1 && ncls >= 1, "wrapper head: bad shape");
  const size_t smem = (size_t)WH_CP * (((head_out + 3) & ~3) + hid + ncls) * sizeof(float);
  YV_REQUIRE(smem <= 48 * 1024, "wrapper head: %zu bytes of shared memory exceed 48 KB", smem);
  YvProf prof(YV_K_HEAD, s, 0.0);
  const int grid = (n + WH_CP - 1) / WH_CP;
  const bool vec = head_out % 4 == 0 && hid % 4 == 0 && (reinterpret_cast<uintptr_t>(w1) & 15) == 0;
  if (vec) wrapper_head_kernel<true><<<grid, WH_T, smem, s>>>(hidden, ld_hidden, w1, b1, w3, b3, n, head_out, hid, ncls, logits, top1);
  else wrapper_head_kernel<false><<<grid, WH_T, smem, s>>>(hidden, ld_hidden, w1, b1, w3, b3, n, head_out, hid, ncls, logits, top1);
  YV_LAUNCH_CHECK();
  return YV_OK;
}

YV_EXPORT int yv_layernorm_bf16(const void *x, const float *gamma, const float *beta, void *y, int rows, int dim,
                                float eps, yv_stream_t stream) {
  return yv_ln_launch(x, 0, (size_t)dim, nullptr, 0, false, gamma, beta, (__nv_bfloat16 *)y, rows, dim, eps, (cudaStream_t)stream);
}
