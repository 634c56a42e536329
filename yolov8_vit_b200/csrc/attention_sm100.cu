// attention_sm100.cu -- K7: softmax(Q K^T / sqrt(dh)) V on tcgen05 / TMEM, fed by TMA.
//
// timm Attention.forward (vit_base_patch{16,8}_224; reached by the reference through
// utils/utils.py:71-72): qkv [n, T, 3, heads, 64] -> per head softmax(q k^T * 64^-0.5) v.
//
// attention_pipe_kernel (below): the kernel for sequences whose whole score row fits twice in tensor memory
// (16 T' <= 224 keys: ViT-B/16, T = 197).  Every other length -- the reference-literal ViT-B/8 with 785 tokens in
// particular -- runs on the flash-style multi-key-block kernel of attention_fa.cu; both are tcgen05 / TMEM / TMA
// kernels (the round-1 mma.sync kernel for long sequences is gone).  attention_fa.cu also has a layout for the
// ViT-B/16 shape (two query tiles in flight); measured 0.43 ms per 1072-crop launch against 0.38 ms for this
// kernel, so this one stays the default there (YV_ATTN_B16=fa selects the other, A/B timing).
#include <cstdlib>
#include "common.cuh"
#include "gemm_sm100.cuh"
#include "sm100_ptx.cuh"
#include "vit_ops.cuh"

using namespace sm100;

namespace {

constexpr int AT_DH = 64;
constexpr int AT_Q_BYTES = 128 * AT_DH * 2;         // 16 KB
constexpr int AT_KV_BYTES = 256 * AT_DH * 2;        // 32 KB reserved each (box is NK rows)

// MN-major B operand (V: keys x 64, 128-byte swizzled rows of dh): 8-key groups are 1024 B apart
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr) {
  const uint32_t lo = ((smem_addr >> 4) & 0x3FFFu) | ((1024u >> 4) << 16);
  const uint32_t hi = (1024u >> 4) | (1u << 14) | (2u << 29);
  return ((uint64_t)hi << 32) | lo;
}
// kind::f16, D=f32, A=bf16 K-major, B=bf16 MN-major
__host__ __device__ constexpr uint32_t umma_idesc_bf16_bmn(int M, int N) {
  return umma_idesc_bf16(M, N) | (1u << 16);
}

__device__ __forceinline__ float ex2_approx(float x) {   // one MUFU.EX2; arguments are <= 0 here
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t *>(&h);
}

// =================================================================================================
// v3: software-pipelined variant for NK <= 224 keys (ViT-B/16: T = 197 -> NK = 208).
//
// One persistent CTA per SM, 24 warps.  A work unit is a (crop, head): K and V are loaded ONCE and
// shared by its NG query tiles; the 256-row Q box covers both tiles.  Tensor memory holds TWO score
// buffers (S0 at [0, NK), S1 at [NK, 2NK)) and a separate O accumulator at [2NK, 2NK + 64):
//   MMA warp      S(t+1) is issued BEFORE it waits for P(t), so by the time the softmax warps finish
//                 tile t the scores of tile t+1 are already complete -- they never wait on an MMA;
//   softmax       16 warps = 4 lane quarters x 4 column quarters (<= 64 columns per thread); the four
//                 threads of a row exchange their maxima through shared memory (one 128-thread named
//                 barrier per lane quarter); partial row sums go to shared memory for the output warps;
//                 P overwrites the head of each thread's own part of S as packed bf16 (A-from-TMEM);
//   output        4 warps drain O after the P V MMA, normalise with the summed partial sums and store,
//                 off the softmax warps' critical path.
// Warps whose 32 query rows are all past T (the tail of the last tile) skip the arithmetic.
// Measured (tools/attn_trace.py, clock64 stamps of CTA 0): the softmax warps run back to back at ~3.7k
// cycles per tile and never wait for S; a variant with dedicated row-max warps moved the max into the
// MMA -> max -> softmax dependency chain and was no faster, and moving half of the exp2 to an FMA-pipe
// polynomial was slower (the MUFU pipe is only 38 % busy).  The binding resource is the TMEM READ port:
// tcgen05.ld moves 64 B/clk per SM, and one tile reads S twice (max pass + exp pass, 2 x 106 KB) plus O
// (32 KB) = 245 KB = 3.8k cycles -- exactly the measured period, independent of the number of softmax
// warps (2/3/4/5/6 column groups: 121/115/118/136/152 us).  Reading S once (scores held in registers
// between the passes) cuts the softmax to 3.1k cycles but needs 96+ registers per thread; with the output
// duty folded into the same warps it spilled and was slower.  A clean single-read version was then built with
// warpgroup register re-allocation (setmaxnreg: TMA/MMA group 40, output group 56, the four softmax groups 96
// registers = the 61 440 registers a 768-thread CTA owns; the inc must sit at the top of the role branch for
// ptxas to honour it, and it can only draw on the CTA's own pool) and the tail-key mask applied once as -inf:
// correct, no spills, 118 us per 268-crop launch -- the same as this two-pass kernel.  So the TMEM read port is
// NOT what bounds the softmax; what remains per 128 x 208 tile and SM sub-partition is 208 MUFU.EX2 warp
// instructions at 8 cycles (1664 cycles) plus the bf16 packs and ~1000 issue slots in a dependent chain
// (max -> smem exchange -> exp -> pack -> tcgen05.st).  The simpler two-pass kernel is kept.
constexpr int AP_THREADS = 24 * 32;              // TMA, MMA, alloc, spare, 16 softmax, 4 output
constexpr int AP_Q_BYTES = 256 * AT_DH * 2;         // 32 KB: both query tiles of a unit
constexpr int AP_STAGE = AP_Q_BYTES + 2 * AT_KV_BYTES;   // 96 KB
constexpr int AP_STAGES = 2;

struct ApBars {
  uint64_t qk_full[AP_STAGES], v_full[AP_STAGES], qk_empty[AP_STAGES], v_empty[AP_STAGES];
  uint64_t s_full[2], p_full[2], o_full, o_free;
  uint32_t tmem_slot, pad;
  float xmax[2][4][128];      // [processed-tile parity][column quarter][row]
  float xsum[4][4][128];      // [tile & 3][column quarter][row]: partial row sums for the output warps
};
constexpr int AP_SMEM = AP_STAGES * AP_STAGE + (int)sizeof(ApBars) + 1024;

// first 16-key chunk of column part c when nch chunks are split over `parts` parts as evenly as possible
__device__ __forceinline__ int ap_cstart(int c, int nch, int parts) {
  const int base = nch / parts, rem = nch - base * parts;
  return c * base + (c < rem ? c : rem);
}

// STRIP = false: the 16 softmax warps work on ONE tile (4 lane quarters x 4 column quarters, row maximum exchanged through
// shared memory + a named barrier).  The four warps of a lane quarter are the four warps of one SM sub-partition, so a
// sub-partition's warps are phase-locked: all in the max pass (MUFU idle), then all in the exp pass.
// STRIP = true: 8 softmax warps, each owns a whole 32-row x NK-column strip of a tile -- no exchange, no barrier -- and
// the two warps of a sub-partition work on DIFFERENT tiles (even / odd, both score buffers are in flight anyway), so one
// warp's max pass runs under the other's exp pass.
// General form: SLOTS tiles in flight (1 or 2) x PARTS column parts per strip: <1, 4> is the first form, <2, 1> the strip
// form, <2, 2> splits each strip between two warps that exchange the row maximum (a 64-thread barrier).
template <int SLOTS, int PARTS>
__global__ void __launch_bounds__((8 + 4 * SLOTS * PARTS) * 32, 1)
attention_pipe_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                      __nv_bfloat16 *__restrict__ out, int n_units, int T, int heads, float scale_log2e,
                      unsigned long long *__restrict__ trace) {
  constexpr bool STRIP = PARTS == 1;                // a strip per warp: no exchange
  constexpr int NSW = 4 * SLOTS * PARTS;            // softmax warps
  constexpr int OUTW = 4 + NSW;                     // first output warp
  // trace (debug, normally null): CTA 0 stamps clock64() per tile and role, 8 slots per tile
#define AP_STAMP(t, e) do { if (trace && blockIdx.x == 0 && (t) < 64) trace[(t) * 8 + (e)] = clock64(); } while (0)
  extern __shared__ __align__(1024) uint8_t ap_raw[];
  uint8_t *smem = ap_raw;
  if ((smem_u32(smem) & 1023u) != 0u) __trap();
  ApBars *bar = reinterpret_cast<ApBars *>(smem + AP_STAGES * AP_STAGE);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int NK = (T + 15) & ~15;                    // <= 224
  const int NCH = NK >> 4;
  const int NG = (T + 127) >> 7;                    // 1 or 2 query tiles per unit
  const int ng_sh = NG - 1;                         // NG is 1 or 2
  const int D = heads * AT_DH;
  const uint32_t kv_bytes = (uint32_t)NK * AT_DH * 2;
  const uint32_t q_bytes = (uint32_t)NG * 128 * AT_DH * 2;
  const int my_units = (n_units - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int NT = my_units * NG;                     // tiles this CTA walks

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmQ);
    prefetch_tmap(&tmKV);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < AP_STAGES; ++i) {
      mbar_init(&bar->qk_full[i], 1);
      mbar_init(&bar->v_full[i], 1);
      mbar_init(&bar->qk_empty[i], 1);
      mbar_init(&bar->v_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bar->s_full[i], 1);
      mbar_init(&bar->p_full[i], 4 * PARTS);        // the softmax warps of a tile
    }
    mbar_init(&bar->o_full, 1);
    mbar_init(&bar->o_free, 4);                     // the 4 output warps
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(&bar->tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bar->tmem_slot;
  const uint32_t o_col = 2u * (uint32_t)NK;         // O accumulator [o_col, o_col + 64)


  if (warp == 0) {
    // ===================== TMA producer: one unit (Q tiles + K + V) per stage =====================
    if (elect_one()) {
      for (int u = 0; u < my_units; ++u) {
        const int unit = (int)blockIdx.x + u * (int)gridDim.x;
        const int crop = unit / heads, head = unit - crop * heads;
        const int row0 = crop * T;
        const int st = u & 1;
        const uint32_t ph = (uint32_t)(u >> 1) & 1u;
        uint8_t *sQ = smem + st * AP_STAGE, *sK = sQ + AP_Q_BYTES, *sV = sK + AT_KV_BYTES;
        mbar_wait(&bar->qk_empty[st], ph ^ 1u);         // S MMAs of unit u-2 retired
        mbar_arrive_expect_tx(&bar->qk_full[st], q_bytes + kv_bytes);
        tma_load_2d(sQ, &tmQ, &bar->qk_full[st], head * AT_DH, row0);
        tma_load_2d(sK, &tmKV, &bar->qk_full[st], D + head * AT_DH, row0);
        mbar_wait(&bar->v_empty[st], ph ^ 1u);          // P V MMAs of unit u-2 retired
        mbar_arrive_expect_tx(&bar->v_full[st], kv_bytes);
        tma_load_2d(sV, &tmKV, &bar->v_full[st], 2 * D + head * AT_DH, row0);
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer: S runs one tile ahead of P V =====================
    if (elect_one()) {
      const uint32_t idesc_s = umma_idesc_bf16(128, NK);
      const uint32_t idesc_o = umma_idesc_bf16_bmn(128, AT_DH);
      auto issue_s = [&](int t) {
        const int u = t >> ng_sh, g = t & (NG - 1), st = u & 1;
        if (g == 0) {
          mbar_wait(&bar->qk_full[st], (uint32_t)(u >> 1) & 1u);
          tc_fence_after();
        }
        const uint32_t aQ = smem_u32(smem + st * AP_STAGE) + (uint32_t)g * AT_Q_BYTES;
        const uint32_t aK = smem_u32(smem + st * AP_STAGE + AP_Q_BYTES);
        const uint64_t dq = umma_desc_k_sw128(aQ), dk = umma_desc_k_sw128(aK);
        const uint32_t d = tmem_base + (uint32_t)(t & 1) * (uint32_t)NK;
#pragma unroll
        for (int k = 0; k < AT_DH / 16; ++k)
          umma_bf16(d, dq + (uint64_t)(k * 2), dk + (uint64_t)(k * 2), idesc_s, (uint32_t)(k != 0));
        umma_commit(&bar->s_full[t & 1]);
        if (g == NG - 1) umma_commit(&bar->qk_empty[st]);
      };
      if (NT > 0) issue_s(0);
      for (int t = 0; t < NT; ++t) {
        // buffer (t+1)&1 was last read by P V(t-1), issued earlier by this thread: in order
        if (t + 1 < NT) issue_s(t + 1);
        const int u = t >> ng_sh, g = t & (NG - 1), st = u & 1;
        if (g == 0) mbar_wait(&bar->v_full[st], (uint32_t)(u >> 1) & 1u);
        mbar_wait(&bar->p_full[t & 1], (uint32_t)(t >> 1) & 1u);
        AP_STAMP(t, 0);
        if (t > 0) mbar_wait(&bar->o_free, (uint32_t)(t - 1) & 1u);   // O(t-1) drained
        tc_fence_after();
        const uint32_t aV = smem_u32(smem + st * AP_STAGE + AP_Q_BYTES + AT_KV_BYTES);
        const uint32_t pbase = tmem_base + (uint32_t)(t & 1) * (uint32_t)NK;
        {
          uint64_t dv = umma_desc_mn_sw128(aV);     // +16 keys = +2048 B = +128 in the (addr >> 4) field
          uint32_t acc = 0;
#pragma unroll
#pragma unroll
          for (int c = 0; c < PARTS; ++c) {         // part c keeps its packed P at the head of its own score columns
            const int s0 = ap_cstart(c, NCH, PARTS), s1 = ap_cstart(c + 1, NCH, PARTS);
            uint32_t pc = pbase + 16u * (uint32_t)s0;
            for (int k = s0; k < s1; ++k, pc += 8, dv += 128) {
              umma_bf16_ts(tmem_base + o_col, pc, dv, idesc_o, acc);
              acc = 1;
            }
          }
        }
        umma_commit(&bar->o_full);
        if (g == NG - 1) umma_commit(&bar->v_empty[st]);
        AP_STAMP(t, 1);
      }
    }
  }

  else if (warp >= 4 && warp < OUTW) {
    // ===================== softmax: 4 lane quarters x (4 column quarters | 2 tile parities) =====================
    const int q = warp & 3, cq = ((warp - 4) >> 2) % PARTS;        // lane quarter, column part
    const int slot = ((warp - 4) >> 2) / PARTS;     // SLOTS == 2: this warp takes the tiles of this parity
    const int trow = q * 32 + lane;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
    const int k_lo = ap_cstart(cq, NCH, PARTS), k_hi = ap_cstart(cq + 1, NCH, PARTS);
    const uint64_t sc2 = pk2(scale_log2e, scale_log2e);
    float *xsum = &bar->xsum[0][cq][trow];
    int np = 0;                                     // tiles this warp has processed (not skipped)
    for (int t = slot; t < NT; t += SLOTS) {
      const int g = t & (NG - 1);
      const int b = t & 1;
      const uint32_t sbuf = lane_addr + (uint32_t)b * (uint32_t)NK;
      mbar_wait(&bar->s_full[b], (uint32_t)(t >> 1) & 1u);    // (also keeps skipping warps in step)
      if (warp == 4 && lane == 0) AP_STAMP(t, 4);
      if (g * 128 + q * 32 >= T) {                  // every row of this warp is padding: nothing to compute
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar->p_full[b]);
        continue;
      }
      tc_fence_after();
      float *xm = &bar->xmax[np & 1][0][trow];      // the 4 warps of a lane quarter skip the same tiles
      ++np;
      // pass 1: maximum over this thread's columns (keys >= T masked)
      float mx = -INFINITY;
      if (STRIP) {
        // the whole row in one thread (13 chunks): two register sets, the next pair of chunks is requested before
        // the current pair is reduced, so a trip costs one reduction, not one tensor-memory round trip
        uint32_t a[16], c[16], a2[16], c2[16];
        auto reduce = [&](int k, const uint32_t (&x)[16], const uint32_t (&y)[16]) {
          if (16 * k + 32 <= T) {
            float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
            for (int j = 0; j < 16; j += 2) {
              m0 = fmaxf(m0, fmaxf(__uint_as_float(x[j]), __uint_as_float(y[j])));
              m1 = fmaxf(m1, fmaxf(__uint_as_float(x[j + 1]), __uint_as_float(y[j + 1])));
            }
            mx = fmaxf(mx, fmaxf(m0, m1));
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              if (16 * k + j < T) mx = fmaxf(mx, __uint_as_float(x[j]));
              if (k + 1 < k_hi && 16 * k + 16 + j < T) mx = fmaxf(mx, __uint_as_float(y[j]));
            }
          }
        };
        tmem_ld_32x16(sbuf, a);
        if (1 < k_hi) tmem_ld_32x16(sbuf + 16, c);
        for (int k = 0; k < k_hi; k += 4) {
          tmem_ld_wait();
          if (k + 2 < k_hi) tmem_ld_32x16(sbuf + 16 * (k + 2), a2);
          if (k + 3 < k_hi) tmem_ld_32x16(sbuf + 16 * (k + 3), c2);
          reduce(k, a, c);
          if (k + 2 < k_hi) {
            tmem_ld_wait();
            if (k + 4 < k_hi) tmem_ld_32x16(sbuf + 16 * (k + 4), a);
            if (k + 5 < k_hi) tmem_ld_32x16(sbuf + 16 * (k + 5), c);
            reduce(k + 2, a2, c2);
          }
        }
      } else
      for (int k = k_lo; k < k_hi; k += 2) {
        uint32_t a[16], c[16];
        const bool two = k + 1 < k_hi;
        tmem_ld_32x16(sbuf + 16 * k, a);
        if (two) tmem_ld_32x16(sbuf + 16 * k + 16, c);
        tmem_ld_wait();
        if (two && 16 * k + 32 <= T) {
#pragma unroll
          for (int j = 0; j < 16; ++j) mx = fmaxf(mx, fmaxf(__uint_as_float(a[j]), __uint_as_float(c[j])));
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (16 * k + j < T) mx = fmaxf(mx, __uint_as_float(a[j]));
          if (two) {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (16 * k + 16 + j < T) mx = fmaxf(mx, __uint_as_float(c[j]));
          }
        }
      }
      if (PARTS > 1) {                              // row maximum over the column parts of this strip
        float *xp = xm + slot * PARTS * 128;
        xp[cq * 128] = mx;
        asm volatile("bar.sync %0, %1;" ::"r"(1 + slot * 4 + q), "n"(32 * PARTS) : "memory");
        mx = xp[0];
#pragma unroll
        for (int i = 1; i < PARTS; ++i) mx = fmaxf(mx, xp[i * 128]);
      }
      const float mneg = -mx * scale_log2e;
      const uint64_t mn2 = pk2(mneg, mneg);
      // pass 2: p = exp2(s * scale - max * scale); partial row sum; P -> TMEM as packed bf16 pairs behind
      // this thread's own read pointer
      uint64_t sum2 = pk2(0.f, 0.f);
      {
      uint32_t r[16];
      if (k_lo < k_hi) tmem_ld_32x16(sbuf + 16 * k_lo, r);
      for (int k = k_lo; k < k_hi; ++k) {
        tmem_ld_wait();
        float p[16];
#pragma unroll
        for (int j = 0; j < 16; j += 2) {
          float x0, x1;
          unpk2(fma2(pk2(__uint_as_float(r[j]), __uint_as_float(r[j + 1])), sc2, mn2), x0, x1);
          p[j] = ex2_approx(x0);
          p[j + 1] = ex2_approx(x1);
        }
        if (k + 1 < k_hi) tmem_ld_32x16(sbuf + 16 * (k + 1), r);
        if (16 * k + 16 > T) {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (16 * k + j >= T) p[j] = 0.f;
        }
#pragma unroll
        for (int j = 0; j < 16; j += 4) sum2 = add2(sum2, add2(pk2(p[j], p[j + 1]), pk2(p[j + 2], p[j + 3])));
        uint32_t w[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) w[j] = pack_bf16x2(p[2 * j], p[2 * j + 1]);
        tmem_st_32x8(sbuf + 16 * k_lo + 8 * (k - k_lo), w);
      }
      }
      float sum, sum_hi;
      unpk2(sum2, sum, sum_hi);
      xsum[(t & 3) * 512] = sum + sum_hi;           // partial row sum for the output warps
      tmem_st_wait();
      tc_fence_before();                            // P stores and S loads precede the MMA that reads P
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar->p_full[b]);
      if (warp == 4 && lane == 0) AP_STAMP(t, 5);
    }
  } else if (warp >= OUTW) {
    // ===================== output: drain O, normalise, store =====================
    const int q = warp & 3;
    const int trow = q * 32 + lane;
    const uint32_t o_addr = tmem_base + ((uint32_t)(q * 32) << 16) + o_col;
    for (int t = 0; t < NT; ++t) {
      const int u = t >> ng_sh, g = t & (NG - 1);
      const int unit = (int)blockIdx.x + u * (int)gridDim.x;
      const int crop = unit / heads, head = unit - crop * heads;
      const int tok = g * 128 + trow;
      mbar_wait_sleep(&bar->o_full, (uint32_t)t & 1u);
      tc_fence_after();
      if (warp == OUTW && lane == 0) AP_STAMP(t, 6);
      const bool live = g * 128 + q * 32 < T;       // warp-uniform
      float inv = 0.f;
      if (live) {
        const float *xs = &bar->xsum[t & 3][0][trow];
        float tot = xs[0];
#pragma unroll
        for (int i = 1; i < PARTS; ++i) tot += xs[i * 128];
        inv = 1.0f / tot;
      }
      auto pack8 = [&](const uint32_t *v) {
        uint4 w;
        w.x = pack_bf16x2(__uint_as_float(v[0]) * inv, __uint_as_float(v[1]) * inv);
        w.y = pack_bf16x2(__uint_as_float(v[2]) * inv, __uint_as_float(v[3]) * inv);
        w.z = pack_bf16x2(__uint_as_float(v[4]) * inv, __uint_as_float(v[5]) * inv);
        w.w = pack_bf16x2(__uint_as_float(v[6]) * inv, __uint_as_float(v[7]) * inv);
        return w;
      };
      uint4 h0[4], h1[4];
      if (live) {
        uint32_t o[32];
        tmem_ld_32x32(o_addr, o);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 4; ++j) h0[j] = pack8(o + 8 * j);
        tmem_ld_32x32(o_addr + 32, o);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 4; ++j) h1[j] = pack8(o + 8 * j);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar->o_free);     // the accumulator may be overwritten by P V(t+1)
      if (warp == OUTW && lane == 0) AP_STAMP(t, 7);
      if (live && tok < T) {
        uint4 *dst = reinterpret_cast<uint4 *>(out + ((size_t)crop * T + tok) * D + (size_t)head * AT_DH);
#pragma unroll
        for (int j = 0; j < 4; ++j) dst[j] = h0[j];
#pragma unroll
        for (int j = 0; j < 4; ++j) dst[4 + j] = h1[j];
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}


}  // namespace

// debug: per-tile clock stamps of CTA 0 (tools/attn_trace.py); off unless yv_debug_attn_trace(1, ...) was called
static unsigned long long *g_trace = nullptr;

YV_EXPORT int yv_debug_attn_trace(int enable, unsigned long long *host_out, int n) {
  if (enable && !g_trace) {
    YV_CUDA(cudaMalloc(&g_trace, 64 * 8 * sizeof(unsigned long long)));
    YV_CUDA(cudaMemset(g_trace, 0, 64 * 8 * sizeof(unsigned long long)));
  }
  if (host_out && g_trace) {
    YV_CUDA(cudaDeviceSynchronize());
    YV_CUDA(cudaMemcpy(host_out, g_trace, (size_t)(n < 512 ? n : 512) * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  }
  if (!enable && g_trace) {
    cudaFree(g_trace);
    g_trace = nullptr;
  }
  return YV_OK;
}

int yv_attention_launch(const __nv_bfloat16 *qkv, __nv_bfloat16 *out, int n, int T, int heads, int dh,
                        cudaStream_t s) {
  YV_REQUIRE(dh == AT_DH, "attention: head dim %d unsupported (need %d)", dh, AT_DH);
  YV_REQUIRE(n >= 1 && heads >= 1 && T >= 1, "attention: bad shape");
  const int NK = (T + 15) & ~15, NG = (T + 127) >> 7;
  const bool pipelined = 2 * NK + AT_DH <= 512 && NG <= 2;      // two score buffers + O fit in tensor memory
  static const bool b16_fa = []() { const char *e = getenv("YV_ATTN_B16"); return e && e[0] == 'f'; }();
  if (!pipelined || b16_fa) return yv_attention_fa_launch(qkv, out, n, T, heads, dh, s);
  YV_REQUIRE(((uintptr_t)qkv % 16) == 0 && ((uintptr_t)out % 16) == 0, "attention: operands must be 16-byte aligned");
  const int D = heads * AT_DH;
  const float scale_log2e = (1.0f / sqrtf((float)dh)) * 1.4426950408889634f;
  CUtensorMap tmQ, tmKV;
  int rc = yv_tmap_bf16_2d(&tmQ, qkv, (uint64_t)n * T, (uint64_t)3 * D, (uint32_t)(128 * NG));
  if (rc) return rc;
  rc = yv_tmap_bf16_2d(&tmKV, qkv, (uint64_t)n * T, (uint64_t)3 * D, (uint32_t)NK);
  if (rc) return rc;
  rc = yv_device_once(YV_INIT_ATTN, []() -> int {
    YV_CUDA((cudaFuncSetAttribute(attention_pipe_kernel<1, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, AP_SMEM)));
    YV_CUDA((cudaFuncSetAttribute(attention_pipe_kernel<2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, AP_SMEM)));
    YV_CUDA((cudaFuncSetAttribute(attention_pipe_kernel<2, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, AP_SMEM)));
    return YV_OK;
  });
  if (rc) return rc;
  YvProf prof(YV_K_ATTN, s, 4.0 * n * heads * (double)T * T * dh);
  const int units = n * heads;
  const int grid = units < yv_attention_sm_budget() ? units : yv_attention_sm_budget();
  // softmax form: strip (default; measured 0.324 ms per 1072 crops against 0.383 for quarters and 0.343 for split);
  // YV_ATTN_B16=quarters | split | strip selects one for A/B
  static const int form = []() { const char *e = getenv("YV_ATTN_B16"); return !e ? 1 : e[0] == 'q' ? 0 : (e[0] == 's' && e[1] == 'p') ? 2 : 1; }();
  if (form == 1) attention_pipe_kernel<2, 1><<<grid, (8 + 8) * 32, AP_SMEM, s>>>(tmQ, tmKV, out, units, T, heads, scale_log2e, g_trace);
  else if (form == 2) attention_pipe_kernel<2, 2><<<grid, (8 + 16) * 32, AP_SMEM, s>>>(tmQ, tmKV, out, units, T, heads, scale_log2e, g_trace);
  else attention_pipe_kernel<1, 4><<<grid, (8 + 16) * 32, AP_SMEM, s>>>(tmQ, tmKV, out, units, T, heads, scale_log2e, g_trace);
  YV_LAUNCH_CHECK();
  return YV_OK;
}

YV_EXPORT int yv_attention_bf16(const void *qkv, void *out, int n, int T, int heads, int dh, yv_stream_t stream) {
  YV_REQUIRE(qkv && out, "yv_attention_bf16: null pointer");
  return yv_attention_launch((const __nv_bfloat16 *)qkv, (__nv_bfloat16 *)out, n, T, heads, dh, (cudaStream_t)stream);
}
