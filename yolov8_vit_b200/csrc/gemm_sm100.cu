// gemm_sm100.cu -- K5: D[M,N] = A[M,K] * W[N,K]^T with fused epilogues, bf16 operands, fp32
// accumulation, hand-written for sm_100a:
//   * operands arrive by TMA (cp.async.bulk.tensor.2d, 128-byte swizzle) into a 6-stage
//     shared-memory ring, completion on mbarriers;
//   * CTA PAIRS (cluster 2x1x1, cta_group::2): ONE elected thread of the leader CTA issues
//     tcgen05.mma kind::f16 256 x 256 x 16 over both SMs; each CTA stages its 128 rows of A and its
//     128 of the 256 rows of W, so a 256 x 256 tile costs 32 KB of L2->SM traffic per CTA per 64-wide
//     k-block instead of 48 KB for a 128 x 256 tile (the 1-CTA version measured L2-feed-bound);
//   * the fp32 accumulator lives in TMEM, double buffered (2 x 256 columns = all 512), so the
//     epilogue of tile i overlaps the MMAs of tile i+1;
//   * 8 epilogue warps pull the accumulator back with tcgen05.ld (32 lanes x 16 columns), transpose
//     it through padded shared memory so that global I/O is row-contiguous, apply bias / tanh-form GELU /
//     residual / position-embedding and store bf16 or fp32;
//   * persistent: one CTA per SM walks the tile list (n fastest, so the A rows of an m-block
//     are served from L2 to its n-tiles while the weights stay L2 resident).
// This is the ViT's only dense contraction: QKV, proj, fc1, fc2,
// patch-embed and the 1000-way head all run through it.
#include "common.cuh"
#include "gemm_sm100.cuh"
#include "sm100_ptx.cuh"

#include <cuda_fp16.h>
#include <cstdio>
#include <cstdlib>
#include <mutex>

using namespace sm100;

namespace {

constexpr int BM = 128;                // accumulator rows per CTA (the pair's MMA is 256 x 256)
constexpr int PM = 2 * BM;             // rows per tile of the CTA pair
constexpr int BN = 256;
constexpr int BNH = BN / 2;            // B rows each CTA of the pair stages
constexpr int BK = 64;                 // 64 bf16 = 128 B = one swizzle atom row
constexpr int MAX_STAGES = 6;
constexpr int UMMA_K = 16;
constexpr int A_BYTES = BM * BK * 2;   // 16 KB
constexpr int B_BYTES = BNH * BK * 2;  // 16 KB
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
// Epilogue forms (template parameter DIRECT):
//   EPI_LSU   TMEM -> registers -> smem transpose -> LSU stores (fp32 / patch-embed / bf16-residual outputs)
//   EPI_TMA   TMEM -> registers -> swizzled smem -> TMA store, bf16 output (qkv, fc1)
//   EPI_TMA_RES  the same with the RESIDUAL ADD fused: the warp's 32 x 64 piece of the fp16 residual stream is
//             fetched by TMA into the staging block while the tile's MMAs still run, added in place and stored
//             back by TMA as fp16 (proj, fc2: x += y W^T + b without a single LSU global access)
//   EPI_TMA_LN   EPI_TMA with the LayerNorm that precedes the Linear FOLDED in: A is the raw fp16 residual stream x, W
//             holds g (.) W_orig as fp16 (kind::f16 with fp16 operands, fp32 accumulation), and the epilogue applies
//             D = rstd_m (acc - mu_m colsum_n) + cfold_n with (mu, rstd) per row from a statistics kernel that only
//             READS the stream (2 B per element instead of the 4 B of a LayerNorm kernel that also writes y)
constexpr int EPI_LSU = 0, EPI_TMA = 1, EPI_TMA_RES = 2, EPI_TMA_LN = 3;
// smem ring depth: the residual form needs 4 KB of staging per epilogue warp (both 32 x 32 boxes of the tile
// resident at once) instead of 2 KB, paid for with one stage of the ring
constexpr int gemm_stages(int direct) { return direct == EPI_TMA_RES ? 5 : 6; }
constexpr int epi_warp_bytes(int direct) { return direct == EPI_TMA_RES ? 4096 : 2048; }
// Epilogue warps per CTA: 8 (two 128-column groups x 4 TMEM lane quarters) or 16 (four 64-column groups).  16 warps
// hide the TMEM-load / smem-transpose latency of a heavy epilogue (GELU) better, at 96 registers per thread.
constexpr int gemm_threads(int new_) { return (4 + new_) * 32; }
constexpr int TMEM_COLS = 512;         // 2 accumulator stages x BN
constexpr int BAR_BYTES = 512;         // mbarriers (ring 2 x 6, accumulator 2 x 2, residual 16) + TMEM slot
constexpr size_t smem_bytes(int new_, int direct) {
  return (size_t)gemm_stages(direct) * STAGE_BYTES + 1024 /*align slack*/ + BAR_BYTES + (size_t)new_ * epi_warp_bytes(direct);
}
// output / residual element types
constexpr int T_BF16 = 0, T_F32 = 1, T_F16 = 2;

// GELU in the epilogue.  The reference's nn.GELU is the erf form x * Phi(x); the first version evaluated it with
// a degree-8 minimax polynomial for erf (|error| < 6e-5, ~10 issue slots per element), which cost 45 us of the
// 197 us fc1 GEMM (K = 768: the epilogue is the critical path; measured with YV_GEMM_DBG).  This version uses
// the tanh form 0.5 x (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3))) on the MUFU pipe: 5 packed fp32x2 operations
// + 2 MUFU.TANH per PAIR.  |tanh form - erf form| <= 4.8e-4 absolute (rms 1.8e-4 for N(0, 1.5) inputs) and
// tanh.approx adds <= 2^-11 relative -- both several times below the bf16 rounding of the stored value (rms
// 1.7e-3 on the same inputs); the ViT goldens (1e-2 relative on the logits, identical top-1) cover it.
__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void gelu2(float &a, float &b) {
  const uint64_t x = pk2(a, b);
  const uint64_t p = fma2(mul2(x, x), pk2(0.0356774081f, 0.0356774081f), pk2(0.7978845608f, 0.7978845608f));
  float u0, u1;
  unpk2(mul2(p, x), u0, u1);
  const uint64_t t = pk2(tanh_approx(u0), tanh_approx(u1));
  const uint64_t hx = mul2(x, pk2(0.5f, 0.5f));
  unpk2(fma2(hx, t, hx), a, b);
}
__device__ __forceinline__ void gelu4(float4 &v) {
  gelu2(v.x, v.y);
  gelu2(v.z, v.w);
}

__device__ __forceinline__ uint32_t pack_h2(float a, float b) {          // fp16 pair, saturating (no inf in the stream)
  const float lim = 65504.f;
  __half2 h = __floats2half2_rn(fminf(fmaxf(a, -lim), lim), fminf(fmaxf(b, -lim), lim));
  return *reinterpret_cast<uint32_t *>(&h);
}
__device__ __forceinline__ uint32_t pack_b2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t *>(&h);
}

// GELU / OUT select the epilogue arithmetic at compile time; MODE 1 compiles in the rarely used position-table
// (patch embedding) and residual operands of the LSU form.
constexpr int EPI_PLAIN = 0, EPI_EXTRA = 1;
// CS = CTAs per cluster: 2 (one CTA pair) or 4 (two pairs on M-adjacent tiles of the same N-tile: the W operand is
// loaded once and MULTICAST to both pairs -- 25 % less L2 -> shared-memory traffic, which is what the kernel's power
// goes into: the operand feed alone, without a single MMA, pins the GPU at its 1000 W cap (tools/gemm_power.py)).
template <bool GELU, int OUT, int MODE, int NEW, int DIRECT = EPI_LSU, int CS = 2>
__global__ void __cluster_dims__(CS, 1, 1) __launch_bounds__(gemm_threads(NEW), 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmD, const __grid_constant__ CUtensorMap tmR,
                 const GemmEpilogue ep, int M, int N, int K) {
  constexpr int EPI_COLS = BN * 4 / NEW;       // columns per epilogue warp
  constexpr int STAGES = gemm_stages(DIRECT);
  static_assert(MODE == EPI_PLAIN || MODE == EPI_EXTRA, "epilogue mode");
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t *epi_smem = smem + (size_t)STAGES * STAGE_BYTES;        // 1024-byte aligned (TMA swizzle atoms)
  uint64_t *bars = (uint64_t *)(epi_smem + (size_t)NEW * epi_warp_bytes(DIRECT));
  uint64_t *full = bars, *empty = bars + MAX_STAGES;
  uint64_t *tfull = bars + 2 * MAX_STAGES, *tempty = bars + 2 * MAX_STAGES + 2;
  uint64_t *rbar = bars + 2 * MAX_STAGES + 4;                     // [16] residual boxes have landed (one per epilogue warp)
  uint32_t *tmem_slot = (uint32_t *)(bars + 2 * MAX_STAGES + 4 + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_tiles = (M + PM - 1) / PM, n_tiles = (N + BN - 1) / BN;
  const int kblocks = K / BK;
  const uint32_t crank = cluster_ctarank();           // rank in the cluster
  const uint32_t rank = crank & 1u;                   // 0 = leader of the CTA pair (issues the MMAs)
  const uint32_t lead = crank & ~1u;                  // cluster rank of this pair's leader
  const int pr = (int)(crank >> 1);                   // which pair of the cluster
  // work units: tiles (CS == 2) or SUPER-tiles of PC M-adjacent tiles (CS == 4 / 6, pair pr takes M-tile PC (u / n_tiles) + pr;
  // an M-tile past the end is all zero-fill loads and clipped stores)
  constexpr int PC = CS / 2;                          // CTA pairs per cluster (1, 2 or 3)
  const int pair = (int)blockIdx.x / CS;              // cluster index = first work unit
  const int npairs = (int)gridDim.x / CS;
  const int num_units = ((m_tiles + PC - 1) / PC) * n_tiles;
  const uint16_t pair_mask = (uint16_t)(3u << lead);

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    if (DIRECT != EPI_LSU) prefetch_tmap(&tmD);
    if (DIRECT == EPI_TMA_RES) prefetch_tmap(&tmR);
  }
  if (warp == 1 && lane == 0) {
    // full: one arrive per producer of the pair (+ the bytes of both); empty / tfull: one multicast
    // tcgen05.commit; tempty (leader's copy is the one waited on): every epilogue warp of both CTAs
    // (CS == 4: a stage is free when BOTH pairs' MMAs have read it -- pair 0 multicasts W into pair 1's slots too)
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 2); mbar_init(&empty[i], PC); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 2 * NEW); }
    for (int i = 0; i < 16; ++i) mbar_init(&rbar[i], 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc_pair(tmem_slot, TMEM_COLS);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                                  // peer barriers / TMEM exist before any remote signal
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer (one elected lane) =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = pair; tile < num_units; tile += npairs) {
        // this CTA stages its 128 rows of A and its 128 (of 256) rows of W
        const int mt = PC * (tile / n_tiles) + pr;
        const int m0 = mt * PM + (int)rank * BM, n0 = (tile % n_tiles) * BN + (int)rank * BNH;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1u);
          uint8_t *sa = smem + (size_t)stage * STAGE_BYTES, *sb = sa + A_BYTES;
          if (rank == 0) mbar_arrive_expect_tx(&full[stage], 2 * STAGE_BYTES);   // bytes of both CTAs
          tma_load_2d_pair(sa, &tmA, &full[stage], kb * BK, m0);
          if (CS == 2) tma_load_2d_pair(sb, &tmB, &full[stage], kb * BK, n0);
          else if (pr == 0) tma_load_2d_pair_mc(sb, &tmB, &full[stage], kb * BK, n0, (uint16_t)((CS == 6 ? 21u : 5u) << rank));   // -> ranks r, r + 2 (, r + 4)
          if (rank != 0) mbar_arrive_remote(&full[stage], lead);
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1 && rank == 0) {
    // ===================== MMA issuer (one elected lane of the leader CTA) =====================
    if (elect_one()) {
      // operand format bits [7,10) / [10,13): 1 = bf16, 0 = f16 (the folded-LayerNorm form multiplies fp16 x fp16)
      constexpr uint32_t idesc = DIRECT == EPI_TMA_LN ? (umma_idesc_bf16(PM, BN) & ~((7u << 7) | (7u << 10))) : umma_idesc_bf16(PM, BN);
      int stage = 0;
      uint32_t phase = 0, it = 0;
      for (int tile = pair; tile < num_units; tile += npairs, ++it) {
        const uint32_t as = it & 1u, aph = (it >> 1) & 1u;
        mbar_wait(&tempty[as], aph ^ 1u);          // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&full[stage], phase);          // TMA bytes have landed
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + (size_t)stage * STAGE_BYTES);
          const uint64_t da = umma_desc_k_sw128(sa), db = umma_desc_k_sw128(sa + A_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            // advance 16 elements (32 B) along K inside the swizzle atom: +2 in the addr>>4 field
            if (ep.dbg != 8)      // (measurement: dbg 8 = operand feed only, no MMAs -- what the L2 -> smem stream costs in power)
              umma_bf16_pair(d_tmem, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (uint32_t)((kb | k) != 0));
          }
          umma_commit_pair(&empty[stage], CS > 2 ? (uint16_t)((1u << CS) - 1u) : pair_mask);   // smem slots free once these MMAs retire
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
        umma_commit_pair(&tfull[as], pair_mask);   // accumulator complete -> both CTAs' epilogues
      }
    }
  } else if (warp >= 4 && DIRECT != EPI_LSU) {
    // ===================== epilogue, TMA-store form: TMEM -> registers -> smem -> TMA =====================
    // tcgen05.ld hands every thread one accumulator ROW.  Bias (a broadcast load per chunk), GELU and the 16-bit
    // conversion happen on that layout; the thread drops its 32 columns (64 B) into the warp's staging block in
    // the 64-byte-swizzle pattern (conflict-free 16-byte accesses) and one lane hands the 32 x 32 block to the TMA
    // unit (cp.async.bulk.tensor store).  No LSU global accesses at all: LSU stores spend 8 line slots per
    // instruction on 32-byte row pieces, and they are what bounds the K = 768 GEMMs (YV_GEMM_DBG=1).  Rows >= M /
    // columns >= N are clipped by the tensor map.
    // EPI_TMA_RES: the residual stream's two 32 x 32 boxes of this warp are requested by TMA (same swizzle, so
    // every thread finds its own row's 64 B exactly where it will write the sum) right after the previous tile's
    // stores have read the block, i.e. while the MMAs of THIS tile still run; sum = acc + bias + residual is
    // written over the residual in shared memory and stored as fp16.  out may alias residual (x += ...): each
    // box is loaded before it is stored by the same warp and no other warp touches it.
    const int ew = warp - 4;
    const int q = warp & 3;
    const int cg = ew >> 2;
    static_assert(DIRECT == EPI_LSU || EPI_COLS == 64, "TMA-store epilogue: two 32-column halves per warp");
    constexpr bool RES = DIRECT == EPI_TMA_RES;
    uint8_t *stg = epi_smem + ew * epi_warp_bytes(DIRECT);
    const uint32_t sw = (uint32_t)(lane >> 1) & 3u;
    uint32_t it = 0, rit = 0;
    for (int tile = pair; tile < num_units; tile += npairs, ++it) {
      const uint32_t as = it & 1u, aph = (it >> 1) & 1u;
      const int m0 = (PC * (tile / n_tiles) + pr) * PM + (int)rank * BM, n0 = (tile % n_tiles) * BN;
      const int row0 = m0 + q * 32;
      const int col0 = n0 + cg * EPI_COLS;
      const bool box0 = row0 < M && col0 < N, box1 = row0 < M && col0 + 32 < N;      // warp-uniform
      if (RES && box0) {
        if (lane == 0) {
          bulk_wait_read0();                       // the previous tile's stores have read both staging boxes
          mbar_arrive_expect_tx(&rbar[ew], box1 ? 4096u : 2048u);
          tma_load_2d(stg, &tmR, &rbar[ew], col0, row0);
          if (box1) tma_load_2d(stg + 2048, &tmR, &rbar[ew], col0 + 32, row0);
        }
        __syncwarp();
      }
      mbar_wait(&tfull[as], aph);
      tc_fence_after();
      const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN + cg * EPI_COLS;
      uint32_t r[2][16];
      if (ep.dbg != 2) tmem_ld_32x16(tbase, r[0]);
      float st_shift = 0.f;                        // EPI_TMA_RES + stats_part: shifted sums of this thread's 64 new values
      uint64_t st_s1 = 0ull, st_s2 = 0ull;         // (packed fp32 pairs: even / odd columns)
      float ln_nmu = 0.f, ln_rstd = 1.f;           // EPI_TMA_LN: -mean and 1 / std of this thread's row
      if (DIRECT == EPI_TMA_LN && row0 + lane < M) {
        const float2 st = __ldg(reinterpret_cast<const float2 *>(ep.row_stats) + row0 + lane);
        ln_nmu = -st.x;
        ln_rstd = st.y;
      }
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) {
        if (ep.dbg == 2) break;
        const int col = col0 + ch * 16;
        float4 b4[4], s4[DIRECT == EPI_TMA_LN ? 4 : 1];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          b4[j] = (ep.bias && col + 4 * j < N) ? __ldg(reinterpret_cast<const float4 *>(ep.bias + col + 4 * j))
                                               : make_float4(0.f, 0.f, 0.f, 0.f);
          if (DIRECT == EPI_TMA_LN)
            s4[j] = col + 4 * j < N ? __ldg(reinterpret_cast<const float4 *>(ep.col_s + col + 4 * j)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        tmem_ld_wait();
        if (ch + 1 < 4) {
          tmem_ld_32x16(tbase + (ch + 1) * 16, r[(ch + 1) & 1]);
        } else {                                   // the accumulator is in registers: hand it back to the MMA warp now
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (rank == 0) mbar_arrive(&tempty[as]);
            else mbar_arrive_remote(&tempty[as], lead);
          }
        }
        float4 v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t *rr = r[ch & 1] + 4 * j;
          if (DIRECT == EPI_TMA_LN) {                // rstd (acc - mu colsum) + cfold
            v[j] = make_float4(fmaf(ln_rstd, fmaf(ln_nmu, s4[j].x, __uint_as_float(rr[0])), b4[j].x),
                               fmaf(ln_rstd, fmaf(ln_nmu, s4[j].y, __uint_as_float(rr[1])), b4[j].y),
                               fmaf(ln_rstd, fmaf(ln_nmu, s4[j].z, __uint_as_float(rr[2])), b4[j].z),
                               fmaf(ln_rstd, fmaf(ln_nmu, s4[j].w, __uint_as_float(rr[3])), b4[j].w));
          } else {
            v[j] = make_float4(__uint_as_float(rr[0]) + b4[j].x, __uint_as_float(rr[1]) + b4[j].y,
                               __uint_as_float(rr[2]) + b4[j].z, __uint_as_float(rr[3]) + b4[j].w);
          }
          if (GELU) gelu4(v[j]);
        }
        uint8_t *blk = stg + (RES ? (ch >> 1) * 2048 : 0);
        if (RES) {
          if (ch == 0 && box0) mbar_wait(&rbar[ew], rit & 1u);      // residual boxes are in shared memory
        } else if ((ch & 1) == 0) {                // the previous store of this warp has read the block
          if (lane == 0) bulk_wait_read0();
          __syncwarp();
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const uint32_t c16 = (uint32_t)((ch & 1) * 2 + j);             // 16-byte piece of the 64-byte row
          uint4 *slot = reinterpret_cast<uint4 *>(blk + lane * 64 + ((c16 ^ sw) << 4));
          uint4 u;
          if (RES) {
            const uint4 x = *slot;
            const float2 x0 = __half22float2(*reinterpret_cast<const __half2 *>(&x.x));
            const float2 x1 = __half22float2(*reinterpret_cast<const __half2 *>(&x.y));
            const float2 x2 = __half22float2(*reinterpret_cast<const __half2 *>(&x.z));
            const float2 x3 = __half22float2(*reinterpret_cast<const __half2 *>(&x.w));
            const float f0 = v[2 * j].x + x0.x, f1 = v[2 * j].y + x0.y, f2 = v[2 * j].z + x1.x, f3 = v[2 * j].w + x1.y;
            const float f4 = v[2 * j + 1].x + x2.x, f5 = v[2 * j + 1].y + x2.y, f6 = v[2 * j + 1].z + x3.x, f7 = v[2 * j + 1].w + x3.y;
            if (ep.stats_part) {                     // statistics of the next LayerNorm, shifted by the row's first value
              if (ch == 0 && j == 0) st_shift = f0;
              const uint64_t sh = pk2(-st_shift, -st_shift);
              const uint64_t d0 = add2(pk2(f0, f1), sh), d1 = add2(pk2(f2, f3), sh), d2 = add2(pk2(f4, f5), sh), d3 = add2(pk2(f6, f7), sh);
              st_s1 = add2(add2(st_s1, add2(d0, d1)), add2(d2, d3));
              st_s2 = fma2(d0, d0, fma2(d1, d1, fma2(d2, d2, fma2(d3, d3, st_s2))));
            }
            u.x = pack_h2(f0, f1);
            u.y = pack_h2(f2, f3);
            u.z = pack_h2(f4, f5);
            u.w = pack_h2(f6, f7);
          } else {
            u.x = pack_b2(v[2 * j].x, v[2 * j].y);
            u.y = pack_b2(v[2 * j].z, v[2 * j].w);
            u.z = pack_b2(v[2 * j + 1].x, v[2 * j + 1].y);
            u.w = pack_b2(v[2 * j + 1].z, v[2 * j + 1].w);
          }
          *slot = u;
        }
        if (ch & 1) {
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0 && ep.dbg != 1 && ((ch >> 1) ? box1 : box0)) {
            tma_store_2d(&tmD, blk, col0 + (ch >> 1) * 32, row0);
            bulk_commit();
          }
        }
      }
      if (RES && box0) ++rit;
      if (RES && ep.stats_part && box0 && row0 + lane < M) {
        float a0, a1, q0, q1;
        unpk2(st_s1, a0, a1);
        unpk2(st_s2, q0, q1);
        const float s1 = a0 + a1, s2 = q0 + q1;    // sum (x - shift), sum (x - shift)^2 over the 64 columns
        reinterpret_cast<float2 *>(ep.stats_part)[(size_t)(col0 >> 6) * M + row0 + lane] =
            make_float2(fmaf(64.f, st_shift, s1), fmaxf(fmaf(-s1 * (1.f / 64.f), s1, s2), 0.f));
      }
      if (ep.dbg == 2) {                           // (measurement mode: nothing was read)
        if (RES && box0) mbar_wait(&rbar[ew], (rit - 1) & 1u);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (rank == 0) mbar_arrive(&tempty[as]);
          else mbar_arrive_remote(&tempty[as], lead);
        }
      }
    }
    if (lane == 0) bulk_wait0();
  } else if (warp >= 4) {
    // ===================== epilogue, LSU form: TMEM -> registers -> smem transpose -> global =====================
    // tcgen05.ld hands every thread one accumulator ROW (32 lanes x 16 columns per load).  Storing
    // that directly would make each warp instruction touch 32 rows; instead the 32 x 16 fp32 block
    // goes through a per-warp XOR-swizzled smem buffer (conflict-free both ways) and comes back with
    // 4 lanes per row, so every global access is a run of 64 contiguous bytes per row and the
    // per-column operand (bias) lives in registers, fetched BEFORE the wait on the accumulator.
    // The TMEM load of chunk c+1 is issued as soon as chunk c is in smem, hiding its latency.
    const int ew = warp - 4;
    const int q = warp & 3;                        // TMEM lane quarter this warp may access
    const int cg = ew >> 2;                        // column group of the tile (EPI_COLS columns)
    float *stg = reinterpret_cast<float *>(epi_smem) + ew * (32 * 16);
    const int trow = lane >> 2, tseg = lane & 3;   // transposed role: row-in-group, 16-byte segment
    constexpr int NCH = EPI_COLS / 16;             // 16-column chunks per warp
    uint32_t it = 0;
    for (int tile = pair; tile < num_units; tile += npairs, ++it) {
      const uint32_t as = it & 1u, aph = (it >> 1) & 1u;
      const int m0 = (PC * (tile / n_tiles) + pr) * PM + (int)rank * BM, n0 = (tile % n_tiles) * BN;
      const int colw = n0 + cg * EPI_COLS + tseg * 4;       // this thread's first column (chunk 0)
      float4 bias4[NCH];
#pragma unroll
      for (int ch = 0; ch < NCH; ++ch) {
        bias4[ch] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ep.bias && colw + ch * 16 < N) bias4[ch] = __ldg(reinterpret_cast<const float4 *>(ep.bias + colw + ch * 16));
      }
      // the 4 rows this thread stores (per chunk)
      size_t obase[4];
      const float *pos_row[4];
      bool rok[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int row = m0 + q * 32 + i * 8 + trow;
        rok[i] = row < M;
        long long drow = row;
        pos_row[i] = nullptr;
        if (MODE == EPI_EXTRA && ep.group_rows > 0) {          // patch-embed: token row = r + r/196 + 1, + pos[1 + r%196]
          const int g = row / ep.group_rows;
          drow = (long long)row + g + 1;
          pos_row[i] = ep.pos + (size_t)(row - g * ep.group_rows + 1) * N;
        }
        obase[i] = (size_t)drow * N + colw;
      }
      mbar_wait(&tfull[as], aph);
      tc_fence_after();
      const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN + cg * EPI_COLS;
      uint32_t r[16];
      if (ep.dbg != 2) tmem_ld_32x16(tbase, r);
#pragma unroll
      for (int ch = 0; ch < NCH; ++ch) {
        if (ep.dbg == 2) break;
        tmem_ld_wait();
        {
          float4 *srow = reinterpret_cast<float4 *>(stg + lane * 16);
          const int sw = (lane >> 1) & 3;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            srow[j ^ sw] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                       __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
        }
        if (ch + 1 < NCH) tmem_ld_32x16(tbase + (ch + 1) * 16, r);
        __syncwarp();
        if (colw + ch * 16 < N) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int rr = i * 8 + trow;
            float4 v = *reinterpret_cast<const float4 *>(stg + rr * 16 + ((tseg ^ ((rr >> 1) & 3)) << 2));
            if (!rok[i]) continue;
            v.x += bias4[ch].x; v.y += bias4[ch].y; v.z += bias4[ch].z; v.w += bias4[ch].w;
            if (GELU) gelu4(v);
            const size_t o = obase[i] + ch * 16;
            if (MODE == EPI_EXTRA) {
              if (pos_row[i]) {
                const float4 p = __ldg(reinterpret_cast<const float4 *>(pos_row[i] + colw + ch * 16));
                v.x += p.x; v.y += p.y; v.z += p.z; v.w += p.w;
              }
              if (ep.residual) {
                if (ep.res_type == T_F32) {
                  const float4 f = *reinterpret_cast<const float4 *>((const float *)ep.residual + o);
                  v.x += f.x; v.y += f.y; v.z += f.z; v.w += f.w;
                } else {
                  const uint2 u = *reinterpret_cast<const uint2 *>((const __nv_bfloat16 *)ep.residual + o);
                  const float2 f0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&u.x));
                  const float2 f1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&u.y));
                  v.x += f0.x; v.y += f0.y; v.z += f1.x; v.w += f1.y;
                }
              }
            }
            if (ep.dbg == 1 && v.x != 12345.678f) continue;      // keeps the arithmetic alive, drops the stores
            if (OUT == T_F32) {
              *reinterpret_cast<float4 *>((float *)ep.out + o) = v;
            } else if (OUT == T_F16) {
              *reinterpret_cast<uint2 *>((__half *)ep.out + o) = make_uint2(pack_h2(v.x, v.y), pack_h2(v.z, v.w));
            } else {
              *reinterpret_cast<uint2 *>((__nv_bfloat16 *)ep.out + o) = make_uint2(pack_b2(v.x, v.y), pack_b2(v.z, v.w));
            }
          }
        }
        __syncwarp();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {                             // this warp is done with the accumulator (leader's barrier)
        if (rank == 0) mbar_arrive(&tempty[as]);
        else mbar_arrive_remote(&tempty[as], lead);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                                  // the peer may still be reading our smem / signalling us
  if (warp == 2) tmem_dealloc_pair(tmem_base, TMEM_COLS);
}

// ---- host side: tensor maps ----------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                    CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                    CUtensorMapFloatOOBfill);
PFN_encodeTiled g_encode = nullptr;
std::once_flag g_once;
int g_init_status = YV_OK;

// process-wide: the driver entry point of cuTensorMapEncodeTiled
void init_once() {
  void *fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
    yv_set_error("cuTensorMapEncodeTiled not available from the driver (err %d, query %d)", (int)e, (int)qres);
    g_init_status = YV_ECUDA;
    return;
  }
  g_encode = (PFN_encodeTiled)fn;
}

// per device (cudaFuncSetAttribute applies to the current device only): dynamic shared memory of every instance
#define YV_GEMM_INSTANCES(X)                                                                                       \
  X(false, T_BF16, 0, 8, EPI_LSU) X(false, T_F32, 0, 8, EPI_LSU) X(true, T_BF16, 0, 8, EPI_LSU) X(true, T_F32, 0, 8, EPI_LSU)     \
  X(false, T_BF16, 0, 16, EPI_LSU) X(false, T_F32, 0, 16, EPI_LSU) X(true, T_BF16, 0, 16, EPI_LSU) X(true, T_F32, 0, 16, EPI_LSU) \
  X(false, T_F16, 0, 16, EPI_LSU)                                                                                  \
  X(false, T_BF16, 0, 16, EPI_TMA) X(true, T_BF16, 0, 16, EPI_TMA) X(false, T_F16, 0, 16, EPI_TMA_RES)             \
  X(false, T_BF16, 0, 16, EPI_TMA_LN) X(true, T_BF16, 0, 16, EPI_TMA_LN)                                              \
  X(false, T_BF16, 1, 8, EPI_LSU) X(false, T_F32, 1, 8, EPI_LSU) X(true, T_BF16, 1, 8, EPI_LSU) X(true, T_F32, 1, 8, EPI_LSU)     \
  X(false, T_F16, 1, 8, EPI_LSU)
// the instances that also exist for clusters of 4 (the ViT block GEMMs)
#define YV_GEMM_INSTANCES4(X)                                                                                      \
  X(false, T_BF16, 0, 16, EPI_TMA) X(true, T_BF16, 0, 16, EPI_TMA) X(false, T_F16, 0, 16, EPI_TMA_RES)             \
  X(false, T_BF16, 0, 16, EPI_TMA_LN) X(true, T_BF16, 0, 16, EPI_TMA_LN)
int init_device() {
  cudaError_t e = cudaSuccess;
#define YV_SET(G, O, X, W, D)                                                                                      \
  if (e == cudaSuccess)                                                                                          \
    e = cudaFuncSetAttribute(gemm_bf16_kernel<G, O, X, W, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes(W, D));
  YV_GEMM_INSTANCES(YV_SET)
#undef YV_SET
#define YV_SET4(G, O, X, W, D)                                                                                     \
  if (e == cudaSuccess)                                                                                          \
    e = cudaFuncSetAttribute(gemm_bf16_kernel<G, O, X, W, D, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes(W, D)); \
  if (e == cudaSuccess)                                                                                          \
    e = cudaFuncSetAttribute(gemm_bf16_kernel<G, O, X, W, D, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes(W, D));
  YV_GEMM_INSTANCES4(YV_SET4)
#undef YV_SET4
  if (e != cudaSuccess) return yv_cuda_fail(e, "cudaFuncSetAttribute(gemm smem)", __FILE__, __LINE__);
  return YV_OK;
}

// row-major 16-bit [rows, cols]; box = box_rows x box_cols, zero OOB fill
// (operands: 64 columns, 128-byte swizzle; TMA epilogues: 32 x 32 box, 64-byte swizzle)
int make_tmap(CUtensorMap *m, const void *ptr, uint64_t rows, uint64_t cols, uint32_t box_rows, uint32_t box_cols = BK,
              CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B, CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16) {
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(m, dt, 2, const_cast<void *>(ptr), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    yv_set_error("cuTensorMapEncodeTiled failed (%d) for [%llu x %llu] 16-bit at %p", (int)r, (unsigned long long)rows,
                 (unsigned long long)cols, ptr);
    return YV_ECUDA;
  }
  return YV_OK;
}

}  // namespace

int yv_tmap_bf16_2d(CUtensorMap *m, const void *ptr, uint64_t rows, uint64_t cols, uint32_t box_rows) {
  std::call_once(g_once, init_once);
  if (g_init_status != YV_OK) return g_init_status;
  return make_tmap(m, ptr, rows, cols, box_rows);
}

int yv_gemm_launch(const void *A, const void *W, int M, int N, int K, const GemmEpilogue &ep, cudaStream_t s) {
  YV_REQUIRE(M >= 1 && N >= 8 && K >= BK, "yv_gemm: bad shape M=%d N=%d K=%d", M, N, K);
  YV_REQUIRE(K % BK == 0, "yv_gemm: K=%d must be a multiple of %d", K, BK);
  YV_REQUIRE(N % 8 == 0, "yv_gemm: N=%d must be a multiple of 8", N);
  YV_REQUIRE(A && W && ep.out, "yv_gemm: null pointer");
  YV_REQUIRE(((uintptr_t)A % 16) == 0 && ((uintptr_t)W % 16) == 0 && ((uintptr_t)ep.out % 16) == 0 &&
                 ((uintptr_t)ep.residual % 16) == 0,
             "yv_gemm: operands must be 16-byte aligned");
  YV_REQUIRE(ep.group_rows == 0 || ep.pos, "yv_gemm: group_rows needs a position table");
  YV_REQUIRE(ep.out_type >= T_BF16 && ep.out_type <= T_F16 && ep.res_type >= T_BF16 && ep.res_type <= T_F16, "yv_gemm: bad element type");
  std::call_once(g_once, init_once);
  if (g_init_status != YV_OK) return g_init_status;
  int rc = yv_device_once(YV_INIT_GEMM, init_device);
  if (rc) return rc;
  CUtensorMap tmA, tmB;
  rc = make_tmap(&tmA, A, (uint64_t)M, (uint64_t)K, BM);
  if (rc) return rc;
  rc = make_tmap(&tmB, W, (uint64_t)N, (uint64_t)K, BNH);
  if (rc) return rc;
  const int tiles = ((M + PM - 1) / PM) * ((N + BN - 1) / BN);
  const int pairs = yv_gemm_sm_budget() / 2;
  const int grid = 2 * (tiles < pairs ? tiles : pairs);
  YvProf prof(YV_K_GEMM, s, 2.0 * M * (double)N * K);
  static const int dbg = []() { const char *e = getenv("YV_GEMM_DBG"); return e ? atoi(e) : 0; }();
  const_cast<GemmEpilogue &>(ep).dbg = dbg;
  // Which epilogue: the fused fp16 residual form (x16 = x16 + A W^T + b, TMA in and out), the bf16 TMA-store
  // form, or the LSU form (fp32 outputs, the patch-embed row map, bf16 / fp32 residuals of the generic C-ABI call).
  const bool tma_res = ep.residual && ep.res_type == T_F16 && ep.out_type == T_F16 && !ep.gelu && ep.group_rows == 0;
  const bool extra = !tma_res && (ep.group_rows > 0 || ep.residual != nullptr);
  YV_REQUIRE(tma_res || !ep.residual || ep.res_type != T_F16, "yv_gemm: an fp16 residual needs an fp16 output and no GELU");
  YV_REQUIRE(ep.out_type != T_F16 || !ep.gelu, "yv_gemm: fp16 output with GELU is not instantiated");
  // epilogue warps of the plain mode: 16 (measured +1.2 % on the whole step against 8; fc1 179 -> 168 us);
  // YV_GEMM_EPI_WARPS=8 selects the 8-warp form, YV_GEMM_TMA_STORE=0 turns the TMA-store epilogue of bf16 outputs off (A/B timing).  The patch-embed mode keeps 8 (it spills at 96 registers).
  static const int force = []() { const char *e = getenv("YV_GEMM_EPI_WARPS"); return e ? atoi(e) : 0; }();
  const bool w16 = force != 8 || ep.out_type == T_F16;
  static const bool direct = []() { const char *e = getenv("YV_GEMM_TMA_STORE"); return !(e && e[0] == '0'); }();
  YV_REQUIRE(!ep.stats_part || (tma_res && N % 64 == 0), "yv_gemm: partial row statistics need the fused fp16 residual epilogue and N %% 64 == 0");
  const bool lnfold = ep.row_stats != nullptr;       // A and W are fp16 here (the caller packs them so)
  YV_REQUIRE(!lnfold || (ep.col_s && ep.bias && !extra && !tma_res && ep.out_type == T_BF16),
             "yv_gemm: the folded-LayerNorm form needs col_s and bias, a bf16 output and no residual");
  const bool tma_out = (direct || lnfold) && !extra && !tma_res && ep.out_type == T_BF16;
  CUtensorMap tmD = tmA, tmR = tmA;            // placeholders unless a TMA epilogue runs
  if (tma_out && (rc = make_tmap(&tmD, ep.out, (uint64_t)M, (uint64_t)N, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B))) return rc;
  if (tma_res) {
    if ((rc = make_tmap(&tmD, ep.out, (uint64_t)M, (uint64_t)N, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_DATA_TYPE_FLOAT16))) return rc;
    if ((rc = make_tmap(&tmR, ep.residual, (uint64_t)M, (uint64_t)N, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_DATA_TYPE_FLOAT16))) return rc;
  }
#define YV_GO(G, O, X, W, D) gemm_bf16_kernel<G, O, X, W, D><<<grid, gemm_threads(W), smem_bytes(W, D), s>>>(tmA, tmB, tmD, tmR, ep, M, N, K)
  // Clusters of 4 / 6 (W multicast to two / three M-adjacent CTA pairs) for the TMA-epilogue forms: opt-in with
  // YV_GEMM_CLUSTER=4 / 6 (clusters of 6: 22 co-resident = the same 132 SMs, same outcome).
  // Measured (tools/gemm_power.py, profiles/r02_gemm_power.md): 25 % less L2 -> smem traffic raises the power-capped
  // clock by 11 - 24 %, but only 33 clusters of 4 fit the GPCs (132 of 148 SMs) and the sustained rate ends 2 - 3 %
  // BELOW the CTA-pair form, so pairs stay the default.
  static const int cs_env = []() { const char *e = getenv("YV_GEMM_CLUSTER"); return e ? atoi(e) : 0; }();
  const int m_tiles_h = (M + PM - 1) / PM;
  const int cs = (cs_env == 4 || cs_env == 6) && (lnfold || tma_res || tma_out) && m_tiles_h >= cs_env / 2 ? cs_env : 2;
  if (cs > 2) {
    static std::mutex mu;
    static int max_cl[64][2] = {};
    int dev = 0;
    YV_CUDA(cudaGetDevice(&dev));
    YV_REQUIRE(dev >= 0 && dev < 64, "yv_gemm: device %d out of range", dev);
    const int pc = cs / 2, slot = cs == 6;
    const int units = ((m_tiles_h + pc - 1) / pc) * ((N + BN - 1) / BN);
    {
      std::lock_guard<std::mutex> lk(mu);
      if (max_cl[dev][slot] == 0) {
        cudaLaunchConfig_t cfg = {};                 // (the cluster size is compiled into the kernel: no launch attribute)
        cfg.blockDim = dim3(gemm_threads(16));
        cfg.dynamicSmemBytes = smem_bytes(16, EPI_TMA);
        cfg.stream = s;
        cfg.gridDim = dim3((unsigned)(yv_num_sms() / cs * cs));
        int n = 0;
        cudaError_t qe = cs == 6 ? cudaOccupancyMaxActiveClusters(&n, gemm_bf16_kernel<false, T_BF16, 0, 16, EPI_TMA, 6>, &cfg)
                                 : cudaOccupancyMaxActiveClusters(&n, gemm_bf16_kernel<false, T_BF16, 0, 16, EPI_TMA, 4>, &cfg);
        if (qe != cudaSuccess || n < 1) { (void)cudaGetLastError(); n = yv_num_sms() / cs; }
        max_cl[dev][slot] = n;
        if (getenv("YV_GEMM_VERBOSE")) fprintf(stderr, "[yv8vit] clusters of %d co-resident on device %d: %d (%d SMs), query rc %d\n", cs, dev, n, cs * n, (int)qe);
      }
    }
    const int ncl = units < max_cl[dev][slot] ? units : max_cl[dev][slot];
    const int gridc = cs * ncl;
#define YV_GOC(G, O, X, W, D)                                                                                                        \
    do { if (cs == 6) gemm_bf16_kernel<G, O, X, W, D, 6><<<gridc, gemm_threads(W), smem_bytes(W, D), s>>>(tmA, tmB, tmD, tmR, ep, M, N, K); \
         else gemm_bf16_kernel<G, O, X, W, D, 4><<<gridc, gemm_threads(W), smem_bytes(W, D), s>>>(tmA, tmB, tmD, tmR, ep, M, N, K); } while (0)
    if (lnfold) { if (ep.gelu) YV_GOC(true, T_BF16, 0, 16, EPI_TMA_LN); else YV_GOC(false, T_BF16, 0, 16, EPI_TMA_LN); }
    else if (tma_res) YV_GOC(false, T_F16, 0, 16, EPI_TMA_RES);
    else { if (ep.gelu) YV_GOC(true, T_BF16, 0, 16, EPI_TMA); else YV_GOC(false, T_BF16, 0, 16, EPI_TMA); }
#undef YV_GOC
  } else
  if (lnfold) {
    if (ep.gelu) YV_GO(true, T_BF16, 0, 16, EPI_TMA_LN); else YV_GO(false, T_BF16, 0, 16, EPI_TMA_LN);
  } else if (tma_res) {
    YV_GO(false, T_F16, 0, 16, EPI_TMA_RES);
  } else if (tma_out) {
    if (ep.gelu) YV_GO(true, T_BF16, 0, 16, EPI_TMA); else YV_GO(false, T_BF16, 0, 16, EPI_TMA);
  } else if (ep.out_type == T_F16) {
    if (extra) YV_GO(false, T_F16, 1, 8, EPI_LSU); else YV_GO(false, T_F16, 0, 16, EPI_LSU);
  } else if (ep.gelu) {
    if (ep.out_type == T_F32) { if (extra) YV_GO(true, T_F32, 1, 8, EPI_LSU); else if (w16) YV_GO(true, T_F32, 0, 16, EPI_LSU); else YV_GO(true, T_F32, 0, 8, EPI_LSU); }
    else { if (extra) YV_GO(true, T_BF16, 1, 8, EPI_LSU); else if (w16) YV_GO(true, T_BF16, 0, 16, EPI_LSU); else YV_GO(true, T_BF16, 0, 8, EPI_LSU); }
  } else {
    if (ep.out_type == T_F32) { if (extra) YV_GO(false, T_F32, 1, 8, EPI_LSU); else if (w16) YV_GO(false, T_F32, 0, 16, EPI_LSU); else YV_GO(false, T_F32, 0, 8, EPI_LSU); }
    else { if (extra) YV_GO(false, T_BF16, 1, 8, EPI_LSU); else if (w16) YV_GO(false, T_BF16, 0, 16, EPI_LSU); else YV_GO(false, T_BF16, 0, 8, EPI_LSU); }
  }
#undef YV_GO
  YV_LAUNCH_CHECK();
  return YV_OK;
}

YV_EXPORT int yv_gemm_bf16(const void *A, const void *W, const float *bias, const void *residual, void *D, int M,
                           int N, int K, int epi, int out_type, yv_stream_t stream) {
  YV_REQUIRE(epi == YV_EPI_NONE || epi == YV_EPI_GELU, "yv_gemm_bf16: unknown epilogue %d", epi);
  YV_REQUIRE(out_type >= 0 && out_type <= 2, "yv_gemm_bf16: out_type %d (0 bf16, 1 f32, 2 f16)", out_type);
  GemmEpilogue ep;
  ep.out = D;
  ep.bias = bias;
  ep.residual = residual;
  ep.res_type = out_type == T_F16 ? T_F16 : T_BF16;     // an fp16 output takes an fp16 residual (may alias D)
  ep.pos = nullptr;
  ep.group_rows = 0;
  ep.gelu = (epi == YV_EPI_GELU);
  ep.out_type = out_type;
  return yv_gemm_launch(A, W, M, N, K, ep, (cudaStream_t)stream);
}
