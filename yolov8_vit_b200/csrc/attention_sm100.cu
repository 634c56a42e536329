// attention_sm100.cu -- K7: softmax(Q K^T / sqrt(dh)) V on tcgen05 / TMEM, fed by TMA.
//
// timm Attention.forward (vit_base_patch{16,8}_224; reached by the reference through
// utils/utils.py:71-72): qkv [n, T, 3, heads, 64] -> per head softmax(q k^T * 64^-0.5) v.
//
// One persistent CTA per SM walks the (crop, head) work items.  Per item (T <= 256 tokens, NK = T
// rounded up to 16 keys):
//   TMA      Q (256 rows x 64), K (NK x 64), V (NK x 64) boxes straight out of the packed qkv
//            activation (128-byte swizzle = the UMMA operand layout, no staging copies);
//   MMA      S_g = Q_g K^T  (128 x NK x 64, g = query rows 0-127 / 128-255) into TMEM,
//            then O_g = P_g V (128 x 64 x NK) into the TMEM columns S_g just vacated; V is consumed
//            as an MN-major B operand, so no transpose is ever materialised;
//   softmax  16 warps (8 per query tile; a row is split between 2 threads): two passes over S in TMEM
//            (max, then exp2 / sum), P written as bf16 into shared memory in the K-major
//            128B-swizzled A-operand layout; finally O * (1/sum) -> bf16 -> global.
// The fp32 scores never leave the SM; HBM traffic is qkv in, out out.
#include "common.cuh"
#include "gemm_sm100.cuh"
#include "sm100_ptx.cuh"
#include "vit_ops.cuh"

using namespace sm100;

namespace {

constexpr int AT_DH = 64;
constexpr int AT_QROWS = 256;                       // Q box rows (two 128-row query tiles)
constexpr int AT_THREADS = 20 * 32;                 // TMA, MMA, TMEM-alloc, spare + 16 softmax warps
constexpr int AT_Q_BYTES = AT_QROWS * AT_DH * 2;    // 32 KB
constexpr int AT_KV_BYTES = 256 * AT_DH * 2;        // 32 KB reserved each (box is NK rows)
constexpr int AT_P_BYTES = 4 * 128 * 128;           // per query tile: 4 chunks of 64 keys x 128 rows x 128 B
constexpr int AT_SMEM = AT_Q_BYTES + 2 * AT_KV_BYTES + 2 * AT_P_BYTES + 2304;   // + barriers / exchange
constexpr int AT_TMEM_COLS = 512;                   // S_0 at column 0, S_1 at column 256

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

struct AttnBars {
  uint64_t qk_full, v_full;
  uint64_t s_full[2], p_full[2], o_full[2], t_free[2];
  uint32_t tmem_slot;
  float xch[2][2][128];       // [tile][slot][row]: max / sum exchange between the two column halves
};

__global__ void __launch_bounds__(AT_THREADS, 1)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                    __nv_bfloat16 *__restrict__ out, int n_items, int T, int heads, float scale_log2e) {
  extern __shared__ __align__(1024) uint8_t at_raw[];
  uint8_t *smem = at_raw;                           // 128B-swizzled tiles need 1024-byte alignment
  if ((smem_u32(smem) & 1023u) != 0u) __trap();
  uint8_t *sQ = smem, *sK = sQ + AT_Q_BYTES, *sV = sK + AT_KV_BYTES, *sP = sV + AT_KV_BYTES;
  AttnBars *bar = reinterpret_cast<AttnBars *>(sP + 2 * AT_P_BYTES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int NK = (T + 15) & ~15;                    // keys per MMA (multiple of 16, <= 256)
  const int NG = (T + 127) >> 7;                    // query tiles in use (1 or 2)
  const int D = heads * AT_DH;
  const uint32_t kv_bytes = (uint32_t)NK * AT_DH * 2;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmQ);
    prefetch_tmap(&tmKV);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(&bar->qk_full, 1);
    mbar_init(&bar->v_full, 1);
    for (int g = 0; g < 2; ++g) {
      mbar_init(&bar->s_full[g], 1);
      mbar_init(&bar->p_full[g], 8);                // the 8 softmax warps of the tile
      mbar_init(&bar->o_full[g], 1);
      mbar_init(&bar->t_free[g], 8);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(&bar->tmem_slot, AT_TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bar->tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      uint32_t it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
        const int crop = item / heads, head = item - crop * heads;
        const int row0 = crop * T;
        const uint32_t prev = (it & 1u) ^ 1u;       // parity of the previous item's completions
        // Q, K are free once the previous item's last S MMA retired; V once its last P V retired
        if (it > 0) mbar_wait(&bar->s_full[NG - 1], prev);
        mbar_arrive_expect_tx(&bar->qk_full, AT_Q_BYTES + kv_bytes);
        tma_load_2d(sQ, &tmQ, &bar->qk_full, head * AT_DH, row0);
        tma_load_2d(sK, &tmKV, &bar->qk_full, D + head * AT_DH, row0);
        if (it > 0) mbar_wait(&bar->o_full[NG - 1], prev);
        mbar_arrive_expect_tx(&bar->v_full, kv_bytes);
        tma_load_2d(sV, &tmKV, &bar->v_full, 2 * D + head * AT_DH, row0);
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (elect_one()) {
      const uint32_t idesc_s = umma_idesc_bf16(128, NK);
      const uint32_t idesc_o = umma_idesc_bf16_bmn(128, AT_DH);
      const uint32_t aQ = smem_u32(sQ), aK = smem_u32(sK), aV = smem_u32(sV), aP = smem_u32(sP);
      uint32_t it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
        const uint32_t ph = it & 1u;
        mbar_wait(&bar->qk_full, ph);
        tc_fence_after();
        for (int g = 0; g < NG; ++g) {              // S_g = Q_g K^T
          mbar_wait(&bar->t_free[g], ph ^ 1u);      // softmax warps have drained O_g of the previous item
          tc_fence_after();
          const uint64_t dq = umma_desc_k_sw128(aQ + g * (128 * 128)), dk = umma_desc_k_sw128(aK);
#pragma unroll
          for (int k = 0; k < AT_DH / 16; ++k)
            umma_bf16(tmem_base + g * 256, dq + (uint64_t)(k * 2), dk + (uint64_t)(k * 2), idesc_s, (uint32_t)(k != 0));
          umma_commit(&bar->s_full[g]);
        }
        mbar_wait(&bar->v_full, ph);
        for (int g = 0; g < NG; ++g) {              // O_g = P_g V, into the columns S_g occupied
          mbar_wait(&bar->p_full[g], ph);
          tc_fence_after();
          for (int k = 0; k < NK / 16; ++k) {
            const uint64_t dp = umma_desc_k_sw128(aP + g * AT_P_BYTES + (k >> 2) * (128 * 128) + (k & 3) * 32);
            const uint64_t dv = umma_desc_mn_sw128(aV + k * (16 * 128));
            umma_bf16(tmem_base + g * 256, dp, dv, idesc_o, (uint32_t)(k != 0));
          }
          umma_commit(&bar->o_full[g]);
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== softmax + output =====================
    // 16 warps: query tile g, TMEM lane quarter q (one thread per query row), column half hf.  The two
    // warps that share a row exchange their partial max / sum through shared memory (named barrier per
    // tile).  4 softmax warps per scheduler hide the MUFU / tcgen05.ld latencies that 2 could not.
    const int sw_id = warp - 4;
    const int q = warp & 3, g = (sw_id >> 2) & 1, hf = sw_id >> 3;
    if (g < NG) {
      const int trow = q * 32 + lane;               // row inside the tile
      const int tok = g * 128 + trow;               // token index inside the crop
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + g * 256;
      uint8_t *prow = sP + g * AT_P_BYTES + trow * 128;
      const int swz = trow & 7;
      const int H0 = ((NK / 16 + 1) / 2) * 16;      // first half: ceil(chunks / 2) 16-key chunks
      const int c_lo = hf ? H0 : 0, c_hi = hf ? NK : H0;
      // slot[hf] carries my max to the partner; once I have read the partner's max from slot[hf ^ 1]
      // that slot is mine to carry my sum back (each slot has one writer and one reader per phase)
      float *xmine = bar->xch[g][hf] + trow, *xother = bar->xch[g][hf ^ 1] + trow;
      uint32_t it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
        const uint32_t ph = it & 1u;
        const int crop = item / heads, head = item - crop * heads;
        mbar_wait(&bar->s_full[g], ph);
        tc_fence_after();
        // pass 1: maximum of this warp's half of the row (keys >= T are masked)
        float mx = -INFINITY;
        for (int c = c_lo; c < c_hi; c += 32) {
          uint32_t a[16], b[16];
          const bool two = c + 16 < c_hi;
          tmem_ld_32x16(taddr + c, a);
          if (two) tmem_ld_32x16(taddr + c + 16, b);
          tmem_ld_wait();
          if (c + 32 <= T) {
#pragma unroll
            for (int j = 0; j < 16; ++j) mx = fmaxf(mx, fmaxf(__uint_as_float(a[j]), __uint_as_float(b[j])));
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (c + j < T) mx = fmaxf(mx, __uint_as_float(a[j]));
            if (two) {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (c + 16 + j < T) mx = fmaxf(mx, __uint_as_float(b[j]));
            }
          }
        }
        *xmine = mx;
        asm volatile("bar.sync %0, 256;" ::"r"(1 + g) : "memory");
        mx = fmaxf(mx, *xother);
        const float mneg = -mx * scale_log2e;
        // pass 2: p = exp2(s * scale - max * scale); row sum; P -> smem (bf16, K-major, 128B swizzle)
        float sum = 0.f;
        uint32_t r[16];
        tmem_ld_32x16(taddr + c_lo, r);
        for (int c = c_lo; c < c_hi; c += 16) {
          tmem_ld_wait();
          float p[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) p[j] = ex2_approx(fmaf(__uint_as_float(r[j]), scale_log2e, mneg));
          if (c + 16 < c_hi) tmem_ld_32x16(taddr + c + 16, r);
          if (c + 16 > T) {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (c + j >= T) p[j] = 0.f;
          }
#pragma unroll
          for (int j = 0; j < 16; j += 4) sum += (p[j] + p[j + 1]) + (p[j + 2] + p[j + 3]);
          uint8_t *chunk = prow + (c >> 6) * (128 * 128);
          const int u0 = (c & 63) >> 3;             // 16-byte unit (8 keys) inside the 128-byte row
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            uint4 w;
            __nv_bfloat162 h0 = __floats2bfloat162_rn(p[8 * u + 0], p[8 * u + 1]);
            __nv_bfloat162 h1 = __floats2bfloat162_rn(p[8 * u + 2], p[8 * u + 3]);
            __nv_bfloat162 h2 = __floats2bfloat162_rn(p[8 * u + 4], p[8 * u + 5]);
            __nv_bfloat162 h3 = __floats2bfloat162_rn(p[8 * u + 6], p[8 * u + 7]);
            w.x = *reinterpret_cast<uint32_t *>(&h0);
            w.y = *reinterpret_cast<uint32_t *>(&h1);
            w.z = *reinterpret_cast<uint32_t *>(&h2);
            w.w = *reinterpret_cast<uint32_t *>(&h3);
            *reinterpret_cast<uint4 *>(chunk + (((u0 + u) ^ swz) << 4)) = w;
          }
        }
        fence_proxy_async_smem();                   // P (generic-proxy stores) -> visible to the UMMA reads
        tc_fence_before();                          // our tcgen05.ld of S_g precede the MMA that overwrites it
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar->p_full[g]);
        *xother = sum;
        asm volatile("bar.sync %0, 256;" ::"r"(1 + g) : "memory");
        const float inv = 1.0f / (sum + *xmine);

        // O_g = P_g V is in TMEM columns [0, 64) of the tile: this warp normalises and stores 32 of them
        mbar_wait(&bar->o_full[g], ph);
        tc_fence_after();
        uint32_t o0[16], o1[16];
        tmem_ld_32x16(taddr + hf * 32, o0);
        tmem_ld_32x16(taddr + hf * 32 + 16, o1);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar->t_free[g]);   // the accumulator columns may be reused
        if (tok < T) {
          uint4 *dst = reinterpret_cast<uint4 *>(out + ((size_t)crop * T + tok) * D + (size_t)head * AT_DH + hf * 32);
          auto pack8 = [&](const uint32_t *v) {
            uint4 w;
            __nv_bfloat162 h0 = __floats2bfloat162_rn(__uint_as_float(v[0]) * inv, __uint_as_float(v[1]) * inv);
            __nv_bfloat162 h1 = __floats2bfloat162_rn(__uint_as_float(v[2]) * inv, __uint_as_float(v[3]) * inv);
            __nv_bfloat162 h2 = __floats2bfloat162_rn(__uint_as_float(v[4]) * inv, __uint_as_float(v[5]) * inv);
            __nv_bfloat162 h3 = __floats2bfloat162_rn(__uint_as_float(v[6]) * inv, __uint_as_float(v[7]) * inv);
            w.x = *reinterpret_cast<uint32_t *>(&h0);
            w.y = *reinterpret_cast<uint32_t *>(&h1);
            w.z = *reinterpret_cast<uint32_t *>(&h2);
            w.w = *reinterpret_cast<uint32_t *>(&h3);
            return w;
          };
          dst[0] = pack8(o0); dst[1] = pack8(o0 + 8);
          dst[2] = pack8(o1); dst[3] = pack8(o1 + 8);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, AT_TMEM_COLS);
}

}  // namespace

int yv_attention_launch(const __nv_bfloat16 *qkv, __nv_bfloat16 *out, int n, int T, int heads, int dh,
                        cudaStream_t s) {
  YV_REQUIRE(dh == AT_DH, "attention: head dim %d unsupported (need %d)", dh, AT_DH);
  YV_REQUIRE(n >= 1 && heads >= 1 && T >= 1, "attention: bad shape");
  YV_REQUIRE(T <= 256, "attention: T=%d exceeds 256 tokens (ViT-B/8's 785 tokens need the multi-block variant)", T);
  YV_REQUIRE(((uintptr_t)qkv % 16) == 0 && ((uintptr_t)out % 16) == 0, "attention: operands must be 16-byte aligned");
  const int NK = (T + 15) & ~15;
  const int D = heads * AT_DH;
  CUtensorMap tmQ, tmKV;
  int rc = yv_tmap_bf16_2d(&tmQ, qkv, (uint64_t)n * T, (uint64_t)3 * D, AT_QROWS);
  if (rc) return rc;
  rc = yv_tmap_bf16_2d(&tmKV, qkv, (uint64_t)n * T, (uint64_t)3 * D, (uint32_t)NK);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    YV_CUDA(cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
    attr_set = true;
  }
  const int items = n * heads;
  const int grid = items < yv_num_sms() ? items : yv_num_sms();
  const float scale_log2e = (1.0f / sqrtf((float)dh)) * 1.4426950408889634f;
  YvProf prof(YV_K_ATTN, s, 4.0 * n * heads * (double)T * T * dh);
  attention_tc_kernel<<<grid, AT_THREADS, AT_SMEM, s>>>(tmQ, tmKV, out, items, T, heads, scale_log2e);
  YV_LAUNCH_CHECK();
  return YV_OK;
}

YV_EXPORT int yv_attention_bf16(const void *qkv, void *out, int n, int T, int heads, int dh, yv_stream_t stream) {
  YV_REQUIRE(qkv && out, "yv_attention_bf16: null pointer");
  return yv_attention_launch((const __nv_bfloat16 *)qkv, (__nv_bfloat16 *)out, n, T, heads, dh, (cudaStream_t)stream);
}
