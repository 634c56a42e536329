// gemm_sm100.cu -- K5: D[M,N] = A[M,K] * W[N,K]^T with fused epilogues, bf16 operands, fp32
// accumulation, hand-written for sm_100a:
//   * operands arrive by TMA (cp.async.bulk.tensor.2d, 128-byte swizzle) into a 4-stage
//     shared-memory ring, completion on mbarriers;
//   * ONE elected thread issues tcgen05.mma (cta_group::1, kind::f16, 128 x 256 x 16) reading
//     both operands through UMMA shared-memory descriptors;
//   * the fp32 accumulator lives in TMEM, double buffered (2 x 256 columns = all 512), so the
//     epilogue of tile i overlaps the MMAs of tile i+1;
//   * 8 epilogue warps pull the accumulator back with tcgen05.ld (32 lanes x 32 columns),
//     apply bias / erf-GELU / residual / position-embedding and store bf16 or fp32;
//   * persistent: one CTA per SM walks the tile list (n fastest, so the A rows of an m-block
//     are served from L2 to its n-tiles while the weights stay L2 resident).
// This is the ViT's only dense contraction: QKV, proj, fc1, fc2,
// patch-embed and the 1000-way head all run through it.
#include "common.cuh"
#include "gemm_sm100.cuh"
#include "sm100_ptx.cuh"

#include <mutex>

using namespace sm100;

namespace {

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 64;                 // 64 bf16 = 128 B = one swizzle atom row
constexpr int STAGES = 4;
constexpr int UMMA_K = 16;
constexpr int A_BYTES = BM * BK * 2;   // 16 KB
constexpr int B_BYTES = BN * BK * 2;   // 32 KB
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int NUM_EPI_WARPS = 8;
constexpr int GEMM_THREADS = (4 + NUM_EPI_WARPS) * 32;
constexpr int TMEM_COLS = 512;         // 2 accumulator stages x BN
constexpr size_t SMEM_BYTES = (size_t)STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;

__device__ __forceinline__ float gelu_erf(float x) {
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}

__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const GemmEpilogue ep, int M, int N, int K) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t *bars = (uint64_t *)(smem + (size_t)STAGES * STAGE_BYTES);
  uint64_t *full = bars, *empty = bars + STAGES;
  uint64_t *tfull = bars + 2 * STAGES, *tempty = bars + 2 * STAGES + 2;
  uint32_t *tmem_slot = (uint32_t *)(bars + 2 * STAGES + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_tiles = (M + BM - 1) / BM, n_tiles = (N + BN - 1) / BN;
  const int num_tiles = m_tiles * n_tiles, kblocks = K / BK;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], NUM_EPI_WARPS); }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer (one elected lane) =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m0 = (tile / n_tiles) * BM, n0 = (tile % n_tiles) * BN;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1u);
          uint8_t *sa = smem + (size_t)stage * STAGE_BYTES, *sb = sa + A_BYTES;
          mbar_arrive_expect_tx(&full[stage], STAGE_BYTES);
          tma_load_2d(sa, &tmA, &full[stage], kb * BK, m0);
          tma_load_2d(sb, &tmB, &full[stage], kb * BK, n0);
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one elected lane) =====================
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
      int stage = 0;
      uint32_t phase = 0, it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
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
            umma_bf16(d_tmem, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (uint32_t)((kb | k) != 0));
          }
          umma_commit(&empty[stage]);              // smem slot free once these MMAs retire
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
        umma_commit(&tfull[as]);                   // accumulator complete -> epilogue
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: TMEM -> registers -> global =====================
    const int ew = warp - 4;
    const int q = warp & 3;                        // TMEM lane quarter this warp may access
    const int half = ew >> 2;                      // column half of the tile
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const uint32_t as = it & 1u, aph = (it >> 1) & 1u;
      const int m0 = (tile / n_tiles) * BM, n0 = (tile % n_tiles) * BN;
      mbar_wait(&tfull[as], aph);
      tc_fence_after();
      const int row = m0 + q * 32 + lane;
      const bool row_ok = row < M;
      long long drow = row;
      const float *pos_row = nullptr;
      if (ep.group_rows > 0) {                     // patch-embed: token row = r + r/196 + 1, + pos[1 + r%196]
        const int g = row / ep.group_rows;
        drow = (long long)row + g + 1;
        pos_row = ep.pos + (size_t)(row - g * ep.group_rows + 1) * N;
      }
#pragma unroll 1
      for (int ch = 0; ch < BN / 2 / 32; ++ch) {
        const int cbase = half * (BN / 2) + ch * 32;
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + as * BN + cbase, r);
        tmem_ld_wait();
        const int col0 = n0 + cbase;
        if (row_ok && col0 < N) {
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
          const int ncols = min(32, N - col0);     // multiple of 8
          if (ep.bias) {
#pragma unroll
            for (int j = 0; j < 32; ++j) if (j < ncols) v[j] += __ldg(ep.bias + col0 + j);
          }
          if (ep.gelu) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
          }
          if (pos_row) {
#pragma unroll
            for (int j = 0; j < 32; ++j) if (j < ncols) v[j] += __ldg(pos_row + col0 + j);
          }
          if (ep.residual && ep.res_f32) {
            const float4 *rp = reinterpret_cast<const float4 *>((const float *)ep.residual + (size_t)drow * N + col0);
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              if (g * 4 < ncols) {
                const float4 f = rp[g];
                v[4 * g] += f.x; v[4 * g + 1] += f.y; v[4 * g + 2] += f.z; v[4 * g + 3] += f.w;
              }
            }
          } else if (ep.residual) {
            const uint4 *rp = reinterpret_cast<const uint4 *>((const __nv_bfloat16 *)ep.residual + (size_t)drow * N + col0);
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              if (g * 8 < ncols) {
                const uint4 u = rp[g];
                const __nv_bfloat162 *h = reinterpret_cast<const __nv_bfloat162 *>(&u);
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const float2 f = __bfloat1622float2(h[e]);
                  v[g * 8 + 2 * e] += f.x;
                  v[g * 8 + 2 * e + 1] += f.y;
                }
              }
            }
          }
          if (ep.out_f32) {
            float4 *op = reinterpret_cast<float4 *>((float *)ep.out + (size_t)drow * N + col0);
#pragma unroll
            for (int g = 0; g < 8; ++g)
              if (g * 4 < ncols) op[g] = make_float4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
          } else {
            uint4 *op = reinterpret_cast<uint4 *>((__nv_bfloat16 *)ep.out + (size_t)drow * N + col0);
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              if (g * 8 < ncols) {
                uint4 u;
                __nv_bfloat162 h0 = __floats2bfloat162_rn(v[8 * g], v[8 * g + 1]);
                __nv_bfloat162 h1 = __floats2bfloat162_rn(v[8 * g + 2], v[8 * g + 3]);
                __nv_bfloat162 h2 = __floats2bfloat162_rn(v[8 * g + 4], v[8 * g + 5]);
                __nv_bfloat162 h3 = __floats2bfloat162_rn(v[8 * g + 6], v[8 * g + 7]);
                u.x = *reinterpret_cast<unsigned *>(&h0);
                u.y = *reinterpret_cast<unsigned *>(&h1);
                u.z = *reinterpret_cast<unsigned *>(&h2);
                u.w = *reinterpret_cast<unsigned *>(&h3);
                op[g] = u;
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[as]);     // this warp is done with the accumulator
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, TMEM_COLS);
}

// ---- host side: tensor maps ----------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                    CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                    CUtensorMapFloatOOBfill);
PFN_encodeTiled g_encode = nullptr;
std::once_flag g_once;
int g_init_status = YV_OK;

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
  e = cudaFuncSetAttribute(gemm_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
  if (e != cudaSuccess) {
    g_init_status = yv_cuda_fail(e, "cudaFuncSetAttribute(gemm smem)", __FILE__, __LINE__);
  }
}

// row-major bf16 [rows, cols]; box = box_rows x 64 columns, 128-byte swizzle, zero OOB fill
int make_tmap(CUtensorMap *m, const void *ptr, uint64_t rows, uint64_t cols, uint32_t box_rows) {
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(ptr), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    yv_set_error("cuTensorMapEncodeTiled failed (%d) for [%llu x %llu] bf16 at %p", (int)r, (unsigned long long)rows,
                 (unsigned long long)cols, ptr);
    return YV_ECUDA;
  }
  return YV_OK;
}

}  // namespace

int yv_gemm_launch(const void *A, const void *W, int M, int N, int K, const GemmEpilogue &ep, cudaStream_t s) {
  YV_REQUIRE(M >= 1 && N >= 8 && K >= BK, "yv_gemm: bad shape M=%d N=%d K=%d", M, N, K);
  YV_REQUIRE(K % BK == 0, "yv_gemm: K=%d must be a multiple of %d", K, BK);
  YV_REQUIRE(N % 8 == 0, "yv_gemm: N=%d must be a multiple of 8", N);
  YV_REQUIRE(A && W && ep.out, "yv_gemm: null pointer");
  YV_REQUIRE(((uintptr_t)A % 16) == 0 && ((uintptr_t)W % 16) == 0 && ((uintptr_t)ep.out % 16) == 0 &&
                 ((uintptr_t)ep.residual % 16) == 0,
             "yv_gemm: operands must be 16-byte aligned");
  YV_REQUIRE(ep.group_rows == 0 || ep.pos, "yv_gemm: group_rows needs a position table");
  std::call_once(g_once, init_once);
  if (g_init_status != YV_OK) return g_init_status;
  CUtensorMap tmA, tmB;
  int rc = make_tmap(&tmA, A, (uint64_t)M, (uint64_t)K, BM);
  if (rc) return rc;
  rc = make_tmap(&tmB, W, (uint64_t)N, (uint64_t)K, BN);
  if (rc) return rc;
  const int tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
  const int grid = tiles < yv_num_sms() ? tiles : yv_num_sms();
  YvProf prof(YV_K_GEMM, s, 2.0 * M * (double)N * K);
  gemm_bf16_kernel<<<grid, GEMM_THREADS, SMEM_BYTES, s>>>(tmA, tmB, ep, M, N, K);
  YV_LAUNCH_CHECK();
  return YV_OK;
}

YV_EXPORT int yv_gemm_bf16(const void *A, const void *W, const float *bias, const void *residual, void *D, int M,
                           int N, int K, int epi, int out_f32, yv_stream_t stream) {
  YV_REQUIRE(epi == YV_EPI_NONE || epi == YV_EPI_GELU, "yv_gemm_bf16: unknown epilogue %d", epi);
  GemmEpilogue ep;
  ep.out = D;
  ep.bias = bias;
  ep.residual = residual;
  ep.res_f32 = false;
  ep.pos = nullptr;
  ep.group_rows = 0;
  ep.gelu = (epi == YV_EPI_GELU);
  ep.out_f32 = out_f32 != 0;
  return yv_gemm_launch(A, W, M, N, K, ep, (cudaStream_t)stream);
}
