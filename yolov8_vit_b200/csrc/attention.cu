// attention.cu -- K7 for LONG sequences (T > 256: ViT-B/8's 785 tokens, row f4): softmax(Q K^T / sqrt(dh)) V.
//
// timm Attention.forward (vit_base_patch8_224, the reference-literal model of utils/class_config.py:21):
// qkv [n, T, 3, heads, dh] -> per head softmax(q k^T * dh^-0.5) v.
//
// mma.sync.m16n8k16 (bf16 x bf16 -> f32) flash-style kernel: a CTA owns up to 160 query rows of one
// (crop, head) -- one warp per 16 rows, Q staged once in shared memory -- and walks ALL keys in blocks of 64 with
// the online-softmax recurrence (running max / sum in registers, exp2 domain); every K / V block is staged
// through padded shared memory (ldmatrix conflict-free), P stays in registers as the A operand of P V.
// K / V blocks go through a 3-deep cp.async ring (blocks kb+1, kb+2 in flight while block kb is in the tensor cores,
// one __syncthreads per block); the query
// rows are split evenly over the CTAs of a (crop, head) (785 tokens: 5 CTAs x 10 warps = 800 rows, not 4 x 256) and
// warps whose rows are all past T only help with the staging.  T <= 256 never comes here: the
// tcgen05 kernels of attention_sm100.cu handle it.  This is the functional path for the reference-literal
// B/8 model; a tcgen05 multi-block variant is the "next" item of DESIGN.md.
#include "common.cuh"
#include "vit_ops.cuh"

namespace {

constexpr int ATT_DH = 64;
constexpr int ATT_LD = 72;         // padded smem row (bf16 elements): 144 B, ldmatrix conflict-free
constexpr int ATT_KB = 64;         // keys per block

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const void *p) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], const void *p) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t *>(&h);
}

__device__ __forceinline__ float ex2_approx(float x) {      // MUFU.EX2; -inf -> +0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void cp_async16(void *dst, const void *src) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(a), "l"(src) : "memory");
}

constexpr int ATT_MAXW = 10;       // warps per CTA: 2 CTAs x 10 warps per SM at <= 102 registers

__global__ void __launch_bounds__(ATT_MAXW * 32, 2)
attention_long_kernel(const __nv_bfloat16 *__restrict__ qkv, __nv_bfloat16 *__restrict__ out, int T, int heads,
                      float scale_log2e) {
  extern __shared__ __align__(16) uint8_t att_smem[];
  const int nw = blockDim.x >> 5;                 // warps = 16-row query groups of this CTA
  const int nkb = (T + ATT_KB - 1) / ATT_KB;
  const int qrows = nw * 16;
  __nv_bfloat16 *Qs = reinterpret_cast<__nv_bfloat16 *>(att_smem);
  __nv_bfloat16 *KV = Qs + (size_t)qrows * ATT_LD;          // 3 buffers x (K block, V block) of 64 keys
  constexpr size_t KV_BUF = (size_t)2 * ATT_KB * ATT_LD;

  // the query blocks of one (crop, head) are adjacent in launch order, so its K / V stay in L2 for all of them
  const int head = blockIdx.y, crop = blockIdx.z, q0 = blockIdx.x * qrows;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int D = heads * ATT_DH;
  const __nv_bfloat16 *base = qkv + (size_t)crop * T * 3 * D + (size_t)head * ATT_DH;

  // stage this CTA's Q rows: 8 x 16-byte chunks per row; rows past T are zero
  const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
  for (int i = tid; i < qrows * 8; i += blockDim.x) {
    const int r = i >> 3, c = i & 7;
    uint4 v = zero;
    if (q0 + r < T) v = *reinterpret_cast<const uint4 *>(base + (size_t)(q0 + r) * 3 * D + c * 8);
    *reinterpret_cast<uint4 *>(Qs + (size_t)r * ATT_LD + c * 8) = v;
  }
  __syncthreads();

  // Q fragments are re-read from shared memory at every k-step (4 extra ldmatrix per key block) instead of living
  // in 16 registers: that is what lets two CTAs share an SM
  const __nv_bfloat16 *qsrc = Qs + (size_t)(warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * ATT_LD + (lane >> 4) * 8;

  float o[8][4];
#pragma unroll
  for (int j = 0; j < 8; ++j) { o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f; }
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  const bool active = q0 + warp * 16 < T;         // warp-uniform: warps past the last token only stage

  auto stage = [&](int kb) {                      // K / V block kb -> buffer kb % 3 (asynchronous copies)
    __nv_bfloat16 *Kd = KV + (size_t)(kb % 3) * KV_BUF, *Vd = Kd + (size_t)ATT_KB * ATT_LD;
    for (int i = tid; i < ATT_KB * 8; i += blockDim.x) {
      const int r = i >> 3, c = i & 7, key = kb * ATT_KB + r;
      if (key < T) {
        cp_async16(Kd + (size_t)r * ATT_LD + c * 8, base + (size_t)key * 3 * D + D + c * 8);
        cp_async16(Vd + (size_t)r * ATT_LD + c * 8, base + (size_t)key * 3 * D + 2 * D + c * 8);
      } else {
        *reinterpret_cast<uint4 *>(Kd + (size_t)r * ATT_LD + c * 8) = zero;
        *reinterpret_cast<uint4 *>(Vd + (size_t)r * ATT_LD + c * 8) = zero;
      }
    }
  };
  stage(0);
  asm volatile("cp.async.commit_group;" ::: "memory");
  if (nkb > 1) stage(1);
  asm volatile("cp.async.commit_group;" ::: "memory");

  for (int kb = 0; kb < nkb; ++kb) {
    asm volatile("cp.async.wait_group 1;" ::: "memory");      // this thread's copies of block kb have landed
    __syncthreads();          // ... and everybody else's; block kb-1 is no longer read: its buffer takes block kb+2
    if (kb + 2 < nkb) stage(kb + 2);
    asm volatile("cp.async.commit_group;" ::: "memory");
    if (!active) continue;
    const __nv_bfloat16 *Ks = KV + (size_t)(kb % 3) * KV_BUF, *Vs = Ks + (size_t)ATT_KB * ATT_LD;
    float s[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) { s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f; }
    const __nv_bfloat16 *Kb = Ks;
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      uint32_t qf[4];
      ldsm_x4(qf, qsrc + ks * 16);
#pragma unroll
      for (int np = 0; np < 4; ++np) {          // n-tile pairs (2 x 8 keys)
        uint32_t kf[4];
        const int r = np * 16 + (lane & 7) + (lane >> 4) * 8;
        const int c = ks * 16 + ((lane >> 3) & 1) * 8;
        ldsm_x4(kf, Kb + (size_t)r * ATT_LD + c);
        mma_bf16(s[2 * np], qf, kf[0], kf[1]);
        mma_bf16(s[2 * np + 1], qf, kf[2], kf[3]);
      }
    }
    // keys >= T exist only in the last block: mask there (warp-uniform branch); the block maximum is taken on the raw
    // scores (scale > 0) and the scale is folded into the exponent's FMA
    if (kb == nkb - 1) {
      const int key0 = kb * ATT_KB + (lane & 3) * 2;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int k = key0 + j * 8;
        if (k >= T) { s[j][0] = -INFINITY; s[j][2] = -INFINITY; }
        if (k + 1 >= T) { s[j][1] = -INFINITY; s[j][3] = -INFINITY; }
      }
    }
    float bm0 = -INFINITY, bm1 = -INFINITY;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      bm0 = fmaxf(bm0, fmaxf(s[j][0], s[j][1]));
      bm1 = fmaxf(bm1, fmaxf(s[j][2], s[j][3]));
    }
    bm0 = fmaxf(bm0, __shfl_xor_sync(0xffffffffu, bm0, 1));
    bm0 = fmaxf(bm0, __shfl_xor_sync(0xffffffffu, bm0, 2));
    bm1 = fmaxf(bm1, __shfl_xor_sync(0xffffffffu, bm1, 1));
    bm1 = fmaxf(bm1, __shfl_xor_sync(0xffffffffu, bm1, 2));
    const float mn0 = fmaxf(m0, bm0 * scale_log2e), mn1 = fmaxf(m1, bm1 * scale_log2e);
    const float corr0 = ex2_approx(m0 - mn0), corr1 = ex2_approx(m1 - mn1);
    m0 = mn0;
    m1 = mn1;
    float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s[j][0] = ex2_approx(fmaf(s[j][0], scale_log2e, -mn0));
      s[j][1] = ex2_approx(fmaf(s[j][1], scale_log2e, -mn0));
      s[j][2] = ex2_approx(fmaf(s[j][2], scale_log2e, -mn1));
      s[j][3] = ex2_approx(fmaf(s[j][3], scale_log2e, -mn1));
      rs0 += s[j][0] + s[j][1];
      rs1 += s[j][2] + s[j][3];
    }
    l0 = l0 * corr0 + rs0;
    l1 = l1 * corr1 + rs1;
    if (__any_sync(0xffffffffu, corr0 != 1.f || corr1 != 1.f)) {   // the running maxima settle after a few blocks
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        o[j][0] *= corr0; o[j][1] *= corr0;
        o[j][2] *= corr1; o[j][3] *= corr1;
      }
    }
    // O += P V
    const __nv_bfloat16 *Vb = Vs;
#pragma unroll
    for (int t = 0; t < 4; ++t) {               // k16 steps over the 64 keys of the block
      uint32_t pf[4];
      pf[0] = pack_bf16(s[2 * t][0], s[2 * t][1]);
      pf[1] = pack_bf16(s[2 * t][2], s[2 * t][3]);
      pf[2] = pack_bf16(s[2 * t + 1][0], s[2 * t + 1][1]);
      pf[3] = pack_bf16(s[2 * t + 1][2], s[2 * t + 1][3]);
#pragma unroll
      for (int dp = 0; dp < 4; ++dp) {          // d n-tile pairs
        uint32_t vf[4];
        const int r = t * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
        const int c = dp * 16 + (lane >> 4) * 8;
        ldsm_x4_t(vf, Vb + (size_t)r * ATT_LD + c);
        mma_bf16(o[2 * dp], pf, vf[0], vf[1]);
        mma_bf16(o[2 * dp + 1], pf, vf[2], vf[3]);
      }
    }
  }

  if (!active) return;
  // finalise: row sums across the quad, normalise, store bf16
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float inv0 = 1.f / l0, inv1 = 1.f / l1;
  const int r0 = q0 + warp * 16 + (lane >> 2), r1 = r0 + 8;
  __nv_bfloat16 *ob = out + (size_t)crop * T * D + (size_t)head * ATT_DH + (lane & 3) * 2;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (r0 < T) *reinterpret_cast<uint32_t *>(ob + (size_t)r0 * D + j * 8) = pack_bf16(o[j][0] * inv0, o[j][1] * inv0);
    if (r1 < T) *reinterpret_cast<uint32_t *>(ob + (size_t)r1 * D + j * 8) = pack_bf16(o[j][2] * inv1, o[j][3] * inv1);
  }
}

}  // namespace

int yv_attention_long_launch(const __nv_bfloat16 *qkv, __nv_bfloat16 *out, int n, int T, int heads, int dh,
                             cudaStream_t s) {
  YV_REQUIRE(dh == ATT_DH, "attention: head dim %d unsupported (need %d)", dh, ATT_DH);
  YV_REQUIRE(n >= 1 && n <= 65535 && heads >= 1 && T >= 1, "attention: bad shape");
  const int groups = (T + 15) / 16;                       // 16-row query groups, spread evenly over the CTAs
  const int qblocks = (groups + ATT_MAXW - 1) / ATT_MAXW;
  const int nw = (groups + qblocks - 1) / qblocks;        // <= ATT_MAXW warps
  const size_t smem = (size_t)(nw * 16 + 6 * ATT_KB) * ATT_LD * sizeof(__nv_bfloat16);
  int rc = yv_device_once(YV_INIT_ATTN_LONG, []() -> int {
    YV_CUDA(cudaFuncSetAttribute(attention_long_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    return YV_OK;
  });
  if (rc) return rc;
  dim3 grid(qblocks, heads, n);
  const float scale_log2e = (1.0f / sqrtf((float)dh)) * 1.4426950408889634f;
  YvProf prof(YV_K_ATTN, s, 4.0 * n * heads * (double)T * T * dh);
  attention_long_kernel<<<grid, nw * 32, smem, s>>>(qkv, out, T, heads, scale_log2e);
  YV_LAUNCH_CHECK();
  return YV_OK;
}
