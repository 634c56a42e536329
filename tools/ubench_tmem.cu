// ubench_tmem.cu -- micro-benchmarks that size the attention kernel's softmax stage on B200 (sm_100a):
//   * tcgen05.ld throughput per SM (TMEM -> registers) for 4 / 8 / 16 warps and the x8 / x16 / x32 / x64 shapes
//   * tcgen05.st throughput (registers -> TMEM)
//   * MUFU ex2.approx throughput per SM, FFMA2 throughput, 3-input fmax throughput
// One CTA per SM, clock64() around a long unrolled loop; prints bytes (or ops) per clock per SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_tmem tools/ubench_tmem.cu && ./tools/ubench_tmem
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int X>
__device__ __forceinline__ void tmem_ld(uint32_t taddr, uint32_t &acc) {
  if constexpr (X == 8) {
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 8; ++i) acc ^= r[i];
  } else if constexpr (X == 16) {
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) acc ^= r[i];
  } else {
    uint32_t r[32];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
                 "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                   "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                   "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) acc ^= r[i];
  }
}

// two loads in flight before one wait (the pattern of the softmax loops)
template <int X>
__device__ __forceinline__ void tmem_ld2(uint32_t t0, uint32_t t1, uint32_t &acc) {
  static_assert(X == 16, "x16 only");
  uint32_t r[16], q[16];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                 "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(t0) : "memory");
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(q[0]), "=r"(q[1]), "=r"(q[2]), "=r"(q[3]), "=r"(q[4]), "=r"(q[5]), "=r"(q[6]), "=r"(q[7]),
                 "=r"(q[8]), "=r"(q[9]), "=r"(q[10]), "=r"(q[11]), "=r"(q[12]), "=r"(q[13]), "=r"(q[14]), "=r"(q[15])
               : "r"(t1) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) acc ^= r[i] ^ q[i];
}

__device__ __forceinline__ void tmem_st8(uint32_t taddr, uint32_t v) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(taddr), "r"(v) : "memory");
}

// mode 0: ld x8, 1: ld x16, 2: ld x32, 3: two x16 in flight, 4: st x8, 5: ex2, 6: ffma2, 7: fmax3, 8: fmax (2-input),
// 9: cvt.rn.bf16x2.f32 (F2FP, 32 per iteration), 10: 32 ex2 + 16 cvt.rn.bf16x2.f32 interleaved (do they share a pipe?)
template <int MODE>
__global__ void __launch_bounds__(1024, 1) bench(long long *cycles, uint32_t *sink, int iters) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t acc = lane;
  float f0 = 1.0f + lane * 1e-3f, f1 = 0.5f, f2 = 0.25f, f3 = 0.125f;
  __syncthreads();
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    const uint32_t col = (uint32_t)((i * 32 + (warp >> 2) * 64) & 255);
    if (MODE == 0) { tmem_ld<8>(base + col, acc); tmem_ld<8>(base + col + 8, acc); tmem_ld<8>(base + col + 16, acc); tmem_ld<8>(base + col + 24, acc); }
    if (MODE == 1) { tmem_ld<16>(base + col, acc); tmem_ld<16>(base + col + 16, acc); }
    if (MODE == 2) { tmem_ld<32>(base + col, acc); }
    if (MODE == 3) { tmem_ld2<16>(base + col, base + col + 16, acc); }
    if (MODE == 4) { tmem_st8(base + col, acc); tmem_st8(base + col + 8, acc); tmem_st8(base + col + 16, acc); tmem_st8(base + col + 24, acc); }
    if (MODE == 5) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f0)); asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f1));
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f2)); asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f3));
      }
    }
    if (MODE == 6) {
      uint64_t a, b;
      asm volatile("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(f0), "f"(f1));
      asm volatile("mov.b64 %0, {%1, %2};" : "=l"(b) : "f"(f2), "f"(f3));
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(a) : "l"(b));
        asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(b) : "l"(a));
      }
      asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(f0), "=f"(f1) : "l"(a));
      asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(f2), "=f"(f3) : "l"(b));
    }
    if (MODE == 7) {
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(f0) : "f"(f1), "f"(f2));
        asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(f3) : "f"(f1), "f"(f2));
      }
    }
    if (MODE == 9) {
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        uint32_t w0, w1;
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w0) : "f"(f0), "f"(f1));
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w1) : "f"(f2), "f"(f3));
        acc ^= w0 ^ w1;
      }
    }
    if (MODE == 10) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        uint32_t w0, w1;
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f0)); asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f1));
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f2)); asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f3));
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w0) : "f"(f0), "f"(f1));
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w1) : "f"(f2), "f"(f3));
        acc ^= w0 ^ w1;
      }
    }
    if (MODE == 8) {
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        asm volatile("max.f32 %0, %0, %1;" : "+f"(f0) : "f"(f1));
        asm volatile("max.f32 %0, %0, %1;" : "+f"(f3) : "f"(f2));
      }
    }
  }
  if (MODE == 4) asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  const long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc ^ __float_as_uint(f0 + f1 + f2 + f3);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(slot), "r"(512u) : "memory");
}

template <int MODE>
double run(int warps, int iters, long long *d_cyc, uint32_t *d_sink) {
  bench<MODE><<<148, warps * 32>>>(d_cyc, d_sink, iters);
  bench<MODE><<<148, warps * 32>>>(d_cyc, d_sink, iters);
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("CUDA error mode %d: %s\n", MODE, cudaGetErrorString(cudaGetLastError())); return -1; }
  long long h[148];
  cudaMemcpy(h, d_cyc, sizeof(h), cudaMemcpyDeviceToHost);
  double s = 0;
  for (int i = 0; i < 148; ++i) s += (double)h[i];
  return s / 148.0;
}

int main() {
  long long *d_cyc; uint32_t *d_sink;
  cudaMalloc(&d_cyc, 148 * sizeof(long long));
  cudaMalloc(&d_sink, 148 * 1024 * sizeof(uint32_t));
  const int iters = 4000;
  printf("{");
  for (int warps : {4, 8, 16, 32}) {
    const double per_iter_bytes = (double)warps * 32 * 32 * 4;      // every mode 0-4 moves 32 columns x 32 lanes x 4 B per warp per iteration
    double c;
    c = run<0>(warps, iters, d_cyc, d_sink); printf("\"ld_x8_w%d_B_per_clk\": %.1f, ", warps, per_iter_bytes * iters / c);
    c = run<1>(warps, iters, d_cyc, d_sink); printf("\"ld_x16_w%d_B_per_clk\": %.1f, ", warps, per_iter_bytes * iters / c);
    c = run<2>(warps, iters, d_cyc, d_sink); printf("\"ld_x32_w%d_B_per_clk\": %.1f, ", warps, per_iter_bytes * iters / c);
    c = run<3>(warps, iters, d_cyc, d_sink); printf("\"ld_2x16_w%d_B_per_clk\": %.1f, ", warps, per_iter_bytes * iters / c);
    c = run<4>(warps, iters, d_cyc, d_sink); printf("\"st_x8_w%d_B_per_clk\": %.1f, ", warps, per_iter_bytes * iters / c);
    c = run<5>(warps, iters, d_cyc, d_sink); printf("\"ex2_w%d_per_clk\": %.2f, ", warps, (double)warps * 32 * 32 * iters / c);
    c = run<6>(warps, iters, d_cyc, d_sink); printf("\"ffma2_w%d_lanes_per_clk\": %.2f, ", warps, (double)warps * 32 * 32 * 2 * iters / c);
    c = run<7>(warps, iters, d_cyc, d_sink); printf("\"fmax3_w%d_per_clk\": %.2f, ", warps, (double)warps * 32 * 32 * iters / c);
    c = run<8>(warps, iters, d_cyc, d_sink); printf("\"fmax2_w%d_per_clk\": %.2f, ", warps, (double)warps * 32 * 32 * iters / c);
    c = run<9>(warps, iters, d_cyc, d_sink); printf("\"cvt_bf16x2_w%d_per_clk\": %.2f, ", warps, (double)warps * 32 * 32 * iters / c);
    c = run<10>(warps, iters, d_cyc, d_sink); printf("\"ex2_32_plus_cvt_16_w%d_clk_per_iter_per_smsp_warpgroup\": %.1f, ", warps, c / iters);
  }
  printf("\"iters\": %d}\n", iters);
  return 0;
}
