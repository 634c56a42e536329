// ubench_coissue.cu -- do two warps of one SM sub-partition overlap?  (What the strip attention's two softmax warps per
// sub-partition rely on: one in its max pass -- FMNMX3 -- while the other is in its exp pass -- MUFU.EX2 + FFMA2.)
// One CTA per SM, 8 warps: warps 0-3 run stream A, warps 4-7 stream B; warp w and warp w + 4 share sub-partition w % 4.
// Each stream is an unrolled loop of independent instructions of one kind; clock64() per warp; the table gives clocks per
// warp-instruction of each stream alone and with the other stream running on the same sub-partition.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_coissue tools/ubench_coissue.cu && ./tools/ubench_coissue
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

enum Kind { NONE = 0, MUFU = 1, FMNMX3 = 2, FFMA2 = 3, FADD = 4, F2FP = 5 };

template <int K>
__device__ __forceinline__ void body(float (&v)[16], float s) {
  if constexpr (K == MUFU) {
#pragma unroll
    for (int i = 0; i < 16; ++i) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(v[i]));
  } else if constexpr (K == FMNMX3) {
#pragma unroll
    for (int i = 0; i < 16; ++i) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(v[i]) : "f"(v[(i + 5) & 15]), "f"(s));
  } else if constexpr (K == FFMA2) {
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
      uint64_t a, b;
      asm volatile("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(v[i]), "f"(v[i + 1]));
      asm volatile("mov.b64 %0, {%1, %1};" : "=l"(b) : "f"(s));
      asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(a) : "l"(b));
      asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(v[i]), "=f"(v[i + 1]) : "l"(a));
    }
  } else if constexpr (K == FADD) {
#pragma unroll
    for (int i = 0; i < 16; ++i) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(v[i]) : "f"(s));
  } else if constexpr (K == F2FP) {
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
      uint32_t w;
      asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w) : "f"(v[i]), "f"(v[i + 1]));
      v[i] = __uint_as_float(w);
    }
  }
}
template <int K> constexpr int ops_per_body() { return K == FFMA2 || K == F2FP ? 8 : 16; }

template <int KA, int KB>
__global__ void __launch_bounds__(256, 1) coissue(float *out, long long *clk, int iters, float s) {
  const int warp = threadIdx.x >> 5;
  float v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = -0.001f * (float)(threadIdx.x + i);
  __syncthreads();
  const long long t0 = clock64();
  if (warp < 4) {
    if constexpr (KA != NONE)
      for (int it = 0; it < iters; ++it) body<KA>(v, s);
  } else {
    if constexpr (KB != NONE)
      for (int it = 0; it < iters; ++it) body<KB>(v, s);
  }
  const long long t1 = clock64();
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) acc += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if ((threadIdx.x & 31) == 0) clk[blockIdx.x * 8 + warp] = t1 - t0;
}

static const char *kname(int k) {
  const char *n[] = {"none", "MUFU.EX2", "FMNMX3", "FFMA2", "FADD", "F2FP.BF16x2"};
  return n[k];
}

template <int KA, int KB>
static void run(float *out, long long *clk, long long *h, int nsm) {
  const int iters = 4000;
  coissue<KA, KB><<<nsm, 256>>>(out, clk, 100, -0.5f);
  coissue<KA, KB><<<nsm, 256>>>(out, clk, iters, -0.5f);
  cudaDeviceSynchronize();
  cudaMemcpy(h, clk, sizeof(long long) * nsm * 8, cudaMemcpyDeviceToHost);
  double a = 0, b = 0;
  for (int i = 0; i < nsm; ++i)
    for (int w = 0; w < 8; ++w) (w < 4 ? a : b) += (double)h[i * 8 + w];
  a /= nsm * 4.0;
  b /= nsm * 4.0;
  printf("{\"A\": \"%s\", \"B\": \"%s\", \"A_clk_per_instr\": %.2f, \"B_clk_per_instr\": %.2f}\n", kname(KA), kname(KB),
         KA ? a / ((double)iters * ops_per_body<KA>()) : 0.0, KB ? b / ((double)iters * ops_per_body<KB>()) : 0.0);
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int nsm = p.multiProcessorCount;
  float *out;
  long long *clk, *h = new long long[nsm * 8];
  cudaMalloc(&out, sizeof(float) * nsm * 256);
  cudaMalloc(&clk, sizeof(long long) * nsm * 8);
  run<MUFU, NONE>(out, clk, h, nsm);
  run<FMNMX3, NONE>(out, clk, h, nsm);
  run<FFMA2, NONE>(out, clk, h, nsm);
  run<FADD, NONE>(out, clk, h, nsm);
  run<F2FP, NONE>(out, clk, h, nsm);
  run<MUFU, MUFU>(out, clk, h, nsm);
  run<FMNMX3, MUFU>(out, clk, h, nsm);
  run<FFMA2, MUFU>(out, clk, h, nsm);
  run<FADD, MUFU>(out, clk, h, nsm);
  run<F2FP, MUFU>(out, clk, h, nsm);
  run<FMNMX3, FFMA2>(out, clk, h, nsm);
  run<FADD, FADD>(out, clk, h, nsm);
  run<FMNMX3, FMNMX3>(out, clk, h, nsm);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 1; }
  return 0;
}
