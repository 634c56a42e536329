// common.cuh -- shared host/device helpers for libyv8vit.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/yv8vit.h"

#ifndef YV_EXPORT
#define YV_EXPORT extern "C" __attribute__((visibility("default")))
#endif

// ---- error plumbing (api.cu owns the thread-local buffer) --------------------
void yv_set_error(const char *fmt, ...);
int yv_cuda_fail(cudaError_t e, const char *what, const char *file, int line);

#define YV_CUDA(call)                                                          \
  do {                                                                         \
    cudaError_t e__ = (call);                                                  \
    if (e__ != cudaSuccess) return yv_cuda_fail(e__, #call, __FILE__, __LINE__); \
  } while (0)

#define YV_LAUNCH_CHECK()                                                      \
  do {                                                                         \
    cudaError_t e__ = cudaGetLastError();                                      \
    if (e__ != cudaSuccess)                                                    \
      return yv_cuda_fail(e__, "kernel launch", __FILE__, __LINE__);           \
  } while (0)

#define YV_REQUIRE(cond, ...)                                                  \
  do {                                                                         \
    if (!(cond)) {                                                             \
      yv_set_error(__VA_ARGS__);                                               \
      return YV_EINVAL;                                                        \
    }                                                                          \
  } while (0)

static inline size_t yv_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
int yv_num_sms();                                  // of the CURRENT device (cached per device)
// SMs a GEMM / an attention launch may occupy (yv_tune_sm_budget; 0 = all): lets the two run side by side on
// disjoint SMs from two streams
int yv_gemm_sm_budget();
int yv_attention_sm_budget();
// one-time setup per (slot, current device), thread-safe: cudaFuncSetAttribute and friends (api.cu)
enum YvInitSlot { YV_INIT_GEMM = 0, YV_INIT_ATTN, YV_INIT_ATTN_FA, YV_INIT_COUNT };
int yv_device_once(int slot, int (*fn)());

// ---- launch accounting / optional CUDA-event profiling (api.cu) ------------------------------
enum YvKernelKind { YV_K_DECODE = 0, YV_K_NMS, YV_K_BUILD, YV_K_CROP, YV_K_GEMM, YV_K_LN, YV_K_ATTN, YV_K_HEAD, YV_K_MISC, YV_K_COUNT };
void yv_prof_begin(int kind, cudaStream_t s, double work);
void yv_prof_end(int kind, cudaStream_t s);
// RAII: counts one launch of `kind` (always) and, when profiling is on, brackets it with CUDA
// events on the launching stream.  `work` = algorithmic bytes (HBM-bound kernels) or flops (GEMM).
struct YvProf {
  int kind; cudaStream_t s;
  YvProf(int k, cudaStream_t st, double work) : kind(k), s(st) { yv_prof_begin(k, st, work); }
  ~YvProf() { yv_prof_end(kind, s); }
};

// ---- device helpers ------------------------------------------------------------
__device__ __forceinline__ float4 ld_stream_f4(const float *p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ float ld_stream_f1(const float *p) {
  float r;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}
