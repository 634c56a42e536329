// api.cu -- library-level entry points and error plumbing of libyv8vit.so.
#include "common.cuh"
#include <string.h>
#include <atomic>
#include <mutex>
#include <nvtx3/nvToolsExt.h>   // header-only NVTX v3: a no-op unless a profiler injects itself

static thread_local char g_err[512] = "";

void yv_set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int yv_cuda_fail(cudaError_t e, const char *what, const char *file, int line) {
  yv_set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
  return YV_ECUDA;
}

// ---- per-device state: one process may drive several GPUs (ViTModel(device='cuda:1')) ------------------
constexpr int YV_MAX_DEVICES = 64;

int yv_num_sms() {
  static std::atomic<int> sms[YV_MAX_DEVICES];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= YV_MAX_DEVICES) return 148;
  int v = sms[dev].load(std::memory_order_relaxed);
  if (v == 0) {
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    sms[dev].store(v, std::memory_order_relaxed);
  }
  return v;
}

// Runs `fn` once per (slot, current device) -- cudaFuncSetAttribute(MaxDynamicSharedMemorySize) applies to the
// current device only -- under a lock, so concurrent first use from several threads is safe.  A failing `fn`
// is retried by the next caller.
int yv_device_once(int slot, int (*fn)()) {
  static std::mutex mu;
  static bool done[YV_INIT_COUNT][YV_MAX_DEVICES];
  int dev = 0;
  YV_CUDA(cudaGetDevice(&dev));
  YV_REQUIRE(slot >= 0 && slot < YV_INIT_COUNT && dev >= 0 && dev < YV_MAX_DEVICES, "yv_device_once: slot %d / device %d out of range", slot, dev);
  std::lock_guard<std::mutex> lk(mu);
  if (done[slot][dev]) return YV_OK;
  const int rc = fn();
  if (rc == YV_OK) done[slot][dev] = true;
  return rc;
}

YV_EXPORT int yv_version(void) { return 100; }

YV_EXPORT const char *yv_last_error(void) { return g_err; }

YV_EXPORT int yv_check_device(int dev) {
  int major = 0, minor = 0;
  YV_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  YV_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
  if (major != 10) {
    yv_set_error("device %d is sm_%d%d; libyv8vit is built for sm_100a only", dev, major, minor);
    return YV_ENODEV;
  }
  return YV_OK;
}

// ---- launch accounting + optional event profiling + NVTX ranges -------------------------------------
#include <vector>
namespace {
struct ProfState {
  std::mutex mu;
  bool on = false;
  long long launches[YV_K_COUNT] = {0};
  double work[YV_K_COUNT] = {0};
  std::vector<cudaEvent_t> ev[YV_K_COUNT];     // start, end, start, end, ...
  std::vector<cudaEvent_t> pool;
} g_prof;
cudaEvent_t prof_event() {
  if (!g_prof.pool.empty()) { cudaEvent_t e = g_prof.pool.back(); g_prof.pool.pop_back(); return e; }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}
}  // namespace

void yv_prof_begin(int kind, cudaStream_t s, double work) {
  nvtxRangePushA(yv_prof_kind_name(kind));       // one NVTX range per launch, named after the kernel class
  std::lock_guard<std::mutex> lk(g_prof.mu);
  g_prof.launches[kind] += 1;
  g_prof.work[kind] += work;
  if (g_prof.on) {
    cudaEvent_t e = prof_event();
    cudaEventRecord(e, s);
    g_prof.ev[kind].push_back(e);
  }
}
void yv_prof_end(int kind, cudaStream_t s) {
  {
    std::lock_guard<std::mutex> lk(g_prof.mu);
    if (g_prof.on && (g_prof.ev[kind].size() & 1)) {
      cudaEvent_t e = prof_event();
      cudaEventRecord(e, s);
      g_prof.ev[kind].push_back(e);
    }
  }
  nvtxRangePop();
}

// on != 0: reset counters and start bracketing every kernel launch with CUDA events; 0: stop.
YV_EXPORT int yv_prof_enable(int on) {
  std::lock_guard<std::mutex> lk(g_prof.mu);
  for (int k = 0; k < YV_K_COUNT; ++k) {
    for (cudaEvent_t e : g_prof.ev[k]) g_prof.pool.push_back(e);
    g_prof.ev[k].clear();
    g_prof.launches[k] = 0;
    g_prof.work[k] = 0;
  }
  g_prof.on = on != 0;
  return YV_OK;
}
YV_EXPORT int yv_prof_kinds(void) { return YV_K_COUNT; }
YV_EXPORT const char *yv_prof_kind_name(int kind) {
  static const char *names[YV_K_COUNT] = {"decode_filter", "nms", "build_crops", "crop_resize", "gemm_tcgen05",
                                          "layernorm", "attention", "wrapper_head", "misc"};
  return (kind >= 0 && kind < YV_K_COUNT) ? names[kind] : "?";
}
// Synchronises the device, then fills per-kind totals since yv_prof_enable: device milliseconds
// (sum of event-bracketed launch durations; 0 when profiling was off), launches, work.
YV_EXPORT int yv_prof_read(double *ms, long long *launches, double *work) {
  YV_CUDA(cudaDeviceSynchronize());
  std::lock_guard<std::mutex> lk(g_prof.mu);
  for (int k = 0; k < YV_K_COUNT; ++k) {
    double t = 0;
    for (size_t i = 0; i + 1 < g_prof.ev[k].size(); i += 2) {
      float f = 0;
      if (cudaEventElapsedTime(&f, g_prof.ev[k][i], g_prof.ev[k][i + 1]) == cudaSuccess) t += f;
    }
    if (ms) ms[k] = t;
    if (launches) launches[k] = g_prof.launches[k];
    if (work) work[k] = g_prof.work[k];
  }
  return YV_OK;
}

// ---- SM budgets (measurement / scheduling knob) ----------------------------------------------------------------
namespace { int g_gemm_sms = 0, g_attn_sms = 0; }
int yv_gemm_sm_budget() { const int n = yv_num_sms(); return g_gemm_sms > 0 && g_gemm_sms < n ? g_gemm_sms : n; }
int yv_attention_sm_budget() { const int n = yv_num_sms(); return g_attn_sms > 0 && g_attn_sms < n ? g_attn_sms : n; }
YV_EXPORT int yv_tune_sm_budget(int gemm_sms, int attention_sms) {
  YV_REQUIRE(gemm_sms >= 0 && attention_sms >= 0 && gemm_sms % 2 == 0, "yv_tune_sm_budget: bad budget %d / %d (GEMM SMs must be even)", gemm_sms, attention_sms);
  g_gemm_sms = gemm_sms;
  g_attn_sms = attention_sms;
  return YV_OK;
}
