// api.cu -- library-level entry points and error plumbing of libyv8vit.so.
#include "common.cuh"
#include <string.h>

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

int yv_num_sms() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
  }
  return sms;
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
