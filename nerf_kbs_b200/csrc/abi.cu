// C-ABI plumbing shared by every entry point: version, per-thread error string, device query.
#include <stdarg.h>
#include <string.h>
#include "common.cuh"

#define NKBS_VERSION 200

static thread_local char g_err[512] = "";

void nkbs_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int nkbs_num_sms() {
  static thread_local int cached_dev = -1;
  static thread_local int cached_sms = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
    cached_dev = dev;
    cached_sms = sms;
  }
  return cached_sms;
}

extern "C" int nkbs_version(void) { return NKBS_VERSION; }
extern "C" const char* nkbs_last_error(void) { return g_err; }

// memset node on the caller's stream (a CUDA-graph memset node when captured, not a kernel launch)
extern "C" int nkbs_zero_async(void* dev_ptr, int64_t bytes, void* stream) {
  NKBS_CHECK_ARG(bytes >= 0 && (dev_ptr || bytes == 0), "nkbs_zero_async: bad arguments");
  if (bytes == 0) return 0;
  cudaError_t e = cudaMemsetAsync(dev_ptr, 0, (size_t)bytes, (cudaStream_t)stream);
  NKBS_CHECK_ARG(e == cudaSuccess, "nkbs_zero_async: %s", cudaGetErrorString(e));
  return 0;
}
