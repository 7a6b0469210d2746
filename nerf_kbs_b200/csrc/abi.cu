// C-ABI plumbing shared by every entry point: version, per-thread error string, device query.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <mutex>
#include <vector>
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

bool nkbs_pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("NKBS_PDL");
    return e && e[0] == '1';
  }();
  return on;
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

// ---- deterministic gradient accumulation: registry of (float buffer, fixed-point shadow) pairs ----
static std::mutex g_fixed_mu;
static std::vector<FixedMap> g_fixed;

FixedMap nkbs_fixed_map_for(const void* p) {
  FixedMap none{nullptr, nullptr, 0, 0.f};
  if (!p) return none;
  std::lock_guard<std::mutex> lk(g_fixed_mu);
  const float* q = static_cast<const float*>(p);
  for (const FixedMap& m : g_fixed)
    if (q >= m.base && q < m.base + m.n) return m;
  return none;
}

extern "C" int nkbs_fixed_point_register(float* base, int64_t n, int64_t* shadow, int32_t log2_scale) {
  NKBS_CHECK_ARG(base && n >= 0, "nkbs_fixed_point_register: bad arguments");
  NKBS_CHECK_ARG(!shadow || (log2_scale >= 16 && log2_scale <= 60), "nkbs_fixed_point_register: log2_scale must be in [16, 60]");
  std::lock_guard<std::mutex> lk(g_fixed_mu);
  for (size_t i = 0; i < g_fixed.size(); ++i)
    if (g_fixed[i].base == base) {
      g_fixed.erase(g_fixed.begin() + i);
      break;
    }
  if (shadow && n > 0) g_fixed.push_back(FixedMap{base, reinterpret_cast<long long*>(shadow), (long long)n, ldexpf(1.f, log2_scale)});
  return 0;
}

// Mostly-zero input (a step touches a few percent of a hash table): 16-byte loads, four in flight per thread, and the
// float buffer / the shadow are only written where the shadow is not zero.
__global__ void __launch_bounds__(256) fixed_point_flush_kernel(long long* acc, float* dst, int64_t n, double inv_scale) {
  nkbs_pdl_sync();
  const int64_t pairs = n >> 1, stride = (int64_t)gridDim.x * blockDim.x;
  longlong2* acc2 = reinterpret_cast<longlong2*>(acc);
  for (int64_t i0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i0 < pairs; i0 += 4 * stride) {
    longlong2 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t i = i0 + u * stride;
      v[u] = i < pairs ? acc2[i] : make_longlong2(0, 0);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (v[u].x == 0 && v[u].y == 0) continue;
      const int64_t i = i0 + u * stride;
      float2* d = reinterpret_cast<float2*>(dst) + i;
      float2 g = *d;
      g.x += (float)((double)v[u].x * inv_scale);
      g.y += (float)((double)v[u].y * inv_scale);
      *d = g;
      acc2[i] = make_longlong2(0, 0);
    }
  }
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0 && acc[n - 1] != 0) {
    dst[n - 1] += (float)((double)acc[n - 1] * inv_scale);
    acc[n - 1] = 0;
  }
}

extern "C" int nkbs_fixed_point_flush(float* dst, int64_t n, void* stream) {
  NKBS_CHECK_ARG(n >= 0 && (n == 0 || dst), "nkbs_fixed_point_flush: bad arguments");
  if (n == 0) return 0;
  const FixedMap fm = nkbs_fixed_map_for(dst);
  NKBS_CHECK_ARG(fm.shadow && dst + n <= fm.base + fm.n, "nkbs_fixed_point_flush: [dst, dst + n) is not inside a registered buffer");
  NKBS_CHECK_ARG(((dst - fm.base) & 1) == 0 && (reinterpret_cast<uintptr_t>(dst) & 7u) == 0 &&
                     (reinterpret_cast<uintptr_t>(fm.shadow) & 15u) == 0,
                 "nkbs_fixed_point_flush: dst must start an even number of elements into a 16-byte aligned shadow and be 8-byte aligned");
  const int64_t blocks = (n / 8 + 255) / 256 + 1, cap = (int64_t)nkbs_num_sms() * 8;
  nkbs_launch(fixed_point_flush_kernel, (int)(blocks < cap ? blocks : cap), 256, 0, (cudaStream_t)stream, 
      fm.shadow + (dst - fm.base), dst, n, 1.0 / (double)fm.scale);
  NKBS_CHECK_LAUNCH("nkbs_fixed_point_flush");
  return 0;
}
