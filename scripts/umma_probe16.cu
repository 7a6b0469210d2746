// Probe of 16-bit (bf16 / fp16) UMMA operands in the chunk-major no-swizzle layout, K-major and MN-major views.
// tile element (r,k): (k/8)*cs + r*16 + (k%8)*2 bytes, cs = R*16+16
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include "umma.cuh"
using namespace umma;

__host__ __device__ inline int tix(int r, int k, int R) { return (k >> 3) * (R * 8 + 8) + r * 8 + (k & 7); }
__host__ __device__ inline int tile_elems(int R, int K) { return (K >> 3) * (R * 8 + 8); }

struct Args {
  const float *a, *b;
  float* d;
  int Ra, Ka, Rb, Kb, M, N, a_mn, b_mn, Mi, a_fmt, b_fmt;
};

__device__ inline uint16_t cvt(float x, int fmt) {
  if (fmt) { __nv_bfloat16 h = __float2bfloat16_rn(x); return *reinterpret_cast<uint16_t*>(&h); }
  __half h = __float2half_rn(x); return *reinterpret_cast<uint16_t*>(&h);
}

__global__ void __launch_bounds__(128) probe(Args p) {
  extern __shared__ __align__(128) uint16_t smem16[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base;
  uint16_t* sa = smem16;
  uint16_t* sb = smem16 + ((tile_elems(p.Ra, p.Ka) + 63) & ~63);
  for (int e = threadIdx.x; e < p.Ra * p.Ka; e += 128) sa[tix(e / p.Ka, e % p.Ka, p.Ra)] = cvt(p.a[e], p.a_fmt);
  for (int e = threadIdx.x; e < p.Rb * p.Kb; e += 128) sb[tix(e / p.Kb, e % p.Kb, p.Rb)] = cvt(p.b[e], p.b_fmt);
  if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
  if (threadIdx.x < 32) tmem_alloc(&tmem_base, 128);
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tm = tmem_base;
  const uint32_t csa = p.Ra * 16 + 16, csb = p.Rb * 16 + 16;
  if (threadIdx.x == 0) {
    uint32_t id = make_idesc_f16(p.Mi, p.N, p.a_mn, p.b_mn, p.a_fmt, p.b_fmt);
    int Kc = p.a_mn ? p.Ra : p.Ka;
    for (int k = 0; k < Kc / 16; ++k) {
      uint64_t da = p.a_mn ? make_desc(smem_u32(sa) + k * 256, 128, csa) : make_desc(smem_u32(sa) + k * 2 * csa, csa, 128);
      uint64_t db = p.b_mn ? make_desc(smem_u32(sb) + k * 256, 128, csb) : make_desc(smem_u32(sb) + k * 2 * csb, csb, 128);
      mma_f16(tm, da, db, id, k > 0);
    }
    commit(&bar);
  }
  mbar_wait(&bar, 0);
  fence_after_sync();
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int row = threadIdx.x;
  bool live = true;
  if (p.Mi == 64) { row = 16 * w + lane; live = lane < 16; }
  for (int c0 = 0; c0 < p.N; c0 += 8) {
    float v[8];
    tmem_ld8(tm + ((uint32_t)(32 * w) << 16) + c0, v);
    if (live && row < p.M)
      for (int j = 0; j < 8; ++j) p.d[row * p.N + c0 + j] = v[j];
  }
  fence_before_sync();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tm, 128);
}

static float rnd(float x, int fmt) {
  if (fmt) return __bfloat162float(__float2bfloat16_rn(x));
  return __half2float(__float2half_rn(x));
}

int run(int test, int a_mn, int b_mn, int Mi, int Ra, int Ka, int Rb, int Kb, int a_fmt, int b_fmt) {
  int M = a_mn ? Ka : Ra, N = b_mn ? Kb : Rb;
  if (M > Mi) M = Mi;
  int Kc = a_mn ? Ra : Ka;
  std::vector<float> a(Ra * Ka), b(Rb * Kb), d(M * N, -777.f), want(M * N, 0.f);
  srand(test);
  for (auto& x : a) x = (rand() % 2001 - 1000) / 500.f;
  for (auto& x : b) x = (rand() % 2001 - 1000) / 500.f;
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) {
      double s = 0;
      for (int k = 0; k < Kc; ++k) {
        float av = a_mn ? a[k * Ka + m] : a[m * Ka + k];
        float bv = b_mn ? b[k * Kb + n] : b[n * Kb + k];
        s += (double)rnd(av, a_fmt) * rnd(bv, b_fmt);
      }
      want[m * N + n] = (float)s;
    }
  float *da, *db, *dd;
  cudaMalloc(&da, a.size() * 4); cudaMalloc(&db, b.size() * 4); cudaMalloc(&dd, d.size() * 4);
  cudaMemcpy(da, a.data(), a.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(db, b.data(), b.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dd, d.data(), d.size() * 4, cudaMemcpyHostToDevice);
  Args p{da, db, dd, Ra, Ka, Rb, Kb, M, N, a_mn, b_mn, Mi, a_fmt, b_fmt};
  size_t smem = (size_t)(((tile_elems(Ra, Ka) + 63) & ~63) + tile_elems(Rb, Kb) + 64) * 2;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  probe<<<1, 128, smem>>>(p);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("T%d: CUDA error %s\n", test, cudaGetErrorString(e)); return 1; }
  cudaMemcpy(d.data(), dd, d.size() * 4, cudaMemcpyDeviceToHost);
  double worst = 0, scale = 0; int nz = 0;
  for (size_t i = 0; i < d.size(); ++i) {
    worst = fmax(worst, fabs((double)d[i] - want[i]));
    scale = fmax(scale, fabs((double)want[i]));
    nz += d[i] != 0.f;
  }
  bool ok = worst <= 1e-4 * scale;
  printf("T%d a_mn=%d b_mn=%d fmt=%d/%d M=%d N=%d Kc=%d: max abs err %.3e (max |want| %.3e, nonzero %d/%zu) %s\n", test, a_mn, b_mn,
         a_fmt, b_fmt, Mi, N, Kc, worst, scale, nz, d.size(), ok ? "OK" : "MISMATCH");
  if (!ok) for (int i = 0; i < 3; ++i) printf("   d[%d]=%f want %f   d[%d]=%f want %f\n", i, d[i], want[i], N + i, d[N + i], want[N + i]);
  return ok ? 0 : 1;
}

int main() {
  int bad = 0;
  //          id a_mn b_mn Mi   Ra  Ka  Rb  Kb  fa fb
  bad += run(1, 0, 0, 128, 128, 64, 64, 64, 1, 1);    // fwd: Y = X W^T
  bad += run(2, 0, 0, 128, 128, 80, 16, 80, 0, 0);    // fp16, K=80 (bias-augmented), N=16
  bad += run(3, 0, 1, 128, 128, 64, 64, 80, 1, 1);    // dX = dY W  (W viewed MN-major), N=80
  bad += run(4, 1, 1, 64, 128, 64, 128, 80, 1, 1);    // dW = dY^T X (both MN-major), M=64
  bad += run(5, 1, 1, 64, 128, 64, 128, 80, 1, 0);    // mixed: dY bf16, X fp16
  bad += run(6, 0, 0, 128, 128, 64, 64, 64, 0, 1);    // mixed K-major
  bad += run(7, 1, 1, 64, 128, 64, 128, 16, 0, 0);    // fp16 dW, N=16
  bad += run(8, 0, 1, 128, 128, 16, 16, 64, 1, 1);    // single MMA, B MN-major
  printf(bad ? "PROBE16 FAILED (%d)\n" : "PROBE16 OK\n", bad);
  return 0;
}
