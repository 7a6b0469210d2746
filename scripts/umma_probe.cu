// Stand-alone probe of the UMMA descriptors used by csrc/mlp_tc.cu (run on the B200 box):
//   T1  D[128x64] = A[128x32] * B[64x32]^T        A, B K-major
//   T2  D[128x48] = A[128x64] * W[64x48]          A K-major, B = W viewed MN-major
//   T3  D[64x40]  = G[128x64]^T * X[128x40]       A = G viewed MN-major (M=64), B = X viewed MN-major
// nvcc -gencode arch=compute_100a,code=sm_100a -I nerf_kbs_b200/csrc scripts/umma_probe.cu -o /tmp/umma_probe
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include "umma.cuh"

using namespace umma;

// tile element (r,k) -> float index inside a chunk-major tile of R rows
__host__ __device__ inline int tix(int r, int k, int R) { return (k >> 2) * (R * 4 + 4) + r * 4 + (k & 3); }
__host__ __device__ inline int tile_floats(int R, int K) { return (K >> 2) * (R * 4 + 4); }

struct Args {
  const float* a;  // row-major [Ra x Ka]
  const float* b;  // row-major [Rb x Kb]
  float* d;        // row-major [M x N]
  int Ra, Ka, Rb, Kb, M, N, test;
  int a_mn, b_mn, swap, Mi;
};

__global__ void __launch_bounds__(128) probe(Args p) {
  extern __shared__ __align__(128) float smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base;
  float* sa = smem;
  float* sb = smem + ((tile_floats(p.Ra, p.Ka) + 31) & ~31);
  for (int e = threadIdx.x; e < p.Ra * p.Ka; e += 128) sa[tix(e / p.Ka, e % p.Ka, p.Ra)] = to_tf32(p.a[e]);
  for (int e = threadIdx.x; e < p.Rb * p.Kb; e += 128) sb[tix(e / p.Kb, e % p.Kb, p.Rb)] = to_tf32(p.b[e]);
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    mbar_fence_init();
  }
  if (threadIdx.x < 32) tmem_alloc(&tmem_base, 64);
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tm = tmem_base;
  const uint32_t csa = p.Ra * 16 + 16, csb = p.Rb * 16 + 16;  // chunk strides (bytes)
  if (threadIdx.x == 0) {
    // generic: a_mn/b_mn choose the view of each tile; contraction length Kc
    uint32_t id = make_idesc_tf32(p.Mi, p.N, p.a_mn, p.b_mn);
    int Kc = p.a_mn ? p.Ra : p.Ka;
    for (int k = 0; k < Kc / 8; ++k) {
      uint64_t da, db;
      if (!p.a_mn) da = make_desc(smem_u32(sa) + k * 2 * csa, csa, 128);
      else da = p.swap ? make_desc(smem_u32(sa) + k * 128, csa, 128) : make_desc(smem_u32(sa) + k * 128, 128, csa);
      if (!p.b_mn) db = make_desc(smem_u32(sb) + k * 2 * csb, csb, 128);
      else db = p.swap ? make_desc(smem_u32(sb) + k * 128, csb, 128) : make_desc(smem_u32(sb) + k * 128, 128, csb);
      mma_tf32(tm, da, db, id, k > 0);
    }
    commit(&bar);
  }
  mbar_wait(&bar, 0);
  fence_after_sync();
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int row = threadIdx.x;
  bool live = true;
  if (p.Mi == 64) {
    row = 16 * w + lane;
    live = lane < 16;
  }
  for (int c0 = 0; c0 < p.N; c0 += 8) {
    float v[8];
    tmem_ld8(tm + ((uint32_t)(32 * w) << 16) + c0, v);
    if (live && row < p.M)
      for (int j = 0; j < 8; ++j) p.d[row * p.N + c0 + j] = v[j];
  }
  fence_before_sync();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tm, 64);
}

static float tf32r(float x) {  // round to nearest, ties away (cvt.rna)
  uint32_t u;
  memcpy(&u, &x, 4);
  u += 0x1000u;
  u &= 0xFFFFE000u;
  float y;
  memcpy(&y, &u, 4);
  return y;
}

int run(int test, int a_mn, int b_mn, int swap, int Mi, int Ra, int Ka, int Rb, int Kb) {
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
        s += (double)tf32r(av) * tf32r(bv);
      }
      want[m * N + n] = (float)s;
    }
  float *da, *db, *dd;
  cudaMalloc(&da, a.size() * 4);
  cudaMalloc(&db, b.size() * 4);
  cudaMalloc(&dd, d.size() * 4);
  cudaMemcpy(da, a.data(), a.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(db, b.data(), b.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dd, d.data(), d.size() * 4, cudaMemcpyHostToDevice);
  Args p{da, db, dd, Ra, Ka, Rb, Kb, M, N, test, a_mn, b_mn, swap, Mi};
  size_t smem = (size_t)(((tile_floats(Ra, Ka) + 31) & ~31) + tile_floats(Rb, Kb) + 64) * 4;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  probe<<<1, 128, smem>>>(p);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("T%d: CUDA error %s\n", test, cudaGetErrorString(e));
    return 1;
  }
  cudaMemcpy(d.data(), dd, d.size() * 4, cudaMemcpyDeviceToHost);
  double worst = 0, scale = 0;
  int nz = 0;
  for (size_t i = 0; i < d.size(); ++i) {
    worst = fmax(worst, fabs((double)d[i] - want[i]));
    scale = fmax(scale, fabs((double)want[i]));
    nz += d[i] != 0.f;
  }
  bool ok = worst <= 1e-4 * scale;
  printf("T%d a_mn=%d b_mn=%d swap=%d M=%d N=%d Kc=%d: max abs err %.3e (max |want| %.3e, nonzero %d/%zu) %s\n", test, a_mn, b_mn,
         swap, Mi, N, Kc, worst, scale, nz, d.size(), ok ? "OK" : "MISMATCH");
  if (!ok)
    for (int i = 0; i < 3; ++i) printf("   d[%d]=%f want %f   d[%d]=%f want %f\n", i, d[i], want[i], N + i, d[N + i], want[N + i]);
  return ok ? 0 : 1;
}

int main() {
  int bad = 0;
  //            id a_mn b_mn swap Mi  Ra  Ka  Rb  Kb
  bad += run(1, 0, 0, 0, 128, 128, 32, 64, 32);
  bad += run(2, 0, 0, 0, 64, 64, 32, 64, 32);     // M=64 read-back layout
  bad += run(3, 0, 1, 0, 128, 128, 64, 64, 48);   // B MN-major, N=48
  bad += run(4, 0, 1, 0, 128, 128, 64, 64, 64);   // B MN-major, N=64
  bad += run(5, 0, 1, 1, 128, 128, 64, 64, 64);   // swapped LBO/SBO
  bad += run(6, 1, 0, 0, 64, 128, 64, 40, 128);   // A MN-major only (B K-major: [40 x 128])
  bad += run(7, 1, 0, 1, 64, 128, 64, 40, 128);
  bad += run(8, 1, 1, 0, 64, 128, 64, 128, 40);
  bad += run(9, 1, 1, 1, 64, 128, 64, 128, 40);
  bad += run(10, 0, 1, 0, 128, 128, 8, 8, 64);    // single MMA, B MN-major
  bad += run(11, 0, 1, 1, 128, 128, 8, 8, 64);
  printf(bad ? "PROBE FAILED (%d)\n" : "PROBE OK\n", bad);
  return 0;
}
