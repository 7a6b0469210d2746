// K8: the proposal density field in ONE pass per direction: ray point -> L-inf contraction -> hash-grid
// encoding -> MLP (2L -> H -> 1, ReLU) -> trunc_exp * selector, with nothing but the density leaving the SM.
// The unfused route (K1 + K3) writes and re-reads 2L+4 floats per sample; at the proposal levels that is 23 M
// samples per 65 536-ray step.  H is 16 (nerfacto) or 64 (semantic-NeRF-W defaults): an MMA of N = 16 would use a
// sixteenth of a tcgen05 tile, so these two layers run on the CUDA cores, weights broadcast from shared memory,
// one thread per sample (adjacent lanes = adjacent samples of one ray, as in encode.cu).
// Backward recomputes the forward, scatters the feature gradient straight into the table gradient (warp-merged
// on the coarse levels) and reduces the weight gradients warp -> CTA -> global with a transposed shared-memory
// sum (no shuffles: 32 rows x 32 lanes per pass).
// Replaces HashMLPDensityField.get_density / density_fn reached from ref:nerfacto.py:299-311,325-332 and
// ref:semantic_nerfw.py:119-137.
#include <stdlib.h>
#include "common.cuh"
#include "encode.cuh"
#include "sampler.cuh"

namespace {

constexpr int kThreads = 256;
#ifndef NKBS_DENSITY_BWD_CTAS
#define NKBS_DENSITY_BWD_CTAS 2
#endif
constexpr int kWarps = kThreads / 32;
constexpr int kRedStride = 36;  // floats per row of the transposed-sum scratch: 16-byte rows, conflict-free both ways

struct DensityArgs : SampleSource {
  int64_t N;
  LevelParams lv;
  const float2* table;
  const float* W1;  // [H, 2L]
  const float* b1;  // [H]
  const float* W2;  // [1, H]
  const float* b2;  // [1]
  float* density;   // [N]            (forward)
  float* feat_out;  // forward, optional (nkbs_density_fwd_save): hash-grid features [N, 2L],
  float* xn_out;    //   normalised positions x selector [N, 3]
  float* sel_out;   //   and selector [N] of every sample, for a backward pass through nkbs_mlp_*_bwd + nkbs_encode_bwd
  const float* ddensity;  // [N]      (backward)
  const float* feat_in;   // backward, optional (nkbs_density_bwd_saved): the forward's features [N, 2L] and
  const float* sel_in;    //   selectors [N] — no positions, no gathers
  float* dtable;    // [L*T, 2]
  float* dW1;
  float* db1;
  float* dW2;
  float* db2;
  int dedup_levels;
  float v4_min_scale;  // see encode.cuh
  FixedMap fm;         // backward, deterministic mode (common.cuh): shadow of the gradient buffer, or a null map
  float* df_out;       // backward, optional [N, 2L]: the feature gradients are written here INSTEAD of being scattered
                       // (the ray-major scatter nkbs_encode_bwd_rays follows)
};

// shared weights: W1 rows padded to KP floats (16-byte rows -> LDS.128 broadcasts), then b1, W2, b2
template <int H, int L>
struct Smem {
  static constexpr int K = 2 * L;
  static constexpr int KP = (K + 3) & ~3;
  float w1[H * KP];
  float b1[H];
  float w2[H];
  float b2;
};

template <int H, int L>
__device__ __forceinline__ void stage(const DensityArgs& a, Smem<H, L>& s) {
  constexpr int K = 2 * L, KP = Smem<H, L>::KP;
  for (int e = threadIdx.x; e < H * KP; e += kThreads) {
    int j = e / KP, i = e - j * KP;
    s.w1[e] = i < K ? __ldg(a.W1 + j * K + i) : 0.f;
  }
  for (int j = threadIdx.x; j < H; j += kThreads) {
    s.b1[j] = __ldg(a.b1 + j);
    s.w2[j] = __ldg(a.W2 + j);
  }
  if (threadIdx.x == 0) s.b2 = __ldg(a.b2);
}

template <int L>
__device__ __forceinline__ void encode_sample(const DensityArgs& a, const float* s_scale, float3 x, float (&f)[2 * L]) {
  const uint32_t mask = a.lv.table_mask;
#pragma unroll
  for (int l = 0; l < L; ++l) {
    Cell cl = nkbs_cell(x, s_scale[l]);
    uint32_t h[8];
    nkbs_corner_hashes(cl, mask, h);
    const float2* tl = a.table + ((size_t)l << a.lv.log2_T);
    float2 v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = __ldg(tl + h[k]);
    float v0[8], v1[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      v0[k] = v[k].x;
      v1[k] = v[k].y;
    }
    f[2 * l] = nkbs_trilerp(v0, cl.o[0], cl.o[1], cl.o[2]);
    f[2 * l + 1] = nkbs_trilerp(v1, cl.o[0], cl.o[1], cl.o[2]);
  }
}

// z_j = b1_j + sum_i W1[j][i] f_i  (i ascending, fused multiply-adds), hidden unit j of the first layer
template <int H, int L>
__device__ __forceinline__ float hidden_pre(const Smem<H, L>& s, const float (&f)[2 * L], int j) {
  constexpr int K = 2 * L, KP = Smem<H, L>::KP;
  float z = s.b1[j];
  const float4* w = reinterpret_cast<const float4*>(s.w1 + j * KP);
#pragma unroll
  for (int q = 0; q < KP / 4; ++q) {
    float4 wv = w[q];
    if (4 * q + 0 < K) z = fmaf(wv.x, f[4 * q + 0 < K ? 4 * q + 0 : 0], z);
    if (4 * q + 1 < K) z = fmaf(wv.y, f[4 * q + 1 < K ? 4 * q + 1 : 0], z);
    if (4 * q + 2 < K) z = fmaf(wv.z, f[4 * q + 2 < K ? 4 * q + 2 : 0], z);
    if (4 * q + 3 < K) z = fmaf(wv.w, f[4 * q + 3 < K ? 4 * q + 3 : 0], z);
  }
  return z;
}

template <int H, int L, bool SAVE>
__global__ void __launch_bounds__(kThreads) density_fwd_kernel(const DensityArgs a) {
  nkbs_pdl_sync();
  __shared__ float s_scale[NKBS_MAX_LEVELS];
  __shared__ __align__(16) Smem<H, L> s;
  if (threadIdx.x < NKBS_MAX_LEVELS) s_scale[threadIdx.x] = a.lv.scale[threadIdx.x];
  stage<H, L>(a, s);
  __syncthreads();
  for (int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; n < a.N; n += (int64_t)gridDim.x * blockDim.x) {
    float sel;
    float3 x = load_sample_position(a, n, sel);
    float f[2 * L];
    encode_sample<L>(a, s_scale, x, f);
    if constexpr (SAVE) {
      if constexpr (L % 2 == 0) {  // rows of 2L floats are 16-byte aligned
        float4* fo = reinterpret_cast<float4*>(a.feat_out + n * (2 * L));
#pragma unroll
        for (int i = 0; i < L / 2; ++i) fo[i] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
      } else {
        float2* fo = reinterpret_cast<float2*>(a.feat_out + n * (2 * L));
#pragma unroll
        for (int i = 0; i < L; ++i) fo[i] = make_float2(f[2 * i], f[2 * i + 1]);
      }
      if (a.xn_out) {
        a.xn_out[3 * n] = x.x;
        a.xn_out[3 * n + 1] = x.y;
        a.xn_out[3 * n + 2] = x.z;
      }
      a.sel_out[n] = sel;
    }
    float out = s.b2;
#pragma unroll
    for (int j = 0; j < H; ++j) out = fmaf(s.w2[j], fmaxf(hidden_pre<H, L>(s, f, j), 0.f), out);
    a.density[n] = expf(out) * sel;
  }
}

// One pass of the transposed sum: every lane contributes `v` for 32 consecutive gradient rows (row r of the pass
// is written by the caller through put(r, value)); afterwards lane i holds the warp total of row i.
__device__ __forceinline__ float row_total(const float* scratch, int lane) {
  const float4* p = reinterpret_cast<const float4*>(scratch + lane * kRedStride);
  float t = 0.f;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    float4 v = p[q];
    t += (v.x + v.y) + (v.z + v.w);
  }
  return t;
}


// ---- weight gradients of the H = 16 field on the tensor cores (mma.sync m16n8k8, 3xTF32) ----------------------
// The weight gradients are sums over samples of outer products, i.e. GEMMs that contract over the 32 samples of a
// warp: dW1|db1 = GH^T [16 x 32] . [F | 1] [32 x 16]  and  dW2 = RZ^T [16 x 32] . g_out [32 x 1].  Every lane owns one
// sample (one row of GH, F, RZ), so the operands cross lanes through a per-warp shared-memory tile (row = sample, 20
// floats: 16-byte rows, conflict-free 128-bit stores) and come back as MMA fragments; the accumulators stay in
// registers over the whole grid-stride loop.  Each product is split into tf32 high and low parts (hi*hi + hi*lo +
// lo*hi) so the result keeps fp32 accuracy.  Against the transposed shared-memory sum this moves a quarter of the
// shared-memory wavefronts per sample.
#ifndef NKBS_DENSITY_BWD_MMA
#define NKBS_DENSITY_BWD_MMA 1
#endif
#ifndef NKBS_DENSITY_BWD_MMA_MAXH
#define NKBS_DENSITY_BWD_MMA_MAXH 64  // 16: only the nerfacto proposal shape takes the tensor-core route (A/B builds)
#endif
constexpr int kTileStride = 20;

__device__ __forceinline__ void split_tf32(float v, uint32_t& hi, uint32_t& lo) {
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi) : "f"(v));
  const float r = v - __uint_as_float(hi);
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lo) : "f"(r));
}

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// d += A . B with both operands split (three MMAs)
__device__ __forceinline__ void mma_3xtf32(float (&d)[4], const float (&a)[4], const float (&b)[2]) {
  uint32_t ah[4], al[4], bh[2], bl[2];
#pragma unroll
  for (int i = 0; i < 4; ++i) split_tf32(a[i], ah[i], al[i]);
#pragma unroll
  for (int i = 0; i < 2; ++i) split_tf32(b[i], bh[i], bl[i]);
  mma_tf32(d, al, bh);
  mma_tf32(d, ah, bl);
  mma_tf32(d, ah, bh);
}

// Geometry of the tensor-core route for hidden width H (a multiple of 16) and K = 2L inputs: kMT m-tiles of 16 hidden
// units, kNT n-tiles of 8 columns of [F | 1 | 0].  Per warp: tile A [32 samples][kStrideA] (relu(z), then GH) and tile
// B [32 samples][kStrideB].  H = 16 keeps the layout it was tuned with (stride 20).  H = 64: strides = 4 mod 32 floats
// keep the row-wise 128-bit stores conflict-free, and the contraction index k of an MMA step is mapped to the samples
// (0, 2, 4, 6 | 1, 3, 5, 7) of its group of eight — any fixed permutation is legal as long as A and B use the same one —
// so that the four rows a transposed fragment load touches start 8 banks apart: conflict-free as well.
template <int H, int L>
struct MmaGeom {
  static constexpr int K = 2 * L;
  static constexpr bool kMma = NKBS_DENSITY_BWD_MMA && (H % 16 == 0) && H <= NKBS_DENSITY_BWD_MMA_MAXH;
  static constexpr int kMT = H / 16;
  static constexpr int kNT = (K + 1 + 7) / 8;
  static constexpr bool kPermute = H != 16;
  static constexpr int kStrideA = H == 16 ? kTileStride : H + 4;
  static constexpr int kStrideB = H == 16 ? kTileStride : 8 * kNT + 4;
  static constexpr int kScratch = kMma ? 32 * (kStrideA + kStrideB) : 32 * kRedStride;  // floats per warp
  static_assert(!kMma || (kStrideA % 4 == 0 && kStrideB % 4 == 0 && 8 * kNT <= kStrideB), "tile rows are float4 rows");
  // sample (= lane) that contraction index kk (0..7) of k-step ks stands for
  __device__ static __forceinline__ int sample(int ks, int kk) {
    return 8 * ks + (kPermute ? ((kk & 3) * 2 + (kk >> 2)) : kk);
  }
};

// A fragment of the k-step `ks` for m-tile `mt` from tile A read TRANSPOSED (row = hidden unit, contraction index =
// sample): a0 (g, t), a1 (g+8, t), a2 (g, t+4), a3 (g+8, t+4), g = lane/4, t = lane%4
template <typename G>
__device__ __forceinline__ void load_a_transposed(const float* tile, int ks, int mt, int g, int t, float (&a)[4]) {
  const float* r0 = tile + G::sample(ks, t) * G::kStrideA + 16 * mt;
  const float* r1 = tile + G::sample(ks, t + 4) * G::kStrideA + 16 * mt;
  a[0] = r0[g];
  a[1] = r0[g + 8];
  a[2] = r1[g];
  a[3] = r1[g + 8];
}

// resident CTAs per SM the backward is compiled for: the H = 16 field fits three (85 registers per thread)
template <int H>
constexpr int bwd_ctas() { return H <= 16 ? NKBS_DENSITY_BWD_CTAS : 1; }

// FEAT: features and selectors come from the forward pass (feat_in / sel_in) instead of being recomputed from the
// sample positions; the feature gradients go to df_out.
template <int H, int L, bool FX, bool FEAT = false, int MINB = 0>
__global__ void __launch_bounds__(kThreads, MINB ? MINB : bwd_ctas<H>()) density_bwd_kernel(const DensityArgs a) {
  nkbs_pdl_sync();
  constexpr int K = 2 * L, KP = Smem<H, L>::KP;
  // gradient rows, in this order: for each hidden unit j: dW1[j][0..K), db1[j], dW2[j]; then db2
  constexpr int kPerJ = K + 2;
  constexpr int kRows = H * kPerJ + 1;
  constexpr int kPasses = (kRows + 31) / 32;
  __shared__ float s_scale[NKBS_MAX_LEVELS];
  __shared__ __align__(16) Smem<H, L> s;
  using G = MmaGeom<H, L>;
  constexpr bool kMma = G::kMma;
  constexpr int kMT = G::kMT, kNT = G::kNT;
  // transposed-sum scratch, or (tensor-core route) the two operand tiles of every warp: dynamic shared memory
  constexpr int kScratch = G::kScratch;
  extern __shared__ __align__(16) float s_red[];  // [kWarps][kScratch]
  __shared__ float s_tot[kPasses * 32];
  if (threadIdx.x < NKBS_MAX_LEVELS) s_scale[threadIdx.x] = a.lv.scale[threadIdx.x];
  for (int e = threadIdx.x; e < kPasses * 32; e += kThreads) s_tot[e] = 0.f;
  stage<H, L>(a, s);
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float* red = s_red + wid * kScratch;
  const uint32_t mask = a.lv.table_mask;
  float acc[kPasses];
#pragma unroll
  for (int p = 0; p < kPasses; ++p) acc[p] = 0.f;
  // tensor-core route: D1 = dW1|db1 as kMT x kNT tiles of 16 x 8 (hidden unit x input column), D2 = dW2 in column 0 of
  // one tile per m-tile
  float d1[kMma ? kMT : 1][kMma ? kNT : 1][4];
  float d2[kMma ? kMT : 1][4];
#pragma unroll
  for (int mt = 0; mt < (kMma ? kMT : 1); ++mt) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      d2[mt][i] = 0.f;
#pragma unroll
      for (int nt = 0; nt < (kMma ? kNT : 1); ++nt) d1[mt][nt][i] = 0.f;
    }
  }
  float db2_acc = 0.f;
  const int fg = lane >> 2, ft = lane & 3;
  const int64_t n_round = (a.N + 31) / 32 * 32;  // whole warps reach the collectives
  for (int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; n < n_round; n += (int64_t)gridDim.x * blockDim.x) {
    const bool live = n < a.N;
    float sel = 0.f;
    float3 x = make_float3(0.f, 0.f, 0.f);
    float g_up = 0.f;
    if (live) {
      if constexpr (FEAT) sel = a.sel_in[n];
      else x = load_sample_position(a, n, sel);
      g_up = a.ddensity[n] * sel;
    }
    // a warp whose samples all lie outside the grid (or carry no gradient) has nothing to add
    if (!__any_sync(0xffffffffu, g_up != 0.f)) {
      if (a.df_out && live) {
        float2* o = reinterpret_cast<float2*>(a.df_out + n * K);
#pragma unroll
        for (int i = 0; i < L; ++i) o[i] = make_float2(0.f, 0.f);
      }
      continue;
    }
    float f[K];
    if constexpr (FEAT) {
      const float2* fi = reinterpret_cast<const float2*>(a.feat_in + n * K);
#pragma unroll
      for (int i = 0; i < L; ++i) {
        const float2 v = live ? fi[i] : make_float2(0.f, 0.f);
        f[2 * i] = v.x;
        f[2 * i + 1] = v.y;
      }
    } else {
      encode_sample<L>(a, s_scale, x, f);
    }
    float z[kMma ? 1 : H];
    float out = s.b2;
    float df[K];
#pragma unroll
    for (int i = 0; i < K; ++i) df[i] = 0.f;
    float g_out;
    if constexpr (kMma) {
      // ---- weight gradients on the tensor cores (see mma_3xtf32 above) ----
      float* tA = red;                        // [32][kStrideA]: relu(z), then GH
      float* tB = red + 32 * G::kStrideA;     // [32][kStrideB]: F | 1 | 0
      float4* rowA = reinterpret_cast<float4*>(tA + lane * G::kStrideA);
      float4* rowB = reinterpret_cast<float4*>(tB + lane * G::kStrideB);
      uint64_t positive = 0ull;  // bit j: hidden unit j is active for this sample
#pragma unroll
      for (int q = 0; q < H / 4; ++q) {
        float r[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int j = 4 * q + i;
          const float zj = hidden_pre<H, L>(s, f, j);
          r[i] = fmaxf(zj, 0.f);
          if (zj > 0.f) positive |= 1ull << j;
          out = fmaf(s.w2[j], r[i], out);
        }
        rowA[q] = make_float4(r[0], r[1], r[2], r[3]);
      }
      // trunc_exp backward (A.4): g * exp(clamp(raw, -15, 15))
      g_out = g_up * expf(fminf(fmaxf(out, -15.f), 15.f));
      __syncwarp();
      // dW2 = RZ^T . g_out: B carries g_out in column 0 (k = sample), i.e. on the lanes with g == 0
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const float g0 = __shfl_sync(0xffffffffu, g_out, G::sample(ks, ft));
        const float g1 = __shfl_sync(0xffffffffu, g_out, G::sample(ks, ft + 4));
        const float bf[2] = {fg == 0 ? g0 : 0.f, fg == 0 ? g1 : 0.f};
#pragma unroll
        for (int mt = 0; mt < kMT; ++mt) {
          float af[4];
          load_a_transposed<G>(tA, ks, mt, fg, ft, af);
          mma_3xtf32(d2[mt], af, bf);
        }
      }
      __syncwarp();
#pragma unroll
      for (int q = 0; q < H / 4; ++q) {
        float gq[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int j = 4 * q + i;
          gq[i] = (positive >> j) & 1ull ? g_out * s.w2[j] : 0.f;
          const float4* w = reinterpret_cast<const float4*>(s.w1 + j * KP);
#pragma unroll
          for (int qq = 0; qq < KP / 4; ++qq) {
            float4 wv = w[qq];
            if (4 * qq + 0 < K) df[4 * qq + 0 < K ? 4 * qq + 0 : 0] = fmaf(gq[i], wv.x, df[4 * qq + 0 < K ? 4 * qq + 0 : 0]);
            if (4 * qq + 1 < K) df[4 * qq + 1 < K ? 4 * qq + 1 : 0] = fmaf(gq[i], wv.y, df[4 * qq + 1 < K ? 4 * qq + 1 : 0]);
            if (4 * qq + 2 < K) df[4 * qq + 2 < K ? 4 * qq + 2 : 0] = fmaf(gq[i], wv.z, df[4 * qq + 2 < K ? 4 * qq + 2 : 0]);
            if (4 * qq + 3 < K) df[4 * qq + 3 < K ? 4 * qq + 3 : 0] = fmaf(gq[i], wv.w, df[4 * qq + 3 < K ? 4 * qq + 3 : 0]);
          }
        }
        rowA[q] = make_float4(gq[0], gq[1], gq[2], gq[3]);
      }
#pragma unroll
      for (int q = 0; q < 2 * kNT; ++q) {
        float e[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int c = 4 * q + i;
          e[i] = c < K ? f[c < K ? c : 0] : (c == K ? 1.f : 0.f);
        }
        rowB[q] = make_float4(e[0], e[1], e[2], e[3]);
      }
      __syncwarp();
      // dW1 | db1 = GH^T . [F | 1]
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        // B (k = sample, n = input column): b0 (t, g), b1 (t+4, g); the fragments of all n-tiles stay in registers
        float bf[kNT][2];
#pragma unroll
        for (int nt = 0; nt < kNT; ++nt) {
          bf[nt][0] = tB[G::sample(ks, ft) * G::kStrideB + 8 * nt + fg];
          bf[nt][1] = tB[G::sample(ks, ft + 4) * G::kStrideB + 8 * nt + fg];
        }
#pragma unroll
        for (int mt = 0; mt < kMT; ++mt) {
          float af[4];
          load_a_transposed<G>(tA, ks, mt, fg, ft, af);
#pragma unroll
          for (int nt = 0; nt < kNT; ++nt) mma_3xtf32(d1[mt][nt], af, bf[nt]);
        }
      }
      __syncwarp();
      db2_acc += g_out;
    } else {
#pragma unroll
      for (int j = 0; j < H; ++j) {
        z[j] = hidden_pre<H, L>(s, f, j);
        out = fmaf(s.w2[j], fmaxf(z[j], 0.f), out);
      }
      // trunc_exp backward (A.4): g * exp(clamp(raw, -15, 15))
      g_out = g_up * expf(fminf(fmaxf(out, -15.f), 15.f));
    // ---- weight gradients: rows are produced unit by unit and summed 32 rows at a time ----
    int row = 0;  // compile-time after unrolling
    int pass = 0;
#pragma unroll
    for (int j = 0; j < H; ++j) {
      const float gh = z[j] > 0.f ? g_out * s.w2[j] : 0.f;
      const float4* w = reinterpret_cast<const float4*>(s.w1 + j * KP);
#pragma unroll
      for (int q = 0; q < KP / 4; ++q) {
        float4 wv = w[q];
        if (4 * q + 0 < K) df[4 * q + 0 < K ? 4 * q + 0 : 0] = fmaf(gh, wv.x, df[4 * q + 0 < K ? 4 * q + 0 : 0]);
        if (4 * q + 1 < K) df[4 * q + 1 < K ? 4 * q + 1 : 0] = fmaf(gh, wv.y, df[4 * q + 1 < K ? 4 * q + 1 : 0]);
        if (4 * q + 2 < K) df[4 * q + 2 < K ? 4 * q + 2 : 0] = fmaf(gh, wv.z, df[4 * q + 2 < K ? 4 * q + 2 : 0]);
        if (4 * q + 3 < K) df[4 * q + 3 < K ? 4 * q + 3 : 0] = fmaf(gh, wv.w, df[4 * q + 3 < K ? 4 * q + 3 : 0]);
      }
#pragma unroll
      for (int c = 0; c < kPerJ; ++c) {
        const float v = c < K ? gh * f[c < K ? c : 0] : (c == K ? gh : g_out * fmaxf(z[j], 0.f));
        red[(row & 31) * kRedStride + lane] = v;
        ++row;
        if ((row & 31) == 0) {
          __syncwarp();
          acc[pass] += row_total(red, lane);
          ++pass;
          __syncwarp();
        }
      }
    }
    red[(row & 31) * kRedStride + lane] = g_out;  // db2
    ++row;
#pragma unroll
    for (int r = row & 31; r < 32 && (row & 31) != 0; ++r) red[r * kRedStride + lane] = 0.f;
    __syncwarp();
    acc[kPasses - 1] += row_total(red, lane);
    __syncwarp();
    }
    // ---- feature gradient -> table gradient ----
    if (FEAT || a.df_out) {
      if (live) {
        float2* o = reinterpret_cast<float2*>(a.df_out + n * K);
#pragma unroll
        for (int i = 0; i < L; ++i) o[i] = make_float2(df[2 * i], df[2 * i + 1]);
      }
      continue;
    }
#pragma unroll
    for (int l = 0; l < L; ++l) {
      Cell cl = nkbs_cell(x, s_scale[l]);
      uint32_t h[8];
      nkbs_corner_hashes(cl, mask, h);
      float w[8];
      nkbs_corner_weights(cl.o[0], cl.o[1], cl.o[2], w);
      float* base = a.dtable + ((size_t)l << (a.lv.log2_T + 1));
      long long* fixed = FX ? a.fm.at(base) : nullptr;
      if (l < a.dedup_levels) {
        scatter_level_runs<FX>(base, cl, h, w, df[2 * l], df[2 * l + 1], lane, s_scale[l] >= a.v4_min_scale, fixed, a.fm.scale);
      } else {
        float vx[8], vy[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          vx[k] = w[k] * df[2 * l];
          vy[k] = w[k] * df[2 * l + 1];
        }
        scatter_cell<FX>(base, cl, h, vx, vy, s_scale[l] >= a.v4_min_scale, fixed, a.fm.scale);
      }
    }
  }
  // ---- warp totals -> CTA totals -> global ----
  // row r of the totals -> its gradient element
  auto dest = [&](int r) -> float* {
    if (r == kRows - 1) return a.db2;
    const int j = r / kPerJ, c = r - j * kPerJ;
    return c < K ? a.dW1 + j * K + c : (c == K ? a.db1 + j : a.dW2 + j);
  };
  // deterministic mode: the float atomics on the CTA totals depend on the order of the warps, so every warp adds its
  // totals to the fixed-point shadow directly
  constexpr bool det = FX;
  auto add = [&](int r, float v) {
    if (v == 0.f) return;
    if constexpr (det) {
      if (r < kRows) nkbs_acc(dest(r), a.fm, v);
    } else {
      atomicAdd(&s_tot[r], v);
    }
  };
  if constexpr (kMma) {
    // D1 tile (mt, nt): d0 (j = 16mt + g, c = 8nt + 2t), d1 (j, c + 1), d2 (j + 8, c), d3 (j + 8, c + 1); columns c <= K are real
#pragma unroll
    for (int mt = 0; mt < kMT; ++mt) {
#pragma unroll
      for (int nt = 0; nt < kNT; ++nt) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int j = 16 * mt + fg + (i >= 2 ? 8 : 0), c = 8 * nt + 2 * ft + (i & 1);
          if (c <= K) add(j * kPerJ + c, d1[mt][nt][i]);
        }
      }
      if (ft == 0) {  // column 0 of D2: dW2[16mt + g], dW2[16mt + g + 8]
        add((16 * mt + fg) * kPerJ + K + 1, d2[mt][0]);
        add((16 * mt + fg + 8) * kPerJ + K + 1, d2[mt][2]);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) db2_acc += __shfl_xor_sync(0xffffffffu, db2_acc, o);
    if (lane == 0) add(kRows - 1, db2_acc);
  }
#pragma unroll
  for (int p = 0; p < kPasses; ++p) add(p * 32 + lane, acc[p]);
  if constexpr (det) return;
  __syncthreads();
  for (int r = threadIdx.x; r < kRows; r += kThreads) {
    const float v = s_tot[r];
    if (v != 0.f) atomicAdd(dest(r), v);
  }
}

template <typename Kern>
int grid_for(Kern kern, int64_t N, size_t smem = 0) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
  int64_t blocks = (N + kThreads - 1) / kThreads;
  int64_t cap = (int64_t)nkbs_num_sms() * per_sm;
  return (int)(blocks < cap ? blocks : cap);
}

int fill(DensityArgs& a, const char* who, int mode, const float* x, const float* origins, const float* directions,
         const float* starts, const float* ends, int bin_stride, int64_t R, int32_t S, int contraction, const float* aabb,
         const float* table, const float* scalings, int L, int log2_T, const float* W1, const float* b1,
         const float* W2, const float* b2, int H) {
  NKBS_CHECK_ARG(mode == 1 || mode == 2, "%s: mode must be 1 (positions) or 2 (frustums), got %d", who, mode);
  NKBS_CHECK_ARG(R >= 0 && S >= 1, "%s: bad sizes R=%lld S=%d", who, (long long)R, S);
  NKBS_CHECK_ARG((H == 16 && L == 5) || (H == 64 && L == 8),
                 "%s: fused shapes are (L=5,H=16) and (L=8,H=64); got L=%d H=%d (use encode + mlp)", who, L, H);
  NKBS_CHECK_ARG(table && scalings && W1 && b1 && W2 && b2, "%s: null table/scalings/weights", who);
  if (mode == 2)
    NKBS_CHECK_ARG(origins && directions && starts && ends, "%s: frustum mode needs origins/directions/starts/ends", who);
  else
    NKBS_CHECK_ARG(x != nullptr || R == 0, "%s: null positions", who);
  a.mode = mode;
  a.x = x;
  a.origins = origins;
  a.directions = directions;
  a.starts = starts;
  a.ends = ends;
  a.S = S;
  NKBS_CHECK_ARG(bin_stride == 0 || bin_stride >= S, "%s: bin_stride %d < S", who, bin_stride);
  a.bstride = bin_stride ? bin_stride : S;
  a.N = R * (int64_t)S;
  a.norm.contraction = contraction;
  for (int i = 0; i < 3; ++i) {
    a.norm.aabb_min[i] = aabb ? aabb[i] : 0.f;
    a.norm.aabb_len[i] = aabb ? aabb[3 + i] - aabb[i] : 1.f;
  }
  NKBS_CHECK_ARG(fill_levels(a.lv, scalings, L, log2_T) == 0, "%s: unsupported grid L=%d log2_T=%d", who, L, log2_T);
  a.table = reinterpret_cast<const float2*>(table);
  a.W1 = W1;
  a.b1 = b1;
  a.W2 = W2;
  a.b2 = b2;
  a.density = nullptr;
  a.ddensity = nullptr;
  a.feat_in = a.sel_in = nullptr;
  a.dtable = a.dW1 = a.db1 = a.dW2 = a.db2 = nullptr;
  a.dedup_levels = 0;
  a.v4_min_scale = v4_min_scale(1e30f);
  a.df_out = nullptr;
  a.fm = FixedMap{nullptr, nullptr, 0, 0.f};
  return 0;
}


// ---- one proposal level in ONE ray-major pass (forward) ---------------------------------------------------------
// [bins of the level] -> density of every sample (the fused field above) -> alpha / transmittance / weights and the
// median depth (RaySamples.get_weights + DepthRenderer("median")) -> inverse-CDF resampling of the next level's bins
// (PDFSampler).  A CTA of G warps owns G rays at a time: all threads evaluate the field for the G*S samples (adjacent
// lanes = adjacent samples of one ray), then warp g composites and resamples ray g out of shared memory.  Only what
// later kernels read leaves the SM: the level's bins (when they are generated here: the initial sampler of level 0),
// density, weights, median depth and the next level's bins.  Replaces, per level, nkbs_initial_sampler (level 0) +
// nkbs_density_fwd + nkbs_composite_fwd + nkbs_pdf_sample and the [R,S] round trips between them
// (ref:nerfacto.py:325-332,397; ProposalNetworkSampler.generate_ray_samples).
struct LevelArgs {
  DensityArgs d;          // field parameters, origins / directions; S = samples of this level; N = R * S
  int64_t R;
  int gen_bins;           // 1: the level's bins are the initial sampler's (written to sbins / ebins); 0: read them
  const float* lin;       // [S+1] torch.linspace(0, 1, S+1)               (gen_bins)
  const float* t_rand;    // [R] per-ray jitter or null (evaluation)        (gen_bins)
  const float* nears;     // [R]
  const float* fars;      // [R]
  float* sbins;           // [R,S+1] spacing-domain bin edges of this level
  float* ebins;           // [R,S+1] euclidean bin edges of this level
  float* density;         // [R,S]
  float* weights;         // [R,S]
  float* depth_med;       // [R] or null
  // resampling of the next level
  int nb;                 // bins of the next level = its samples + 1
  float anneal;
  const float* anneal_dev;
  float hist_pad, eps;
  const float* u_base;    // [nb]
  const float* rand;      // [R] per-ray jitter or null (evaluation: centred)
  float* out_sbins;       // [R,nb]
  float* out_ebins;       // [R,nb]
};

#define LVL_PI(s) ((s) + ((s) >> 5))
__host__ __device__ constexpr int lvl_strip(int S) { return S + (S >> 5) + 1; }
// floats of dynamic shared memory per ray: euclid bins, spacing bins, strip (delta*sigma -> weights -> pdf values), cdf
__host__ __device__ constexpr int lvl_ray_floats(int S) { return 2 * (S + 1) + lvl_strip(S) + (S + 1); }

__device__ __forceinline__ float lvl_nan_to_num(float w) {
  if (w != w) return 0.f;
  if (w > FLT_MAX) return FLT_MAX;
  if (w < -FLT_MAX) return -FLT_MAX;
  return w;
}

// G rays per group, W >= G warps per CTA: all W warps evaluate the field, warps 0..G-1 then own one ray each
template <int H, int L, int G, int W>
__global__ void __launch_bounds__(32 * W) proposal_level_fwd_kernel(const LevelArgs a) {
  nkbs_pdl_sync();
  __shared__ float s_scale[NKBS_MAX_LEVELS];
  __shared__ __align__(16) Smem<H, L> s;
  extern __shared__ float s_dyn[];
  constexpr int kT = 32 * W;
  const DensityArgs& d = a.d;
  if (threadIdx.x < NKBS_MAX_LEVELS) s_scale[threadIdx.x] = d.lv.scale[threadIdx.x];
  {
    constexpr int K = 2 * L, KP = Smem<H, L>::KP;
    for (int e = threadIdx.x; e < H * KP; e += kT) {
      int j = e / KP, i = e - j * KP;
      s.w1[e] = i < K ? __ldg(d.W1 + j * K + i) : 0.f;
    }
    for (int j = threadIdx.x; j < H; j += kT) {
      s.b1[j] = __ldg(d.b1 + j);
      s.w2[j] = __ldg(d.W2 + j);
    }
    if (threadIdx.x == 0) s.b2 = __ldg(d.b2);
  }
  const int S = d.S, nbin = S + 1;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int per_ray = lvl_ray_floats(S);
  const PdfParams pp{S, a.nb, a.anneal_dev ? *a.anneal_dev : a.anneal, a.hist_pad, a.eps, a.u_base};
  __syncthreads();
  for (int64_t g0 = (int64_t)blockIdx.x * G; g0 < a.R; g0 += (int64_t)gridDim.x * G) {
    const int rays = a.R - g0 < G ? (int)(a.R - g0) : G;
    // ---- bins of the group -> shared memory (generated here for level 0) ----
    for (int e = threadIdx.x; e < rays * nbin; e += kT) {
      const int g = e / nbin, i = e - g * nbin;
      const int64_t r = g0 + g;
      float* ray = s_dyn + g * per_ray;
      float sb, eb;
      if (a.gen_bins) {
        sb = nkbs_initial_bin(a.lin, i, S, a.t_rand != nullptr, a.t_rand ? a.t_rand[r] : 0.f);
        eb = nkbs_to_euclid(sb, nkbs_spacing_fn(a.nears[r]), nkbs_spacing_fn(a.fars[r]));
        a.sbins[r * nbin + i] = sb;
        a.ebins[r * nbin + i] = eb;
      } else {
        sb = a.sbins[r * nbin + i];
        eb = a.ebins[r * nbin + i];
      }
      ray[i] = eb;
      ray[nbin + i] = sb;
    }
    __syncthreads();
    // ---- the field on every sample of the group ----
    for (int e = threadIdx.x; e < rays * S; e += kT) {
      const int g = e / S, sidx = e - g * S;
      const int64_t r = g0 + g;
      float* ray = s_dyn + g * per_ray;
      const float st = ray[sidx], en = ray[sidx + 1];
      const float3 o = make_float3(d.origins[3 * r], d.origins[3 * r + 1], d.origins[3 * r + 2]);
      const float3 dir = make_float3(d.directions[3 * r], d.directions[3 * r + 1], d.directions[3 * r + 2]);
      float sel;
      const float3 x = nkbs_normalise(nkbs_frustum_position(o, dir, st, en), d.norm, sel);
      float f[2 * L];
      encode_sample<L>(d, s_scale, x, f);
      float out = s.b2;
#pragma unroll
      for (int j = 0; j < H; ++j) out = fmaf(s.w2[j], fmaxf(hidden_pre<H, L>(s, f, j), 0.f), out);
      const float sigma = expf(out) * sel;
      a.density[r * S + sidx] = sigma;
      ray[2 * nbin + LVL_PI(sidx)] = (en - st) * sigma;
    }
    __syncthreads();
    // ---- warp g: weights, median depth and the next level's bins of ray g ----
    if (wid < rays) {
      const int64_t r = g0 + wid;
      float* ray = s_dyn + wid * per_ray;
      const float* eb = ray;
      const float* sb = ray + nbin;
      float* sw = ray + 2 * nbin;
      float* scdf = sw + lvl_strip(S);
      const int per = (S + 31) >> 5;
      const int r0 = min(lane * per, S), r1 = min(r0 + per, S);
      // transmittance: serial prefix inside the lane's contiguous run + one warp scan of the run totals
      // (same arithmetic as composite_fwd_kernel)
      float tot = 0.f;
      for (int q = r0; q < r1; ++q) tot += sw[LVL_PI(q)];
      float incl = tot;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
      }
      float run = incl - tot;
      for (int q = r0; q < r1; ++q) {
        const float x = sw[LVL_PI(q)];
        sw[LVL_PI(q)] = lvl_nan_to_num((1.f - expf(-x)) * expf(-run));
        run += x;
      }
      __syncwarp();
      for (int q = lane; q < S; q += 32) a.weights[r * S + q] = sw[LVL_PI(q)];
      if (a.depth_med) {
        // first sample whose inclusive cumulative weight (double accumulator rounded per element) reaches 0.5
        double local = 0.0;
        for (int q = r0; q < r1; ++q) local += (double)sw[LVL_PI(q)];
        double inc = local;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const double t = __shfl_up_sync(0xffffffffu, inc, o);
          if (lane >= o) inc += t;
        }
        double cum = inc - local;
        int idx = S;
        for (int q = r0; q < r1; ++q) {
          cum += (double)sw[LVL_PI(q)];
          if ((float)cum >= 0.5f) {
            idx = q;
            break;
          }
        }
        const unsigned hit = __ballot_sync(0xffffffffu, idx < S);
        idx = hit ? __shfl_sync(0xffffffffu, idx, __ffs(hit) - 1) : S - 1;
        if (lane == 0) a.depth_med[r] = (eb[idx] + eb[idx + 1]) / 2.f;
      }
      __syncwarp();
      // resampling: the pdf values overwrite the weights in the strip (sv and the weight source are the same
      // padded strip: element s is read before it is rewritten by the same lane)
      struct Strip {
        float* p;
        __device__ __forceinline__ float& operator[](int q) const { return p[LVL_PI(q)]; }
      };
      Strip sv{sw};
      nkbs_pdf_ray_warp(pp, [&](int q) { return sw[LVL_PI(q)]; }, [&](int k) { return sb[k]; }, sv, scdf, lane,
                        a.rand != nullptr, a.rand ? a.rand[r] : 0.f, nullptr, a.nears[r], a.fars[r],
                        a.out_sbins + r * a.nb, a.out_ebins + r * a.nb, nullptr);
    }
    __syncthreads();
  }
}

}  // namespace

static int density_fwd_launch(const char* who, int32_t mode, const float* x, const float* origins, const float* directions,
                              const float* starts, const float* ends, int32_t bin_stride, int64_t R, int32_t S,
                              int32_t contraction, const float* aabb, const float* table, const float* scalings, int32_t L,
                              int32_t log2_T, const float* W1, const float* b1, const float* W2, const float* b2,
                              int32_t H, float* density, float* feat_out, float* xn_out, float* sel_out, void* stream) {
  DensityArgs a;
  if (fill(a, who, mode, x, origins, directions, starts, ends, bin_stride, R, S, contraction, aabb, table, scalings,
           L, log2_T, W1, b1, W2, b2, H))
    return -1;
  if (a.N == 0) return 0;
  NKBS_CHECK_ARG(density, "%s: null output", who);
  a.density = density;
  a.feat_out = feat_out;
  a.xn_out = xn_out;
  a.sel_out = sel_out;
  cudaStream_t st = (cudaStream_t)stream;
  if (feat_out) {
    NKBS_CHECK_ARG(sel_out, "%s: feat_out and sel_out come together", who);
    NKBS_CHECK_ARG((reinterpret_cast<uintptr_t>(feat_out) & 15u) == 0, "%s: feat_out must be 16-byte aligned", who);
    if (H == 16)
      nkbs_launch(density_fwd_kernel<16, 5, true>, grid_for(density_fwd_kernel<16, 5, true>, a.N), kThreads, 0, st, a);
    else
      nkbs_launch(density_fwd_kernel<64, 8, true>, grid_for(density_fwd_kernel<64, 8, true>, a.N), kThreads, 0, st, a);
  } else {
    if (H == 16)
      nkbs_launch(density_fwd_kernel<16, 5, false>, grid_for(density_fwd_kernel<16, 5, false>, a.N), kThreads, 0, st, a);
    else
      nkbs_launch(density_fwd_kernel<64, 8, false>, grid_for(density_fwd_kernel<64, 8, false>, a.N), kThreads, 0, st, a);
  }
  NKBS_CHECK_LAUNCH(who);
  return 0;
}

extern "C" int nkbs_density_fwd(int32_t mode, const float* x, const float* origins, const float* directions,
                                const float* starts, const float* ends, int32_t bin_stride, int64_t R, int32_t S,
                                int32_t contraction,
                                const float* aabb, const float* table, const float* scalings, int32_t L,
                                int32_t log2_T, const float* W1, const float* b1, const float* W2, const float* b2,
                                int32_t H, float* density, void* stream) {
  return density_fwd_launch("nkbs_density_fwd", mode, x, origins, directions, starts, ends, bin_stride, R, S, contraction, aabb,
                            table, scalings, L, log2_T, W1, b1, W2, b2, H, density, nullptr, nullptr, nullptr, stream);
}

extern "C" int nkbs_density_fwd_save(int32_t mode, const float* x, const float* origins, const float* directions,
                                     const float* starts, const float* ends, int32_t bin_stride, int64_t R, int32_t S,
                                     int32_t contraction, const float* aabb, const float* table, const float* scalings,
                                     int32_t L, int32_t log2_T, const float* W1, const float* b1, const float* W2,
                                     const float* b2, int32_t H, float* density, float* feat_out, float* xn_out,
                                     float* sel_out, void* stream) {
  NKBS_CHECK_ARG(feat_out && sel_out, "nkbs_density_fwd_save: null outputs");
  return density_fwd_launch("nkbs_density_fwd_save", mode, x, origins, directions, starts, ends, bin_stride, R, S, contraction,
                            aabb, table, scalings, L, log2_T, W1, b1, W2, b2, H, density, feat_out, xn_out, sel_out, stream);
}

extern "C" int nkbs_density_bwd(int32_t mode, const float* x, const float* origins, const float* directions,
                                const float* starts, const float* ends, int32_t bin_stride, int64_t R, int32_t S,
                                int32_t contraction,
                                const float* aabb, const float* table, const float* scalings, int32_t L,
                                int32_t log2_T, const float* W1, const float* b1, const float* W2, const float* b2,
                                int32_t H, const float* ddensity, int32_t dedup_levels, float* dtable, float* dW1,
                                float* db1, float* dW2, float* db2, float* dfeat_out, void* stream) {
  DensityArgs a;
  if (fill(a, "nkbs_density_bwd", mode, x, origins, directions, starts, ends, bin_stride, R, S, contraction, aabb, table, scalings,
           L, log2_T, W1, b1, W2, b2, H))
    return -1;
  if (a.N == 0) return 0;
  NKBS_CHECK_ARG(ddensity && (dtable || dfeat_out) && dW1 && db1 && dW2 && db2, "nkbs_density_bwd: null gradient buffers");
  NKBS_CHECK_ARG((reinterpret_cast<uintptr_t>(dtable) & 15u) == 0, "nkbs_density_bwd: dtable must be 16-byte aligned");
  NKBS_CHECK_ARG((reinterpret_cast<uintptr_t>(dfeat_out) & 7u) == 0, "nkbs_density_bwd: dfeat_out must be 8-byte aligned");
  a.df_out = dfeat_out;
  a.fm = nkbs_fixed_map_for(dW1 ? (const void*)dW1 : (const void*)dtable);
  a.ddensity = ddensity;
  a.dtable = dtable;
  a.dW1 = dW1;
  a.db1 = db1;
  a.dW2 = dW2;
  a.db2 = db2;
  a.dedup_levels = dedup_levels < 0 ? L : dedup_levels;
  cudaStream_t st = (cudaStream_t)stream;
  if (H == 16) {
    constexpr size_t smem = (size_t)kWarps * MmaGeom<16, 5>::kScratch * sizeof(float);
    static_assert(smem <= 48 * 1024, "default dynamic shared memory limit");
    if (a.fm.shadow) nkbs_launch(density_bwd_kernel<16, 5, true>, grid_for(density_bwd_kernel<16, 5, true>, a.N, smem), kThreads, smem, st, a);
    else nkbs_launch(density_bwd_kernel<16, 5, false>, grid_for(density_bwd_kernel<16, 5, false>, a.N, smem), kThreads, smem, st, a);
  } else {
    constexpr size_t smem = (size_t)kWarps * MmaGeom<64, 8>::kScratch * sizeof(float);
    if (smem > 48 * 1024) {
      cudaError_t e = cudaFuncSetAttribute(density_bwd_kernel<64, 8, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(density_bwd_kernel<64, 8, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      NKBS_CHECK_ARG(e == cudaSuccess, "nkbs_density_bwd: %zu bytes of shared memory refused (%s)", smem, cudaGetErrorString(e));
    }
    if (a.fm.shadow) nkbs_launch(density_bwd_kernel<64, 8, true>, grid_for(density_bwd_kernel<64, 8, true>, a.N, smem), kThreads, smem, st, a);
    else nkbs_launch(density_bwd_kernel<64, 8, false>, grid_for(density_bwd_kernel<64, 8, false>, a.N, smem), kThreads, smem, st, a);
  }
  NKBS_CHECK_LAUNCH("nkbs_density_bwd");
  return 0;
}

extern "C" int nkbs_density_bwd_saved(int64_t N, int32_t L, int32_t H, const float* feat_in, const float* sel_in,
                                      const float* ddensity, const float* W1, const float* b1, const float* W2,
                                      const float* b2, float* dW1, float* db1, float* dW2, float* db2, float* dfeat_out,
                                      void* stream) {
  NKBS_CHECK_ARG(N >= 0, "nkbs_density_bwd_saved: negative N");
  NKBS_CHECK_ARG(H == 16 && L == 5, "nkbs_density_bwd_saved: built for the (L=5, H=16) field; got L=%d H=%d", L, H);
  if (N == 0) return 0;
  NKBS_CHECK_ARG(feat_in && sel_in && ddensity && W1 && b1 && W2 && b2 && dW1 && db1 && dW2 && db2 && dfeat_out,
                 "nkbs_density_bwd_saved: null pointers");
  NKBS_CHECK_ARG(((reinterpret_cast<uintptr_t>(feat_in) | reinterpret_cast<uintptr_t>(dfeat_out)) & 7u) == 0,
                 "nkbs_density_bwd_saved: feat_in / dfeat_out must be 8-byte aligned");
  DensityArgs a = {};
  a.mode = 1;
  a.S = 1;
  a.bstride = 1;
  a.N = N;
  a.lv.num_levels = L;
  a.W1 = W1;
  a.b1 = b1;
  a.W2 = W2;
  a.b2 = b2;
  a.feat_in = feat_in;
  a.sel_in = sel_in;
  a.ddensity = ddensity;
  a.dW1 = dW1;
  a.db1 = db1;
  a.dW2 = dW2;
  a.db2 = db2;
  a.df_out = dfeat_out;
  a.fm = nkbs_fixed_map_for(dW1);
  cudaStream_t st = (cudaStream_t)stream;
  constexpr size_t smem = (size_t)kWarps * MmaGeom<16, 5>::kScratch * sizeof(float);
  // resident CTAs per SM the kernel is compiled for (registers 128 / 80 / 64 at 2 / 3 / 4): without the gathers the
  // kernel streams, and more warps in flight win slightly — 4096-ray step 0.7466 / 0.7463 / 0.7445 ms, 65 536 rays
  // 7.538 / 7.535 / 7.482 ms.  NKBS_SAVED_BWD_CTAS = 2, 3 overrides.
  static const int minb = [] { const char* e = getenv("NKBS_SAVED_BWD_CTAS"); return e ? atoi(e) : 4; }();
  if (a.fm.shadow)
    nkbs_launch(density_bwd_kernel<16, 5, true, true>, grid_for(density_bwd_kernel<16, 5, true, true>, N, smem), kThreads, smem, st, a);
  else if (minb == 3)
    nkbs_launch(density_bwd_kernel<16, 5, false, true, 3>, grid_for(density_bwd_kernel<16, 5, false, true, 3>, N, smem), kThreads, smem, st, a);
  else if (minb == 4)
    nkbs_launch(density_bwd_kernel<16, 5, false, true, 4>, grid_for(density_bwd_kernel<16, 5, false, true, 4>, N, smem), kThreads, smem, st, a);
  else
    nkbs_launch(density_bwd_kernel<16, 5, false, true>, grid_for(density_bwd_kernel<16, 5, false, true>, N, smem), kThreads, smem, st, a);
  NKBS_CHECK_LAUNCH("nkbs_density_bwd_saved");
  return 0;
}

extern "C" int nkbs_proposal_level_fwd(int64_t R, int32_t S, const float* origins, const float* directions,
                                       const float* nears, const float* fars, int32_t gen_bins, const float* lin,
                                       const float* t_rand, float* sbins, float* ebins, int32_t contraction,
                                       const float* aabb, const float* table, const float* scalings, int32_t L,
                                       int32_t log2_T, const float* W1, const float* b1, const float* W2,
                                       const float* b2, int32_t H, float* density, float* weights, float* depth_med,
                                       int32_t num_bins, float anneal, const float* anneal_dev, float histogram_padding,
                                       float eps, const float* u_base, const float* rand, float* out_sbins,
                                       float* out_ebins, void* stream) {
  LevelArgs a;
  // the bin arrays double as the frustum source: starts = ebins[:, :-1], ends = ebins[:, 1:]
  if (fill(a.d, "nkbs_proposal_level_fwd", 2, nullptr, origins, directions, ebins, ebins ? ebins + 1 : nullptr, S + 1, R, S,
           contraction, aabb, table, scalings, L, log2_T, W1, b1, W2, b2, H))
    return -1;
  if (R == 0) return 0;
  NKBS_CHECK_ARG(S <= 1024 && num_bins >= 2, "nkbs_proposal_level_fwd: S=%d (max 1024), bins=%d", S, num_bins);
  NKBS_CHECK_ARG(nears && fars && sbins && density && weights && u_base && out_sbins && out_ebins,
                 "nkbs_proposal_level_fwd: null argument");
  NKBS_CHECK_ARG(!gen_bins || lin, "nkbs_proposal_level_fwd: the initial sampler needs the linspace table");
  a.R = R;
  a.gen_bins = gen_bins;
  a.lin = lin;
  a.t_rand = t_rand;
  a.nears = nears;
  a.fars = fars;
  a.sbins = sbins;
  a.ebins = ebins;
  a.density = density;
  a.weights = weights;
  a.depth_med = depth_med;
  a.nb = num_bins;
  a.anneal = anneal;
  a.anneal_dev = anneal_dev;
  a.hist_pad = histogram_padding;
  a.eps = eps;
  a.u_base = u_base;
  a.rand = rand;
  a.out_sbins = out_sbins;
  a.out_ebins = out_ebins;
  // CTA shape: as many threads as the group has samples where possible (one field evaluation per thread keeps the
  // gathers of a whole wave in flight, like nkbs_density_fwd), G rays per group.  NKBS_LEVEL_SHAPE=<G><W> overrides.
  static const int forced = [] { const char* e = getenv("NKBS_LEVEL_SHAPE"); return e ? atoi(e) : 0; }();
  int shape = forced ? forced : (S >= 224 ? 18 : (S >= 80 && S <= 96 ? 26 : 44));
  const int G = shape / 10, Wn = shape % 10;
  NKBS_CHECK_ARG((G == 1 && Wn == 8) || (G == 2 && Wn == 6) || (G == 4 && Wn == 4) || (G == 2 && Wn == 8),
                 "nkbs_proposal_level_fwd: unsupported NKBS_LEVEL_SHAPE %d", shape);
  const size_t smem = (size_t)G * lvl_ray_floats(S) * sizeof(float);
  int64_t groups = (R + G - 1) / G;
  const int64_t cap = (int64_t)nkbs_num_sms() * 32;
  const int grid = (int)(groups < cap ? groups : cap);
  cudaStream_t st = (cudaStream_t)stream;
  auto launch = [&](auto kernel) -> int {
    if (smem > 48 * 1024) {
      cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      NKBS_CHECK_ARG(e == cudaSuccess, "nkbs_proposal_level_fwd: %zu bytes of shared memory refused (%s)", smem, cudaGetErrorString(e));
    }
    nkbs_launch(kernel, grid, 32 * Wn, smem, st, a);
    return 0;
  };
  int rc;
#define NKBS_LVL(HH, LL)                                                        \
  (shape == 18   ? launch(proposal_level_fwd_kernel<HH, LL, 1, 8>)              \
   : shape == 26 ? launch(proposal_level_fwd_kernel<HH, LL, 2, 6>)              \
   : shape == 28 ? launch(proposal_level_fwd_kernel<HH, LL, 2, 8>)              \
                 : launch(proposal_level_fwd_kernel<HH, LL, 4, 4>))
  rc = H == 16 ? NKBS_LVL(16, 5) : NKBS_LVL(64, 8);
#undef NKBS_LVL
  if (rc) return rc;
  NKBS_CHECK_LAUNCH("nkbs_proposal_level_fwd");
  return 0;
}
