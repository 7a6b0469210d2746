// Device pieces of the samplers shared by sampler.cu (stand-alone K5 kernels) and density.cu (the fused
// proposal-level pass): the spacing <-> euclidean maps of UniformLinDispPiecewiseSampler and one warp's inverse-CDF
// resampling of one ray (PDFSampler).  Bit-exactness rules are those of sampler.cu: explicit round-to-nearest IEEE ops
// in the torch path's order, sums of float terms held in double (exact for these magnitudes, so the reduction order
// does not matter), linspace tables from the host.
#pragma once
#include <float.h>
#include "common.cuh"

__device__ __forceinline__ float nkbs_spacing_fn(float x) {
  // where(x < 1, x / 2, 1 - 1 / (2 * x))
  return x < 1.f ? __fmul_rn(x, 0.5f) : __fsub_rn(1.f, __fdiv_rn(1.f, __fmul_rn(2.f, x)));
}
__device__ __forceinline__ float nkbs_spacing_inv(float x) {
  // where(x < 0.5, 2 * x, 1 / (2 - 2 * x))
  return x < 0.5f ? __fmul_rn(2.f, x) : __fdiv_rn(1.f, __fsub_rn(2.f, __fmul_rn(2.f, x)));
}
__device__ __forceinline__ float nkbs_to_euclid(float s, float s_near, float s_far) {
  // spacing_inv(x * s_far + (1 - x) * s_near)
  return nkbs_spacing_inv(__fadd_rn(__fmul_rn(s, s_far), __fmul_rn(__fsub_rn(1.f, s), s_near)));
}

// bin edge i of the initial sampler: linspace value, stratified by the ray's jitter when there is one
__device__ __forceinline__ float nkbs_initial_bin(const float* lin, int i, int S, bool jitter, float tr) {
  float b = lin[i];
  if (jitter) {
    // centers = (bins[1:] + bins[:-1]) / 2 ; upper = cat[centers, last] ; lower = cat[first, centers]
    const float upper = (i < S) ? __fmul_rn(__fadd_rn(lin[i + 1], b), 0.5f) : b;
    const float lower = (i > 0) ? __fmul_rn(__fadd_rn(b, lin[i - 1]), 0.5f) : b;
    b = __fadd_rn(lower, __fmul_rn(__fsub_rn(upper, lower), tr));
  }
  return b;
}

__device__ __forceinline__ double nkbs_warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

struct PdfParams {
  int S;            // existing samples per ray
  int nb;           // new bins = num_samples + 1
  float anneal;
  float hist_pad;   // 0.01
  float eps;        // 1e-5
  const float* u_base;  // [nb] torch.linspace(0, 1-1/nb, nb)
};

// One warp resamples one ray.  `W(s)` returns the ray's weight s (any memory), `EB(k)` its existing spacing bin k.
// `sv` [S] (anything indexable, may alias the weights: element s is read before it is written, by the same lane) and
// `scdf` [S+1] are the warp's scratch (shared memory).  `have_rand`: u = u_base + rand / nb (per-ray
// jitter `rand_ray`, or `rand_bins[i]` when given), else the centred evaluation points.  Writes the new spacing and
// euclidean bins (and the searchsorted indices when `out_inds` is not null).
template <typename WF, typename BF, typename SV>
__device__ __forceinline__ void nkbs_pdf_ray_warp(const PdfParams& p, WF W, BF EB, SV sv, float* scdf, int lane,
                                                  bool have_rand, float rand_ray, const float* rand_bins, float near,
                                                  float far, float* out_sbins, float* out_ebins, int64_t* out_inds) {
  const int S = p.S, nb = p.nb;
  const bool do_pow = p.anneal != 1.0f;
  const float fS = (float)S, fnb = (float)nb;
  const float eval_off = (float)(1.0 / (2.0 * (double)nb));
  // pass 1: v = pow(w, anneal); weights_sum = sum(v + histogram_padding)
  double dsum = 0.0;
  for (int s = lane; s < S; s += 32) {
    float v = W(s);
    if (do_pow) v = powf(v, p.anneal);
    sv[s] = v;
    dsum += (double)__fadd_rn(v, p.hist_pad);
  }
  dsum = nkbs_warp_sum_d(dsum);
  float sum = (float)dsum;
  const float padding = fmaxf(__fsub_rn(p.eps, sum), 0.f);
  const float pad_each = __fdiv_rn(padding, fS);
  sum = __fadd_rn(sum, padding);
  // pass 2: cdf[k] = min(1, float(sum_{s<k} pdf_s)), running sum in double, 32 samples per round
  double carry = 0.0;
  if (lane == 0) scdf[0] = 0.f;
  __syncwarp();
  for (int s0 = 0; s0 < S; s0 += 32) {
    const int s = s0 + lane;
    double x = 0.0;
    if (s < S) x = (double)__fdiv_rn(__fadd_rn(__fadd_rn(sv[s], p.hist_pad), pad_each), sum);
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      double n = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += n;
    }
    x += carry;
    if (s < S) scdf[s + 1] = fminf(1.f, (float)x);
    carry = __shfl_sync(0xffffffffu, x, 31);
  }
  __syncwarp();
  const float s_near = nkbs_spacing_fn(near), s_far = nkbs_spacing_fn(far);
  float rj = 0.f;
  if (have_rand && !rand_bins) rj = __fdiv_rn(rand_ray, fnb);
  for (int i = lane; i < nb; i += 32) {
    float u = p.u_base[i];
    if (have_rand) {
      if (rand_bins) rj = __fdiv_rn(rand_bins[i], fnb);
      u = __fadd_rn(u, rj);
    } else {
      u = __fadd_rn(u, eval_off);
    }
    // idx = number of entries of cdf[0..S] that are <= u  (side="right")
    int lo = 0, hi = S + 1;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (scdf[mid] <= u) lo = mid + 1;
      else hi = mid;
    }
    const int idx = lo;
    const int below = idx - 1 < 0 ? 0 : (idx - 1 > S ? S : idx - 1);
    const int above = idx > S ? S : idx;
    const float g0 = scdf[below], g1 = scdf[above];
    const float b0 = EB(below), b1 = EB(above);
    float t = __fdiv_rn(__fsub_rn(u, g0), __fsub_rn(g1, g0));
    if (t != t) t = 0.f;                    // nan_to_num(nan=0)
    else if (t > FLT_MAX) t = FLT_MAX;      // +inf -> max
    else if (t < -FLT_MAX) t = -FLT_MAX;
    t = fminf(fmaxf(t, 0.f), 1.f);
    const float b = __fadd_rn(b0, __fmul_rn(t, __fsub_rn(b1, b0)));
    out_sbins[i] = b;
    out_ebins[i] = nkbs_to_euclid(b, s_near, s_far);
    if (out_inds) out_inds[i] = idx;
  }
  __syncwarp();
}
