// K5: initial piecewise lin-disparity sampler and inverse-CDF (PDF) resampling.
// Replaces nerfstudio UniformLinDispPiecewiseSampler and PDFSampler inside ProposalNetworkSampler,
// reached from ref:nerfacto.py:325-332,397 and ref:semantic_nerfw.py:132-137,187.
//
// Bit-exactness: every op is an explicit round-to-nearest IEEE op in the torch path's order; sums and
// the cumulative distribution are accumulated in sample order; the cdf uses a double accumulator rounded
// to float per entry, which is what torch.cumsum does on the CPU (the normaliser torch.sum uses a
// vectorised float order that no kernel can reproduce: there the kernel sums floats in sample order).  The
// linspace tables come from the host (torch.linspace on the CPU) so their rounding is the oracle's.
#include <float.h>
#include "common.cuh"
#include "sampler.cuh"

namespace {

constexpr int kThreads = 128;

__device__ __forceinline__ float spacing_fn(float x) { return nkbs_spacing_fn(x); }
__device__ __forceinline__ float to_euclid(float s, float s_near, float s_far) { return nkbs_to_euclid(s, s_near, s_far); }

struct InitArgs {
  int64_t R;
  int S;
  const float* lin;     // [S+1] torch.linspace(0,1,S+1)
  const float* t_rand;  // [R] or [R,S+1] or null (eval: no jitter)
  int rand_per_bin;
  const float* nears;   // [R]
  const float* fars;    // [R]
  float* sbins;         // [R,S+1] spacing-domain bins
  float* ebins;         // [R,S+1] euclidean bins
};

// One warp per ray: the two spacing_fn divisions of the ray's near / far planes are evaluated once per ray (a thread per
// bin edge re-evaluated them, and a 64-bit division for its ray index, for each of the S + 1 edges), lanes stride over the
// edges (coalesced stores).
__global__ void __launch_bounds__(256) initial_sampler_kernel(const InitArgs a) {
  nkbs_pdl_sync();
  const int nb = a.S + 1;
  const int lane = threadIdx.x & 31;
  const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < a.R; r += nwarps) {
    const float s_near = spacing_fn(a.nears[r]), s_far = spacing_fn(a.fars[r]);
    const float tr = (a.t_rand && !a.rand_per_bin) ? a.t_rand[r] : 0.f;
    float* sb = a.sbins + r * nb;
    float* eb = a.ebins + r * nb;
    for (int i = lane; i < nb; i += 32) {
      const float t = (a.t_rand && a.rand_per_bin) ? a.t_rand[r * nb + i] : tr;
      const float b = nkbs_initial_bin(a.lin, i, a.S, a.t_rand != nullptr, t);
      sb[i] = b;
      eb[i] = to_euclid(b, s_near, s_far);
    }
  }
}

struct PdfArgs {
  int64_t R;
  int S;                // existing samples per ray
  int nb;               // new bins = num_samples + 1
  const float* weights; // [R,S]
  float anneal;
  const float* anneal_dev; // optional device scalar overriding `anneal` (CUDA-graph replay)
  float hist_pad;       // 0.01
  float eps;            // 1e-5
  const float* sbins;   // [R,S+1] existing spacing bins
  const float* u_base;  // [nb] torch.linspace(0, 1-1/nb, nb)
  const float* rand;    // [R] or [R,nb] or null (eval: centred)
  int rand_per_bin;
  const float* nears;
  const float* fars;
  float* out_sbins;     // [R,nb]
  float* out_ebins;     // [R,nb]
  int64_t* out_inds;    // [R,nb] searchsorted result, optional
};

// One thread per ray.  The query points u are increasing, so searchsorted(cdf, u, right) is a merge
// walk: the cdf is produced on the fly in sample order and never stored.
__global__ void __launch_bounds__(kThreads) pdf_sample_kernel(const PdfArgs a) {
  nkbs_pdl_sync();
  const int S = a.S, nb = a.nb;
  const float anneal = a.anneal_dev ? *a.anneal_dev : a.anneal;
  const float fS = (float)S, fnb = (float)nb;
  const float eval_off = (float)(1.0 / (2.0 * (double)nb));
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < a.R; r += (int64_t)gridDim.x * blockDim.x) {
    const float* w = a.weights + r * S;
    const float* eb = a.sbins + r * (S + 1);
    const bool do_pow = anneal != 1.0f;
    // pass 1: weights_sum = sum(pow(w, anneal) + histogram_padding)
    // The normaliser is accumulated in double and rounded once: torch.sum's vectorised float order on the
    // CPU is machine dependent, the correctly rounded sum is the closest well-defined stand-in.
    double dsum = 0.0;
    for (int s = 0; s < S; ++s) {
      float v = w[s];
      if (do_pow) v = powf(v, anneal);
      dsum += (double)__fadd_rn(v, a.hist_pad);
    }
    float sum = (float)dsum;
    float padding = fmaxf(__fsub_rn(a.eps, sum), 0.f);
    float pad_each = __fdiv_rn(padding, fS);
    sum = __fadd_rn(sum, padding);
    float s_near = spacing_fn(a.nears[r]), s_far = spacing_fn(a.fars[r]);
    // merge walk state: idx = number of cdf entries known to be <= u so far
    int idx = 0;            // candidate searchsorted result
    double run = 0.0;       // unclipped cumsum of pdf up to entry idx (double accumulator, as torch.cumsum on the CPU)
    float c_cur = 0.f;      // cdf[idx]   (cdf[0] = 0)
    float c_prev = 0.f;     // cdf[idx-1] (unused while idx == 0)
    float rj = 0.f;
    if (a.rand && !a.rand_per_bin) rj = __fdiv_rn(a.rand[r], fnb);
    for (int i = 0; i < nb; ++i) {
      float u = a.u_base[i];
      if (a.rand) {
        if (a.rand_per_bin) rj = __fdiv_rn(a.rand[r * nb + i], fnb);
        u = __fadd_rn(u, rj);
      } else {
        u = __fadd_rn(u, eval_off);
      }
      // advance while cdf[idx] <= u  (side="right": first index whose value is > u)
      while (idx <= S && c_cur <= u) {
        c_prev = c_cur;
        ++idx;
        if (idx <= S) {
          float v = w[idx - 1];
          if (do_pow) v = powf(v, anneal);
          v = __fadd_rn(__fadd_rn(v, a.hist_pad), pad_each);
          run += (double)__fdiv_rn(v, sum);
          c_cur = fminf(1.f, (float)run);
        }
      }
      int below = idx - 1 < 0 ? 0 : (idx - 1 > S ? S : idx - 1);
      int above = idx > S ? S : idx;
      float g0, g1;
      if (idx == 0) {
        g0 = c_cur;  // cdf[0]
        g1 = c_cur;
      } else if (idx > S) {
        g0 = c_prev;  // cdf[S]
        g1 = c_prev;
      } else {
        g0 = c_prev;
        g1 = c_cur;
      }
      float b0 = eb[below], b1 = eb[above];
      float t = __fdiv_rn(__fsub_rn(u, g0), __fsub_rn(g1, g0));
      if (t != t) t = 0.f;                     // nan_to_num(nan=0)
      else if (t > FLT_MAX) t = FLT_MAX;       // +inf -> max
      else if (t < -FLT_MAX) t = -FLT_MAX;
      t = fminf(fmaxf(t, 0.f), 1.f);
      float b = __fadd_rn(b0, __fmul_rn(t, __fsub_rn(b1, b0)));
      a.out_sbins[r * nb + i] = b;
      a.out_ebins[r * nb + i] = to_euclid(b, s_near, s_far);
      if (a.out_inds) a.out_inds[r * nb + i] = idx;
    }
  }
}

// One warp per ray (S <= kWarpPdfMaxS).  Same arithmetic as pdf_sample_kernel: the normaliser and the running cdf
// are sums of float terms held in double, which are exact for these magnitudes (terms >= histogram_padding / sum,
// totals < 2^9: fewer than 53 significant bits), so the warp-parallel reduction / scan produce the very bits of the
// sequential accumulation.  searchsorted(cdf, u, right) becomes a binary search in the warp's shared-memory cdf.
constexpr int kWarpPdfMaxS = 1024;
constexpr int kPdfWarps = 8;

__global__ void __launch_bounds__(kPdfWarps * 32) pdf_sample_warp_kernel(const PdfArgs a) {
  nkbs_pdl_sync();
  extern __shared__ float s_pdf[];  // per warp: v[S] then cdf[S+1]
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int S = a.S, nb = a.nb;
  float* sv = s_pdf + wid * (2 * S + 1);
  float* scdf = sv + S;
  const PdfParams p{S, nb, a.anneal_dev ? *a.anneal_dev : a.anneal, a.hist_pad, a.eps, a.u_base};
  for (int64_t r = blockIdx.x * (int64_t)kPdfWarps + wid; r < a.R; r += (int64_t)gridDim.x * kPdfWarps) {
    const float* w = a.weights + r * S;
    const float* eb = a.sbins + r * (S + 1);
    nkbs_pdf_ray_warp(p, [&](int s) { return w[s]; }, [&](int k) { return eb[k]; }, sv, scdf, lane, a.rand != nullptr,
                      (a.rand && !a.rand_per_bin) ? a.rand[r] : 0.f, (a.rand && a.rand_per_bin) ? a.rand + r * nb : nullptr,
                      a.nears[r], a.fars[r], a.out_sbins + r * nb, a.out_ebins + r * nb,
                      a.out_inds ? a.out_inds + r * nb : nullptr);
  }
}

}  // namespace

extern "C" int nkbs_initial_sampler(int64_t R, int32_t S, const float* lin, const float* t_rand,
                                    int32_t rand_per_bin, const float* nears, const float* fars, float* sbins,
                                    float* ebins, void* stream) {
  NKBS_CHECK_ARG(R >= 0 && S >= 1, "nkbs_initial_sampler: bad sizes R=%lld S=%d", (long long)R, S);
  if (R == 0) return 0;
  NKBS_CHECK_ARG(lin && nears && fars && sbins && ebins, "nkbs_initial_sampler: null argument");
  InitArgs a{R, S, lin, t_rand, rand_per_bin, nears, fars, sbins, ebins};
  int64_t blocks = (R + 7) / 8;  // eight rays (warps) per CTA
  int64_t cap = (int64_t)nkbs_num_sms() * 16;
  nkbs_launch(initial_sampler_kernel, (int)(blocks < cap ? blocks : cap), 256, 0, (cudaStream_t)stream, a);
  NKBS_CHECK_LAUNCH("nkbs_initial_sampler");
  return 0;
}

extern "C" int nkbs_pdf_sample(int64_t R, int32_t S, int32_t num_bins, const float* weights, float anneal,
                               const float* anneal_dev,
                               float histogram_padding, float eps, const float* sbins, const float* u_base,
                               const float* rand, int32_t rand_per_bin, const float* nears, const float* fars,
                               float* out_sbins, float* out_ebins, int64_t* out_inds, void* stream) {
  NKBS_CHECK_ARG(R >= 0 && S >= 1 && num_bins >= 2, "nkbs_pdf_sample: bad sizes R=%lld S=%d bins=%d", (long long)R, S, num_bins);
  if (R == 0) return 0;
  NKBS_CHECK_ARG(weights && sbins && u_base && nears && fars && out_sbins && out_ebins, "nkbs_pdf_sample: null argument");
  PdfArgs a{R, S, num_bins, weights, anneal, anneal_dev, histogram_padding, eps, sbins, u_base, rand, rand_per_bin,
            nears, fars, out_sbins, out_ebins, out_inds};
  if (S <= kWarpPdfMaxS) {
    int64_t blocks = (R + kPdfWarps - 1) / kPdfWarps;
    int64_t cap = (int64_t)nkbs_num_sms() * 8;
    size_t smem = (size_t)kPdfWarps * (2 * S + 1) * sizeof(float);
    static thread_local size_t configured = 48 * 1024;
    if (smem > configured) {
      cudaFuncSetAttribute(pdf_sample_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      configured = smem;
    }
    nkbs_launch(pdf_sample_warp_kernel, (int)(blocks < cap ? blocks : cap), kPdfWarps * 32, smem, (cudaStream_t)stream, a);
  } else {
    int64_t blocks = (R + kThreads - 1) / kThreads;
    nkbs_launch(pdf_sample_kernel, (int)blocks, kThreads, 0, (cudaStream_t)stream, a);
  }
  NKBS_CHECK_LAUNCH("nkbs_pdf_sample");
  return 0;
}
