// K6/K7: ray-major volume compositing, forward and backward, one warp per ray.
// Replaces RaySamples.get_weights and the RGB / Accumulation / Depth(median, expected) / Semantic /
// Uncertainty renderers reached from ref:nerfacto.py:399-449 and ref:semantic_nerfw.py:190-229
// (SURVEY a11, a13-a18, A.8).
#include <float.h>
#include "common.cuh"
#include "nkbs.h"
#include "sampler.cuh"

namespace {

constexpr int kWarpsPerCta = 8;
constexpr int kThreads = kWarpsPerCta * 32;
constexpr int kMaxS = 512;  // samples per ray held in a warp's shared-memory strip

// Per-warp strips are indexed through PI: one pad word per 32 samples, so that lanes walking contiguous runs of
// 2, 4, 8 or 16 samples (stride = run length) and lanes striding by 32 both hit distinct banks.
#define PI(s) ((s) + ((s) >> 5))
__host__ __device__ constexpr int strip_len(int S) { return S + (S >> 5) + 1; }

struct CompArgs {
  int64_t R;
  int S;
  // weights source: density (compute alpha/transmittance) or precomputed weights
  const float* density;   // [R,S] or null
  const float* weights_in;// [R,S] or null
  const float* starts;    // [R,S] euclidean bin starts (needed for density and depth), row stride bstride
  const float* ends;      // [R,S]
  int bstride;            // floats between consecutive rays in starts / ends (S, or S+1 for views of bin edges)
  const float* rgb;       // [R,S,3] or null
  const float* sem;       // [R,S,C] or null
  int C;
  int bg_mode;            // 0 last_sample, 1 black, 2 white
  int eval_mode;          // nan_to_num rgb in, clamp out
  // outputs (nullable)
  float* weights;         // [R,S]
  float* rgb_out;         // [R,3]
  float* acc_out;         // [R]
  float* depth_exp;       // [R] unclipped expected depth
  float* depth_med;       // [R]
  float* sem_out;         // [R,C]
  int* minmax;            // [2] ordered-int keys of min/max step (for the expected-depth clip)
  // optional tail: inverse-CDF resampling of the NEXT level's bins from this level's weights (PDFSampler), done by the
  // warp that has just computed the ray's weights, straight from its shared-memory strip (pdf_nb = 0: none)
  int pdf_nb;
  float pdf_anneal;
  const float* pdf_anneal_dev;
  float pdf_hist_pad, pdf_eps;
  const float* pdf_sbins;     // [R,S+1] spacing-domain bin edges of this level
  const float* pdf_u;         // [pdf_nb] torch.linspace(0, 1 - 1/nb, nb)
  const float* pdf_rand;      // [R] per-ray jitter or null (centred evaluation points)
  const float* nears;         // [R]
  const float* fars;          // [R]
  float* pdf_out_sbins;       // [R,pdf_nb]
  float* pdf_out_ebins;       // [R,pdf_nb]
};

__device__ __forceinline__ int float_key(float f) {
  int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float key_float(int k) { return __int_as_float(k >= 0 ? k : k ^ 0x7fffffff); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_incl_scan(float v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    float n = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += n;
  }
  return v;
}
__device__ __forceinline__ float nan_to_num(float w) {
  if (w != w) return 0.f;
  if (w > FLT_MAX) return FLT_MAX;
  if (w < -FLT_MAX) return -FLT_MAX;
  return w;
}

__global__ void __launch_bounds__(kThreads) composite_fwd_kernel(const CompArgs a) {
  nkbs_pdl_sync();
  extern __shared__ float s_w[];  // [kWarpsPerCta][strip (+ S + 1 floats of cdf when resampling)]
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int S = a.S;
  float* sw = s_w + wid * (strip_len(S) + (a.pdf_nb ? S + 1 : 0));
  const PdfParams pp{S, a.pdf_nb, (a.pdf_nb && a.pdf_anneal_dev) ? *a.pdf_anneal_dev : a.pdf_anneal, a.pdf_hist_pad, a.pdf_eps, a.pdf_u};
  float lmin = FLT_MAX, lmax = -FLT_MAX;
  for (int64_t r = blockIdx.x * (int64_t)kWarpsPerCta + wid; r < a.R; r += (int64_t)gridDim.x * kWarpsPerCta) {
    const int64_t base = r * S;
    const int64_t bb = r * a.bstride;
    // Three passes.  (A) coalesced, independent loads: delta*sigma of every sample into the strip.  (B) every lane
    // owns a contiguous run of `per` samples: serial prefix inside the run plus ONE warp scan of the run totals
    // (instead of a scan per 32-sample chunk).  (C) coalesced stores and the per-ray sums.
    const int per = (S + 31) >> 5;
    const int r0 = min(lane * per, S), r1 = min(r0 + per, S);
    float acc = 0.f, dsum = 0.f;
    if (a.density) {
      for (int s = lane; s < S; s += 32) sw[PI(s)] = (a.ends[bb + s] - a.starts[bb + s]) * a.density[base + s];
      __syncwarp();
      float tot = 0.f;
      for (int s = r0; s < r1; ++s) tot += sw[PI(s)];
      float run = warp_incl_scan(tot, lane) - tot;  // delta*sigma accumulated before this lane's run
      for (int s = r0; s < r1; ++s) {
        const float x = sw[PI(s)];
        // alphas * transmittance, nan_to_num as in get_weights
        sw[PI(s)] = nan_to_num((1.f - expf(-x)) * expf(-run));
        run += x;
      }
      __syncwarp();
    }
    for (int s = lane; s < S; s += 32) {
      float st = 0.f, en = 0.f, w;
      if (a.starts) {
        st = a.starts[bb + s];
        en = a.ends[bb + s];
      }
      if (a.density) {
        w = sw[PI(s)];
      } else {
        w = a.weights_in[base + s];
        sw[PI(s)] = w;
      }
      if (a.weights) a.weights[base + s] = w;
      const float step = (st + en) / 2.f;
      acc += w;
      dsum += w * step;
      lmin = fminf(lmin, step);
      lmax = fmaxf(lmax, step);
    }
    acc = warp_sum(acc);
    dsum = warp_sum(dsum);
    __syncwarp();
    if (a.acc_out && lane == 0) a.acc_out[r] = acc;
    if (a.depth_exp && lane == 0) a.depth_exp[r] = dsum / (acc + 1e-10f);
    if (a.depth_med) {
      // first index whose inclusive cumulative weight (double accumulator rounded to float per element, what
      // torch.cumsum does on the CPU) reaches 0.5.  Every lane owns a contiguous run of samples: run totals are
      // scanned across the warp in double, then each lane walks its own run.
      const int per = (S + 31) >> 5;
      const int s0 = min(lane * per, S), s1 = min(s0 + per, S);
      double local = 0.0;
      for (int s = s0; s < s1; ++s) local += (double)sw[PI(s)];
      double incl = local;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const double t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      double cum = incl - local;
      int idx = S;
      for (int s = s0; s < s1; ++s) {
        cum += (double)sw[PI(s)];
        if ((float)cum >= 0.5f) {
          idx = s;
          break;
        }
      }
      const unsigned hit = __ballot_sync(0xffffffffu, idx < S);
      idx = hit ? __shfl_sync(0xffffffffu, idx, __ffs(hit) - 1) : S - 1;
      if (lane == 0) a.depth_med[r] = (a.starts[bb + idx] + a.ends[bb + idx]) / 2.f;
    }
    if (a.rgb_out) {
      // lanes 0..2 = colour channel accumulators would idle 29 lanes; instead stride the flat [S*3] strip
      float c0 = 0.f, c1 = 0.f, c2 = 0.f;
      const float* rgb = a.rgb + base * 3;
      for (int s = lane; s < S; s += 32) {
        float w = sw[PI(s)];
        float r0 = rgb[3 * s], g0 = rgb[3 * s + 1], b0 = rgb[3 * s + 2];
        if (a.eval_mode) {
          r0 = nan_to_num(r0);
          g0 = nan_to_num(g0);
          b0 = nan_to_num(b0);
        }
        c0 += w * r0;
        c1 += w * g0;
        c2 += w * b0;
      }
      c0 = warp_sum(c0);
      c1 = warp_sum(c1);
      c2 = warp_sum(c2);
      if (lane < 3) {
        float comp = lane == 0 ? c0 : (lane == 1 ? c1 : c2);
        float bg = 0.f;
        if (a.bg_mode == 0) {
          bg = rgb[3 * (S - 1) + lane];
          if (a.eval_mode) bg = nan_to_num(bg);
        } else if (a.bg_mode == 2) {
          bg = 1.f;
        }
        comp = comp + bg * (1.0f - acc);
        if (a.eval_mode) comp = fminf(fmaxf(comp, 0.f), 1.f);
        a.rgb_out[r * 3 + lane] = comp;
      }
    }
    if (a.sem_out) {
      // lanes over classes: each sample's C logits are one contiguous read
      const float* sem = a.sem + base * a.C;
      for (int c0 = 0; c0 < a.C; c0 += 32) {
        int c = c0 + lane;
        float v = 0.f;
        if (c < a.C)
          for (int s = 0; s < S; ++s) v += sw[PI(s)] * sem[(int64_t)s * a.C + c];
        if (c < a.C) a.sem_out[r * a.C + c] = v;
      }
    }
    if (a.pdf_nb) {
      // last use of the strip for this ray: the pdf values overwrite the weights in place (element s is read and then
      // written by the same lane)
      __syncwarp();
      struct Strip {
        float* p;
        __device__ __forceinline__ float& operator[](int q) const { return p[PI(q)]; }
      };
      const float* sb = a.pdf_sbins + r * (S + 1);
      nkbs_pdf_ray_warp(pp, [&](int q) { return sw[PI(q)]; }, [&](int k) { return sb[k]; }, Strip{sw}, sw + strip_len(S),
                        lane, a.pdf_rand != nullptr, a.pdf_rand ? a.pdf_rand[r] : 0.f, nullptr, a.nears[r], a.fars[r],
                        a.pdf_out_sbins + r * a.pdf_nb, a.pdf_out_ebins + r * a.pdf_nb, nullptr);
    }
    __syncwarp();
  }
  if (a.minmax) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lmin = fminf(lmin, __shfl_xor_sync(0xffffffffu, lmin, o));
      lmax = fmaxf(lmax, __shfl_xor_sync(0xffffffffu, lmax, o));
    }
    if (lane == 0 && lmin <= lmax) {
      atomicMin(a.minmax, float_key(lmin));
      atomicMax(a.minmax + 1, float_key(lmax));
    }
  }
}

__global__ void minmax_init_kernel(int* mm) {
  nkbs_pdl_sync();
  mm[0] = float_key(FLT_MAX);
  mm[1] = float_key(-FLT_MAX);
}

// depth = clip(raw, steps.min(), steps.max()); mask = 1 where the clip passed the value through
__global__ void clip_depth_kernel(const float* raw, const int* mm, int64_t R, float* out, float* mask) {
  nkbs_pdl_sync();
  float lo = key_float(mm[0]), hi = key_float(mm[1]);
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < R; r += (int64_t)gridDim.x * blockDim.x) {
    float v = raw[r];
    out[r] = fminf(fmaxf(v, lo), hi);
    if (mask) mask[r] = (v >= lo && v <= hi) ? 1.f : 0.f;
  }
}

struct CompBwdArgs {
  int64_t R;
  int S;
  const float* density;    // [R,S] or null (then d_weights is produced instead of d_density)
  const float* weights;    // [R,S] saved forward weights
  const float* starts;
  const float* ends;
  int bstride;
  const float* rgb;        // [R,S,3] or null
  const float* sem;        // [R,S,C] or null
  int C;
  int bg_mode;
  int sem_weight_grad;     // pass_semantic_gradients
  const float* acc;        // [R] saved accumulation
  const float* depth_raw;  // [R] saved unclipped expected depth (needed when g_depth given)
  // upstream gradients (nullable)
  const float* g_weights;  // [R,S]
  const float* g_rgb;      // [R,3]
  const float* g_acc;      // [R]
  const float* g_depth;    // [R] gradient w.r.t. the unclipped expected depth
  const float* g_sem;      // [R,C]
  // outputs
  float* d_density;        // [R,S]
  float* d_weights;        // [R,S]
  float* d_rgb;            // [R,S,3]
  float* d_sem;            // [R,S,C]
};

__global__ void __launch_bounds__(kThreads) composite_bwd_kernel(const CompBwdArgs a) {
  nkbs_pdl_sync();
  extern __shared__ float s_buf[];  // [kWarpsPerCta][3][S]: weights, dL/dw, exp(-incl)
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int S = a.S;
  float* sw = s_buf + wid * 3 * strip_len(S);
  float* sg = sw + strip_len(S);
  float* sE = sg + strip_len(S);
  for (int64_t r = blockIdx.x * (int64_t)kWarpsPerCta + wid; r < a.R; r += (int64_t)gridDim.x * kWarpsPerCta) {
    const int64_t base = r * S;
    const int64_t bb = r * a.bstride;
    const float acc = a.acc ? a.acc[r] : 0.f;
    float gr[3] = {0.f, 0.f, 0.f}, bg[3] = {0.f, 0.f, 0.f};
    if (a.g_rgb) {
      gr[0] = a.g_rgb[r * 3];
      gr[1] = a.g_rgb[r * 3 + 1];
      gr[2] = a.g_rgb[r * 3 + 2];
      if (a.bg_mode == 0) {
        bg[0] = a.rgb[(base + S - 1) * 3];
        bg[1] = a.rgb[(base + S - 1) * 3 + 1];
        bg[2] = a.rgb[(base + S - 1) * 3 + 2];
      } else if (a.bg_mode == 2) {
        bg[0] = bg[1] = bg[2] = 1.f;
      }
    }
    // expected depth = dsum / (acc + eps):  d/d dsum = g/(acc+eps),  d/d acc = -g * depth / (acc+eps)
    float g_dsum = 0.f, g_acc = a.g_acc ? a.g_acc[r] : 0.f;
    if (a.g_depth) {
      float g = a.g_depth[r], den = acc + 1e-10f;
      g_dsum = g / den;
      g_acc += -g * a.depth_raw[r] / den;
    }
    // pass A: gradient w.r.t. each weight
    for (int s = lane; s < S; s += 32) {
      float w = a.weights[base + s];
      float g = g_acc + (a.g_weights ? a.g_weights[base + s] : 0.f);
      if (a.g_rgb) {
        const float* c = a.rgb + (base + s) * 3;
        g += gr[0] * (c[0] - bg[0]) + gr[1] * (c[1] - bg[1]) + gr[2] * (c[2] - bg[2]);
      }
      if (a.g_depth) g += g_dsum * ((a.starts[bb + s] + a.ends[bb + s]) / 2.f);
      if (a.g_sem && a.sem_weight_grad) {
        const float* q = a.sem + (base + s) * a.C;
        const float* gq = a.g_sem + r * a.C;
        float d = 0.f;
        for (int c = 0; c < a.C; ++c) d += gq[c] * q[c];
        g += d;
      }
      sw[PI(s)] = w;
      sg[PI(s)] = g;
      if (a.d_weights) a.d_weights[base + s] = g;
      if (a.d_rgb) {
        float* o = a.d_rgb + (base + s) * 3;
        float extra = (a.bg_mode == 0 && s == S - 1) ? (1.f - acc) : 0.f;
        o[0] = gr[0] * (w + extra);
        o[1] = gr[1] * (w + extra);
        o[2] = gr[2] * (w + extra);
      }
    }
    __syncwarp();
    if (a.d_sem) {
      float* o = a.d_sem + base * a.C;
      for (int c0 = 0; c0 < a.C; c0 += 32) {
        int c = c0 + lane;
        if (c < a.C) {
          float g = a.g_sem ? a.g_sem[r * a.C + c] : 0.f;
          for (int s = 0; s < S; ++s) o[(int64_t)s * a.C + c] = sw[PI(s)] * g;
        }
      }
    }
    if (a.d_density) {
      // w_i = (1-exp(-x_i)) exp(-sum_{j<i} x_j):  dL/dx_i = g_i exp(-incl_i) - sum_{j>i} g_j w_j
      // forward sweep: E_i = exp(-inclusive prefix of x); reverse sweep: exclusive suffix sum of g_j w_j
      // same three passes as the forward: coalesced loads, lane-contiguous runs with one scan, coalesced stores
      const int per = (S + 31) >> 5;
      const int r0 = min(lane * per, S), r1 = min(r0 + per, S);
      for (int s = lane; s < S; s += 32) sE[PI(s)] = (a.ends[bb + s] - a.starts[bb + s]) * a.density[base + s];
      __syncwarp();
      float tot = 0.f;
      for (int s = r0; s < r1; ++s) tot += sE[PI(s)];
      float run = warp_incl_scan(tot, lane) - tot;
      float qtot = 0.f;
      for (int s = r0; s < r1; ++s) {
        run += sE[PI(s)];
        sE[PI(s)] = expf(-run);
        qtot += sg[PI(s)] * sw[PI(s)];
      }
      // exclusive suffix of the run totals of g_j w_j (runs of the lanes above this one)
      float suf = qtot;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float n = __shfl_down_sync(0xffffffffu, suf, o);
        if (lane + o < 32) suf += n;
      }
      float after = suf - qtot;  // sum of g_j w_j over all later samples
      for (int s = r1 - 1; s >= r0; --s) {
        const float q = sg[PI(s)] * sw[PI(s)];
        sE[PI(s)] = sg[PI(s)] * sE[PI(s)] - after;
        after += q;
      }
      __syncwarp();
      for (int s = lane; s < S; s += 32) a.d_density[base + s] = (a.ends[bb + s] - a.starts[bb + s]) * sE[PI(s)];
    }
    __syncwarp();
  }
}

int comp_grid(int64_t R) {
  int64_t blocks = (R + kWarpsPerCta - 1) / kWarpsPerCta;
  int64_t cap = (int64_t)nkbs_num_sms() * 8;
  if (blocks < 1) blocks = 1;
  return (int)(blocks < cap ? blocks : cap);
}

}  // namespace

extern "C" int nkbs_composite_fwd(int64_t R, int32_t S, const float* density, const float* weights_in,
                                  const float* starts, const float* ends, int32_t bin_stride, const float* rgb,
                                  const float* sem, int32_t C, int32_t bg_mode, int32_t eval_mode, float* weights, float* rgb_out,
                                  float* acc_out, float* depth_exp_raw, float* depth_exp, float* depth_mask,
                                  float* depth_med, float* sem_out, int32_t* minmax_ws, void* stream) {
  NKBS_CHECK_ARG(R >= 0 && S >= 1 && S <= kMaxS, "nkbs_composite_fwd: S=%d outside 1..%d", S, kMaxS);
  if (R == 0) return 0;
  NKBS_CHECK_ARG((density != nullptr) != (weights_in != nullptr), "nkbs_composite_fwd: give exactly one of density / weights");
  NKBS_CHECK_ARG(!density || (starts && ends), "nkbs_composite_fwd: density needs starts/ends");
  NKBS_CHECK_ARG(!(depth_exp_raw || depth_med) || (starts && ends), "nkbs_composite_fwd: depth needs starts/ends");
  NKBS_CHECK_ARG(!rgb_out || rgb, "nkbs_composite_fwd: rgb_out without rgb");
  NKBS_CHECK_ARG(!sem_out || (sem && C >= 1), "nkbs_composite_fwd: sem_out without sem");
  NKBS_CHECK_ARG(!depth_exp || (depth_exp_raw && minmax_ws), "nkbs_composite_fwd: clipped depth needs raw + workspace");
  NKBS_CHECK_ARG(bg_mode >= 0 && bg_mode <= 2, "nkbs_composite_fwd: unknown background mode %d", bg_mode);
  if (R == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  NKBS_CHECK_ARG(bin_stride == 0 || bin_stride >= S, "nkbs_composite_fwd: bin_stride %d < S", bin_stride);
  CompArgs a{R, S, density, weights_in, starts, ends, bin_stride ? bin_stride : S, rgb, sem, C, bg_mode, eval_mode,
             weights, rgb_out, acc_out, depth_exp_raw, depth_med, sem_out, depth_exp ? minmax_ws : nullptr};
  a.pdf_nb = 0;
  a.pdf_anneal_dev = nullptr;
  if (depth_exp) nkbs_launch(minmax_init_kernel, 1, 1, 0, st, minmax_ws);
  size_t smem = (size_t)kWarpsPerCta * strip_len(S) * sizeof(float);
  nkbs_launch(composite_fwd_kernel, comp_grid(R), kThreads, smem, st, a);
  NKBS_CHECK_LAUNCH("nkbs_composite_fwd");
  if (depth_exp) {
    nkbs_launch(clip_depth_kernel, (int)((R + 255) / 256), 256, 0, st, depth_exp_raw, minmax_ws, R, depth_exp, depth_mask);
    NKBS_CHECK_LAUNCH("nkbs_composite_fwd(clip)");
  }
  return 0;
}

extern "C" int nkbs_composite_bwd(int64_t R, int32_t S, const float* density, const float* weights,
                                  const float* starts, const float* ends, int32_t bin_stride, const float* rgb,
                                  const float* sem, int32_t C, int32_t bg_mode, int32_t sem_weight_grad, const float* acc,
                                  const float* depth_raw, const float* g_weights, const float* g_rgb,
                                  const float* g_acc, const float* g_depth, const float* g_sem, float* d_density,
                                  float* d_weights, float* d_rgb, float* d_sem, void* stream) {
  NKBS_CHECK_ARG(R >= 0 && S >= 1 && S <= kMaxS, "nkbs_composite_bwd: S=%d outside 1..%d", S, kMaxS);
  if (R == 0) return 0;
  NKBS_CHECK_ARG(weights, "nkbs_composite_bwd: null weights");
  NKBS_CHECK_ARG(!d_density || (density && starts && ends), "nkbs_composite_bwd: d_density needs density/starts/ends");
  NKBS_CHECK_ARG(!g_rgb || (rgb && acc), "nkbs_composite_bwd: g_rgb needs rgb and accumulation");
  NKBS_CHECK_ARG(!g_depth || (depth_raw && acc && starts && ends), "nkbs_composite_bwd: g_depth needs depth_raw/acc/starts/ends");
  NKBS_CHECK_ARG(!(g_sem && sem_weight_grad) || sem, "nkbs_composite_bwd: semantic weight gradient needs sem");
  NKBS_CHECK_ARG(!d_rgb || acc || bg_mode != 0, "nkbs_composite_bwd: d_rgb with last_sample background needs accumulation");
  if (R == 0) return 0;
  NKBS_CHECK_ARG(bin_stride == 0 || bin_stride >= S, "nkbs_composite_bwd: bin_stride %d < S", bin_stride);
  CompBwdArgs a{R, S, density, weights, starts, ends, bin_stride ? bin_stride : S, rgb, sem, C, bg_mode, sem_weight_grad, acc, depth_raw,
                g_weights, g_rgb, g_acc, g_depth, g_sem, d_density, d_weights, d_rgb, d_sem};
  size_t smem = (size_t)kWarpsPerCta * 3 * strip_len(S) * sizeof(float);
  if (smem > 48 * 1024) {  // S >= 496: opt in to more than the default 48 KB of dynamic shared memory
    cudaError_t e = cudaFuncSetAttribute(composite_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    NKBS_CHECK_ARG(e == cudaSuccess, "nkbs_composite_bwd: %zu bytes of shared memory refused (%s)", smem, cudaGetErrorString(e));
  }
  nkbs_launch(composite_bwd_kernel, comp_grid(R), kThreads, smem, (cudaStream_t)stream, a);
  NKBS_CHECK_LAUNCH("nkbs_composite_bwd");
  return 0;
}

// ---- expected depth without any allocation by anybody (no counterpart in the reference) -------
// DepthRenderer("expected") needs two pieces of scratch besides its output: the unclipped depth of every ray and the
// batch-wide minimum / maximum sample position it is clipped to.  nkbs_composite_fwd takes them from the caller; this
// variant keeps them in a file-scope device array, so a caller that may not allocate (CUDA-graph capture, a tight
// serving loop, a harness that forbids cudaMalloc) passes inputs and the output only.  The array is part of the
// library image: every GPU that loads libnkbs.so holds its 4 MiB + 256 B until the process exits.  All calls share it
// and it keeps the last call's contents: the min / max keys are re-initialised by every call and the raw depths of rays
// [0, R) are overwritten before they are read; two calls that overlap on the GPU (different streams) race on it.
namespace {
constexpr int64_t kNoAllocMaxRays = NKBS_NOALLOC_MAX_RAYS;  // include/nkbs.h
constexpr int kNoAllocHeader = 256;  // bytes: the two ordered-int min / max keys, padded to the array's alignment
constexpr int kNoAllocMaxDevices = 64;
}  // namespace
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocHeader + kNoAllocMaxRays * sizeof(float)];

namespace {
// Device address of g_noalloc_scratch on the CURRENT device.  In host code the array's name is only a stand-in for
// the symbol, not a device pointer: the address comes from cudaGetSymbolAddress, asked once per device with that device
// current, and is cached per device ordinal (racing first calls store the same value).
int noalloc_scratch(unsigned char** out) {
  static void* cache[kNoAllocMaxDevices] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  NKBS_CHECK_ARG(e == cudaSuccess, "nkbs_expected_depth_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
  NKBS_CHECK_ARG(dev >= 0 && dev < kNoAllocMaxDevices, "nkbs_expected_depth_noalloc: device ordinal %d not supported", dev);
  void* p = __atomic_load_n(&cache[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    NKBS_CHECK_ARG(e == cudaSuccess && p, "nkbs_expected_depth_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
    __atomic_store_n(&cache[dev], p, __ATOMIC_RELEASE);
  }
  *out = static_cast<unsigned char*>(p);
  return 0;
}
}  // namespace

extern "C" int nkbs_expected_depth_noalloc(int64_t R, int32_t S, const float* weights, const float* starts,
                                           const float* ends, int32_t bin_stride, float* depth, void* stream) {
  NKBS_CHECK_ARG(R >= 0 && S >= 1 && S <= kMaxS, "nkbs_expected_depth_noalloc: S=%d outside 1..%d", S, kMaxS);
  NKBS_CHECK_ARG(R <= kNoAllocMaxRays,
                 "nkbs_expected_depth_noalloc: %lld rays exceed the built-in scratch of %lld rays (split the batch, or use "
                 "nkbs_composite_fwd with caller-owned scratch)", (long long)R, (long long)kNoAllocMaxRays);
  if (R == 0) return 0;
  NKBS_CHECK_ARG(weights && starts && ends && depth, "nkbs_expected_depth_noalloc: null argument");
  NKBS_CHECK_ARG(bin_stride == 0 || bin_stride >= S, "nkbs_expected_depth_noalloc: bin_stride %d < S", bin_stride);
  unsigned char* scratch = nullptr;
  if (noalloc_scratch(&scratch)) return -1;
  int* mm = reinterpret_cast<int*>(scratch);
  float* raw = reinterpret_cast<float*>(scratch + kNoAllocHeader);
  cudaStream_t st = (cudaStream_t)stream;
  CompArgs a{R, S, nullptr, weights, starts, ends, bin_stride ? bin_stride : S, nullptr, nullptr, 0, 0, 0,
             nullptr, nullptr, nullptr, raw, nullptr, nullptr, mm};
  a.pdf_nb = 0;
  a.pdf_anneal_dev = nullptr;
  nkbs_launch(minmax_init_kernel, 1, 1, 0, st, mm);
  const size_t smem = (size_t)kWarpsPerCta * strip_len(S) * sizeof(float);
  nkbs_launch(composite_fwd_kernel, comp_grid(R), kThreads, smem, st, a);
  NKBS_CHECK_LAUNCH("nkbs_expected_depth_noalloc");
  nkbs_launch(clip_depth_kernel, (int)((R + 255) / 256), 256, 0, st, raw, mm, R, depth, nullptr);
  NKBS_CHECK_LAUNCH("nkbs_expected_depth_noalloc(clip)");
  return 0;
}

// ---- get_weights + median depth of one proposal level and PDFSampler of the next, one ray-major launch ----------------
extern "C" int nkbs_weights_resample_fwd(int64_t R, int32_t S, const float* density, const float* ebins,
                                         const float* sbins, const float* nears, const float* fars, int32_t num_bins,
                                         float anneal, const float* anneal_dev, float histogram_padding, float eps,
                                         const float* u_base, const float* rand, float* weights, float* depth_med,
                                         float* out_sbins, float* out_ebins, void* stream) {
  NKBS_CHECK_ARG(R >= 0 && S >= 1 && S <= kMaxS && num_bins >= 2, "nkbs_weights_resample_fwd: S=%d outside 1..%d or bins=%d < 2",
                 S, kMaxS, num_bins);
  if (R == 0) return 0;
  NKBS_CHECK_ARG(density && ebins && sbins && nears && fars && u_base && weights && out_sbins && out_ebins,
                 "nkbs_weights_resample_fwd: null argument");
  CompArgs a{R, S, density, nullptr, ebins, ebins + 1, S + 1, nullptr, nullptr, 0, 0, 0,
             weights, nullptr, nullptr, nullptr, depth_med, nullptr, nullptr};
  a.pdf_nb = num_bins;
  a.pdf_anneal = anneal;
  a.pdf_anneal_dev = anneal_dev;
  a.pdf_hist_pad = histogram_padding;
  a.pdf_eps = eps;
  a.pdf_sbins = sbins;
  a.pdf_u = u_base;
  a.pdf_rand = rand;
  a.nears = nears;
  a.fars = fars;
  a.pdf_out_sbins = out_sbins;
  a.pdf_out_ebins = out_ebins;
  const size_t smem = (size_t)kWarpsPerCta * (strip_len(S) + S + 1) * sizeof(float);
  nkbs_launch(composite_fwd_kernel, comp_grid(R), kThreads, smem, (cudaStream_t)stream, a);
  NKBS_CHECK_LAUNCH("nkbs_weights_resample_fwd");
  return 0;
}
