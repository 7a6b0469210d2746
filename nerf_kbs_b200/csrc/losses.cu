// N1: the per-ray training losses and their gradients in ONE ray-major pass, one warp per ray:
//   * interlevel (proposal) loss  lossfun_outer  of each proposal level against the final level's histogram
//   * distortion loss of the final level (O(S^2) pairwise term kept in shared memory)
//   * MSE of the rendered colour, cross-entropy (ignore_index) of the rendered semantic logits
// The values land in a 5-float device vector; the gradients w.r.t. the weights of every level, the rendered colour
// and the logits are written in the same pass (the losses are scalar sums over rays, so each ray's share of every
// gradient is known as soon as the ray has been read).  The torch route runs ~130 elementwise / scan / gather /
// searchsorted launches and materialises an [R,S,S] tensor for the distortion term.
// Replaces nerfstudio.model_components.losses.interlevel_loss / distortion_loss and the MSE / CrossEntropy modules
// called at ref:nerfacto.py:468,483-489,505-515 and ref:semantic_nerfw.py:253-279.
#include "common.cuh"

namespace {

constexpr int kWarpsPerCta = 8;
constexpr int kThreads = kWarpsPerCta * 32;
constexpr int kMaxProp = 4;
constexpr int kMaxS = 512;

struct LossArgs {
  int64_t R;
  int S;                        // final-level samples per ray
  const float* sdist;           // [R,S+1] spacing-space bin edges of the final level
  const float* w;               // [R,S]   final weights
  float* dw;                    // [R,S]   d(distortion)/dw, null = skip the distortion term
  int n_prop;
  int Sp[kMaxProp];
  const float* sdist_p[kMaxProp];  // [R,Sp+1]
  const float* w_p[kMaxProp];      // [R,Sp]
  float* dw_p[kMaxProp];           // [R,Sp] d(interlevel)/dw_p
  const float* rgb;             // [R,3] or null
  const float* image;           // [R,3]
  float* drgb;                  // [R,3]
  const float* logits;          // [R,C] or null
  const int64_t* labels;        // [R]
  int C;
  int64_t ignore_index;
  float* dlogits;               // [R,C]  softmax - onehot, zero rows for ignored labels (caller divides by the count)
  float* out;                   // [5] rgb mse, interlevel, distortion, CE sum, CE count  (accumulated: caller zeroes)
  float inv_R, inv_RS, inv_R3;  // normalisers of the reported terms
  float g_R, g_RS, g_R3;        // normalisers x term weights, applied to the gradients (weighted entry point)
  float tw[3];                  // term weights (rgb, interlevel, distortion)
  float* total;                 // optional [1]: weighted sum of the three terms, accumulated
  float* scratch;               // optional [8], zero between launches: the CTAs accumulate here instead, and the last
                                // CTA to finish WRITES out[0..4], total[0] and total[1..3] (the weighted terms) and
                                // zeroes the scratch again: out / total need no zero fill before the launch
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
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

// number of elements of the ascending array a[0..n) that are <= v   (torch.searchsorted(..., side="right"))
__device__ __forceinline__ int upper_bound(const float* a, int n, float v) {
  int lo = 0, hi = n;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (a[mid] <= v) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

__global__ void __launch_bounds__(kThreads) ray_losses_kernel(const LossArgs a, int smax) {
  nkbs_pdl_sync();
  extern __shared__ float s_all[];
  __shared__ float s_out[5];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int S = a.S;
  // per-warp strips: c [S+1], w [S], ut [S], cp [smax+1], cy [smax+1], D [smax+2]
  float* sc = s_all + (size_t)wid * ((S + 1) + 2 * S + 3 * (smax + 2));
  float* sw = sc + (S + 1);
  float* sut = sw + S;
  float* scp = sut + S;
  float* scy = scp + (smax + 2);
  float* sD = scy + (smax + 2);
  if (threadIdx.x < 5) s_out[threadIdx.x] = 0.f;
  __syncthreads();
  float t_rgb = 0.f, t_inter = 0.f, t_dist = 0.f, t_ce = 0.f, t_cnt = 0.f;  // this lane's partial sums
  for (int64_t r = blockIdx.x * (int64_t)kWarpsPerCta + wid; r < a.R; r += (int64_t)gridDim.x * kWarpsPerCta) {
    // ---- final level into shared memory ----
    if (a.sdist) {
      const float* c = a.sdist + r * (S + 1);
      const float* w = a.w + r * S;
      for (int s = lane; s <= S; s += 32) sc[s] = c[s];
      for (int s = lane; s < S; s += 32) sw[s] = w[s];
      __syncwarp();
      for (int s = lane; s < S; s += 32) sut[s] = (sc[s + 1] + sc[s]) / 2.f;
      __syncwarp();
    }
    // ---- distortion: sum_ij w_i w_j |u_i - u_j| + sum_i w_i^2 (c_{i+1} - c_i) / 3 ----
    if (a.dw) {
      for (int i = lane; i < S; i += 32) {
        const float ui = sut[i], wi = sw[i];
        float inner = 0.f;
        for (int j = 0; j < S; ++j) inner += sw[j] * fabsf(ui - sut[j]);
        const float delta = sc[i + 1] - sc[i];
        t_dist += wi * inner + wi * wi * delta / 3.f;
        a.dw[r * S + i] = (2.f * inner + 2.f * wi * delta / 3.f) * a.g_R;
      }
    }
    // ---- interlevel: for every proposal level, the final histogram must lie under the proposal's envelope ----
    for (int p = 0; p < a.n_prop; ++p) {
      const int Sp = a.Sp[p];
      const float* cp = a.sdist_p[p] + r * (Sp + 1);
      const float* wp = a.w_p[p] + r * Sp;
      for (int k = lane; k <= Sp; k += 32) {
        scp[k] = cp[k];
        sD[k] = 0.f;
      }
      // cy[0] = 0, cy[k+1] = wp[0] + .. + wp[k]
      float carry = 0.f;
      for (int k0 = 0; k0 < Sp; k0 += 32) {
        const int k = k0 + lane;
        const float v = k < Sp ? wp[k] : 0.f;
        const float incl = warp_incl_scan(v, lane);
        if (k < Sp) scy[k + 1] = carry + incl;
        carry += __shfl_sync(0xffffffffu, incl, 31);
      }
      if (lane == 0) scy[0] = 0.f;
      __syncwarp();
      for (int s = lane; s < S; s += 32) {
        int lo = upper_bound(scp, Sp, sc[s]) - 1;          // t1_starts = cp[:-1]
        int hi = upper_bound(scp + 1, Sp, sc[s + 1]);      // t1_ends   = cp[1:]
        lo = min(max(lo, 0), Sp - 1);
        hi = min(max(hi, 0), Sp - 1);
        const float w_outer = scy[hi + 1] - scy[lo];
        const float ws = sw[s];
        const float d = fmaxf(ws - w_outer, 0.f);
        const float den = ws + 1e-7f;
        t_inter += d * d / den;
        const float g = -2.f * d / den * a.g_RS;  // d loss / d w_outer
        if (g != 0.f) {
          atomicAdd(&sD[lo], g);
          atomicAdd(&sD[hi + 1], -g);
        }
      }
      __syncwarp();
      // d loss / d wp[k] = inclusive prefix sum of the difference array
      float* out = a.dw_p[p] + r * Sp;
      carry = 0.f;
      for (int k0 = 0; k0 < Sp; k0 += 32) {
        const int k = k0 + lane;
        const float v = k < Sp ? sD[k] : 0.f;
        const float incl = warp_incl_scan(v, lane);
        if (k < Sp) out[k] = carry + incl;
        carry += __shfl_sync(0xffffffffu, incl, 31);
      }
      __syncwarp();
    }
    // ---- colour MSE ----
    if (a.rgb && lane < 3) {
      const float e = a.rgb[r * 3 + lane] - a.image[r * 3 + lane];
      t_rgb += e * e;
      a.drgb[r * 3 + lane] = 2.f * e * a.g_R3;
    }
    // ---- semantic cross-entropy ----
    if (a.logits) {
      const int64_t lab = a.labels[r];
      const bool valid = lab != a.ignore_index && lab >= 0 && lab < a.C;
      const float* x = a.logits + r * a.C;
      float m = -INFINITY;
      for (int c = lane; c < a.C; c += 32) m = fmaxf(m, x[c]);
      m = warp_max(m);
      float se = 0.f;
      for (int c = lane; c < a.C; c += 32) se += expf(x[c] - m);
      se = warp_sum(se);
      const float lse = logf(se) + m;
      for (int c = lane; c < a.C; c += 32) {
        const float sm = expf(x[c] - lse);
        a.dlogits[r * a.C + c] = valid ? sm - (c == (int)lab ? 1.f : 0.f) : 0.f;
      }
      if (valid && lane == 0) {
        t_ce += lse - x[lab];
        t_cnt += 1.f;
      }
    }
    __syncwarp();
  }
  t_rgb = warp_sum(t_rgb);
  t_inter = warp_sum(t_inter);
  t_dist = warp_sum(t_dist);
  t_ce = warp_sum(t_ce);
  t_cnt = warp_sum(t_cnt);
  if (lane == 0) {
    atomicAdd(&s_out[0], t_rgb);
    atomicAdd(&s_out[1], t_inter);
    atomicAdd(&s_out[2], t_dist);
    atomicAdd(&s_out[3], t_ce);
    atomicAdd(&s_out[4], t_cnt);
  }
  __syncthreads();
  float* acc = a.scratch ? a.scratch : a.out;
  if (threadIdx.x < 5) {
    const float scale = threadIdx.x == 0 ? a.inv_R3 : (threadIdx.x == 1 ? a.inv_RS : (threadIdx.x == 2 ? a.inv_R : 1.f));
    const float v = s_out[threadIdx.x] * scale;
    if (v != 0.f) atomicAdd(acc + threadIdx.x, v);
  }
  if (a.total && threadIdx.x == 0) {
    const float v = a.tw[0] * (s_out[0] * a.inv_R3) + a.tw[1] * (s_out[1] * a.inv_RS) + a.tw[2] * (s_out[2] * a.inv_R);
    if (v != 0.f) atomicAdd(a.scratch ? a.scratch + 5 : a.total, v);
  }
  if (a.scratch) {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      unsigned* counter = reinterpret_cast<unsigned*>(a.scratch + 7);
      if (atomicAdd(counter, 1u) == gridDim.x - 1) {  // every CTA's sums have arrived
        __threadfence();
        float t[6];
#pragma unroll
        for (int i = 0; i < 6; ++i) t[i] = atomicExch(a.scratch + i, 0.f);  // read and re-arm
#pragma unroll
        for (int i = 0; i < 5; ++i) a.out[i] = t[i];
        if (a.total) {
          a.total[0] = t[5];
#pragma unroll
          for (int i = 0; i < 3; ++i) a.total[1 + i] = a.tw[i] * t[i];
        }
        *counter = 0u;
      }
    }
  }
}

}  // namespace

static int ray_losses_launch(int64_t R, int32_t S, const float* sdist, const float* w, float* dw, int32_t n_prop,
                             const int32_t* S_prop, const float* const* sdist_prop, const float* const* w_prop,
                             float* const* dw_prop, const float* rgb, const float* image, float* drgb,
                             const float* logits, const int64_t* labels, int32_t C, int64_t ignore_index,
                             float* dlogits, float* out, const float* term_weights, float* total, float* scratch,
                             void* stream) {
  NKBS_CHECK_ARG(R >= 0 && out, "nkbs_ray_losses: bad arguments");
  NKBS_CHECK_ARG(n_prop >= 0 && n_prop <= kMaxProp, "nkbs_ray_losses: at most %d proposal levels (got %d)", kMaxProp, n_prop);
  const bool hist = dw != nullptr || n_prop > 0;
  NKBS_CHECK_ARG(!hist || (sdist && w && S >= 1 && S <= kMaxS), "nkbs_ray_losses: final level needs sdist/w and 1 <= S <= %d", kMaxS);
  NKBS_CHECK_ARG(!rgb || (image && drgb), "nkbs_ray_losses: rgb needs image and drgb");
  NKBS_CHECK_ARG(!logits || (labels && dlogits && C >= 1), "nkbs_ray_losses: logits need labels, dlogits and C");
  if (R == 0) return 0;
  LossArgs a;
  a.R = R;
  a.S = hist ? S : 1;
  a.sdist = hist ? sdist : nullptr;
  a.w = w;
  a.dw = dw;
  a.n_prop = n_prop;
  int smax = 0;
  for (int p = 0; p < n_prop; ++p) {
    NKBS_CHECK_ARG(S_prop[p] >= 1 && S_prop[p] <= kMaxS && sdist_prop[p] && w_prop[p] && dw_prop[p],
                   "nkbs_ray_losses: bad proposal level %d", p);
    a.Sp[p] = S_prop[p];
    a.sdist_p[p] = sdist_prop[p];
    a.w_p[p] = w_prop[p];
    a.dw_p[p] = dw_prop[p];
    if (S_prop[p] > smax) smax = S_prop[p];
  }
  a.rgb = rgb;
  a.image = image;
  a.drgb = drgb;
  a.logits = logits;
  a.labels = labels;
  a.C = C;
  a.ignore_index = ignore_index;
  a.dlogits = dlogits;
  a.out = out;
  a.inv_R = 1.f / (float)R;
  a.inv_RS = 1.f / ((float)R * (float)a.S);
  a.inv_R3 = 1.f / ((float)R * 3.f);
  for (int i = 0; i < 3; ++i) a.tw[i] = term_weights ? term_weights[i] : 1.f;
  a.g_R3 = a.inv_R3 * a.tw[0];
  a.g_RS = a.inv_RS * a.tw[1];
  a.g_R = a.inv_R * a.tw[2];
  a.total = total;
  a.scratch = scratch;
  const size_t smem = (size_t)kWarpsPerCta * ((a.S + 1) + 2 * a.S + 3 * (smax + 2)) * sizeof(float);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(ray_losses_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    NKBS_CHECK_ARG(e == cudaSuccess, "nkbs_ray_losses: %zu bytes of shared memory refused", smem);
  }
  int64_t blocks = (R + kWarpsPerCta - 1) / kWarpsPerCta;
  int64_t cap = (int64_t)nkbs_num_sms() * 8;
  int grid = (int)(blocks < cap ? blocks : cap);
  nkbs_launch(ray_losses_kernel, grid, kThreads, smem, (cudaStream_t)stream, a, smax);
  NKBS_CHECK_LAUNCH("nkbs_ray_losses");
  return 0;
}

extern "C" int nkbs_ray_losses(int64_t R, int32_t S, const float* sdist, const float* w, float* dw, int32_t n_prop,
                               const int32_t* S_prop, const float* const* sdist_prop, const float* const* w_prop,
                               float* const* dw_prop, const float* rgb, const float* image, float* drgb,
                               const float* logits, const int64_t* labels, int32_t C, int64_t ignore_index,
                               float* dlogits, float* out, void* stream) {
  return ray_losses_launch(R, S, sdist, w, dw, n_prop, S_prop, sdist_prop, w_prop, dw_prop, rgb, image, drgb, logits,
                           labels, C, ignore_index, dlogits, out, nullptr, nullptr, nullptr, stream);
}

extern "C" int nkbs_ray_losses_weighted(int64_t R, int32_t S, const float* sdist, const float* w, float* dw,
                                        int32_t n_prop, const int32_t* S_prop, const float* const* sdist_prop,
                                        const float* const* w_prop, float* const* dw_prop, const float* rgb,
                                        const float* image, float* drgb, const float* logits, const int64_t* labels,
                                        int32_t C, int64_t ignore_index, float* dlogits, float* out,
                                        const float* term_weights, float* total, float* scratch, void* stream) {
  NKBS_CHECK_ARG(term_weights != nullptr, "nkbs_ray_losses_weighted: null term_weights");
  return ray_losses_launch(R, S, sdist, w, dw, n_prop, S_prop, sdist_prop, w_prop, dw_prop, rgb, image, drgb, logits,
                           labels, C, ignore_index, dlogits, out, term_weights, total, scratch, stream);
}
