// N2: multi-tensor Adam over the model's parameter tensors, one launch per <= 24 tensors.
// Replaces torch.optim.Adam(lr=1e-3, eps=1e-15) that the reference configures for both parameter
// groups (ref:train.py:85-98; groups ref:nerfacto.py:359-363) [nerfstudio Optimizers, NS-recalled].
// HBM streaming kernel: 4 words read + 3 written per parameter, 128-bit accesses.
#include "common.cuh"

namespace {

constexpr int kMaxSegs = 24;
constexpr int kThreads = 256;
constexpr int kPerBlock = kThreads * 16;  // floats per CTA

struct AdamSeg {
  float* p;
  const float* g;
  float* m;
  float* v;
  int64_t n;
};

struct AdamArgs {
  AdamSeg seg[kMaxSegs];
  int block_start[kMaxSegs + 1];  // first CTA of every segment
  int nseg;
  float lr, beta1, beta2, eps, weight_decay;
  const float* step_dev;   // device scalar: step count t >= 1 (float), read by every thread
  const float* inv_scale;  // optional device scalar multiplied into the gradient (GradScaler unscale)
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float b1, float b2, float eps,
                                         float step_size, float sqrt_bc2, float wd) {
  if (wd != 0.f) g = fmaf(wd, p, g);
  m = fmaf(b1, m, (1.f - b1) * g);      // exp_avg.lerp_(grad, 1-beta1)
  v = fmaf(b2, v, (1.f - b2) * g * g);  // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1-beta2)
  float denom = sqrtf(v) / sqrt_bc2 + eps;
  p -= step_size * (m / denom);
}

__global__ void __launch_bounds__(kThreads) adam_kernel(const AdamArgs a) {
  int s = 0;
#pragma unroll 1
  while (s + 1 < a.nseg && (int)blockIdx.x >= a.block_start[s + 1]) ++s;
  const AdamSeg sg = a.seg[s];
  const float t = *a.step_dev;
  const float bc1 = 1.f - powf(a.beta1, t);
  const float bc2 = 1.f - powf(a.beta2, t);
  const float step_size = a.lr / bc1;
  const float sqrt_bc2 = sqrtf(bc2);
  const float gs = a.inv_scale ? *a.inv_scale : 1.f;
  const int64_t base = (int64_t)(blockIdx.x - a.block_start[s]) * kPerBlock;
  const bool vec = ((reinterpret_cast<uintptr_t>(sg.p) | reinterpret_cast<uintptr_t>(sg.g) |
                     reinterpret_cast<uintptr_t>(sg.m) | reinterpret_cast<uintptr_t>(sg.v)) & 15) == 0;
  if (vec && base + kPerBlock <= sg.n) {
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      int64_t i = base + (int64_t)(it * kThreads + threadIdx.x) * 4;
      float4 p = *reinterpret_cast<float4*>(sg.p + i);
      float4 g = __ldcs(reinterpret_cast<const float4*>(sg.g + i));
      float4 m = *reinterpret_cast<float4*>(sg.m + i);
      float4 v = *reinterpret_cast<float4*>(sg.v + i);
      adam_one(p.x, g.x * gs, m.x, v.x, a.beta1, a.beta2, a.eps, step_size, sqrt_bc2, a.weight_decay);
      adam_one(p.y, g.y * gs, m.y, v.y, a.beta1, a.beta2, a.eps, step_size, sqrt_bc2, a.weight_decay);
      adam_one(p.z, g.z * gs, m.z, v.z, a.beta1, a.beta2, a.eps, step_size, sqrt_bc2, a.weight_decay);
      adam_one(p.w, g.w * gs, m.w, v.w, a.beta1, a.beta2, a.eps, step_size, sqrt_bc2, a.weight_decay);
      *reinterpret_cast<float4*>(sg.p + i) = p;
      *reinterpret_cast<float4*>(sg.m + i) = m;
      *reinterpret_cast<float4*>(sg.v + i) = v;
    }
  } else {
    int64_t end = base + kPerBlock < sg.n ? base + kPerBlock : sg.n;
    for (int64_t i = base + threadIdx.x; i < end; i += kThreads) {
      float p = sg.p[i], m = sg.m[i], v = sg.v[i];
      adam_one(p, sg.g[i] * gs, m, v, a.beta1, a.beta2, a.eps, step_size, sqrt_bc2, a.weight_decay);
      sg.p[i] = p;
      sg.m[i] = m;
      sg.v[i] = v;
    }
  }
}

}  // namespace

extern "C" int nkbs_adam_step(int32_t num_tensors, float* const* params, const float* const* grads,
                              float* const* exp_avg, float* const* exp_avg_sq, const int64_t* numel, float lr,
                              float beta1, float beta2, float eps, float weight_decay, const float* step_dev,
                              const float* inv_scale_dev, void* stream) {
  NKBS_CHECK_ARG(num_tensors >= 0 && step_dev, "nkbs_adam_step: bad arguments");
  int i = 0;
  while (i < num_tensors) {
    AdamArgs a;
    a.nseg = 0;
    int blocks = 0;
    while (i < num_tensors && a.nseg < kMaxSegs) {
      if (numel[i] > 0) {
        NKBS_CHECK_ARG(params[i] && grads[i] && exp_avg[i] && exp_avg_sq[i], "nkbs_adam_step: null tensor %d", i);
        int64_t nb = (numel[i] + kPerBlock - 1) / kPerBlock;
        NKBS_CHECK_ARG(blocks + nb < (1ll << 30), "nkbs_adam_step: tensor %d too large", i);
        a.seg[a.nseg] = AdamSeg{params[i], grads[i], exp_avg[i], exp_avg_sq[i], numel[i]};
        a.block_start[a.nseg] = blocks;
        blocks += (int)nb;
        ++a.nseg;
      }
      ++i;
    }
    if (a.nseg == 0) break;
    a.block_start[a.nseg] = blocks;
    a.lr = lr;
    a.beta1 = beta1;
    a.beta2 = beta2;
    a.eps = eps;
    a.weight_decay = weight_decay;
    a.step_dev = step_dev;
    a.inv_scale = inv_scale_dev;
    adam_kernel<<<blocks, kThreads, 0, (cudaStream_t)stream>>>(a);
    NKBS_CHECK_LAUNCH("nkbs_adam_step");
  }
  return 0;
}
