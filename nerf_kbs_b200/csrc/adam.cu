// N2: multi-tensor Adam over the model's parameter tensors, one launch per <= 24 tensors.
// Replaces torch.optim.Adam(lr=1e-3, eps=1e-15) that the reference configures for both parameter
// groups (ref:train.py:85-98; groups ref:nerfacto.py:359-363) [nerfstudio Optimizers, NS-recalled].
// HBM streaming kernel: 4 words read + 3 written per parameter, 128-bit accesses.
#include <stdlib.h>
#include "common.cuh"

namespace {

constexpr int kMaxSegs = 24;
constexpr int kThreads = 256;
constexpr int kPerBlock = kThreads * 16;  // floats per CTA

struct AdamSeg {
  float* p;
  float* g;
  float* m;
  float* v;
  int64_t n;
};

struct AdamArgs {
  AdamSeg seg[kMaxSegs];
  int block_start[kMaxSegs + 1];  // first CTA of every segment
  int nseg;
  float lr, beta1, beta2, eps, weight_decay;
  float* step_dev;         // device scalar: step count t >= 1 (float), read by every thread
  unsigned* counter;       // null: *step_dev is this step's count.  Else *step_dev counts the steps BEFORE this one: the
                           // kernel uses t = *step_dev + 1 and the last CTA to finish stores it (no separate add launch)
  const float* inv_scale;  // optional device scalar multiplied into the gradient (GradScaler unscale / 1/world)
  int zero_grads;          // write 0 back over every gradient read: the buffer is ready for the next backward pass
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float b1, float b2, float eps,
                                         float step_size, float sqrt_bc2, float wd) {
  if (wd != 0.f) g = fmaf(wd, p, g);
  m = fmaf(b1, m, (1.f - b1) * g);      // exp_avg.lerp_(grad, 1-beta1)
  v = fmaf(b2, v, (1.f - b2) * g * g);  // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1-beta2)
  float denom = sqrtf(v) / sqrt_bc2 + eps;
  p -= step_size * (m / denom);
}

// Every CTA has read the old count before it arrives here, and the CTA that arrives last knows that all have: it
// publishes the new count and re-arms the counter for the next launch (graph replays included).
__device__ __forceinline__ void advance_step(float* step_dev, unsigned* counter, float t) {
  if (!counter) return;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(counter, 1u) == gridDim.x - 1) {
      *step_dev = t;
      *counter = 0u;
    }
  }
}

// The denominator and the quotient with the hardware approximations (sqrt.approx, div.approx: <= 3 ulp on the update
// term, which is ~lr relative to the parameter: the stored parameter differs from the IEEE kernel's by at most one ulp
// on ~1/5 of the elements after 5 steps, 3e-8 relative L2).  The IEEE sequences cost ~50 instructions per element with
// their range checks and slow paths and made the kernel issue-bound (issue slots 49 % busy at 58 % of DRAM throughput,
// profiles/r02_top_kernels_ncu_full.csv): 108.5 -> 96.3 us on the main field's 16.8 M parameters (0.77 -> 0.87 of the
// measured copy bandwidth, scripts/adam_bench.py).  NKBS_ADAM_IEEE=1 selects the IEEE kernel.
__device__ __forceinline__ void adam_one_fast(float& p, float g, float& m, float& v, float b1, float b2, float eps,
                                              float step_size, float inv_sqrt_bc2, float wd) {
  if (wd != 0.f) g = fmaf(wd, p, g);
  m = fmaf(b1, m, (1.f - b1) * g);
  v = fmaf(b2, v, (1.f - b2) * g * g);
  float r;
  asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(v));
  const float denom = fmaf(r, inv_sqrt_bc2, eps);
  p -= step_size * __fdividef(m, denom);
}

template <bool FAST>
__global__ void __launch_bounds__(kThreads) adam_kernel(const AdamArgs a) {
  nkbs_pdl_sync();
  int s = 0;
#pragma unroll 1
  while (s + 1 < a.nseg && (int)blockIdx.x >= a.block_start[s + 1]) ++s;
  const AdamSeg sg = a.seg[s];
  const float t = a.counter ? *a.step_dev + 1.f : *a.step_dev;
  const float bc1 = 1.f - powf(a.beta1, t);
  const float bc2 = 1.f - powf(a.beta2, t);
  const float step_size = a.lr / bc1;
  const float sqrt_bc2 = sqrtf(bc2);
  const float c2 = FAST ? 1.f / sqrt_bc2 : sqrt_bc2;  // what the per-element routine takes
  const float gs = a.inv_scale ? *a.inv_scale : 1.f;
  auto one = [&](float& p, float g, float& m, float& v) {
    if (FAST) adam_one_fast(p, g * gs, m, v, a.beta1, a.beta2, a.eps, step_size, c2, a.weight_decay);
    else adam_one(p, g * gs, m, v, a.beta1, a.beta2, a.eps, step_size, c2, a.weight_decay);
  };
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
      one(p.x, g.x, m.x, v.x);
      one(p.y, g.y, m.y, v.y);
      one(p.z, g.z, m.z, v.z);
      one(p.w, g.w, m.w, v.w);
      *reinterpret_cast<float4*>(sg.p + i) = p;
      *reinterpret_cast<float4*>(sg.m + i) = m;
      *reinterpret_cast<float4*>(sg.v + i) = v;
      if (a.zero_grads) __stcs(reinterpret_cast<float4*>(sg.g + i), make_float4(0.f, 0.f, 0.f, 0.f));
    }
  } else {
    int64_t end = base + kPerBlock < sg.n ? base + kPerBlock : sg.n;
    for (int64_t i = base + threadIdx.x; i < end; i += kThreads) {
      float p = sg.p[i], m = sg.m[i], v = sg.v[i];
      one(p, sg.g[i], m, v);
      sg.p[i] = p;
      sg.m[i] = m;
      sg.v[i] = v;
      if (a.zero_grads) sg.g[i] = 0.f;
    }
  }
  advance_step(a.step_dev, a.counter, t);
}

// ---- data-parallel step over NVLink peer memory: reduce-scatter + Adam + all-gather in ONE kernel ----
// Every rank owns the slice [lo, lo+n) of a flat parameter buffer that exists at the same offsets on every GPU
// (symmetric memory).  For its slice a rank reads the gradient from all peers (P2P loads), averages, applies Adam
// with moments it alone keeps (a 1/world shard), and stores the new parameters into every peer's buffer (P2P
// stores).  Link traffic equals an allreduce's (reduce-scatter + all-gather) but the separate Adam pass, its
// 7 words/parameter of HBM traffic on every GPU and the gradient scaling pass are gone.
constexpr int kMaxPeers = 16;
struct DpAdamArgs {
  const float* g[kMaxPeers];  // flat gradient buffer of every rank
  float* p[kMaxPeers];        // flat parameter buffer of every rank
  int world, rank;
  int64_t lo, n;              // this rank's slice (elements; lo and n are multiples of 4)
  float* m;                   // [n] first moments of the slice
  float* v;                   // [n]
  float lr, beta1, beta2, eps, weight_decay;
  float* step_dev;
  unsigned* counter;          // as AdamArgs::counter
  const float* g_mc;          // multicast (NVLS) address of the gradient buffer, or null
  float* p_mc;                // multicast address of the parameter buffer, or null
};

// WORLD > 0: number of ranks known at compile time, the peer loads of one element are all in flight together
// (NVLink round trips are ~2 us: the loop over ranks must not serialise them); WORLD = 0: any rank count.
template <int WORLD>
__global__ void __launch_bounds__(kThreads) dp_adam_kernel(const DpAdamArgs a) {
  nkbs_pdl_sync();
  const float t = a.counter ? *a.step_dev + 1.f : *a.step_dev;
  const float bc1 = 1.f - powf(a.beta1, t);
  const float bc2 = 1.f - powf(a.beta2, t);
  const float step_size = a.lr / bc1;
  const float sqrt_bc2 = sqrtf(bc2);
  const int world = WORLD > 0 ? WORLD : a.world;
  const float inv_world = 1.f / (float)world;
  const int64_t nvec = a.n >> 2;
  for (int64_t q = blockIdx.x * (int64_t)kThreads + threadIdx.x; q < nvec; q += (int64_t)gridDim.x * kThreads) {
    const int64_t i = a.lo + 4 * q;
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
    if (WORLD > 0) {
      float4 x[WORLD > 0 ? WORLD : 1];
#pragma unroll
      for (int r = 0; r < WORLD; ++r) x[r] = *reinterpret_cast<const float4*>(a.g[r] + i);
#pragma unroll
      for (int r = 0; r < WORLD; ++r) {  // rank order on every rank: all ranks compute the same sum
        g.x += x[r].x;
        g.y += x[r].y;
        g.z += x[r].z;
        g.w += x[r].w;
      }
    } else {
#pragma unroll 1
      for (int r = 0; r < world; ++r) {
        const float4 x = *reinterpret_cast<const float4*>(a.g[r] + i);
        g.x += x.x;
        g.y += x.y;
        g.z += x.z;
        g.w += x.w;
      }
    }
    float4 p = *reinterpret_cast<const float4*>(a.p[a.rank] + i);
    float4 m = *reinterpret_cast<const float4*>(a.m + 4 * q);
    float4 v = *reinterpret_cast<const float4*>(a.v + 4 * q);
    adam_one(p.x, g.x * inv_world, m.x, v.x, a.beta1, a.beta2, a.eps, step_size, sqrt_bc2, a.weight_decay);
    adam_one(p.y, g.y * inv_world, m.y, v.y, a.beta1, a.beta2, a.eps, step_size, sqrt_bc2, a.weight_decay);
    adam_one(p.z, g.z * inv_world, m.z, v.z, a.beta1, a.beta2, a.eps, step_size, sqrt_bc2, a.weight_decay);
    adam_one(p.w, g.w * inv_world, m.w, v.w, a.beta1, a.beta2, a.eps, step_size, sqrt_bc2, a.weight_decay);
    *reinterpret_cast<float4*>(a.m + 4 * q) = m;
    *reinterpret_cast<float4*>(a.v + 4 * q) = v;
    if (WORLD > 0) {
#pragma unroll
      for (int r = 0; r < WORLD; ++r) *reinterpret_cast<float4*>(a.p[r] + i) = p;
    } else {
#pragma unroll 1
      for (int r = 0; r < world; ++r) *reinterpret_cast<float4*>(a.p[r] + i) = p;
    }
  }
  advance_step(a.step_dev, a.counter, t);
}


// The same step through the NVSwitch's multicast objects (NVLS): `multimem.ld_reduce` returns the SUM over all ranks of
// one 16-byte piece of the gradient — the reduction happens inside the switch, this GPU receives one slice instead of
// `world` of them — and `multimem.st` hands one store of the new parameters to the switch, which replicates it into
// every rank's buffer (this rank's included).  NVLink traffic per GPU drops from 2 (world - 1) / world to 2 / world of
// the parameter bytes.  The sum's order inside the switch is not specified, but only the owner of a slice reduces it and
// every rank receives the owner's result: ranks stay bit-identical.
__device__ __forceinline__ float4 multimem_ld_reduce_add(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

__global__ void __launch_bounds__(kThreads) dp_adam_multimem_kernel(const DpAdamArgs a) {
  nkbs_pdl_sync();
  const float t = a.counter ? *a.step_dev + 1.f : *a.step_dev;
  const float bc1 = 1.f - powf(a.beta1, t);
  const float bc2 = 1.f - powf(a.beta2, t);
  const float step_size = a.lr / bc1;
  const float sqrt_bc2 = sqrtf(bc2);
  const float inv_world = 1.f / (float)a.world;
  const int64_t nvec = a.n >> 2;
  // kU pieces per thread and pass, all reductions issued before the first is used: a switch round trip is several
  // microseconds, the loop must keep many of them in flight
  constexpr int kU = 4;
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  for (int64_t q0 = blockIdx.x * (int64_t)kThreads + threadIdx.x; q0 < nvec; q0 += kU * stride) {
    float4 g[kU], p[kU], m[kU], v[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t q = q0 + u * stride;
      if (q < nvec) g[u] = multimem_ld_reduce_add(a.g_mc + a.lo + 4 * q);
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t q = q0 + u * stride;
      if (q < nvec) {
        p[u] = *reinterpret_cast<const float4*>(a.p[a.rank] + a.lo + 4 * q);
        m[u] = *reinterpret_cast<const float4*>(a.m + 4 * q);
        v[u] = *reinterpret_cast<const float4*>(a.v + 4 * q);
      }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t q = q0 + u * stride;
      if (q < nvec) {
        adam_one(p[u].x, g[u].x * inv_world, m[u].x, v[u].x, a.beta1, a.beta2, a.eps, step_size, sqrt_bc2, a.weight_decay);
        adam_one(p[u].y, g[u].y * inv_world, m[u].y, v[u].y, a.beta1, a.beta2, a.eps, step_size, sqrt_bc2, a.weight_decay);
        adam_one(p[u].z, g[u].z * inv_world, m[u].z, v[u].z, a.beta1, a.beta2, a.eps, step_size, sqrt_bc2, a.weight_decay);
        adam_one(p[u].w, g[u].w * inv_world, m[u].w, v[u].w, a.beta1, a.beta2, a.eps, step_size, sqrt_bc2, a.weight_decay);
        *reinterpret_cast<float4*>(a.m + 4 * q) = m[u];
        *reinterpret_cast<float4*>(a.v + 4 * q) = v[u];
        multimem_st(a.p_mc + a.lo + 4 * q, p[u]);
      }
    }
  }
  advance_step(a.step_dev, a.counter, t);
}

}  // namespace

extern "C" int nkbs_adam_step(int32_t num_tensors, float* const* params, float* const* grads,
                              float* const* exp_avg, float* const* exp_avg_sq, const int64_t* numel, float lr,
                              float beta1, float beta2, float eps, float weight_decay, float* step_dev,
                              const float* inv_scale_dev, int32_t zero_grads, uint32_t* step_counter, void* stream) {
  NKBS_CHECK_ARG(num_tensors >= 0 && step_dev, "nkbs_adam_step: bad arguments");
  int64_t total_blocks = 0;
  for (int j = 0; j < num_tensors; ++j) total_blocks += (numel[j] + kPerBlock - 1) / kPerBlock;
  NKBS_CHECK_ARG(!step_counter || (num_tensors <= kMaxSegs && total_blocks > 0),
                 "nkbs_adam_step: the in-kernel step count needs a single launch (at most %d tensors, at least one element)", kMaxSegs);
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
    a.zero_grads = zero_grads;
    a.counter = step_counter;
    const char* ev = getenv("NKBS_ADAM_IEEE");
    if (ev && ev[0] == '1') nkbs_launch(adam_kernel<false>, blocks, kThreads, 0, (cudaStream_t)stream, a);
    else nkbs_launch(adam_kernel<true>, blocks, kThreads, 0, (cudaStream_t)stream, a);
    NKBS_CHECK_LAUNCH("nkbs_adam_step");
  }
  return 0;
}

extern "C" int nkbs_dp_adam_step(int32_t world, int32_t rank, const float* const* grad_peers, float* const* param_peers,
                                 int64_t lo, int64_t n, float* exp_avg, float* exp_avg_sq, float lr, float beta1,
                                 float beta2, float eps, float weight_decay, float* step_dev, uint32_t* step_counter,
                                 const float* grad_multicast, float* param_multicast, void* stream) {
  NKBS_CHECK_ARG(world >= 1 && world <= kMaxPeers && rank >= 0 && rank < world, "nkbs_dp_adam_step: bad world/rank %d/%d", world, rank);
  NKBS_CHECK_ARG(lo >= 0 && n >= 0 && (lo & 3) == 0 && (n & 3) == 0, "nkbs_dp_adam_step: slice must be 4-element aligned");
  if (n == 0) return 0;
  NKBS_CHECK_ARG(grad_peers && param_peers && exp_avg && exp_avg_sq && step_dev, "nkbs_dp_adam_step: null pointers");
  DpAdamArgs a;
  for (int r = 0; r < world; ++r) {
    NKBS_CHECK_ARG(grad_peers[r] && param_peers[r], "nkbs_dp_adam_step: null peer buffer %d", r);
    NKBS_CHECK_ARG(((reinterpret_cast<uintptr_t>(grad_peers[r]) | reinterpret_cast<uintptr_t>(param_peers[r])) & 15) == 0,
                   "nkbs_dp_adam_step: peer buffers must be 16-byte aligned");
    a.g[r] = grad_peers[r];
    a.p[r] = param_peers[r];
  }
  a.world = world;
  a.rank = rank;
  a.lo = lo;
  a.n = n;
  a.m = exp_avg;
  a.v = exp_avg_sq;
  a.lr = lr;
  a.beta1 = beta1;
  a.beta2 = beta2;
  a.eps = eps;
  a.weight_decay = weight_decay;
  a.step_dev = step_dev;
  a.counter = step_counter;
  NKBS_CHECK_ARG((grad_multicast == nullptr) == (param_multicast == nullptr), "nkbs_dp_adam_step: both multicast addresses or none");
  NKBS_CHECK_ARG(((reinterpret_cast<uintptr_t>(grad_multicast) | reinterpret_cast<uintptr_t>(param_multicast)) & 15) == 0,
                 "nkbs_dp_adam_step: multicast addresses must be 16-byte aligned");
  a.g_mc = grad_multicast;
  a.p_mc = param_multicast;
  int64_t blocks = ((n >> 2) + kThreads - 1) / kThreads;
  int64_t cap = (int64_t)nkbs_num_sms() * 8;
  const int grid = (int)(blocks < cap ? blocks : cap);
  cudaStream_t st = (cudaStream_t)stream;
  if (a.g_mc) {
    nkbs_launch(dp_adam_multimem_kernel, grid, kThreads, 0, st, a);
    NKBS_CHECK_LAUNCH("nkbs_dp_adam_step(multimem)");
    return 0;
  }
  switch (world) {
    case 2: nkbs_launch(dp_adam_kernel<2>, grid, kThreads, 0, st, a); break;
    case 4: nkbs_launch(dp_adam_kernel<4>, grid, kThreads, 0, st, a); break;
    case 8: nkbs_launch(dp_adam_kernel<8>, grid, kThreads, 0, st, a); break;
    default: nkbs_launch(dp_adam_kernel<0>, grid, kThreads, 0, st, a); break;
  }
  NKBS_CHECK_LAUNCH("nkbs_dp_adam_step");
  return 0;
}
