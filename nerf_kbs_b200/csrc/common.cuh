// Shared device helpers for the nkbs sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>

#define NKBS_MAX_LEVELS 16
#define NKBS_WARP 32

// error plumbing (abi.cu owns the storage)
void nkbs_set_error(const char* fmt, ...);
#define NKBS_CHECK_ARG(cond, ...)            \
  do {                                       \
    if (!(cond)) {                           \
      nkbs_set_error(__VA_ARGS__);           \
      return -1;                             \
    }                                        \
  } while (0)
#define NKBS_CHECK_LAUNCH(name)                                               \
  do {                                                                        \
    cudaError_t e__ = cudaGetLastError();                                     \
    if (e__ != cudaSuccess) {                                                 \
      nkbs_set_error("%s: launch failed: %s", name, cudaGetErrorString(e__)); \
      return -2;                                                              \
    }                                                                         \
  } while (0)

int nkbs_num_sms();

// ---- launches with programmatic dependent launch (PDL), opt-in: NKBS_PDL=1 ----
// Every kernel of the library starts with nkbs_pdl_sync(): griddepcontrol.wait (returns once the kernel launched before
// it on the stream has completed and its writes are visible; a no-op for a launch without the attribute) followed by
// griddepcontrol.launch_dependents.  With the launch attribute the NEXT kernel's CTAs are scheduled onto SMs as soon as
// every CTA of this kernel has started and resources free up, and sit at their own wait, so that launch latency and CTA
// scheduling overlap the predecessor's tail.  Nothing a kernel reads or writes is touched before the wait: the
// semantics are those of ordinary stream order (the GPU test suite passes in both modes).
// MEASURED on the B200 (same build, back to back, captured-graph steps): it does not pay here — 4096-ray step
// 0.752-0.756 ms without, 0.769-0.771 ms with (trigger after or before the wait alike), render 21.9 -> 21.0 M rays/s,
// 8192-ray semantic step and 65 536-ray step unchanged (+-0.3 %).  Inside a graph the kernel-to-kernel edge is already
// cheap, and the early-resident CTAs take SM slots from the kernels of the step's other streams (proposal backward,
// deferred Adam).  Hence off by default.
bool nkbs_pdl_enabled();
#ifdef __CUDACC__
__device__ __forceinline__ void nkbs_pdl_sync() {
  asm volatile("griddepcontrol.wait;\n\tgriddepcontrol.launch_dependents;" ::: "memory");
}
template <typename... P, typename... A>
inline void nkbs_launch(void (*kern)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, A&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  if (nkbs_pdl_enabled()) {
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
  }
  cudaLaunchKernelEx(&cfg, kern, static_cast<P>(args)...);  // an error stays in cudaGetLastError for NKBS_CHECK_LAUNCH
}
#endif

// ---- deterministic gradient accumulation (opt-in, nkbs_fixed_point_register) ----
// A registered gradient buffer [base, base + n) has a 64-bit fixed-point shadow (2^log2_scale units per 1.0).  Kernels
// that would add a float partial sum to an address inside the buffer with an atomic add it to the shadow instead:
// integer additions commute exactly, so the total does not depend on the order in which CTAs and warps arrive.
// nkbs_fixed_point_flush adds the shadow into the float buffer (one thread per element, no atomics) and re-zeroes it.
// Addresses outside every registered buffer (and all addresses when nothing is registered) take the float atomics.
struct FixedMap {
  const float* base;
  long long* shadow;  // null: not registered
  long long n;
  float scale;        // 2^log2_scale
  __device__ __forceinline__ long long* at(const float* p) const {
    if (!shadow) return nullptr;
    const long long i = p - base;
    return (i >= 0 && i < n) ? shadow + i : nullptr;
  }
};
FixedMap nkbs_fixed_map_for(const void* p);  // host side (abi.cu): the registered buffer that holds p, or a null map

#ifdef __CUDACC__
__device__ __forceinline__ void nkbs_fixed_add(long long* addr, float v, float scale) {
  if (v != 0.f) atomicAdd(reinterpret_cast<unsigned long long*>(addr), (unsigned long long)__float2ll_rn(v * scale));
}
// float partial sum v -> *p: fixed-point shadow when p lies in a registered buffer, float atomic otherwise
__device__ __forceinline__ void nkbs_acc(float* p, const FixedMap& fm, float v) {
  long long* f = fm.at(p);
  if (f) nkbs_fixed_add(f, v, fm.scale);
  else atomicAdd(p, v);
}
#endif

struct LevelParams {
  float scale[NKBS_MAX_LEVELS];
  int num_levels;
  uint32_t table_mask;  // T-1
  uint32_t log2_T;
};

// ---- position front end: ray point, L-inf scene contraction, [0,1] normalisation, selector ----
// Every step is an explicit IEEE op in the same order as the torch path so that the cell a sample
// lands in (and hence every hash index) is bit-identical to the oracle: no FMA contraction here.
__device__ __forceinline__ float3 nkbs_frustum_position(float3 o, float3 d, float start, float end) {
  // origins + directions * (starts + ends) / 2   (torch evaluates (d*(s+e))/2 left to right)
  float se = __fadd_rn(start, end);
  float3 p;
  p.x = __fadd_rn(o.x, __fmul_rn(__fmul_rn(d.x, se), 0.5f));
  p.y = __fadd_rn(o.y, __fmul_rn(__fmul_rn(d.y, se), 0.5f));
  p.z = __fadd_rn(o.z, __fmul_rn(__fmul_rn(d.z, se), 0.5f));
  return p;
}

__device__ __forceinline__ float3 nkbs_contract_linf(float3 p) {
  float mag = fmaxf(fmaxf(fabsf(p.x), fabsf(p.y)), fabsf(p.z));
  if (mag < 1.0f) return p;
  float s = __fsub_rn(2.0f, __fdiv_rn(1.0f, mag));
  float3 q;
  q.x = __fmul_rn(s, __fdiv_rn(p.x, mag));
  q.y = __fmul_rn(s, __fdiv_rn(p.y, mag));
  q.z = __fmul_rn(s, __fdiv_rn(p.z, mag));
  return q;
}

struct NormParams {
  int contraction;     // 1: L-inf contraction then (x+2)/4 ; 0: (x-aabb_min)/(aabb_max-aabb_min)
  float aabb_min[3];
  float aabb_len[3];
};

// returns normalised position already multiplied by the selector; sel = 1.f inside (0,1)^3 else 0.f
__device__ __forceinline__ float3 nkbs_normalise(float3 p, const NormParams& np, float& sel) {
  float3 q;
  if (np.contraction) {
    q = nkbs_contract_linf(p);
    q.x = __fmul_rn(__fadd_rn(q.x, 2.0f), 0.25f);
    q.y = __fmul_rn(__fadd_rn(q.y, 2.0f), 0.25f);
    q.z = __fmul_rn(__fadd_rn(q.z, 2.0f), 0.25f);
  } else {
    q.x = __fdiv_rn(__fsub_rn(p.x, np.aabb_min[0]), np.aabb_len[0]);
    q.y = __fdiv_rn(__fsub_rn(p.y, np.aabb_min[1]), np.aabb_len[1]);
    q.z = __fdiv_rn(__fsub_rn(p.z, np.aabb_min[2]), np.aabb_len[2]);
  }
  bool in = (q.x > 0.f) && (q.x < 1.f) && (q.y > 0.f) && (q.y < 1.f) && (q.z > 0.f) && (q.z < 1.f);
  sel = in ? 1.f : 0.f;
  // torch multiplies by the boolean selector: NaN*0 = NaN is preserved by doing the multiply
  q.x = __fmul_rn(q.x, sel);
  q.y = __fmul_rn(q.y, sel);
  q.z = __fmul_rn(q.z, sel);
  return q;
}

// ---- hash ----
// ((x*1) ^ (y*2654435761) ^ (z*805459861)) mod T in int64 == low log2(T) bits of the uint32 product
// chain for non-negative coordinates (T is a power of two).
__device__ __forceinline__ uint32_t nkbs_hash(uint32_t x, uint32_t y, uint32_t z, uint32_t mask) {
  return (x ^ (y * 2654435761u) ^ (z * 805459861u)) & mask;
}

// torch: ceil(s).int32 / floor(s).int32 ; a NaN converts to INT_MIN on x86, cvt.rzi gives 0 on the
// GPU.  NaN positions only arise from NaN rays; both sides then produce NaN features via offset.
struct Cell {
  uint32_t c[3], f[3];
  float o[3];
};
__device__ __forceinline__ Cell nkbs_cell(float3 x, float scale) {
  Cell cl;
  float s0 = __fmul_rn(x.x, scale), s1 = __fmul_rn(x.y, scale), s2 = __fmul_rn(x.z, scale);
  float f0 = floorf(s0), f1 = floorf(s1), f2 = floorf(s2);
  cl.c[0] = (uint32_t)(int)ceilf(s0);
  cl.c[1] = (uint32_t)(int)ceilf(s1);
  cl.c[2] = (uint32_t)(int)ceilf(s2);
  cl.f[0] = (uint32_t)(int)f0;
  cl.f[1] = (uint32_t)(int)f1;
  cl.f[2] = (uint32_t)(int)f2;
  cl.o[0] = __fsub_rn(s0, f0);
  cl.o[1] = __fsub_rn(s1, f1);
  cl.o[2] = __fsub_rn(s2, f2);
  return cl;
}

// corner order of the torch path: h0=(c,c,c) h1=(c,f,c) h2=(f,f,c) h3=(f,c,c) h4=(c,c,f) h5=(c,f,f)
// h6=(f,f,f) h7=(f,c,f)
__device__ __forceinline__ void nkbs_corner_hashes(const Cell& cl, uint32_t mask, uint32_t h[8]) {
  uint32_t yc = cl.c[1] * 2654435761u, yf = cl.f[1] * 2654435761u;
  uint32_t zc = cl.c[2] * 805459861u, zf = cl.f[2] * 805459861u;
  h[0] = (cl.c[0] ^ yc ^ zc) & mask;
  h[1] = (cl.c[0] ^ yf ^ zc) & mask;
  h[2] = (cl.f[0] ^ yf ^ zc) & mask;
  h[3] = (cl.f[0] ^ yc ^ zc) & mask;
  h[4] = (cl.c[0] ^ yc ^ zf) & mask;
  h[5] = (cl.c[0] ^ yf ^ zf) & mask;
  h[6] = (cl.f[0] ^ yf ^ zf) & mask;
  h[7] = (cl.f[0] ^ yc ^ zf) & mask;
}

// trilinear blend in the torch path's association order
__device__ __forceinline__ float nkbs_trilerp(const float v[8], float ox, float oy, float oz) {
  float nx = 1.f - ox, ny = 1.f - oy, nz = 1.f - oz;
  float f03 = v[0] * ox + v[3] * nx;
  float f12 = v[1] * ox + v[2] * nx;
  float f56 = v[5] * ox + v[6] * nx;
  float f47 = v[4] * ox + v[7] * nx;
  float f0312 = f03 * oy + f12 * ny;
  float f4756 = f47 * oy + f56 * ny;
  return f0312 * oz + f4756 * nz;
}

// weight of corner k in the blend above (used by the backward scatter)
__device__ __forceinline__ void nkbs_corner_weights(float ox, float oy, float oz, float w[8]) {
  float nx = 1.f - ox, ny = 1.f - oy, nz = 1.f - oz;
  w[0] = ox * oy * oz;
  w[3] = nx * oy * oz;
  w[1] = ox * ny * oz;
  w[2] = nx * ny * oz;
  w[5] = ox * ny * nz;
  w[6] = nx * ny * nz;
  w[4] = ox * oy * nz;
  w[7] = nx * oy * nz;
}

__device__ __forceinline__ void nkbs_red_add_v2(float* addr, float a, float b) {
  // one 8-byte vector reduction instead of two scalar atomics (sm_90+)
  asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(addr), "f"(a), "f"(b) : "memory");
}

__device__ __forceinline__ void nkbs_red_add_v4(float* addr, float a, float b, float c, float d) {
  // one 16-byte vector reduction (sm_90+); addr must be 16-byte aligned
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__device__ __forceinline__ float nkbs_sigmoid(float x) { return 1.f / (1.f + expf(-x)); }
__device__ __forceinline__ float nkbs_softplus(float x) {
  // torch softplus: beta=1, threshold=20
  return x > 20.f ? x : log1pf(expf(x));
}
