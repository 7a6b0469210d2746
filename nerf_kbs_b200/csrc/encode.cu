// K1/K2: position front end + multiresolution hash-grid encoding, forward and backward.
// Replaces nerfstudio HashEncoding.pytorch_fwd and its autograd, reached from
// ref:nerfacto.py:275-311 and ref:semantic_nerfw.py:102-126.
#include <cstdlib>

#include "common.cuh"
#include "encode.cuh"

namespace {

constexpr int kEncThreads = 256;
// x edges of a cell as (ceil-x corner, floor-x corner) in the corner order of nkbs_corner_hashes
__device__ constexpr int kEdgeC[4] = {0, 1, 4, 5};
__device__ constexpr int kEdgeF[4] = {3, 2, 7, 6};

struct EncodeArgs : SampleSource {
  int64_t N;
  LevelParams lv;
  const float2* table;
  float* xn;
  float* selector;
  float* features;
  int64_t* indices;
};

// One thread per sample; the level loop keeps 4 16-byte (+ up to 4 8-byte) independent gathers in flight per level and
// the compiler overlaps two levels (unroll 2).  Level scales live in shared memory.
template <bool kWriteIdx>
__global__ void __launch_bounds__(kEncThreads) encode_fwd_kernel(const EncodeArgs a) {
  nkbs_pdl_sync();
  __shared__ float s_scale[NKBS_MAX_LEVELS];
  if (threadIdx.x < NKBS_MAX_LEVELS) s_scale[threadIdx.x] = a.lv.scale[threadIdx.x];
  __syncthreads();
  const int L = a.lv.num_levels;
  const uint32_t mask = a.lv.table_mask;
  // gridDim.y slices of the level range (small batches: more CTAs than one per 256 samples, see level_slices)
  const int l_begin = (int)((blockIdx.y * L) / gridDim.y), l_end = (int)(((blockIdx.y + 1) * L) / gridDim.y);
  for (int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; n < a.N; n += (int64_t)gridDim.x * blockDim.x) {
    float sel;
    float3 x = load_sample_position(a, n, sel);
    if (a.xn && blockIdx.y == 0) {
      a.xn[3 * n] = x.x;
      a.xn[3 * n + 1] = x.y;
      a.xn[3 * n + 2] = x.z;
    }
    if (a.selector && blockIdx.y == 0) a.selector[n] = sel;
    float2* out = reinterpret_cast<float2*>(a.features + n * (2 * L));
#pragma unroll 2
    for (int l = l_begin; l < l_end; ++l) {
      Cell cl = nkbs_cell(x, s_scale[l]);
      uint32_t h[8];
      nkbs_corner_hashes(cl, mask, h);
      const float2* tl = a.table + ((size_t)l << a.lv.log2_T);
      // The x prime of the hash is 1, so the two corners of an x edge are rows h and h ^ 1 of one 16-byte pair whenever
      // floor(x) is even (or ceil == floor): one 16-byte gather at the pair base returns both, and the second gather of
      // the edge is issued only when floor(x) is odd.  Same rows and values as eight 8-byte gathers, a quarter fewer
      // sector requests on the levels whose gathers do not coalesce.
      const bool same = cl.c[0] == cl.f[0];
      const bool in_pair = same || !(cl.f[0] & 1u);
      float4 q[4];
      float2 vc[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) q[j] = __ldg(reinterpret_cast<const float4*>(tl + (h[kEdgeF[j]] & ~1u)));
#pragma unroll
      for (int j = 0; j < 4; ++j) vc[j] = in_pair ? make_float2(0.f, 0.f) : __ldg(tl + h[kEdgeC[j]]);
      float2 v[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const bool hi = h[kEdgeF[j]] & 1u;
        const float2 lo2 = make_float2(q[j].x, q[j].y), hi2 = make_float2(q[j].z, q[j].w);
        const float2 vf = hi ? hi2 : lo2, other = hi ? lo2 : hi2;
        v[kEdgeF[j]] = vf;
        v[kEdgeC[j]] = same ? vf : (in_pair ? other : vc[j]);
      }
      float v0[8], v1[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        v0[k] = v[k].x;
        v1[k] = v[k].y;
      }
      out[l] = make_float2(nkbs_trilerp(v0, cl.o[0], cl.o[1], cl.o[2]), nkbs_trilerp(v1, cl.o[0], cl.o[1], cl.o[2]));
      if (kWriteIdx) {
        int64_t* ip = a.indices + (n * L + l) * 8;
#pragma unroll
        for (int k = 0; k < 8; ++k) ip[k] = (int64_t)h[k] + ((int64_t)l << a.lv.log2_T);
      }
    }
  }
}

struct EncodeBwdArgs {
  const float* xn;        // [N,3] normalised positions (selector already applied)
  const float* dfeat;     // [N,2L]
  int64_t N;
  LevelParams lv;
  float* dtable;          // [L*T,2], accumulated into
  int dedup_levels;       // levels below this index use warp-level duplicate merging
  float v4_min_scale;     // levels of at least this resolution pair their reductions (16-byte red)
  FixedMap fm;            // deterministic mode (common.cuh): shadow of the buffer that holds dtable, or a null map
};

template <bool FX>
__global__ void __launch_bounds__(kEncThreads) encode_bwd_kernel(const EncodeBwdArgs a) {
  nkbs_pdl_sync();
  __shared__ float s_scale[NKBS_MAX_LEVELS];
  if (threadIdx.x < NKBS_MAX_LEVELS) s_scale[threadIdx.x] = a.lv.scale[threadIdx.x];
  __syncthreads();
  const int L = a.lv.num_levels;
  const uint32_t mask = a.lv.table_mask;
  const int lane = threadIdx.x & 31;
  const int l_begin = (int)((blockIdx.y * L) / gridDim.y), l_end = (int)(((blockIdx.y + 1) * L) / gridDim.y);
  // grid-stride over whole warps so that every lane of a warp reaches the collectives
  const int64_t n_round = (a.N + 31) / 32 * 32;
  for (int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; n < n_round; n += (int64_t)gridDim.x * blockDim.x) {
    const bool live = n < a.N;
    float3 x = make_float3(0.f, 0.f, 0.f);
    if (live) x = make_float3(a.xn[3 * n], a.xn[3 * n + 1], a.xn[3 * n + 2]);
    const float2* g = reinterpret_cast<const float2*>(a.dfeat + (live ? n : 0) * (2 * L));
    for (int l = l_begin; l < l_end; ++l) {
      Cell cl = nkbs_cell(x, s_scale[l]);
      uint32_t h[8];
      nkbs_corner_hashes(cl, mask, h);
      float w[8];
      nkbs_corner_weights(cl.o[0], cl.o[1], cl.o[2], w);
      float2 gl = live ? __ldg(g + l) : make_float2(0.f, 0.f);
      float* base = a.dtable + ((size_t)l << (a.lv.log2_T + 1));
      long long* fixed = FX ? a.fm.at(base) : nullptr;
      if (l < a.dedup_levels) {
        scatter_level_runs<FX>(base, cl, h, w, gl.x, gl.y, lane, s_scale[l] >= a.v4_min_scale, fixed, a.fm.scale);
      } else {
        float vx[8], vy[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          vx[k] = w[k] * gl.x;
          vy[k] = w[k] * gl.y;
        }
        scatter_cell<FX>(base, cl, h, vx, vy, s_scale[l] >= a.v4_min_scale, fixed, a.fm.scale);
      }
    }
  }
}

// ---- ray-major gradient scatter (proposal grids: 96-256 samples per ray on 5-8 coarse levels) -------------------------
// One warp per ray, every lane owns a CONTIGUOUS run of ceil(S / 32) samples and walks it serially, level by level:
// while consecutive samples stay in one cell their 8 x 2 corner gradients are summed in registers, and only a change of
// cell (or the end of the lane's run) sends the 8 rows to the table.  On these grids neighbouring samples of a ray share
// their cell on most levels, so this replaces the warp-wide segmented shuffle reduction of scatter_level_runs — up to
// 5 steps x 16 shuffles per level and sample, the dominant instruction cost of the sample-major scatter there — by 16
// FMAs per sample and one flush per cell crossing (plus one per lane boundary).
struct RayScatterArgs : SampleSource {
  int64_t R;
  LevelParams lv;
  const float* dfeat;   // [R*S, 2L], zero for samples outside the grid
  float* dtable;        // [L*T, 2], accumulated into
  float v4_min_scale;
  FixedMap fm;          // deterministic mode, see EncodeBwdArgs
};

template <int CMAX, bool FX>
__global__ void __launch_bounds__(kEncThreads) encode_bwd_rays_kernel(const RayScatterArgs a) {
  nkbs_pdl_sync();
  __shared__ float s_scale[NKBS_MAX_LEVELS];
  if (threadIdx.x < NKBS_MAX_LEVELS) s_scale[threadIdx.x] = a.lv.scale[threadIdx.x];
  __syncthreads();
  const int L = a.lv.num_levels, S = a.S;
  const uint32_t mask = a.lv.table_mask;
  const int lane = threadIdx.x & 31;
  const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int C = (S + 31) >> 5;
  for (int64_t r = warp; r < a.R; r += nwarps) {
    const float3 o = make_float3(a.origins[3 * r], a.origins[3 * r + 1], a.origins[3 * r + 2]);
    const float3 d = make_float3(a.directions[3 * r], a.directions[3 * r + 1], a.directions[3 * r + 2]);
    float3 xs[CMAX];
#pragma unroll
    for (int c = 0; c < CMAX; ++c) {
      const int sidx = lane * C + c;
      xs[c] = make_float3(0.f, 0.f, 0.f);
      if (c < C && sidx < S) {
        const int64_t b = r * a.bstride + sidx;
        float sel;
        xs[c] = nkbs_normalise(nkbs_frustum_position(o, d, a.starts[b], a.ends[b]), a.norm, sel);
      }
    }
    const float2* g_ray = reinterpret_cast<const float2*>(a.dfeat + (r * S) * (int64_t)(2 * L));
    for (int l = 0; l < L; ++l) {
      const float scale = s_scale[l];
      const bool v4 = scale >= a.v4_min_scale;
      float* base = a.dtable + ((size_t)l << (a.lv.log2_T + 1));
      long long* fixed = FX ? a.fm.at(base) : nullptr;
      uint32_t ck0 = 0xffffffffu, ck1 = 0xffffffffu;  // no open run
      Cell cc;
      float ax[8], ay[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) ax[k] = ay[k] = 0.f;
      auto flush = [&]() {
        uint32_t h[8];
        nkbs_corner_hashes(cc, mask, h);
        scatter_cell<FX>(base, cc, h, ax, ay, v4, fixed, a.fm.scale);
      };
#pragma unroll
      for (int c = 0; c < CMAX; ++c) {
        const int sidx = lane * C + c;
        if (c < C && sidx < S) {
          const float2 g = __ldg(g_ray + (int64_t)sidx * L + l);
          if (g.x != 0.f || g.y != 0.f) {
            const Cell cl = nkbs_cell(xs[c], scale);
            const uint32_t k0 = (cl.f[0] & 0xffffu) | (cl.f[1] << 16);
            const uint32_t k1 = (cl.f[2] & 0xffffu) | ((cl.c[0] - cl.f[0]) << 16) | ((cl.c[1] - cl.f[1]) << 17) |
                                ((cl.c[2] - cl.f[2]) << 18);
            if (k0 != ck0 || k1 != ck1) {
              if (ck1 != 0xffffffffu) flush();
              cc = cl;
              ck0 = k0;
              ck1 = k1;
#pragma unroll
              for (int k = 0; k < 8; ++k) ax[k] = ay[k] = 0.f;
            }
            float w[8];
            nkbs_corner_weights(cl.o[0], cl.o[1], cl.o[2], w);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              ax[k] = fmaf(w[k], g.x, ax[k]);
              ay[k] = fmaf(w[k], g.y, ay[k]);
            }
          }
        }
      }
      if (ck1 != 0xffffffffu) flush();
    }
  }
}

// A batch of fewer CTAs than two waves (148 SMs x 8 resident CTAs) leaves SMs idle or unevenly loaded: the level range
// is then cut into slices along gridDim.y, each CTA looping over L / slices levels of its 256 samples, so that coarse
// (shuffle-merged) and fine (L2-bound) levels also run side by side.  NKBS_ENCODE_SLICES overrides (a divisor of L).
int level_slices(int64_t blocks, int L) {
  static const int forced = [] {
    const char* e = getenv("NKBS_ENCODE_SLICES");
    return e ? atoi(e) : 0;
  }();
  int s = forced > 0 ? forced : (blocks < (int64_t)nkbs_num_sms() * 16 ? 4 : 1);
  if (s > L) s = L;
  while (s > 1 && L % s != 0) --s;
  return s;
}

}  // namespace

extern "C" int nkbs_encode_fwd(int mode, const float* x, const float* origins, const float* directions,
                               const float* starts, const float* ends, int32_t bin_stride, int64_t R, int32_t S, int contraction,
                               const float* aabb, const float* table, const float* scalings, int32_t L,
                               int32_t log2_T, int32_t F, float* xn, float* selector, float* features,
                               int64_t* indices, void* stream) {
  NKBS_CHECK_ARG(mode >= 0 && mode <= 2, "nkbs_encode_fwd: mode must be 0, 1 or 2 (got %d)", mode);
  NKBS_CHECK_ARG(F == 2, "nkbs_encode_fwd: features_per_level must be 2 (got %d)", F);
  NKBS_CHECK_ARG(R >= 0 && S >= 1, "nkbs_encode_fwd: bad sizes R=%lld S=%d", (long long)R, S);
  NKBS_CHECK_ARG(table && features && scalings, "nkbs_encode_fwd: null table/features/scalings");
  NKBS_CHECK_ARG((reinterpret_cast<uintptr_t>(table) & 15u) == 0, "nkbs_encode_fwd: table must be 16-byte aligned");
  if (mode == 2)
    NKBS_CHECK_ARG(origins && directions && starts && ends, "nkbs_encode_fwd: frustum mode needs origins/directions/starts/ends");
  else
    NKBS_CHECK_ARG(x != nullptr || R == 0, "nkbs_encode_fwd: null positions");
  EncodeArgs a;
  a.mode = mode;
  a.x = x;
  a.origins = origins;
  a.directions = directions;
  a.starts = starts;
  a.ends = ends;
  a.N = R * (int64_t)S;
  a.S = S;
  NKBS_CHECK_ARG(bin_stride == 0 || bin_stride >= S, "nkbs_encode_fwd: bin_stride %d < S", bin_stride);
  a.bstride = bin_stride ? bin_stride : S;
  a.norm.contraction = contraction;
  for (int i = 0; i < 3; ++i) {
    a.norm.aabb_min[i] = aabb ? aabb[i] : 0.f;
    a.norm.aabb_len[i] = aabb ? aabb[3 + i] - aabb[i] : 1.f;
  }
  NKBS_CHECK_ARG(fill_levels(a.lv, scalings, L, log2_T) == 0, "nkbs_encode_fwd: unsupported grid L=%d log2_T=%d", L, log2_T);
  a.table = reinterpret_cast<const float2*>(table);
  a.xn = xn;
  a.selector = selector;
  a.features = features;
  a.indices = indices;
  if (a.N == 0) return 0;
  int64_t blocks = (a.N + kEncThreads - 1) / kEncThreads;
  int64_t cap = (int64_t)nkbs_num_sms() * 32;
  int grid = (int)(blocks < cap ? blocks : cap);
  const dim3 grid2(grid, level_slices(blocks, L));
  if (indices)
    nkbs_launch(encode_fwd_kernel<true>, grid2, kEncThreads, 0, (cudaStream_t)stream, a);
  else
    nkbs_launch(encode_fwd_kernel<false>, grid2, kEncThreads, 0, (cudaStream_t)stream, a);
  NKBS_CHECK_LAUNCH("nkbs_encode_fwd");
  return 0;
}

extern "C" int nkbs_encode_bwd(const float* xn, const float* dfeat, int64_t N, const float* scalings, int32_t L,
                               int32_t log2_T, int32_t F, int32_t dedup_levels, float* dtable, void* stream) {
  NKBS_CHECK_ARG(F == 2, "nkbs_encode_bwd: features_per_level must be 2 (got %d)", F);
  NKBS_CHECK_ARG(N >= 0 && dtable && scalings, "nkbs_encode_bwd: bad arguments");
  NKBS_CHECK_ARG((reinterpret_cast<uintptr_t>(dtable) & 15u) == 0, "nkbs_encode_bwd: dtable must be 16-byte aligned");
  EncodeBwdArgs a;
  a.xn = xn;
  a.dfeat = dfeat;
  a.N = N;
  NKBS_CHECK_ARG(fill_levels(a.lv, scalings, L, log2_T) == 0, "nkbs_encode_bwd: unsupported grid L=%d log2_T=%d", L, log2_T);
  a.dtable = dtable;
  a.dedup_levels = dedup_levels < 0 ? L : dedup_levels;
  a.v4_min_scale = v4_min_scale(0.f);
  a.fm = nkbs_fixed_map_for(dtable);
  if (N == 0) return 0;
  NKBS_CHECK_ARG(xn && dfeat, "nkbs_encode_bwd: null xn/dfeat");
  int64_t blocks = (N + kEncThreads - 1) / kEncThreads;
  int64_t cap = (int64_t)nkbs_num_sms() * 32;
  int grid = (int)(blocks < cap ? blocks : cap);
  if (a.fm.shadow) nkbs_launch(encode_bwd_kernel<true>, dim3(grid, level_slices(blocks, L)), kEncThreads, 0, (cudaStream_t)stream, a);
  else nkbs_launch(encode_bwd_kernel<false>, dim3(grid, level_slices(blocks, L)), kEncThreads, 0, (cudaStream_t)stream, a);
  NKBS_CHECK_LAUNCH("nkbs_encode_bwd");
  return 0;
}

extern "C" int nkbs_encode_bwd_rays(int64_t R, int32_t S, const float* origins, const float* directions, const float* starts,
                                    const float* ends, int32_t bin_stride, int32_t contraction, const float* aabb,
                                    const float* dfeat, const float* scalings, int32_t L, int32_t log2_T, float* dtable,
                                    void* stream) {
  NKBS_CHECK_ARG(R >= 0 && S >= 1 && S <= 256, "nkbs_encode_bwd_rays: S=%d outside 1..256 (use nkbs_encode_bwd)", S);
  NKBS_CHECK_ARG(dtable && scalings, "nkbs_encode_bwd_rays: bad arguments");
  NKBS_CHECK_ARG((reinterpret_cast<uintptr_t>(dtable) & 15u) == 0, "nkbs_encode_bwd_rays: dtable must be 16-byte aligned");
  if (R == 0) return 0;
  NKBS_CHECK_ARG(origins && directions && starts && ends && dfeat, "nkbs_encode_bwd_rays: null argument");
  NKBS_CHECK_ARG((reinterpret_cast<uintptr_t>(dfeat) & 7u) == 0, "nkbs_encode_bwd_rays: dfeat must be 8-byte aligned");
  NKBS_CHECK_ARG(bin_stride == 0 || bin_stride >= S, "nkbs_encode_bwd_rays: bin_stride %d < S", bin_stride);
  RayScatterArgs a;
  a.mode = 2;
  a.x = nullptr;
  a.origins = origins;
  a.directions = directions;
  a.starts = starts;
  a.ends = ends;
  a.S = S;
  a.bstride = bin_stride ? bin_stride : S;
  a.norm.contraction = contraction;
  for (int i = 0; i < 3; ++i) {
    a.norm.aabb_min[i] = aabb ? aabb[i] : 0.f;
    a.norm.aabb_len[i] = aabb ? aabb[3 + i] - aabb[i] : 1.f;
  }
  a.R = R;
  NKBS_CHECK_ARG(fill_levels(a.lv, scalings, L, log2_T) == 0, "nkbs_encode_bwd_rays: unsupported grid L=%d log2_T=%d", L, log2_T);
  a.dfeat = dfeat;
  a.dtable = dtable;
  a.v4_min_scale = v4_min_scale(1e30f);
  a.fm = nkbs_fixed_map_for(dtable);
  const int64_t blocks = (R + kEncThreads / 32 - 1) / (kEncThreads / 32);
  const int64_t cap = (int64_t)nkbs_num_sms() * 8;
  const int grid = (int)(blocks < cap ? blocks : cap);
  cudaStream_t st = (cudaStream_t)stream;
  if (a.fm.shadow) {
    if (S <= 64) nkbs_launch(encode_bwd_rays_kernel<2, true>, grid, kEncThreads, 0, st, a);
    else if (S <= 128) nkbs_launch(encode_bwd_rays_kernel<4, true>, grid, kEncThreads, 0, st, a);
    else nkbs_launch(encode_bwd_rays_kernel<8, true>, grid, kEncThreads, 0, st, a);
  } else {
    if (S <= 64) nkbs_launch(encode_bwd_rays_kernel<2, false>, grid, kEncThreads, 0, st, a);
    else if (S <= 128) nkbs_launch(encode_bwd_rays_kernel<4, false>, grid, kEncThreads, 0, st, a);
    else nkbs_launch(encode_bwd_rays_kernel<8, false>, grid, kEncThreads, 0, st, a);
  }
  NKBS_CHECK_LAUNCH("nkbs_encode_bwd_rays");
  return 0;
}
