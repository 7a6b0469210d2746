// Pieces shared by the encoding kernels (encode.cu) and the fused proposal-density kernels (density.cu):
// where a sample's position comes from, the warp-merged gradient scatter, and the level table.
#pragma once
#include "common.cuh"

struct SampleSource {
  int mode;  // 0: x normalised, 1: x raw positions, 2: frustum
  const float* x;
  const float* origins;
  const float* directions;
  const float* starts;
  const float* ends;
  int S;
  int bstride;  // floats between consecutive rays in starts / ends (S, or S+1 for views of an [R,S+1] bin array)
  NormParams norm;
};

__device__ __forceinline__ float3 load_sample_position(const SampleSource& a, int64_t n, float& sel) {
  float3 p;
  if (a.mode == 2) {
    int64_t r = n / a.S;
    float3 o = make_float3(a.origins[3 * r], a.origins[3 * r + 1], a.origins[3 * r + 2]);
    float3 d = make_float3(a.directions[3 * r], a.directions[3 * r + 1], a.directions[3 * r + 2]);
    const int64_t b = r * a.bstride + (n - r * a.S);
    p = nkbs_frustum_position(o, d, a.starts[b], a.ends[b]);
  } else {
    p = make_float3(a.x[3 * n], a.x[3 * n + 1], a.x[3 * n + 2]);
  }
  if (a.mode == 0) {
    sel = 1.f;
    return p;
  }
  return nkbs_normalise(p, a.norm, sel);
}

// Gradient scatter of one level with duplicate merging.  Adjacent lanes are adjacent samples of one ray, so the
// samples of a warp that fall into the same cell (all 8 corner rows equal) form RUNS of neighbouring lanes.  A run
// is summed with a segmented shuffle-down reduction (5 fixed steps for the 8 corners x 2 features), after which
// only the first lane of each run issues the 8 vector reductions: coarse levels drop from 32 to 1-4 reductions
// per corner and warp, at a cost that does not depend on the run length.  The whole warp must be converged.
__device__ __forceinline__ void scatter_level_runs(float* base, const Cell& cl, const uint32_t h[8], const float w[8],
                                                   float gx, float gy, int lane) {
  // cell key: floor coordinates and the ceil-floor bits (scale < 65536 is checked on the host)
  const uint32_t k0 = (cl.f[0] & 0xffffu) | (cl.f[1] << 16);
  const uint32_t k1 = (cl.f[2] & 0xffffu) | ((cl.c[0] - cl.f[0]) << 16) | ((cl.c[1] - cl.f[1]) << 17) |
                      ((cl.c[2] - cl.f[2]) << 18);
  const uint32_t p0 = __shfl_up_sync(0xffffffffu, k0, 1), p1 = __shfl_up_sync(0xffffffffu, k1, 1);
  const bool head = lane == 0 || p0 != k0 || p1 != k1;
  const uint32_t heads = __ballot_sync(0xffffffffu, head);
  const uint32_t after = heads & ~((2u << lane) - 1u);  // run heads strictly above this lane
  const int end = after ? (__ffs(after) - 2) : 31;      // last lane of this lane's run
  float vx[8], vy[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    vx[k] = w[k] * gx;
    vy[k] = w[k] * gy;
  }
  if (heads != 0xffffffffu) {  // some run is longer than one lane
    // start of this lane's run = highest head at or below the lane; the longest run of the warp bounds the
    // number of reduction steps (fine levels hold runs of 2-4 samples: 1-2 steps instead of 5)
    const int start = 31 - __clz(heads & ((2u << lane) - 1u));
    const int longest = __reduce_max_sync(0xffffffffu, end - start + 1);
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      if (off >= longest) break;
      const bool take = lane + off <= end;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float ox = __shfl_down_sync(0xffffffffu, vx[k], off);
        const float oy = __shfl_down_sync(0xffffffffu, vy[k], off);
        if (take) {
          vx[k] += ox;
          vy[k] += oy;
        }
      }
    }
  }
  if (head) {
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (vx[k] != 0.f || vy[k] != 0.f) nkbs_red_add_v2(base + 2 * (size_t)h[k], vx[k], vy[k]);
  }
}

inline int fill_levels(LevelParams& lv, const float* scalings, int L, int log2_T) {
  if (L < 1 || L > NKBS_MAX_LEVELS) return -1;
  if (log2_T < 1 || log2_T > 30) return -1;
  for (int i = 0; i < L; ++i)
    if (!(scalings[i] < 65536.f)) return -1;  // cell keys of scatter_level_runs hold 16-bit coordinates
  for (int i = 0; i < NKBS_MAX_LEVELS; ++i) lv.scale[i] = i < L ? scalings[i] : 0.f;
  lv.num_levels = L;
  lv.log2_T = (uint32_t)log2_T;
  lv.table_mask = (1u << log2_T) - 1u;
  return 0;
}

