// Pieces shared by the encoding kernels (encode.cu) and the fused proposal-density kernels (density.cu):
// where a sample's position comes from, the warp-merged gradient scatter, and the level table.
#pragma once
#include <stdlib.h>
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

// Gradient of the 8 corner rows of one cell -> table gradient.  The x prime of the hash is 1, so the two corners of an x
// edge (ceil-x corner kc, floor-x corner kf) are rows h and h ^ 1 of one 16-byte aligned pair whenever floor(x) is even
// (as in the forward's paired gathers): one 16-byte vector reduction then carries both rows.  A quarter fewer L2
// reduction operations on average (the scatter is bound by their count).  NKBS_SCATTER_V4 = 0: always two 8-byte ones.
#ifndef NKBS_SCATTER_V4
#define NKBS_SCATTER_V4 1
#endif
// Deterministic mode (common.cuh, FixedMap): `fixed` != nullptr is the fixed-point shadow of this level's table gradient
// and takes the additions instead of the float reductions.
// FX is a compile-time switch so that the default kernels carry no trace of the mode.
template <bool FX = false>
__device__ __forceinline__ void scatter_cell(float* base, const Cell& cl, const uint32_t h[8], const float vx[8],
                                             const float vy[8], bool v4, long long* fixed = nullptr, float fscale = 0.f) {
  if constexpr (FX) {
    if (!fixed) {  // a destination outside the registered buffer
      scatter_cell<false>(base, cl, h, vx, vy, v4);
      return;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      nkbs_fixed_add(fixed + 2 * (size_t)h[k], vx[k], fscale);
      nkbs_fixed_add(fixed + 2 * (size_t)h[k] + 1, vy[k], fscale);
    }
    return;
  }
  constexpr int kC[4] = {0, 1, 4, 5}, kF[4] = {3, 2, 7, 6};  // x edges as (ceil-x corner, floor-x corner)
  const bool pair = NKBS_SCATTER_V4 && v4 && !(cl.f[0] & 1u) && cl.c[0] != cl.f[0];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int kc = kC[j], kf = kF[j];
    const bool nzc = vx[kc] != 0.f || vy[kc] != 0.f, nzf = vx[kf] != 0.f || vy[kf] != 0.f;
    if (pair && nzc && nzf) {
      const bool f_hi = h[kf] & 1u;  // the floor-x corner is the odd row of the pair
      float* p = base + 2 * (size_t)(h[kf] & ~1u);
      if (f_hi) nkbs_red_add_v4(p, vx[kc], vy[kc], vx[kf], vy[kf]);
      else nkbs_red_add_v4(p, vx[kf], vy[kf], vx[kc], vy[kc]);
    } else {
      if (nzc) nkbs_red_add_v2(base + 2 * (size_t)h[kc], vx[kc], vy[kc]);
      if (nzf) nkbs_red_add_v2(base + 2 * (size_t)h[kf], vx[kf], vy[kf]);
    }
  }
}

// Gradient scatter of one level with duplicate merging.  Adjacent lanes are adjacent samples of one ray, so the
// samples of a warp that fall into the same cell (all 8 corner rows equal) form RUNS of neighbouring lanes.  A run
// is summed with a segmented shuffle-down reduction (5 fixed steps for the 8 corners x 2 features), after which
// only the first lane of each run issues the 8 vector reductions: coarse levels drop from 32 to 1-4 reductions
// per corner and warp, at a cost that does not depend on the run length.  The whole warp must be converged.
template <bool FX = false>
__device__ __forceinline__ void scatter_level_runs(float* base, const Cell& cl, const uint32_t h[8], const float w[8],
                                                   float gx, float gy, int lane, bool v4, long long* fixed = nullptr,
                                                   float fscale = 0.f) {
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
  if (head) scatter_cell<FX>(base, cl, h, vx, vy, v4, fixed, fscale);
}

// Levels of at least this resolution pair their reductions (scatter_cell).  Pairing pays where the scatter is bound by
// the number of L2 reductions — the main grid, 48 samples per ray: 120 -> 102 us at 4096 rays, 1.55 -> 1.26 ms at
// 65 536 — and loses a little where run merging has already removed most of them and the kernel is bound by its own
// instructions (proposal grids at 256 samples per ray: 90 -> 103 us); profiles/r02_encode_bwd_v4.txt.  Hence on for
// nkbs_encode_bwd, off for the fused proposal field.  NKBS_V4_MIN_SCALE overrides both.
inline float v4_min_scale(float dflt) {
  static const float v = [] {
    const char* e = getenv("NKBS_V4_MIN_SCALE");
    return e ? (float)atof(e) : -1.f;
  }();
  return v >= 0.f ? v : dflt;
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

