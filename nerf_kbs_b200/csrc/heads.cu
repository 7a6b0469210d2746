// K4: assembly of the head-MLP inputs: SH-16 of the view direction, geometry features and the
// per-image embedding row (appearance 32-d for the colour MLP, transient 16-d for the transient MLP),
// plus the backward split (geo gradient + embedding-row scatter).
// Replaces the cat([...]) + SHEncoding + Embedding lookups of NerfactoField.get_outputs reached from
// ref:nerfacto.py:398 and ref:semantic_nerfw.py:188.
#include "common.cuh"
#include "sh.cuh"

namespace {

constexpr int kThreads = 256;

struct AssembleArgs {
  int64_t N;
  int S;                    // samples per ray for per-ray inputs
  int sh_dim;               // 16 (colour input) or 0 (transient input)
  int geo_dim;              // 15
  int emb_dim;              // 32 or 16
  const float* directions;  // [R,3] if dir_per_ray else [N,3]; unused when sh_dim==0
  int dir_per_ray;
  const float* geo;         // [N,geo_dim]
  const int64_t* cam;       // [R] if cam_per_ray else [N]; null -> use emb_const
  int cam_per_ray;
  const float* emb;         // [num_images, emb_dim]
  const float* emb_const;   // [emb_dim] (eval: mean or zeros)
  int64_t num_images;
  float* h;                 // [N, sh_dim+geo_dim+emb_dim]
};

__global__ void __launch_bounds__(kThreads) assemble_fwd_kernel(const AssembleArgs a) {
  nkbs_pdl_sync();
  const int W = a.sh_dim + a.geo_dim + a.emb_dim;
  const int64_t total = a.N * W;
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t n = e / W;
    int c = (int)(e - n * W);
    float v;
    if (c < a.sh_dim) {
      int64_t di = a.dir_per_ray ? n / a.S : n;
      float x = (a.directions[3 * di] + 1.0f) / 2.0f;
      float y = (a.directions[3 * di + 1] + 1.0f) / 2.0f;
      float z = (a.directions[3 * di + 2] + 1.0f) / 2.0f;
      v = sh_component(c, x, y, z);
    } else if (c < a.sh_dim + a.geo_dim) {
      v = a.geo[n * a.geo_dim + (c - a.sh_dim)];
    } else {
      int j = c - a.sh_dim - a.geo_dim;
      if (a.cam) {
        int64_t ci = a.cam[a.cam_per_ray ? n / a.S : n];
        v = a.emb[ci * a.emb_dim + j];
      } else {
        v = a.emb_const[j];
      }
    }
    a.h[e] = v;
  }
}

struct AssembleBwdArgs {
  int64_t N;
  int S;
  int sh_dim, geo_dim, emb_dim;
  const float* dh;     // [N, W]
  float* dgeo;         // [N, geo_dim] or null
  const int64_t* cam;  // null -> no embedding gradient (eval)
  int cam_per_ray;
  float* demb;         // [num_images, emb_dim], accumulated into
  FixedMap fm;         // deterministic mode (common.cuh): shadow of the buffer that holds demb, or a null map
};

__global__ void __launch_bounds__(kThreads) assemble_bwd_geo_kernel(const AssembleBwdArgs a) {
  nkbs_pdl_sync();
  const int W = a.sh_dim + a.geo_dim + a.emb_dim;
  const int64_t total = a.N * a.geo_dim;
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t n = e / a.geo_dim;
    int c = (int)(e - n * a.geo_dim);
    a.dgeo[e] = a.dh[n * W + a.sh_dim + c];
  }
}

// one thread per (group, column): a group is the S samples of one ray when the camera index is per
// ray (they share the embedding row), so the row receives one reduction per ray instead of S.
__global__ void __launch_bounds__(kThreads) assemble_bwd_emb_kernel(const AssembleBwdArgs a) {
  nkbs_pdl_sync();
  const int W = a.sh_dim + a.geo_dim + a.emb_dim;
  const int G = a.cam_per_ray ? a.S : 1;
  const int64_t groups = a.N / G;
  const int64_t total = groups * a.emb_dim;
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t g = e / a.emb_dim;
    int j = (int)(e - g * a.emb_dim);
    const float* src = a.dh + (g * G) * W + a.sh_dim + a.geo_dim + j;
    float s = 0.f;
    for (int i = 0; i < G; ++i) s += src[(int64_t)i * W];
    int64_t ci = a.cam[g];
    if (s != 0.f) nkbs_acc(a.demb + ci * a.emb_dim + j, a.fm, s);
  }
}

int grid_for(int64_t total) {
  int64_t blocks = (total + kThreads - 1) / kThreads;
  int64_t cap = (int64_t)nkbs_num_sms() * 16;
  if (blocks < 1) blocks = 1;
  return (int)(blocks < cap ? blocks : cap);
}

}  // namespace

extern "C" int nkbs_assemble_fwd(int64_t N, int32_t S, int32_t sh_dim, int32_t geo_dim, int32_t emb_dim,
                                 const float* directions, int32_t dir_per_ray, const float* geo,
                                 const int64_t* cam, int32_t cam_per_ray, const float* emb, const float* emb_const,
                                 int64_t num_images, float* h, void* stream) {
  NKBS_CHECK_ARG(N >= 0 && S >= 1, "nkbs_assemble_fwd: bad sizes");
  NKBS_CHECK_ARG(sh_dim == 0 || sh_dim == 16, "nkbs_assemble_fwd: sh_dim must be 0 or 16 (got %d)", sh_dim);
  NKBS_CHECK_ARG(geo_dim >= 0 && emb_dim >= 0 && h, "nkbs_assemble_fwd: bad arguments");
  NKBS_CHECK_ARG(sh_dim == 0 || directions, "nkbs_assemble_fwd: null directions");
  NKBS_CHECK_ARG(emb_dim == 0 || (cam && emb) || emb_const, "nkbs_assemble_fwd: need camera indices + table, or a constant row");
  NKBS_CHECK_ARG(!(dir_per_ray || cam_per_ray) || N % S == 0, "nkbs_assemble_fwd: N not a multiple of S");
  if (N == 0) return 0;
  AssembleArgs a{N, S, sh_dim, geo_dim, emb_dim, directions, dir_per_ray, geo, cam, cam_per_ray, emb, emb_const, num_images, h};
  nkbs_launch(assemble_fwd_kernel, grid_for(N * (sh_dim + geo_dim + emb_dim)), kThreads, 0, (cudaStream_t)stream, a);
  NKBS_CHECK_LAUNCH("nkbs_assemble_fwd");
  return 0;
}

extern "C" int nkbs_assemble_bwd(int64_t N, int32_t S, int32_t sh_dim, int32_t geo_dim, int32_t emb_dim,
                                 const float* dh, float* dgeo, const int64_t* cam, int32_t cam_per_ray,
                                 float* demb, void* stream) {
  NKBS_CHECK_ARG(N >= 0 && S >= 1 && dh, "nkbs_assemble_bwd: bad arguments");
  if (N == 0) return 0;
  AssembleBwdArgs a{N, S, sh_dim, geo_dim, emb_dim, dh, dgeo, cam, cam_per_ray, demb, nkbs_fixed_map_for(demb)};
  if (dgeo && geo_dim > 0) {
    nkbs_launch(assemble_bwd_geo_kernel, grid_for(N * geo_dim), kThreads, 0, (cudaStream_t)stream, a);
    NKBS_CHECK_LAUNCH("nkbs_assemble_bwd(geo)");
  }
  if (cam && demb && emb_dim > 0) {
    NKBS_CHECK_ARG(!cam_per_ray || N % S == 0, "nkbs_assemble_bwd: N not a multiple of S");
    int64_t groups = cam_per_ray ? N / S : N;
    nkbs_launch(assemble_bwd_emb_kernel, grid_for(groups * emb_dim), kThreads, 0, (cudaStream_t)stream, a);
    NKBS_CHECK_LAUNCH("nkbs_assemble_bwd(emb)");
  }
  return 0;
}
