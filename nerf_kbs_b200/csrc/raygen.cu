// N3/N4: the data-side and image-side neighbours of the ray-batch path, on the device:
//   nkbs_generate_rays      pinhole rays (origins, unit directions, directions_norm, pixel_area) from (camera,row,col)
//   nkbs_compact_masked     stable compaction of candidate pixels whose motion mask is non-zero to exactly R rays
//   nkbs_color_to_label     RGB semantic colour -> class label through a palette (replaces a per-pixel Python loop)
//   nkbs_semantic_colormap  argmax over class logits -> palette colour (the rendered semantic image)
//   nkbs_apply_colormap     scalar image in [0,1] -> 256-entry colour table (accumulation image)
//   nkbs_apply_depth_colormap  depth normalised by its own (or given) range -> colour table, blended over white by accumulation
// Replaces nerfstudio RayGenerator / Cameras.generate_rays as driven by ref:dataparser/datamanager_z.py:96,252-278,
// the masked sampling of ref:dataparser/dataset_z.py:180-191, the color2label loop of ref:nerfacto.py:508-512 and
// the argmax + colormap of ref:nerfacto.py:574-578 / ref:semantic_nerfw.py:346-351, and nerfstudio
// utils.colormaps.apply_colormap / apply_depth_colormap as called at ref:nerfacto.py:574-578,608-613 and
// ref:semantic_nerfw.py:305.
#include <float.h>
#include "common.cuh"

namespace {

constexpr int kThreads = 256;

struct RayGenArgs {
  int64_t N;
  const int64_t* idx;     // [N,3] camera, row, col
  const float* c2w;       // [num_cameras,3,4]
  const float* intr;      // [num_cameras,4] fx, fy, cx, cy
  int64_t num_cameras;
  float* origins;         // [N,3]
  float* directions;      // [N,3] unit
  float* dnorm;           // [N] or null
  float* pixel_area;      // [N] or null
  int64_t* cam_out;       // [N] or null
};

__device__ __forceinline__ float3 rotate(const float* R, float x, float y, float z) {
  // sum(dir[None, :] * R[3x3 of the 3x4], -1): row i = R[i,0] x + R[i,1] y + R[i,2] z, left to right
  float3 o;
  o.x = __fadd_rn(__fadd_rn(__fmul_rn(x, R[0]), __fmul_rn(y, R[1])), __fmul_rn(z, R[2]));
  o.y = __fadd_rn(__fadd_rn(__fmul_rn(x, R[4]), __fmul_rn(y, R[5])), __fmul_rn(z, R[6]));
  o.z = __fadd_rn(__fadd_rn(__fmul_rn(x, R[8]), __fmul_rn(y, R[9])), __fmul_rn(z, R[10]));
  return o;
}
__device__ __forceinline__ float norm3(float3 v) {
  return sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(v.x, v.x), __fmul_rn(v.y, v.y)), __fmul_rn(v.z, v.z)));
}

__global__ void __launch_bounds__(kThreads) generate_rays_kernel(const RayGenArgs a) {
  nkbs_pdl_sync();
  for (int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; n < a.N; n += (int64_t)gridDim.x * blockDim.x) {
    int64_t cam = a.idx[3 * n];
    cam = cam < 0 ? 0 : (cam >= a.num_cameras ? a.num_cameras - 1 : cam);
    const float py = (float)a.idx[3 * n + 1] + 0.5f, px = (float)a.idx[3 * n + 2] + 0.5f;
    const float fx = a.intr[4 * cam], fy = a.intr[4 * cam + 1], cx = a.intr[4 * cam + 2], cy = a.intr[4 * cam + 3];
    const float* M = a.c2w + 12 * cam;
    // camera looks down -z, y up (nerfstudio / OpenGL convention); neighbours one pixel right and one pixel down
    const float x0 = (px - cx) / fx, y0 = (py - cy) / fy;
    const float x1 = (px + 1.f - cx) / fx, y1 = (py + 1.f - cy) / fy;
    float3 d = rotate(M, x0, -y0, -1.f);
    float3 dx = rotate(M, x1, -y0, -1.f);
    float3 dy = rotate(M, x0, -y1, -1.f);
    const float n0 = norm3(d), n1 = norm3(dx), n2 = norm3(dy);
    d = make_float3(d.x / n0, d.y / n0, d.z / n0);
    dx = make_float3(dx.x / n1, dx.y / n1, dx.z / n1);
    dy = make_float3(dy.x / n2, dy.y / n2, dy.z / n2);
    a.origins[3 * n] = M[3];
    a.origins[3 * n + 1] = M[7];
    a.origins[3 * n + 2] = M[11];
    a.directions[3 * n] = d.x;
    a.directions[3 * n + 1] = d.y;
    a.directions[3 * n + 2] = d.z;
    if (a.dnorm) a.dnorm[n] = n0;
    if (a.pixel_area) {
      const float ax = norm3(make_float3(dx.x - d.x, dx.y - d.y, dx.z - d.z));
      const float ay = norm3(make_float3(dy.x - d.x, dy.y - d.y, dy.z - d.z));
      a.pixel_area[n] = ax * ay;
    }
    if (a.cam_out) a.cam_out[n] = cam;
  }
}

// Stable compaction by ONE CTA (candidate lists are ~R long, a few thousand): out[k] = k-th candidate whose mask
// byte is non-zero; if fewer than R survive the survivors repeat cyclically (count returned in *n_kept).
__global__ void __launch_bounds__(1024) compact_masked_kernel(int64_t M, const int64_t* img, const int64_t* pix,
                                                              const uint8_t* mask, int64_t pixels_per_image, int64_t R,
                                                              int64_t* out_img, int64_t* out_pix, int32_t* n_kept) {
  nkbs_pdl_sync();
  __shared__ int s_warp[32];
  __shared__ int s_base, s_total;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  for (int64_t m0 = 0; m0 < M; m0 += blockDim.x) {
    const int64_t m = m0 + threadIdx.x;
    bool keep = false;
    if (m < M) keep = mask[img[m] * pixels_per_image + pix[m]] != 0;
    const unsigned b = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_warp[wid] = __popc(b);
    __syncthreads();
    if (wid == 0) {  // exclusive scan of the 32 warp counts
      const int v = s_warp[lane];
      int incl = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      s_warp[lane] = incl - v;
      if (lane == 31) s_total = incl;
    }
    __syncthreads();
    const int base = s_base;
    const int pos = base + s_warp[wid] + __popc(b & ((1u << lane) - 1u));
    if (keep && pos < R) {
      out_img[pos] = img[m];
      out_pix[pos] = pix[m];
    }
    __syncthreads();
    if (threadIdx.x == 0) s_base = base + s_total;
    __syncthreads();
  }
  const int kept = s_base < R ? s_base : (int)R;
  if (threadIdx.x == 0) *n_kept = s_base;
  __syncthreads();  // the survivors are in global memory before they are repeated
  if (kept > 0 && kept < R) {
    for (int64_t k = kept + threadIdx.x; k < R; k += blockDim.x) {
      out_img[k] = out_img[k % kept];
      out_pix[k] = out_pix[k % kept];
    }
  }
}

// ---- the same compaction for candidate lists of any length (the reference loads pixel chunks of 12.8 M,
// ref:dataparser/datamanager_z.py:49): per-CTA counts, one scan of the counts, ordered scatter, cyclic fill ----
constexpr int kCompactCta = 1024;

__device__ __forceinline__ int block_exclusive_scan_1024(int v, int* s_warp, int& total) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_warp[wid] = incl;
  __syncthreads();
  if (wid == 0) {
    const int w = s_warp[lane];
    int wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += t;
    }
    s_warp[lane] = wi - w;
    if (lane == 31) s_warp[32] = wi;
  }
  __syncthreads();
  total = s_warp[32];
  const int r = s_warp[wid] + incl - v;
  __syncthreads();
  return r;
}

__global__ void __launch_bounds__(kCompactCta) compact_count_kernel(int64_t M, const int64_t* img, const int64_t* pix,
                                                                    const uint8_t* mask, int64_t ppi, int32_t* counts) {
  nkbs_pdl_sync();
  __shared__ int s_warp[33];
  const int64_t m = blockIdx.x * (int64_t)kCompactCta + threadIdx.x;
  const int keep = (m < M && mask[img[m] * ppi + pix[m]] != 0) ? 1 : 0;
  int total;
  block_exclusive_scan_1024(keep, s_warp, total);
  if (threadIdx.x == 0) counts[blockIdx.x] = total;
}

// exclusive scan of counts[0..nb) in place, counts[nb] = *n_kept = total
__global__ void __launch_bounds__(kCompactCta) compact_scan_kernel(int nb, int32_t* counts, int32_t* n_kept) {
  nkbs_pdl_sync();
  __shared__ int s_warp[33];
  int carry = 0;
  for (int b0 = 0; b0 < nb; b0 += kCompactCta) {
    const int b = b0 + threadIdx.x;
    const int v = b < nb ? counts[b] : 0;
    int total;
    const int ex = block_exclusive_scan_1024(v, s_warp, total);
    if (b < nb) counts[b] = carry + ex;
    carry += total;
  }
  if (threadIdx.x == 0) {
    counts[nb] = carry;
    *n_kept = carry;
  }
}

__global__ void __launch_bounds__(kCompactCta) compact_scatter_kernel(int64_t M, const int64_t* img, const int64_t* pix,
                                                                      const uint8_t* mask, int64_t ppi, int64_t R,
                                                                      const int32_t* offsets, int64_t* out_img,
                                                                      int64_t* out_pix) {
  nkbs_pdl_sync();
  __shared__ int s_warp[33];
  const int64_t m = blockIdx.x * (int64_t)kCompactCta + threadIdx.x;
  int64_t im = 0, px = 0;
  int keep = 0;
  if (m < M) {
    im = img[m];
    px = pix[m];
    keep = mask[im * ppi + px] != 0;
  }
  int total;
  const int64_t pos = (int64_t)offsets[blockIdx.x] + block_exclusive_scan_1024(keep, s_warp, total);
  if (keep && pos < R) {
    out_img[pos] = im;
    out_pix[pos] = px;
  }
}

__global__ void __launch_bounds__(kThreads) compact_fill_kernel(int64_t R, const int32_t* n_kept, int64_t* out_img,
                                                                int64_t* out_pix) {
  nkbs_pdl_sync();
  const int64_t kept = *n_kept;
  if (kept <= 0 || kept >= R) return;
  for (int64_t k = kept + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < R; k += (int64_t)gridDim.x * blockDim.x) {
    out_img[k] = out_img[k % kept];  // k % kept < kept: written by the scatter kernel, never by this one
    out_pix[k] = out_pix[k % kept];
  }
}

constexpr int kMaxPalette = 256;
struct Palette {
  int n;
  uint32_t key[kMaxPalette];  // r | g<<8 | b<<16
};

__global__ void __launch_bounds__(kThreads) color_to_label_kernel(int64_t N, const uint8_t* rgb, const Palette* pal,
                                                                  int64_t default_label, int64_t* labels) {
  nkbs_pdl_sync();
  __shared__ uint32_t s_key[kMaxPalette];
  const int n = pal->n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s_key[i] = pal->key[i];
  __syncthreads();
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < N; p += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t k = rgb[3 * p] | ((uint32_t)rgb[3 * p + 1] << 8) | ((uint32_t)rgb[3 * p + 2] << 16);
    int64_t lab = default_label;
    for (int i = 0; i < n; ++i)
      if (s_key[i] == k) {
        lab = i;
        break;
      }
    labels[p] = lab;
  }
}

__global__ void __launch_bounds__(kThreads) semantic_colormap_kernel(int64_t N, int C, const float* logits,
                                                                     const Palette* pal, int64_t* labels, uint8_t* rgb) {
  nkbs_pdl_sync();
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < N; p += (int64_t)gridDim.x * blockDim.x) {
    const float* x = logits + p * C;
    int best = 0;
    float bv = x[0];
    for (int c = 1; c < C; ++c) {  // first maximum wins, as torch.argmax
      const float v = x[c];
      if (v > bv || (v != v && !(bv != bv))) {
        bv = v;
        best = c;
      }
    }
    if (labels) labels[p] = best;
    if (rgb) {
      const uint32_t k = best < pal->n ? pal->key[best] : 0u;
      rgb[3 * p] = k & 0xff;
      rgb[3 * p + 1] = (k >> 8) & 0xff;
      rgb[3 * p + 2] = (k >> 16) & 0xff;
    }
  }
}

// ---- scalar colour maps (nerfstudio utils/colormaps.py) ----
__device__ __forceinline__ int cm_float_key(float f) {
  int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float cm_key_float(int k) { return __int_as_float(k >= 0 ? k : k ^ 0x7fffffff); }

__global__ void cm_minmax_init_kernel(int* mm) {
  nkbs_pdl_sync();
  mm[0] = cm_float_key(FLT_MAX);
  mm[1] = cm_float_key(-FLT_MAX);
}

__global__ void __launch_bounds__(kThreads) cm_minmax_kernel(int64_t N, const float* x, int* mm) {
  nkbs_pdl_sync();
  float lo = FLT_MAX, hi = -FLT_MAX;
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < N; p += (int64_t)gridDim.x * blockDim.x) {
    const float v = x[p];
    lo = fminf(lo, v);
    hi = fmaxf(hi, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if ((threadIdx.x & 31) == 0 && lo <= hi) {
    atomicMin(mm, cm_float_key(lo));
    atomicMax(mm + 1, cm_float_key(hi));
  }
}

// colour = lut[(long)(v * 255)]   (torch: (image * 255).long() truncates toward zero; nerfstudio asserts the index
// range 0..255, here it is clamped).  depth mode: v = clip((d - near) / (far - near + 1e-10), 0, 1) with near / far
// python floats in the reference: the subtraction runs in fp32 against fp32(near), the divisor is
// fp32(double(far) - double(near) + 1e-10), the division is a true fp32 division (CPU torch semantics); then
// colour * acc + (1 - acc).
__global__ void __launch_bounds__(kThreads) colormap_kernel(int64_t N, const float* x, const float* acc, const float* lut,
                                                            int depth_mode, const int* mm, float near_in, float far_in,
                                                            int have_near, int have_far, float* out) {
  nkbs_pdl_sync();
  float near = 0.f, den = 1.f;
  if (depth_mode) {
    near = have_near ? near_in : cm_key_float(mm[0]);
    const float far = have_far ? far_in : cm_key_float(mm[1]);
    den = (float)((double)far - (double)near + 1e-10);
  }
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < N; p += (int64_t)gridDim.x * blockDim.x) {
    float v = x[p];
    if (depth_mode) {
      v = __fdiv_rn(__fsub_rn(v, near), den);
      v = fminf(fmaxf(v, 0.f), 1.f);  // torch.clip: NaN propagates
      if (v != v) v = 0.f;
    }
    long long idx = (long long)__fmul_rn(v, 255.f);
    idx = idx < 0 ? 0 : (idx > 255 ? 255 : idx);
    float r = lut[3 * idx], g = lut[3 * idx + 1], b = lut[3 * idx + 2];
    if (acc) {
      const float a = acc[p];
      const float w = __fsub_rn(1.f, a);
      r = __fadd_rn(__fmul_rn(r, a), w);
      g = __fadd_rn(__fmul_rn(g, a), w);
      b = __fadd_rn(__fmul_rn(b, a), w);
    }
    out[3 * p] = r;
    out[3 * p + 1] = g;
    out[3 * p + 2] = b;
  }
}

int grid_for(int64_t n) {
  int64_t blocks = (n + kThreads - 1) / kThreads;
  int64_t cap = (int64_t)nkbs_num_sms() * 8;
  if (blocks < 1) blocks = 1;
  return (int)(blocks < cap ? blocks : cap);
}

}  // namespace

extern "C" int nkbs_generate_rays(int64_t N, const int64_t* ray_indices, const float* c2w, const float* intrinsics,
                                  int64_t num_cameras, float* origins, float* directions, float* directions_norm,
                                  float* pixel_area, int64_t* camera_indices, void* stream) {
  NKBS_CHECK_ARG(N >= 0 && num_cameras >= 1, "nkbs_generate_rays: bad sizes");
  if (N == 0) return 0;
  NKBS_CHECK_ARG(ray_indices && c2w && intrinsics && origins && directions, "nkbs_generate_rays: null pointers");
  RayGenArgs a{N, ray_indices, c2w, intrinsics, num_cameras, origins, directions, directions_norm, pixel_area, camera_indices};
  nkbs_launch(generate_rays_kernel, grid_for(N), kThreads, 0, (cudaStream_t)stream, a);
  NKBS_CHECK_LAUNCH("nkbs_generate_rays");
  return 0;
}

extern "C" int nkbs_compact_masked(int64_t M, const int64_t* image_indices, const int64_t* pixel_indices,
                                   const uint8_t* masks, int64_t pixels_per_image, int64_t R, int64_t* out_image,
                                   int64_t* out_pixel, int32_t* n_kept, int32_t* block_ws, void* stream) {
  NKBS_CHECK_ARG(M >= 0 && R >= 0 && pixels_per_image >= 1, "nkbs_compact_masked: bad sizes");
  NKBS_CHECK_ARG(image_indices && pixel_indices && masks && out_image && out_pixel && n_kept, "nkbs_compact_masked: null pointers");
  cudaStream_t st = (cudaStream_t)stream;
  if (!block_ws) {  // one CTA walks the list: per-step candidate lists of a few thousand entries
    NKBS_CHECK_ARG(M <= (1ll << 22), "nkbs_compact_masked: %lld candidates need the block workspace", (long long)M);
    nkbs_launch(compact_masked_kernel, 1, 1024, 0, st, M, image_indices, pixel_indices, masks, pixels_per_image, R, out_image,
                                              out_pixel, n_kept);
    NKBS_CHECK_LAUNCH("nkbs_compact_masked");
    return 0;
  }
  const int64_t nb = (M + kCompactCta - 1) / kCompactCta;
  NKBS_CHECK_ARG(nb < (1ll << 30), "nkbs_compact_masked: too many candidates");
  if (nb > 0) {
    nkbs_launch(compact_count_kernel, (int)nb, kCompactCta, 0, st, M, image_indices, pixel_indices, masks, pixels_per_image, block_ws);
    NKBS_CHECK_LAUNCH("nkbs_compact_masked(count)");
  }
  nkbs_launch(compact_scan_kernel, 1, kCompactCta, 0, st, (int)nb, block_ws, n_kept);
  NKBS_CHECK_LAUNCH("nkbs_compact_masked(scan)");
  if (nb > 0) {
    nkbs_launch(compact_scatter_kernel, (int)nb, kCompactCta, 0, st, M, image_indices, pixel_indices, masks, pixels_per_image, R,
                                                           block_ws, out_image, out_pixel);
    NKBS_CHECK_LAUNCH("nkbs_compact_masked(scatter)");
  }
  if (R > 0) {
    nkbs_launch(compact_fill_kernel, grid_for(R), kThreads, 0, st, R, n_kept, out_image, out_pixel);
    NKBS_CHECK_LAUNCH("nkbs_compact_masked(fill)");
  }
  return 0;
}

extern "C" int nkbs_color_to_label(int64_t N, const uint8_t* rgb, const void* palette, int64_t default_label,
                                   int64_t* labels, void* stream) {
  NKBS_CHECK_ARG(N >= 0, "nkbs_color_to_label: negative N");
  if (N == 0) return 0;
  NKBS_CHECK_ARG(rgb && palette && labels, "nkbs_color_to_label: null pointers");
  nkbs_launch(color_to_label_kernel, grid_for(N), kThreads, 0, (cudaStream_t)stream, N, rgb, (const Palette*)palette, default_label, labels);
  NKBS_CHECK_LAUNCH("nkbs_color_to_label");
  return 0;
}

extern "C" int nkbs_semantic_colormap(int64_t N, int32_t C, const float* logits, const void* palette, int64_t* labels,
                                      uint8_t* rgb, void* stream) {
  NKBS_CHECK_ARG(N >= 0 && C >= 1, "nkbs_semantic_colormap: bad sizes");
  if (N == 0) return 0;
  NKBS_CHECK_ARG(logits && (labels || rgb) && (!rgb || palette), "nkbs_semantic_colormap: null pointers");
  nkbs_launch(semantic_colormap_kernel, grid_for(N), kThreads, 0, (cudaStream_t)stream, N, C, logits, (const Palette*)palette, labels, rgb);
  NKBS_CHECK_LAUNCH("nkbs_semantic_colormap");
  return 0;
}

extern "C" int nkbs_apply_colormap(int64_t N, const float* image, const float* lut, float* rgb, void* stream) {
  NKBS_CHECK_ARG(N >= 0, "nkbs_apply_colormap: negative N");
  if (N == 0) return 0;
  NKBS_CHECK_ARG(image && lut && rgb, "nkbs_apply_colormap: null pointers");
  nkbs_launch(colormap_kernel, grid_for(N), kThreads, 0, (cudaStream_t)stream, N, image, nullptr, lut, 0, nullptr, 0.f, 0.f, 0, 0, rgb);
  NKBS_CHECK_LAUNCH("nkbs_apply_colormap");
  return 0;
}

extern "C" int nkbs_apply_depth_colormap(int64_t N, const float* depth, const float* accumulation, const float* lut,
                                         int32_t have_near, float near_plane, int32_t have_far, float far_plane,
                                         int32_t* minmax_ws, float* rgb, void* stream) {
  NKBS_CHECK_ARG(N >= 0, "nkbs_apply_depth_colormap: negative N");
  if (N == 0) return 0;
  NKBS_CHECK_ARG(depth && lut && rgb, "nkbs_apply_depth_colormap: null pointers");
  NKBS_CHECK_ARG((have_near && have_far) || minmax_ws, "nkbs_apply_depth_colormap: the depth range needs a workspace");
  cudaStream_t st = (cudaStream_t)stream;
  if (!(have_near && have_far)) {
    nkbs_launch(cm_minmax_init_kernel, 1, 1, 0, st, minmax_ws);
    nkbs_launch(cm_minmax_kernel, grid_for(N), kThreads, 0, st, N, depth, minmax_ws);
    NKBS_CHECK_LAUNCH("nkbs_apply_depth_colormap(range)");
  }
  nkbs_launch(colormap_kernel, grid_for(N), kThreads, 0, st, N, depth, accumulation, lut, 1, minmax_ws, near_plane, far_plane,
                                                    have_near, have_far, rgb);
  NKBS_CHECK_LAUNCH("nkbs_apply_depth_colormap");
  return 0;
}
