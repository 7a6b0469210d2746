// K3 (tensor-core path): fused small MLP, forward and backward, on tcgen05 tensor cores.
//
// A CTA is nwg epilogue warpgroups (128 threads each: thread t <-> sample row t <-> TMEM lane t of the tile the
// warpgroup currently owns) plus one issuer warp whose elected lane issues every tcgen05.mma of the CTA.  The
// warpgroups and the issuer talk through mbarriers only (operands ready -> MMA batch -> tcgen05.commit), so the
// tensor pipe works on one warpgroup's tile while the others run their epilogues.  Weights of
// every layer stay in shared memory for the whole kernel as 16-bit UMMA operand tiles; activations bounce
// TMEM -> registers (bias is folded into the GEMM through a constant-one input column, ReLU and the 16-bit
// conversion happen in registers) -> shared memory, where they are the A operand of the next layer.
// The backward kernel recomputes the forward, then per layer issues
//     dX_l  = dY_l   . W_l        (A = dY_l K-major,            B = W_l viewed MN-major)
//     dW_l += dY_l^T . [X_l | 1]  (A = dY_l viewed MN-major, M=64, B = X_l viewed MN-major; K = 128 samples)
// with the weight/bias gradient accumulated in TMEM across all tiles of the CTA and flushed once.
// Operand tiles use the chunk-major no-swizzle canonical layout of umma.cuh, so no transposed copies exist.
// fmt 0 = fp16 operands, 1 = bf16 operands; accumulation is always fp32.
// Replaces nerfstudio field_components.mlp.MLP (+ the field-head activations) reached from
// ref:nerfacto.py:275-311,398 and ref:semantic_nerfw.py:102-126,188 (under autocast these linears run in
// 16-bit on the reference side too, ref:train.py:50).
#include <stdlib.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include "common.cuh"
#include "sh.cuh"
#include "umma.cuh"

namespace {

using namespace umma;

constexpr int kTile = 128;
constexpr int kThreads = 128;
constexpr int kMaxLayers = 3;
constexpr int kMaxWidth = 64;
constexpr int kActCs = kTile * 16 + 16;  // chunk stride (bytes) of a 128-row activation tile

enum OutAct { ACT_NONE = 0, ACT_SIGMOID = 1, ACT_DENSITY = 2, ACT_TRANSIENT = 3 };

struct TcLayer {
  const float* W;  // [N,K] torch nn.Linear layout
  const float* b;  // [N]
  float* dW;
  float* db;
  int K, N;
  int relu;   // ReLU after this layer
  int Ka;     // K + 1 (bias column) rounded up to 16
  int Np;     // N rounded up to 16
  int w_off;  // byte offset of the weight tile in shared memory
  int w_cs;   // chunk stride (bytes) of the weight tile = Np*16 + 16
  int x_off;  // byte offset of this layer's input tile X_l
  int dw_col; // first TMEM column of this layer's dW accumulator (backward)
};

// Input built in the kernel instead of read as rows (IN = 1, 2): X0 = [SH16((d+1)/2) | geo15 | embedding row | 1],
// the colour head's input (IN = 1: SH 16, appearance 32), or [geo15 | embedding row | 1], the transient head's
// (IN = 2: no SH, transient embedding 16).  Replaces K4 (heads.cu) on the tensor-core path: the [N,63] tensor and
// its gradient never exist in HBM.  IN = 3: [geo15 | 1], the semantics head's input (no embedding).
struct HeadIn {
  const float* geo;        // row n at geo + n * geo_stride (15 floats)
  int geo_stride;
  const float* dirs;       // [R,3] if dir_per_ray else [N,3]
  int dir_per_ray;
  const int64_t* cam;      // [R] if cam_per_ray else [N]; null -> emb_const
  int cam_per_ray;
  const float* emb;        // [num_images, emb_dim]
  const float* emb_const;  // [emb_dim]
  int S;                   // samples per ray
  int geo_vec;             // rows can be fetched as four aligned 16-byte loads starting one float before geo
  int emb_vec;             // embedding rows (and the constant row) are 16-byte aligned
  float* dgeo;             // backward: row n at dgeo + n * dgeo_stride, or null
  int dgeo_stride;
  int dgeo_lead;           // backward: this many floats in front of every dgeo row are set to zero (the density column)
  float* demb;             // backward: [num_images, emb_dim], accumulated into, or null
};

struct TcArgs {
  HeadIn hin;
  int nl;
  TcLayer layer[kMaxLayers];
  const float* x;         // [N,K0]
  const float* selector;  // [N] or null
  float* y;               // [N,Kn]   (forward)
  float* y0;              // [N] or null: column 0 of y once more, contiguous (the density of the base MLP)
  const float* dy;        // [N,Kn]   (backward)
  const float* dy0;       // [N] or null: added to column 0 of dy (gradient of y0)
  float* dx;              // [N,K0] or null (backward)
  FixedMap fm;            // backward, deterministic mode (common.cuh): shadow of the buffer that holds dW / db, or a null map
  FixedMap fm_emb;        // the same for the buffer that holds hin.demb
  int out_act;
  int fmt;                // 0 fp16, 1 bf16
  int64_t N;
  int x_off1;             // forward: second ping-pong activation buffer   (offsets x_off/x_off1/g_off are
  int g_off;              // backward: gradient tile (64 features wide)    relative to the warpgroup's region)
  int dy_off;             // backward: fp32 staging of the upstream-gradient tile [128 x Kn] (cp.async target)
  int stage_off;          // backward: fp32 staging used to write dx coalesced (aliases dead activation tiles)
  int stage_ok;           // 1 if the dx staging fits
  int ystage_ok;          // forward: the output tile is staged in shared memory and stored coalesced
  int recycle;            // backward, three layers: X_2 lives where X_0 was and X_0 is rebuilt before the last batch
  int wg_base;            // byte offset of warpgroup 0's activation region
  int wg_stride;          // bytes per warpgroup region
  int nwg;                // epilogue warpgroups per CTA (each owns one 128-sample tile at a time)
  int tmem_cols;          // power of two >= 32
  int smem_bytes;
};

template <int FMT>
__device__ __forceinline__ uint16_t cvt16(float v) {
  if (FMT == 1) {
    __nv_bfloat16 h = __float2bfloat16_rn(v);
    return *reinterpret_cast<uint16_t*>(&h);
  } else {
    __half h = __float2half_rn(fminf(fmaxf(v, -65504.f), 65504.f));
    return *reinterpret_cast<uint16_t*>(&h);
  }
}
template <int FMT>
__device__ __forceinline__ uint32_t pack2(float a, float b) {  // one F2FP: two floats -> packed 16-bit pair (a low),
  uint32_t d;                                                  // saturating at the format's largest finite value
  if (FMT == 1) asm("cvt.rn.satfinite.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(b), "f"(a));
  else asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(b), "f"(a));
  return d;
}
template <int FMT>
__device__ __forceinline__ uint32_t pack2_relu(float a, float b) {  // max(., 0) folded into the conversion
  uint32_t d;
  if (FMT == 1) asm("cvt.rn.relu.satfinite.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(b), "f"(a));
  else asm("cvt.rn.relu.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(b), "f"(a));
  return d;
}
// packed pair with the halves whose partner in `h` (same format) is not > 0 cleared: dX * relu'(X)
template <int FMT>
__device__ __forceinline__ uint32_t mask_pos2(uint32_t v, uint32_t h) {
  if (FMT == 1) return v & __hgt2_mask(*reinterpret_cast<const __nv_bfloat162*>(&h), __float2bfloat162_rn(0.f));
  return v & __hgt2_mask(*reinterpret_cast<const __half2*>(&h), __float2half2_rn(0.f));
}

// byte offset of element (row, k) inside a tile with chunk stride cs
__device__ __forceinline__ int tile_off(int row, int k, int cs) { return (k >> 3) * cs + row * 16 + (k & 7) * 2; }

// weights of every layer -> 16-bit operand tiles: rows n < Np (zero padded), columns k < Ka with column K = bias
template <int FMT>
__device__ void stage_weights(const TcArgs& a, unsigned char* smem) {
  for (int l = 0; l < a.nl; ++l) {
    const TcLayer& ly = a.layer[l];
    unsigned char* wt = smem + ly.w_off;
    const int total = ly.Np * ly.Ka;
    for (int e = threadIdx.x; e < total; e += blockDim.x) {
      int n = e / ly.Ka, k = e - n * ly.Ka;
      float v = 0.f;
      if (n < ly.N) {
        if (k < ly.K) v = __ldg(ly.W + (size_t)n * ly.K + k);
        else if (k == ly.K) v = __ldg(ly.b + n);
      }
      *reinterpret_cast<uint16_t*>(wt + tile_off(n, k, ly.w_cs)) = cvt16<FMT>(v);
    }
  }
}

// global x[row0 .. row0+127, 0..K) fp32 -> X0 tile (16-bit), column K = 1, columns K+1..Ka-1 = 0
template <int FMT>
__device__ void load_input_tile(const float* __restrict__ x, int64_t row0, int64_t rows_total, int K, int Ka,
                                unsigned char* tile, int tid) {
  int64_t rows = rows_total - row0;
  if (rows > kTile) rows = kTile;
  const float* src = x + row0 * K;
  const int valid = (int)rows * K;
  const int total = kTile * K;
  if ((K & 3) == 0 && ((reinterpret_cast<uintptr_t>(src) & 15) == 0)) {
    for (int e = tid * 4; e < total; e += kThreads * 4) {
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (e < valid) v = __ldg(reinterpret_cast<const float4*>(src + e));
      int r = e / K, k = e - r * K;
      uint2 p = make_uint2(pack2<FMT>(v.x, v.y), pack2<FMT>(v.z, v.w));
      *reinterpret_cast<uint2*>(tile + tile_off(r, k, kActCs)) = p;
    }
  } else {
    for (int e = tid; e < total; e += kThreads) {
      float v = e < valid ? __ldg(src + e) : 0.f;
      int r = e / K, k = e - r * K;
      *reinterpret_cast<uint16_t*>(tile + tile_off(r, k, kActCs)) = cvt16<FMT>(v);
    }
  }
  const int r = tid;
  for (int k = K; k < Ka; ++k)
    *reinterpret_cast<uint16_t*>(tile + tile_off(r, k, kActCs)) = cvt16<FMT>(k == K ? 1.f : 0.f);
}

// X0 tile of the head MLPs, one row per thread (IN = 1: SHD = 16, EMBD = 32; IN = 2: SHD = 0, EMBD = 16).
// Values are produced in fp32 exactly as nkbs_assemble_fwd writes them, then rounded to the operand format.
template <int FMT, int SHD, int EMBD>
__device__ __forceinline__ void build_head_tile(const TcArgs& a, int64_t row0, unsigned char* tile, int tid) {
  constexpr int GEO = 15, K0 = SHD + GEO + EMBD, Ka = (K0 + 1 + 15) & ~15;
  const HeadIn& h = a.hin;
  const int64_t row = row0 + tid;
  if (row >= a.N) {  // rows past the end (last tile only): zeros and the constant-one column
#pragma unroll
    for (int c = 0; c < Ka / 8; ++c) {
      uint4 z = make_uint4(0u, 0u, 0u, 0u);
      if (c == K0 / 8) {
        const uint32_t one = (uint32_t)cvt16<FMT>(1.f) << ((K0 & 1) * 16);
        if (((K0 & 7) >> 1) == 0) z.x = one;
        else if (((K0 & 7) >> 1) == 1) z.y = one;
        else if (((K0 & 7) >> 1) == 2) z.z = one;
        else z.w = one;
      }
      *reinterpret_cast<uint4*>(tile + c * kActCs + tid * 16) = z;
    }
    return;
  }
  const int64_t ray = row / h.S;
  float sh[16];
  if (SHD > 0) {
    const int64_t di = h.dir_per_ray ? ray : row;
    const float x = (__ldg(h.dirs + 3 * di) + 1.0f) / 2.0f;
    const float y = (__ldg(h.dirs + 3 * di + 1) + 1.0f) / 2.0f;
    const float z = (__ldg(h.dirs + 3 * di + 2) + 1.0f) / 2.0f;
    sh16(x, y, z, sh);
  }
  // geometry: g[1..15].  When the features are columns 1..15 of 16-float rows (the base MLP's output, the usual
  // caller) the row is fetched with four 16-byte loads; element 0 (the density column) rides along unused.
  float g[16];
  const float* gp = h.geo + row * (int64_t)h.geo_stride;
  if (h.geo_vec) {
    const float4* q = reinterpret_cast<const float4*>(gp - 1);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float4 t = __ldg(q + i);
      g[4 * i] = t.x, g[4 * i + 1] = t.y, g[4 * i + 2] = t.z, g[4 * i + 3] = t.w;
    }
  } else {
#pragma unroll
    for (int k = 0; k < GEO; ++k) g[k + 1] = __ldg(gp + k);
  }
  float e[EMBD > 0 ? EMBD : 1];
  if (EMBD > 0) {
    const float* ep = h.emb_const;
    if (h.cam) ep = h.emb + __ldg(h.cam + (h.cam_per_ray ? ray : row)) * EMBD;
    if (h.emb_vec) {
#pragma unroll
      for (int i = 0; i < EMBD / 4; ++i) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(ep) + i);
        e[4 * i] = t.x, e[4 * i + 1] = t.y, e[4 * i + 2] = t.z, e[4 * i + 3] = t.w;
      }
    } else {
#pragma unroll
      for (int k = 0; k < EMBD; ++k) e[k] = __ldg(ep + k);
    }
  }
#pragma unroll
  for (int c = 0; c < Ka / 8; ++c) {
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int k = 8 * c + j;  // compile-time column
      if (k < SHD) v[j] = sh[k < 16 ? k : 0];
      else if (k < SHD + GEO) v[j] = g[k - SHD + 1];
      else if (k < K0) v[j] = e[EMBD > 0 ? k - SHD - GEO : 0];
      else v[j] = k == K0 ? 1.f : 0.f;
    }
    *reinterpret_cast<uint4*>(tile + c * kActCs + tid * 16) =
        make_uint4(pack2<FMT>(v[0], v[1]), pack2<FMT>(v[2], v[3]), pack2<FMT>(v[4], v[5]), pack2<FMT>(v[6], v[7]));
  }
}

template <int FMT, int IN>
__device__ __forceinline__ void build_head_tile_in(const TcArgs& a, int64_t row0, unsigned char* tile, int tid) {
  if (IN == 1) build_head_tile<FMT, 16, 32>(a, row0, tile, tid);
  else if (IN == 2) build_head_tile<FMT, 0, 16>(a, row0, tile, tid);
  else build_head_tile<FMT, 0, 0>(a, row0, tile, tid);
}

// Output activation on a group of 16 accumulator columns starting at c0: one warp-uniform branch per group, and only
// the columns an activation actually touches are computed (density: column 0; sigmoid: columns < N).
__device__ __forceinline__ void apply_out_act16(int act, int c0, int N, float* v, float sel) {
  if (act == ACT_SIGMOID) {
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (c0 + 4 * q < N) {
#pragma unroll
        for (int j = 4 * q; j < 4 * q + 4; ++j) v[j] = nkbs_sigmoid(v[j]);
      }
  } else if (act == ACT_DENSITY) {
    if (c0 == 0) v[0] = expf(v[0]) * sel;
  } else if (act == ACT_TRANSIENT) {
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (c0 + 4 * q < N) {
#pragma unroll
        for (int j = 4 * q; j < 4 * q + 4; ++j) {
          const float sg = nkbs_sigmoid(v[j]), sp = nkbs_softplus(v[j]);
          v[j] = (c0 + j >= 1 && c0 + j <= 3) ? sg : sp;
        }
      }
  }
}

__device__ __forceinline__ float out_act_grad(int act, int n, float raw, float sel, float dy) {
  switch (act) {
    case ACT_SIGMOID: {
      float s = nkbs_sigmoid(raw);
      return dy * s * (1.f - s);
    }
    case ACT_DENSITY: return n == 0 ? dy * sel * expf(fminf(fmaxf(raw, -15.f), 15.f)) : dy;
    case ACT_TRANSIENT:
      if (n >= 1 && n <= 3) {
        float s = nkbs_sigmoid(raw);
        return dy * s * (1.f - s);
      }
      return raw > 20.f ? dy : dy * nkbs_sigmoid(raw);
    default: return dy;
  }
}

// The same on a group of 16 columns: v[j] (raw output) -> dL/d raw, zero for columns >= Kn.  One warp-uniform branch
// on the activation per group, work only for the 4-column pieces that hold outputs.
template <typename G>
__device__ __forceinline__ void grad_loop16(int c0, int Kn, float* v, const float* dyr, G g) {
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    if (c0 + 4 * q < Kn) {
#pragma unroll
      for (int j = 4 * q; j < 4 * q + 4; ++j) {
        const int n = c0 + j;
        const bool in = n < Kn;
        const float r = g(n, v[j], in ? dyr[n] : 0.f);
        v[j] = in ? r : 0.f;
      }
    } else {
#pragma unroll
      for (int j = 4 * q; j < 4 * q + 4; ++j) v[j] = 0.f;
    }
  }
}
__device__ __forceinline__ void out_act_grad16(int act, int c0, int Kn, float* v, float sel, const float* dyr) {
  if (act == ACT_SIGMOID) {
    grad_loop16(c0, Kn, v, dyr, [](int, float raw, float dy) {
      const float s = nkbs_sigmoid(raw);
      return dy * s * (1.f - s);
    });
  } else if (act == ACT_DENSITY) {
    grad_loop16(c0, Kn, v, dyr, [sel](int n, float raw, float dy) {
      return n == 0 ? dy * sel * expf(fminf(fmaxf(raw, -15.f), 15.f)) : dy;
    });
  } else if (act == ACT_TRANSIENT) {
    grad_loop16(c0, Kn, v, dyr, [act, sel](int n, float raw, float dy) { return out_act_grad(act, n, raw, sel, dy); });
  } else {
    grad_loop16(c0, Kn, v, dyr, [](int, float, float dy) { return dy; });
  }
}

// D[128 x Np] = X_l [128 x Ka] . W_l^T      (both K-major).  Warp-uniform: see mma_f16_uniform.
__device__ __forceinline__ void issue_forward(const TcArgs& a, int l, uint32_t smem_base, uint32_t tmem, uint32_t xa,
                                              uint32_t leader) {
  const TcLayer& ly = a.layer[l];
  const uint32_t idesc = make_idesc_f16(128, ly.Np, 0, 0, a.fmt, a.fmt);
  const uint64_t da = make_desc(xa, kActCs, 128), db = make_desc(smem_base + ly.w_off, ly.w_cs, 128);
  const uint32_t sa = (2 * kActCs) >> 4, sb = (2 * ly.w_cs) >> 4;  // descriptor start-address steps per K=16
  const int nk = ly.Ka >> 4;
#pragma unroll
  for (int k = 0; k < (kMaxWidth + 16) / 16; ++k)
    if (k < nk) mma_f16_uniform(tmem, da + (uint64_t)(k * sa), db + (uint64_t)(k * sb), idesc, k > 0, leader);
}


#ifdef NKBS_TRACE
__constant__ int c_dbg;  // experiments (NKBS_TC_DBG): bit 0 skips the TMEM loads, bit 1 the operand-tile stores
#define DBG(bit) (c_dbg & (bit))
#else
#define DBG(bit) 0
#endif

// Walk this thread's accumulator row in groups of 16 columns: f(c0, v[16]) for c0 = 0, 16, .. < ncols_total.
// Columns >= ncols_tmem are zeros.  Two tcgen05.ld (32 columns) are in flight per wait.
template <int kMaxGroups, typename F>
__device__ __forceinline__ void tmem_foreach16(uint32_t taddr, int ncols_tmem, int ncols_total, F f, int c_begin = 0) {
#pragma unroll 1
  for (int c0 = c_begin; c0 < ncols_total; c0 += 32) {
    uint32_t b0[16], b1[16];
    const bool h0 = c0 < ncols_tmem && !DBG(1), h1 = c0 + 16 < ncols_tmem && !DBG(1);
    if (h0) tmem_ld16_issue(taddr + c0, b0);
    if (h1) tmem_ld16_issue(taddr + c0 + 16, b1);
    if (h0) tmem_ld_wait();
    float v[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = h0 ? __uint_as_float(b0[j]) : 0.f;
    f(c0, v);
    if (c0 + 16 < ncols_total) {
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = h1 ? __uint_as_float(b1[j]) : 0.f;
      f(c0 + 16, v);
    }
  }
}

template <int FMT>
__device__ __forceinline__ void store16(unsigned char* tile, int c0, int r, const float v[16]) {
  if (DBG(2)) return;
  *reinterpret_cast<uint4*>(tile + (c0 >> 3) * kActCs + r * 16) =
      make_uint4(pack2<FMT>(v[0], v[1]), pack2<FMT>(v[2], v[3]), pack2<FMT>(v[4], v[5]), pack2<FMT>(v[6], v[7]));
  *reinterpret_cast<uint4*>(tile + ((c0 >> 3) + 1) * kActCs + r * 16) =
      make_uint4(pack2<FMT>(v[8], v[9]), pack2<FMT>(v[10], v[11]), pack2<FMT>(v[12], v[13]), pack2<FMT>(v[14], v[15]));
}

// 16 accumulator columns (raw TMEM words) -> two 16-byte chunks of an operand tile, ReLU folded into the conversion
template <int FMT>
__device__ __forceinline__ void store16_act(unsigned char* tile, int c0, int r, const uint32_t (&b)[16], bool relu) {
  uint32_t p[8];
  if (relu) {
#pragma unroll
    for (int j = 0; j < 8; ++j) p[j] = pack2_relu<FMT>(__uint_as_float(b[2 * j]), __uint_as_float(b[2 * j + 1]));
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) p[j] = pack2<FMT>(__uint_as_float(b[2 * j]), __uint_as_float(b[2 * j + 1]));
  }
  if (DBG(2)) return;
  *reinterpret_cast<uint4*>(tile + (c0 >> 3) * kActCs + r * 16) = make_uint4(p[0], p[1], p[2], p[3]);
  *reinterpret_cast<uint4*>(tile + ((c0 >> 3) + 1) * kActCs + r * 16) = make_uint4(p[4], p[5], p[6], p[7]);
}

// hidden-layer epilogue: TMEM accumulator -> ReLU -> 16-bit -> next layer's input tile (+ bias column).
// Interior columns go 32 at a time through one conversion per pair; the groups that hold the tail of the outputs,
// the constant-one column and the zero padding take the per-column path.
template <int FMT>
__device__ __forceinline__ void hidden_epilogue(uint32_t tmem_row, int Np, int N, int relu, int Ka_next,
                                                unsigned char* dst, int r) {
  int c0 = 0;
#pragma unroll 1
  for (; c0 + 32 <= N; c0 += 32) {
    uint32_t b0[16], b1[16];
    tmem_ld16_issue(tmem_row + c0, b0);
    tmem_ld16_issue(tmem_row + c0 + 16, b1);
    tmem_ld_wait();
    store16_act<FMT>(dst, c0, r, b0, relu != 0);
    store16_act<FMT>(dst, c0 + 16, r, b1, relu != 0);
  }
  if ((N & 15) == 0 && c0 == N) {  // nothing but [1 0 0 0 0 0 0 0] and zero chunks left
    for (int c = N >> 3; c < (Ka_next >> 3); ++c)
      *reinterpret_cast<uint4*>(dst + c * kActCs + r * 16) = make_uint4(c == (N >> 3) ? cvt16<FMT>(1.f) : 0u, 0u, 0u, 0u);
    return;
  }
  tmem_foreach16<(kMaxWidth + 16) / 16>(tmem_row, Np, Ka_next, [&](int c0, float v[16]) {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int n = c0 + j;
      v[j] = n < N ? (relu ? fmaxf(v[j], 0.f) : v[j]) : (n == N ? 1.f : 0.f);
    }
    store16<FMT>(dst, c0, r, v);
  }, c0);
}

#ifdef NKBS_TRACE
__device__ long long g_trace[2][4096];
__device__ int g_trace_n[2];
__device__ __forceinline__ void trace(int role, int tag) {
  if (blockIdx.x != 0) return;
  int i = g_trace_n[role];
  if (i < 2047) {
    g_trace[role][2 * i] = tag;
    g_trace[role][2 * i + 1] = clock64();
    g_trace_n[role] = i + 1;
  }
}
#define TRACE_WG(tag) do { if (wg == 0 && tid == 0) trace(0, tag); } while (0)
#define TRACE_IS(tag) do { if (lane == 0) trace(1, tag); } while (0)
#else
#define TRACE_WG(tag)
#define TRACE_IS(tag)
#endif

struct Sh {
  uint64_t ready[4];  // warpgroup g: operand tiles of its next MMA batch are written   (128 arrivals)
  uint64_t done[4];   // issuer: that batch has completed (tcgen05.commit)
  uint32_t tmem_base;
};

__device__ __forceinline__ int64_t tile_of(int i, int wg, int nwg) {
  return ((int64_t)i * gridDim.x + blockIdx.x) * nwg + wg;
}

// ---- input tile through registers: the loads of tile i+1 are in flight while tile i is processed ----
template <int NV>
__device__ __forceinline__ void x_prefetch(float4 (&xr)[NV], const float* __restrict__ x, int64_t row0, int64_t rows_total,
                                           int K, int tid) {
  int64_t rows = rows_total - row0;
  if (rows > kTile) rows = kTile;
  const float* src = x + row0 * K;
  const int valid = (int)rows * K;
  const int nvec = K >> 2;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int e = 4 * (tid + kThreads * j);
    if (j < nvec && e < valid) xr[j] = ldg_nc_v4(src + e);
    else xr[j] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// next tile's input rows -> L2 (one 128-byte line per thread and pass): the tile after this one then loads from L2
// instead of HBM, without holding registers or shared memory for it
__device__ __forceinline__ void x_prefetch_l2(const float* __restrict__ x, int64_t row0, int64_t rows_total, int K, int tid) {
  if (row0 >= rows_total) return;
  int64_t rows = rows_total - row0;
  if (rows > kTile) rows = kTile;
  const char* src = reinterpret_cast<const char*>(x + row0 * K);
  const int bytes = (int)rows * K * 4;
  for (int o = tid * 128; o < bytes; o += kThreads * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(src + o));
}

template <int FMT, int NV>
__device__ __forceinline__ void x_commit(const float4 (&xr)[NV], int K, int kshift, int Ka, unsigned char* tile, int tid) {
  const int nvec = K >> 2;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    if (j < nvec) {
      const int e = 4 * (tid + kThreads * j);
      const int r = kshift >= 0 ? (e >> kshift) : (e / K);
      const int k = e - r * K;
      *reinterpret_cast<uint2*>(tile + tile_off(r, k, kActCs)) =
          make_uint2(pack2<FMT>(xr[j].x, xr[j].y), pack2<FMT>(xr[j].z, xr[j].w));
    }
  }
  if ((K & 7) == 0) {  // the constant-one column opens a 16-byte chunk: [1 0 0 0 0 0 0 0], then zero chunks
    for (int c = K >> 3; c < (Ka >> 3); ++c)
      *reinterpret_cast<uint4*>(tile + c * kActCs + tid * 16) = make_uint4(c == (K >> 3) ? cvt16<FMT>(1.f) : 0u, 0u, 0u, 0u);
  } else {
    for (int k = K; k < Ka; ++k)
      *reinterpret_cast<uint16_t*>(tile + tile_off(tid, k, kActCs)) = cvt16<FMT>(k == K ? 1.f : 0.f);
  }
}

// warpgroup side of one MMA batch: publish the operand tiles, then wait for the batch to complete
__device__ __forceinline__ void wg_round(Sh& sh, int wg, uint32_t& phase) {
  fence_proxy_async();   // st.shared of the operand tiles -> visible to the tensor core's async proxy
  fence_before_sync();   // earlier tcgen05.ld of this thread are ordered before the batch that overwrites TMEM
  mbar_arrive(&sh.ready[wg]);
  mbar_wait(&sh.done[wg], phase);
  phase ^= 1;
  fence_after_sync();
}

__device__ __forceinline__ void kernel_prologue(Sh& sh, bool issuer, int tmem_cols) {
  if (threadIdx.x == 0) {
    for (int g = 0; g < 4; ++g) {
      mbar_init(&sh.ready[g], kThreads);
      mbar_init(&sh.done[g], 1);
    }
    mbar_fence_init();
  }
  if (issuer) tmem_alloc(&sh.tmem_base, tmem_cols);
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
}

// Register rebalancing (setmaxnreg): the CTA is NWG epilogue warpgroups plus one warpgroup of which only the first warp
// works (the MMA issuer).  At launch every warp gets the same 512 / (NWG + 1) registers per lane of its SM
// sub-partition; the issuer warpgroup shrinks to 56 and the epilogue warpgroups grow into what that frees
// (pool = the CTA's launch allocation: 2 x 224 + 56 = 3 x 168, 3 x 152 + 56 = 4 x 128, 4 x 104 + 56 <= 5 x 96), which is what keeps their tile-wide register arrays out of local memory.
template <int NWG>
__device__ __forceinline__ void rebalance_registers(bool issuer_wg) {
  if (NWG < 2) return;
  if (issuer_wg) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
  } else {
    if (NWG == 2) asm volatile("setmaxnreg.inc.sync.aligned.u32 224;");
    else if (NWG == 3) asm volatile("setmaxnreg.inc.sync.aligned.u32 152;");
    else asm volatile("setmaxnreg.inc.sync.aligned.u32 104;");
  }
}

template <int FMT, int NWG, int IN>
__global__ void __launch_bounds__((NWG + 1) * kThreads, 1) mlp_tc_fwd_kernel(const TcArgs a) {
  nkbs_pdl_sync();
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ Sh sh;
  constexpr int nwg = NWG;
  const int wg = threadIdx.x >> 7;
  const int tid = threadIdx.x & (kThreads - 1);
  const bool issuer = threadIdx.x >> 5 == nwg * 4;  // first warp of the last warpgroup; its other warps only idle
  stage_weights<FMT>(a, smem);
  kernel_prologue(sh, issuer, a.tmem_cols);
  const uint32_t tmem = sh.tmem_base;
  const uint32_t smem_base = smem_u32(smem);
  const int64_t ntiles = (a.N + kTile - 1) / kTile;
  if (wg == nwg) {
    rebalance_registers<NWG>(true);
  }
  if (issuer) {
    uint32_t rph = 0;  // bit g: phase of ready[g]
    const uint32_t leader = elect_one();
    for (int i = 0;; ++i) {
      if (tile_of(i, 0, nwg) >= ntiles) break;
      for (int l = 0; l < a.nl; ++l) {
#pragma unroll
        for (int g = 0; g < nwg; ++g) {
          if (tile_of(i, g, nwg) < ntiles) {
            mbar_wait(&sh.ready[g], (rph >> g) & 1u);
            fence_after_sync();
            const uint32_t base = smem_base + a.wg_base + g * a.wg_stride;
            issue_forward(a, l, smem_base, tmem + g * kMaxWidth, base + ((l & 1) ? a.x_off1 : a.layer[0].x_off), leader);
            commit_uniform(&sh.done[g], leader);
            rph ^= 1u << g;
          }
        }
      }
    }
  } else if (wg < nwg) {
    rebalance_registers<NWG>(false);
    unsigned char* base = smem + a.wg_base + wg * a.wg_stride;
    const uint32_t tmem_row = tmem + wg * kMaxWidth + ((uint32_t)(tid & ~31) << 16);
    const TcLayer& last = a.layer[a.nl - 1];
    const int K0 = a.layer[0].K, Ka0 = a.layer[0].Ka;
    const bool vec = (K0 & 3) == 0 && (K0 >> 2) <= (NWG >= 3 ? 8 : 16) && (reinterpret_cast<uintptr_t>(a.x) & 15) == 0;
    const int kshift = (K0 & (K0 - 1)) == 0 ? (31 - __clz(K0)) : -1;
    uint32_t phase = 0;
    constexpr int NV = NWG >= 3 ? 8 : 16;  // register budget: 120-128 registers per thread with three or four warpgroups
    for (int i = 0;; ++i) {
      const int64_t t = tile_of(i, wg, nwg);
      if (t >= ntiles) break;
      const int64_t row0 = t * kTile;
      const int64_t row = row0 + tid;
      const bool live = row < a.N;
      float sel = 1.f;
      if (a.selector && live) sel = __ldg(a.selector + row);
      unsigned char* cur = base + a.layer[0].x_off;
      unsigned char* nxt = base + a.x_off1;
      TRACE_WG(1);
      if (IN != 0) {
        build_head_tile_in<FMT, IN>(a, row0, cur, tid);
      } else if (vec) {
        float4 xr[NV];
        x_prefetch(xr, a.x, row0, a.N, K0, tid);
        x_prefetch_l2(a.x, tile_of(i + 1, wg, nwg) * kTile, a.N, K0, tid);
        x_commit<FMT>(xr, K0, kshift, Ka0, cur, tid);
      } else {
        load_input_tile<FMT>(a.x, row0, a.N, K0, Ka0, cur, tid);
      }
      TRACE_WG(2);
      for (int l = 0; l < a.nl; ++l) {
        wg_round(sh, wg, phase);
        TRACE_WG(10 + l);
        if (l + 1 < a.nl) {
          hidden_epilogue<FMT>(tmem_row, a.layer[l].Np, a.layer[l].N, a.layer[l].relu, a.layer[l + 1].Ka, nxt, tid);
          TRACE_WG(20 + l);
          unsigned char* tmp = cur;
          cur = nxt;
          nxt = tmp;
        }
      }
      // output layer: activation, then rows -> fp32 staging (odd row stride: conflict-free) -> coalesced global stores.
      // The staging lives in the second activation buffer: the last MMA that read it has completed, and its next
      // writer is the next tile's first hidden epilogue, a warpgroup round away.
      if (a.ystage_ok) {
        float* st = reinterpret_cast<float*>(base + a.x_off1);
        const int NO = last.N;
        int64_t rows = a.N - row0;
        if (rows > kTile) rows = kTile;
        float* dst = a.y + row0 * NO;
        if ((NO & 3) == 0) {
          // rows of 16-byte pieces: stride NO + 4 floats keeps both the row-wise stores and the linear reads aligned
          const int stride = NO + 4;
          tmem_foreach16<kMaxWidth / 16>(tmem_row, last.Np, last.Np, [&](int c0, float v[16]) {
            apply_out_act16(a.out_act, c0, NO, v, sel);
            if (c0 == 0 && a.y0 && live) a.y0[row] = v[0];
            float4* p = reinterpret_cast<float4*>(st + tid * stride + c0);
#pragma unroll
            for (int q = 0; q < 4; ++q)
              if (c0 + 4 * q < NO) p[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
          });
          TRACE_WG(31);
          wg_sync(wg);
          TRACE_WG(32);
          const int nv = NO >> 2, total = (int)rows * nv;
          int rr = tid / nv, c4 = tid - rr * nv;
          const int dr = kThreads / nv, dc = kThreads - dr * nv;
          for (int e = tid; e < total; e += kThreads) {
            reinterpret_cast<float4*>(dst)[e] = *reinterpret_cast<const float4*>(st + rr * stride + 4 * c4);
            rr += dr;
            c4 += dc;
            if (c4 >= nv) {
              c4 -= nv;
              ++rr;
            }
          }
        } else {
          const int stride = NO | 1, pad = stride - NO;
          float* srow = st + tid * stride;
          tmem_foreach16<kMaxWidth / 16>(tmem_row, last.Np, last.Np, [&](int c0, float v[16]) {
            apply_out_act16(a.out_act, c0, NO, v, sel);
            if (c0 == 0 && a.y0 && live) a.y0[row] = v[0];
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (c0 + j < NO) srow[c0 + j] = v[j];
          });
          TRACE_WG(31);
          wg_sync(wg);
          TRACE_WG(32);
          const int total = (int)rows * NO;
          int rr = tid / NO, c = tid - rr * NO;
          const int dr = kThreads / NO, dc = kThreads - dr * NO;
          for (int e = tid; e < total; e += kThreads) {
            dst[e] = st[e + rr * pad];
            rr += dr;
            c += dc;
            if (c >= NO) {
              c -= NO;
              ++rr;
            }
          }
        }
      } else {
        float* yo = a.y + row * last.N;
        tmem_foreach16<kMaxWidth / 16>(tmem_row, last.Np, last.Np, [&](int c0, float v[16]) {
          apply_out_act16(a.out_act, c0, last.N, v, sel);
          if (c0 == 0 && a.y0 && live) a.y0[row] = v[0];
          if (live) {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (c0 + j < last.N) yo[c0 + j] = v[j];
          }
        });
      }
      TRACE_WG(30);
    }
  }
  fence_before_sync();
  __syncthreads();
  if (issuer) tmem_dealloc(tmem, a.tmem_cols);
}

template <int FMT, int NWG, int IN>
__global__ void __launch_bounds__((NWG + 1) * kThreads, 1) mlp_tc_bwd_kernel(const TcArgs a) {
  nkbs_pdl_sync();
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ Sh sh;
  constexpr int nwg = NWG;
  const int wg = threadIdx.x >> 7;
  const int tid = threadIdx.x & (kThreads - 1);
  const int lane = threadIdx.x & 31;
  const bool issuer = threadIdx.x >> 5 == nwg * 4;  // first warp of the last warpgroup; its other warps only idle
  stage_weights<FMT>(a, smem);
  kernel_prologue(sh, issuer, a.tmem_cols);
  const uint32_t tmem = sh.tmem_base;
  const uint32_t smem_base = smem_u32(smem);
  const int64_t ntiles = (a.N + kTile - 1) / kTile;
  const int nl = a.nl;
  const bool any_tile = tile_of(0, 0, nwg) < ntiles;
  // gradient w.r.t. the first layer's input is wanted: as rows (dx) or split into geo / embedding parts (IN != 0)
  const bool want_dx0 = a.dx != nullptr || (IN != 0 && (a.hin.dgeo != nullptr || a.hin.demb != nullptr));
  if (wg == nwg) {
    rebalance_registers<NWG>(true);
  }
  if (issuer) {
    uint32_t rph = 0;      // bit g: phase of ready[g]
    uint32_t started = 0;  // bit l: dW_l accumulator holds data
    const uint32_t leader = elect_one();
    for (int i = 0;; ++i) {
      if (tile_of(i, 0, nwg) >= ntiles) break;
      for (int r = 0; r < 2 * nl; ++r) {
#pragma unroll
        for (int g = 0; g < nwg; ++g) {
          if (tile_of(i, g, nwg) < ntiles) {
            mbar_wait(&sh.ready[g], (rph >> g) & 1u);
            fence_after_sync();
            TRACE_IS(100 + r * 4 + g);
            const uint32_t base = smem_base + a.wg_base + g * a.wg_stride;
            const uint32_t work = tmem + g * kMaxWidth;
            if (r < nl) {
              issue_forward(a, r, smem_base, work, base + a.layer[r].x_off, leader);
            } else {
              const int l = 2 * nl - 1 - r;
              const TcLayer& ly = a.layer[l];
              const uint32_t ga = base + (l == nl - 1 ? a.g_off : a.layer[l + 1].x_off);
              const uint32_t xa = base + ly.x_off, wa = smem_base + ly.w_off;
              if (l > 0 || want_dx0) {
                // dX_l [128 x Nx] = G [128 x Np] . W_l   (W_l viewed MN-major: contraction over its rows); issued
                // first and committed on its own so the warpgroup's epilogue does not wait for the dW GEMM
                const int Nx = (ly.K + 15) & ~15;
                const uint32_t id_x = make_idesc_f16(128, Nx, 0, 1, a.fmt, a.fmt);
                const uint64_t dgk = make_desc(ga, kActCs, 128), dw = make_desc(wa, 128, ly.w_cs);
                const int nk = ly.Np >> 4;
#pragma unroll
                for (int k = 0; k < kMaxWidth / 16; ++k)
                  if (k < nk)
                    mma_f16_uniform(work, dgk + (uint64_t)(k * ((2 * kActCs) >> 4)), dw + (uint64_t)(k * 16), id_x, k > 0, leader);
              }
              // dW_l (M = 64 features of dY) += G^T . [X_l | 1]   : contraction over the 128 samples
              const uint32_t id_w = make_idesc_f16(64, ly.Ka, 1, 1, a.fmt, a.fmt);
              const uint32_t acc0 = (started >> l) & 1u;
              const uint64_t dg = make_desc(ga, 128, kActCs), dxd = make_desc(xa, 128, kActCs);
#pragma unroll
              for (int k = 0; k < kTile / 16; ++k)
                mma_f16_uniform(tmem + ly.dw_col, dg + (uint64_t)(k * 16), dxd + (uint64_t)(k * 16), id_w, acc0 | (k > 0), leader);
              started |= 1u << l;
            }
            commit_uniform(&sh.done[g], leader);
            TRACE_IS(200 + r * 4 + g);
            rph ^= 1u << g;
          }
        }
      }
    }
  } else if (wg < nwg) {
    rebalance_registers<NWG>(false);
    unsigned char* base = smem + a.wg_base + wg * a.wg_stride;
    const uint32_t tmem_row = tmem + wg * kMaxWidth + ((uint32_t)(tid & ~31) << 16);
    const TcLayer& last = a.layer[nl - 1];
    const int K0 = a.layer[0].K, Ka0 = a.layer[0].Ka, Kn = last.N;
    const bool vec = (K0 & 3) == 0 && (K0 >> 2) <= (NWG >= 3 ? 8 : 16) && (reinterpret_cast<uintptr_t>(a.x) & 15) == 0;
    const bool vec_dx = a.dx && (K0 & 3) == 0 && (reinterpret_cast<uintptr_t>(a.dx) & 15) == 0;
    const int kshift = (K0 & (K0 - 1)) == 0 ? (31 - __clz(K0)) : -1;
    const int r = tid;
    float* dys = reinterpret_cast<float*>(base + a.dy_off);
    uint32_t phase = 0;
    constexpr int NV = NWG >= 3 ? 8 : 16;  // register budget: 120-128 registers per thread with three or four warpgroups
    // The output-gradient tile G is read 64 features wide (dW of the last layer, M = 64) but only its first Np
    // columns carry data: the padding is zeroed once here, nothing else ever writes this region (the dx staging
    // ends before it, see build_args).
    for (int c = last.Np >> 3; c < kMaxWidth / 8; ++c)
      *reinterpret_cast<uint4*>(base + a.g_off + c * kActCs + r * 16) = make_uint4(0u, 0u, 0u, 0u);
    for (int i = 0;; ++i) {
      const int64_t t = tile_of(i, wg, nwg);
      if (t >= ntiles) break;
      const int64_t row0 = t * kTile;
      const int64_t row = row0 + r;
      const bool live = row < a.N;
      TRACE_WG(1);
      // upstream gradient tile: asynchronous copy, consumed after the forward recompute
      {
        const int64_t total = a.N * Kn, e0 = row0 * Kn;
        for (int j = 0; j < Kn; ++j) {
          const int e = tid + kThreads * j;
          const bool ok = e0 + e < total;
          cp_async4(dys + e, a.dy + (ok ? e0 + e : 0), ok ? 4u : 0u);
        }
        cp_async_commit();
      }
      TRACE_WG(5);
      float sel = 1.f;
      if (a.selector && live) sel = __ldg(a.selector + row);
      auto stage_x0 = [&]() {
        if (IN != 0) {
          build_head_tile_in<FMT, IN>(a, row0, base + a.layer[0].x_off, tid);
        } else if (vec) {
          float4 xr[NV];
          x_prefetch(xr, a.x, row0, a.N, K0, tid);
          x_prefetch_l2(a.x, tile_of(i + 1, wg, nwg) * kTile, a.N, K0, tid);
          x_commit<FMT>(xr, K0, kshift, Ka0, base + a.layer[0].x_off, tid);
        } else {
          load_input_tile<FMT>(a.x, row0, a.N, K0, Ka0, base + a.layer[0].x_off, tid);
        }
      };
      stage_x0();
      TRACE_WG(2);
      // ---- forward recompute, every layer input kept ----
      for (int l = 0; l < nl; ++l) {
        wg_round(sh, wg, phase);
        TRACE_WG(10 + l);
        if (l + 1 < nl)
          hidden_epilogue<FMT>(tmem_row, a.layer[l].Np, a.layer[l].N, a.layer[l].relu, a.layer[l + 1].Ka,
                               base + a.layer[l + 1].x_off, tid);
      }
      TRACE_WG(3);
      // ---- gradient w.r.t. the raw outputs -> G tile (64 features wide, zero padded) ----
      {
        cp_async_wait_all();
        wg_sync(wg);  // every thread's part of the dy tile has landed
        if (a.dy0 && live) dys[r * Kn] += __ldg(a.dy0 + row);  // row r is read by thread r only
        const float* dyr = dys + r * Kn;
        unsigned char* g = base + a.g_off;
        tmem_foreach16<kMaxWidth / 16>(tmem_row, last.Np, last.Np, [&](int c0, float v[16]) {
          if (c0 < Kn && live) {
            out_act_grad16(a.out_act, c0, Kn, v, sel, dyr);
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = 0.f;
          }
          store16<FMT>(g, c0, r, v);
        });
      }
      TRACE_WG(4);
      // ---- backward sweep ----
      for (int l = nl - 1; l >= 0; --l) {
        const TcLayer& ly = a.layer[l];
        const int Nx = (ly.K + 15) & ~15;
        TRACE_WG(30 + l);
        // recycled layout: X_2 / dY_1 occupied X_0's place; dW_0 needs X_0 again (all earlier batches have completed)
        if (l == 0 && a.recycle) stage_x0();
        wg_round(sh, wg, phase);
        TRACE_WG(20 + l);
        if (l > 0) {
          // dY_{l-1} = dX_l * relu'(X_l), written in place over X_l (its MMAs have completed); 64 columns wide
          unsigned char* xt = base + ly.x_off;
          const bool masked = a.layer[l - 1].relu != 0;
          const int K = ly.K, Ka = ly.Ka;
          // interior columns, 32 at a time: one conversion per pair, ReLU mask applied on the packed pair
          int cb = 0;
#pragma unroll 1
          for (; cb + 32 <= K; cb += 32) {
            uint32_t b0[16], b1[16];
            tmem_ld16_issue(tmem_row + cb, b0);
            tmem_ld16_issue(tmem_row + cb + 16, b1);
            tmem_ld_wait();
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const uint32_t* b = q < 2 ? b0 : b1;
              const int o = (q & 1) * 8;
              uint4* slot = reinterpret_cast<uint4*>(xt + ((cb >> 3) + q) * kActCs + r * 16);
              uint32_t p0 = pack2<FMT>(__uint_as_float(b[o]), __uint_as_float(b[o + 1]));
              uint32_t p1 = pack2<FMT>(__uint_as_float(b[o + 2]), __uint_as_float(b[o + 3]));
              uint32_t p2 = pack2<FMT>(__uint_as_float(b[o + 4]), __uint_as_float(b[o + 5]));
              uint32_t p3 = pack2<FMT>(__uint_as_float(b[o + 6]), __uint_as_float(b[o + 7]));
              if (masked) {  // the stored activation chunk is read right where it is needed (four live registers)
                const uint4 m = *slot;
                p0 = mask_pos2<FMT>(p0, m.x);
                p1 = mask_pos2<FMT>(p1, m.y);
                p2 = mask_pos2<FMT>(p2, m.z);
                p3 = mask_pos2<FMT>(p3, m.w);
              }
              *slot = make_uint4(p0, p1, p2, p3);
            }
          }
          tmem_foreach16<kMaxWidth / 16>(tmem_row, Nx, kMaxWidth, [&](int c0, float v[16]) {
            uint4 m0 = make_uint4(0, 0, 0, 0), m1 = make_uint4(0, 0, 0, 0);
            if (c0 < Ka) {
              m0 = *reinterpret_cast<const uint4*>(xt + (c0 >> 3) * kActCs + r * 16);
              m1 = *reinterpret_cast<const uint4*>(xt + ((c0 >> 3) + 1) * kActCs + r * 16);
            }
            const uint32_t mw[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
            if (c0 + 16 <= K) {
              if (masked) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  // stored activation > 0  <=>  16-bit pattern, as a signed 16-bit integer, is > 0
                  const int h = (int)(mw[j >> 1] << ((j & 1) ? 0 : 16)) >> 16;
                  if (h <= 0) v[j] = 0.f;
                }
              }
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const int h = (int)(mw[j >> 1] << ((j & 1) ? 0 : 16)) >> 16;
                const bool pos = !masked || h > 0;
                if (!(pos && (c0 + j) < K)) v[j] = 0.f;
              }
            }
            store16<FMT>(xt, c0, r, v);
          }, cb);
        } else if (IN != 0 && want_dx0) {
          // dX_0 rows -> fp32 staging; the geometry columns go out as rows of dgeo, the embedding columns are summed
          // over the samples of each ray (they share the embedding row) and added to that row's gradient
          constexpr int SHD = IN == 1 ? 16 : 0, GEO = 15, EMBD = IN == 1 ? 32 : (IN == 2 ? 16 : 0);
          float* st = reinterpret_cast<float*>(base + a.stage_off);
          const int stride = Nx + 4;
          tmem_foreach16<kMaxWidth / 16>(tmem_row, Nx, Nx, [&](int c0, float v[16]) {
            float4* p = reinterpret_cast<float4*>(st + r * stride + c0);
            p[0] = make_float4(v[0], v[1], v[2], v[3]);
            p[1] = make_float4(v[4], v[5], v[6], v[7]);
            p[2] = make_float4(v[8], v[9], v[10], v[11]);
            p[3] = make_float4(v[12], v[13], v[14], v[15]);
          });
          wg_sync(wg);
          int64_t rows = a.N - row0;
          if (rows > kTile) rows = kTile;
          if (a.hin.dgeo) {
            const int total = (int)rows * GEO;
            for (int e = tid; e < total; e += kThreads) {
              const int rr = e / GEO, c = e - rr * GEO;
              a.hin.dgeo[(row0 + rr) * (int64_t)a.hin.dgeo_stride + c] = st[rr * stride + SHD + c];
            }
            if (a.hin.dgeo_lead > 0 && tid < (int)rows)
              for (int c = 1; c <= a.hin.dgeo_lead; ++c) a.hin.dgeo[(row0 + tid) * (int64_t)a.hin.dgeo_stride - c] = 0.f;
          }
          if (a.hin.demb && a.hin.cam) {
            const int G = a.hin.cam_per_ray ? a.hin.S : 1;
            const int64_t g0 = row0 / G, g1 = (row0 + rows - 1) / G;
            const int lane_ = tid & 31, warp_ = tid >> 5;
            for (int64_t g = g0 + warp_; g <= g1; g += kThreads / 32) {
              const int lo = (int)max(g * G - row0, (int64_t)0), hi = (int)min((g + 1) * G - row0, rows);
              if (lane_ < EMBD) {
                float sum = 0.f;
                for (int rr = lo; rr < hi; ++rr) sum += st[rr * stride + SHD + GEO + lane_];
                if (sum != 0.f) nkbs_acc(a.hin.demb + __ldg(a.hin.cam + g) * EMBD + lane_, a.fm_emb, sum);
              }
            }
          }
        } else if (a.dx) {
          if (a.stage_ok) {
            // dX_0 rows -> fp32 staging (row stride Nx+4 floats: conflict-free 16-byte stores) -> coalesced global
            float* st = reinterpret_cast<float*>(base + a.stage_off);
            const int stride = Nx + 4;
            tmem_foreach16<kMaxWidth / 16>(tmem_row, Nx, Nx, [&](int c0, float v[16]) {
              float4* p = reinterpret_cast<float4*>(st + r * stride + c0);
              p[0] = make_float4(v[0], v[1], v[2], v[3]);
              p[1] = make_float4(v[4], v[5], v[6], v[7]);
              p[2] = make_float4(v[8], v[9], v[10], v[11]);
              p[3] = make_float4(v[12], v[13], v[14], v[15]);
            });
            wg_sync(wg);
            int64_t rows = a.N - row0;
            if (rows > kTile) rows = kTile;
            float* dst = a.dx + row0 * K0;
            if (vec_dx && kshift >= 2) {
              // K0 a power of two: thread -> (row, 16-byte column) by shifts, rows advance by 128 / (K0/4)
              const int sh = kshift - 2, nv = 1 << sh;
              const int c4 = tid & (nv - 1), step = kThreads >> sh;
              for (int rr = tid >> sh; rr < (int)rows; rr += step)
                reinterpret_cast<float4*>(dst)[rr * nv + c4] = *reinterpret_cast<const float4*>(st + rr * stride + c4 * 4);
            } else if (vec_dx) {
              const int nv = K0 >> 2, total = (int)rows * nv;
              for (int e = tid; e < total; e += kThreads) {
                const int rr = e / nv, c4 = e - rr * nv;
                reinterpret_cast<float4*>(dst)[e] = *reinterpret_cast<const float4*>(st + rr * stride + c4 * 4);
              }
            } else {
              const int total = (int)rows * K0;
              for (int e = tid; e < total; e += kThreads) {
                const int rr = e / K0, c = e - rr * K0;
                dst[e] = st[rr * stride + c];
              }
            }
          } else {
            float* dxr = a.dx + row * K0;
            tmem_foreach16<kMaxWidth / 16>(tmem_row, Nx, Nx, [&](int c0, float v[16]) {
              if (live) {
#pragma unroll
                for (int j = 0; j < 16; ++j)
                  if (c0 + j < K0) dxr[c0 + j] = v[j];
              }
            });
          }
        }
      }
      // recycled layout: the dx staging covers X_0's place, which the next tile's input is written into right away
      if (a.recycle) wg_sync(wg);
    }
    // ---- flush the weight / bias gradients: M=64 accumulator rows live in lanes 0..15 of every warp.  Every
    // warpgroup has waited for its own last batch; the barrier among them covers the others' dW accumulations. ----
    fence_before_sync();
    epilogue_sync<NWG>();
    fence_after_sync();
    if (any_tile) {
      const int w = tid >> 5;
      const int n = 16 * w + lane;
      const uint32_t lane_base = tmem + ((uint32_t)(tid & ~31) << 16);
      for (int l = wg; l < nl; l += nwg) {
        const TcLayer& ly = a.layer[l];
        for (int c0 = 0; c0 < ly.Ka; c0 += 16) {
          float v[16];
          tmem_ld16(lane_base + ly.dw_col + c0, v);
          if (lane < 16 && n < ly.N) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              int k = c0 + j;
              if (k < ly.K) nkbs_acc(ly.dW + (size_t)n * ly.K + k, a.fm, v[j]);
              else if (k == ly.K) nkbs_acc(ly.db + n, a.fm, v[j]);
            }
          }
        }
      }
    }
  }
  fence_before_sync();
  __syncthreads();
  if (issuer) tmem_dealloc(tmem, a.tmem_cols);
}

int round_up(int v, int m) { return (v + m - 1) / m * m; }

constexpr int kSmemLimit = 227 * 1024;

int build_args(TcArgs& a, bool backward, int nl, const int32_t* dims, const float* const* W, const float* const* b,
               int relu_mask, int out_act, int fmt, const char* who) {
  NKBS_CHECK_ARG(nl >= 1 && nl <= kMaxLayers, "%s: 1..%d linear layers supported (got %d)", who, kMaxLayers, nl);
  NKBS_CHECK_ARG(out_act >= 0 && out_act <= 3, "%s: unknown output activation %d", who, out_act);
  NKBS_CHECK_ARG(fmt == 0 || fmt == 1, "%s: fmt must be 0 (fp16) or 1 (bf16)", who);
  for (int l = 0; l <= nl; ++l)
    NKBS_CHECK_ARG(dims[l] >= 1 && dims[l] <= kMaxWidth, "%s: layer width %d outside 1..%d", who, dims[l], kMaxWidth);
  int off = 0, max_ka = 0;
  for (int l = 0; l < nl; ++l) {
    TcLayer& ly = a.layer[l];
    NKBS_CHECK_ARG(W[l] && b[l], "%s: null weight/bias for layer %d", who, l);
    ly.W = W[l];
    ly.b = b[l];
    ly.dW = nullptr;
    ly.db = nullptr;
    a.fm = a.fm_emb = FixedMap{nullptr, nullptr, 0, 0.f};
    ly.K = dims[l];
    ly.N = dims[l + 1];
    ly.relu = (relu_mask >> l) & 1;
    ly.Ka = round_up(ly.K + 1, 16);
    ly.Np = round_up(ly.N, 16);
    ly.w_cs = ly.Np * 16 + 16;
    ly.w_off = off;
    off += (ly.Ka / 8) * ly.w_cs;
    off = round_up(off, 128);
    ly.x_off = 0;
    ly.dw_col = 0;
    if (ly.Ka > max_ka) max_ka = ly.Ka;
  }
  a.nl = nl;
  a.out_act = out_act;
  a.fmt = fmt;
  a.wg_base = off;
  int per_wg = 0, dw_cols = 0;
  if (!backward) {
    int tile_bytes = round_up((max_ka / 8) * kActCs, 128);
    a.layer[0].x_off = 0;
    a.x_off1 = tile_bytes;
    a.g_off = 0;
    a.dy_off = 0;
    a.stage_off = 0;
    a.stage_ok = 0;
    a.recycle = 0;
    per_wg = 2 * tile_bytes;
    a.ystage_ok = nl >= 2 && a.layer[nl - 1].N >= 4 && kTile * (a.layer[nl - 1].N + 4) * 4 <= tile_bytes;
  } else {
    // Two layouts.  Plain: X_0 .. X_{nl-1} | G | dy.  Recycled (three layers): X_2 is written where X_0 was (X_0 is
    // dead between the first forward batch and the last backward batch, before which it is rebuilt), which frees one
    // tile per warpgroup and lets a third warpgroup fit.  The recycled layout is used only when it adds a warpgroup.
    for (int l = 0; l < nl; ++l) dw_cols += a.layer[l].Ka;
    auto tile_of_layer = [&](int l) {
      const int width = a.layer[l].Ka > kMaxWidth ? a.layer[l].Ka : kMaxWidth;  // later reused as a 64-wide gradient tile
      return round_up((width / 8) * kActCs, 128);
    };
    auto lay = [&](bool recycle) {
      int off = 0;
      if (recycle) {
        a.layer[0].x_off = a.layer[2].x_off = 0;
        off += tile_of_layer(0) > tile_of_layer(2) ? tile_of_layer(0) : tile_of_layer(2);
        a.layer[1].x_off = off;
        off += tile_of_layer(1);
      } else {
        for (int l = 0; l < nl; ++l) {
          a.layer[l].x_off = off;
          off += tile_of_layer(l);
        }
      }
      a.g_off = off;
      off += round_up((kMaxWidth / 8) * kActCs, 128);
      a.dy_off = off;
      off += round_up(kTile * a.layer[nl - 1].N * 4, 128);
      // dx staging: after X_0 in the plain layout; over both activation regions in the recycled one (every tile is dead
      // once the last batch has completed).  It must end before the G tile, whose zero padding is written once.
      const int nx0 = round_up(a.layer[0].K, 16);
      a.stage_off = recycle ? 0 : (nl >= 2 ? a.layer[1].x_off : 0);
      a.stage_ok = nl >= 2 && kTile * (nx0 + 4) * 4 <= a.g_off - a.stage_off;
      a.recycle = recycle ? 1 : 0;
      return off;
    };
    auto fit = [&](int bytes) {
      int n = 4;
      while (n > 1 && (a.wg_base + n * bytes > kSmemLimit - 1024 || n * kMaxWidth + dw_cols > 512)) --n;
      return n;
    };
    a.x_off1 = 0;
    a.ystage_ok = 0;
    per_wg = lay(false);
    // Measured on B200 (round 1): a third warpgroup makes the CTA 13 warps, which the register file allocates as 16
    // (128 registers per thread); the three-layer backward then spills ~0.5 KB per thread and loses more than the
    // extra tile in flight gains (colour head 1.56 -> 1.81 ms at 3.1 M samples).  The recycled layout therefore stays
    // off unless NKBS_TC_RECYCLE=1 asks for it (experiments); two-layer nets fit three warpgroups in the plain layout.
    static const bool want_recycle = [] { const char* e = getenv("NKBS_TC_RECYCLE"); return e && e[0] == '1'; }();
    if (nl == 3 && want_recycle) {
      const int plain = fit(per_wg);
      const int rec_bytes = lay(true);
      if (fit(rec_bytes) > plain && a.stage_ok) per_wg = rec_bytes;
      else per_wg = lay(false);
    }
  }
#ifdef NKBS_TRACE
  {
    const char* e = getenv("NKBS_TC_DBG");
    int d = e ? atoi(e) : 0;
    cudaMemcpyToSymbol(c_dbg, &d, sizeof(int));
  }
#endif
  a.wg_stride = per_wg;
  // NKBS_TC_NWG=1..4 caps the warpgroups per CTA (experiments: scripts/tc_trace.py)
  static const int nwg_cap = [] { const char* e = getenv("NKBS_TC_NWG"); return (e && e[0] >= '1' && e[0] <= '4') ? e[0] - '0' : 4; }();
  // Forward: three warpgroups.  A fourth makes the CTA 17 warps, which caps the kernel at 96 registers per thread; the
  // spill traffic that follows costs more than the extra tile in flight brings (measured on B200, full-frame render:
  // colour + semantics heads 9.8 -> 6.5 ms, base MLP 3.9 -> 2.2 ms per frame with three).
  int nwg = backward ? nwg_cap : (nwg_cap < 3 ? nwg_cap : 3);
  while (nwg > 1 && (a.wg_base + nwg * per_wg > kSmemLimit - 1024 || nwg * kMaxWidth + dw_cols > 512)) --nwg;
  NKBS_CHECK_ARG(a.wg_base + nwg * per_wg <= kSmemLimit - 1024 && nwg * kMaxWidth + dw_cols <= 512,
                 "%s: does not fit (smem %d, tmem %d)", who, a.wg_base + per_wg, kMaxWidth + dw_cols);
  a.nwg = nwg;
  int col = nwg * kMaxWidth;
  if (backward)
    for (int l = 0; l < nl; ++l) {
      a.layer[l].dw_col = col;
      col += a.layer[l].Ka;
    }
  int tc = 32;
  while (tc < col) tc <<= 1;
  a.tmem_cols = tc;
  a.smem_bytes = a.wg_base + nwg * per_wg;
  return 0;
}

template <typename K>
int set_smem(K kernel, int bytes, const char* who) {
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  NKBS_CHECK_ARG(e == cudaSuccess, "%s: %d bytes of shared memory refused: %s", who, bytes, cudaGetErrorString(e));
  return 0;
}

int pick_grid(const TcArgs& a, int64_t N) {
  int64_t ntiles = (N + kTile - 1) / kTile;
  int64_t groups = (ntiles + a.nwg - 1) / a.nwg;
  int threads = (a.nwg + 1) * kThreads;
  int per_sm = kSmemLimit / (a.smem_bytes + 1024);
  if (per_sm > 512 / a.tmem_cols) per_sm = 512 / a.tmem_cols;
  if (per_sm > 2048 / threads) per_sm = 2048 / threads;
  if (per_sm < 1) per_sm = 1;
  int64_t cap = (int64_t)nkbs_num_sms() * per_sm;
  return (int)(groups < cap ? groups : cap);
}

// registers per lane the rebalancing hands out (rebalance_registers) vs. what the kernel was launched with: the
// growth blocks until the pool has room, so a kernel that was allocated less than assumed must not be launched
template <typename K>
int check_register_pool(K kernel, int nwg, const char* who) {
  if (nwg < 2) return 0;
  cudaFuncAttributes fa;
  cudaError_t e = cudaFuncGetAttributes(&fa, kernel);
  NKBS_CHECK_ARG(e == cudaSuccess, "%s: cudaFuncGetAttributes: %s", who, cudaGetErrorString(e));
  const int grown = nwg == 2 ? 224 : (nwg == 3 ? 152 : 104);
  NKBS_CHECK_ARG(fa.numRegs * (nwg + 1) >= nwg * grown + 56,
                 "%s: kernel launched with %d registers per thread, the register rebalancing needs %d", who, fa.numRegs,
                 (nwg * grown + 56 + nwg) / (nwg + 1));
  return 0;
}

template <bool BWD, int FMT, int NWG, int IN>
int launch_one(const TcArgs& a, int64_t N, cudaStream_t st, const char* who) {
  const int grid = pick_grid(a, N), threads = (NWG + 1) * kThreads;
  if (BWD) {
    static const int ok = check_register_pool(mlp_tc_bwd_kernel<FMT, NWG, IN>, NWG, who);
    if (ok) return -1;
    if (set_smem(mlp_tc_bwd_kernel<FMT, NWG, IN>, a.smem_bytes, who)) return -1;
    nkbs_launch(mlp_tc_bwd_kernel<FMT, NWG, IN>, grid, threads, a.smem_bytes, st, a);
  } else {
    static const int ok = check_register_pool(mlp_tc_fwd_kernel<FMT, NWG, IN>, NWG, who);
    if (ok) return -1;
    if (set_smem(mlp_tc_fwd_kernel<FMT, NWG, IN>, a.smem_bytes, who)) return -1;
    nkbs_launch(mlp_tc_fwd_kernel<FMT, NWG, IN>, grid, threads, a.smem_bytes, st, a);
  }
  return 0;
}

template <bool BWD, int IN>
int launch_tc(const TcArgs& a, int64_t N, cudaStream_t st, const char* who) {
  const int key = a.fmt * 8 + a.nwg;
  switch (key) {
    case 1: return launch_one<BWD, 0, 1, IN>(a, N, st, who);
    case 2: return launch_one<BWD, 0, 2, IN>(a, N, st, who);
    case 3: return launch_one<BWD, 0, 3, IN>(a, N, st, who);
    case 4: return launch_one<BWD, 0, 4, IN>(a, N, st, who);
    case 9: return launch_one<BWD, 1, 1, IN>(a, N, st, who);
    case 10: return launch_one<BWD, 1, 2, IN>(a, N, st, who);
    case 11: return launch_one<BWD, 1, 3, IN>(a, N, st, who);
    case 12: return launch_one<BWD, 1, 4, IN>(a, N, st, who);
  }
  nkbs_set_error("%s: unsupported warpgroup count %d", who, a.nwg);
  return -1;
}

int fill_head(TcArgs& a, const char* who, int sh_dim, int geo_dim, int emb_dim, const float* geo, int geo_stride,
              const float* dirs, int dir_per_ray, const int64_t* cam, int cam_per_ray, const float* emb,
              const float* emb_const, int S, int64_t N, int& in_mode) {
  NKBS_CHECK_ARG(geo_dim == 15 && ((sh_dim == 16 && emb_dim == 32) || (sh_dim == 0 && (emb_dim == 16 || emb_dim == 0))),
                 "%s: head inputs are [SH16|geo15|emb32], [geo15|emb16] or [geo15] (got %d/%d/%d)", who, sh_dim, geo_dim, emb_dim);
  in_mode = sh_dim == 16 ? 1 : (emb_dim == 16 ? 2 : 3);
  NKBS_CHECK_ARG(a.layer[0].K == sh_dim + geo_dim + emb_dim, "%s: first layer takes %d inputs, the assembled row has %d",
                 who, a.layer[0].K, sh_dim + geo_dim + emb_dim);
  NKBS_CHECK_ARG(geo && geo_stride >= geo_dim && S >= 1, "%s: bad geo / S", who);
  NKBS_CHECK_ARG(sh_dim == 0 || dirs, "%s: null directions", who);
  NKBS_CHECK_ARG(emb_dim == 0 || (cam && emb) || emb_const, "%s: need camera indices + table, or a constant row", who);
  NKBS_CHECK_ARG(!(dir_per_ray || cam_per_ray) || N % S == 0, "%s: N not a multiple of S", who);
  a.hin.geo = geo;
  a.hin.geo_stride = geo_stride;
  a.hin.dirs = dirs;
  a.hin.dir_per_ray = dir_per_ray;
  a.hin.cam = (emb_const || emb_dim == 0) ? nullptr : cam;
  a.hin.cam_per_ray = cam_per_ray;
  a.hin.emb = emb;
  a.hin.emb_const = emb_const;
  a.hin.S = S;
  // a pointer 4 bytes past a 16-byte boundary is never the start of an allocation: the float before it is readable
  a.hin.geo_vec = (geo_stride % 4 == 0) && (reinterpret_cast<uintptr_t>(geo) & 15) == 4;
  a.hin.emb_vec = emb_dim % 4 == 0 && emb_dim > 0 &&
                  (reinterpret_cast<uintptr_t>(a.hin.cam ? (const void*)emb : (const void*)emb_const) & 15) == 0;
  a.hin.dgeo = nullptr;
  a.hin.dgeo_stride = 0;
  a.hin.dgeo_lead = 0;
  a.hin.demb = nullptr;
  return 0;
}

}  // namespace

#ifdef NKBS_TRACE
extern "C" int nkbs_debug_trace(long long* host, int* n) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(host, g_trace, sizeof(long long) * 2 * 4096);
  cudaMemcpyFromSymbol(n, g_trace_n, sizeof(int) * 2);
  int z[2] = {0, 0};
  cudaMemcpyToSymbol(g_trace_n, z, sizeof(z));
  return 0;
}
#endif

extern "C" int nkbs_mlp_tc_fwd(const float* x, int64_t N, int32_t nl, const int32_t* dims, const float* const* W,
                               const float* const* b, int32_t relu_mask, int32_t out_act, int32_t fmt,
                               const float* selector, float* y, float* y0, void* stream) {
  TcArgs a;
  if (build_args(a, false, nl, dims, W, b, relu_mask, out_act, fmt, "nkbs_mlp_tc_fwd")) return -1;
  NKBS_CHECK_ARG(N >= 0, "nkbs_mlp_tc_fwd: negative N");
  if (N == 0) return 0;
  NKBS_CHECK_ARG(x && y, "nkbs_mlp_tc_fwd: null x/y");
  a.hin = HeadIn{};
  a.x = x;
  a.y = y;
  a.y0 = y0;
  a.selector = selector;
  a.dy = nullptr;
  a.dy0 = nullptr;
  a.dx = nullptr;
  a.N = N;
  if (launch_tc<false, 0>(a, N, (cudaStream_t)stream, "nkbs_mlp_tc_fwd")) return -1;
  NKBS_CHECK_LAUNCH("nkbs_mlp_tc_fwd");
  return 0;
}

extern "C" int nkbs_mlp_tc_bwd(const float* x, const float* dy, int64_t N, int32_t nl, const int32_t* dims,
                               const float* const* W, const float* const* b, int32_t relu_mask, int32_t out_act,
                               int32_t fmt, const float* selector, const float* dy0, float* dx, float* const* dW,
                               float* const* db, void* stream) {
  TcArgs a;
  if (build_args(a, true, nl, dims, W, b, relu_mask, out_act, fmt, "nkbs_mlp_tc_bwd")) return -1;
  NKBS_CHECK_ARG(N >= 0, "nkbs_mlp_tc_bwd: negative N");
  if (N == 0) return 0;
  NKBS_CHECK_ARG(x && dy, "nkbs_mlp_tc_bwd: null x/dy");
  for (int l = 0; l < nl; ++l) {
    NKBS_CHECK_ARG(dW[l] && db[l], "nkbs_mlp_tc_bwd: null dW/db for layer %d", l);
    a.layer[l].dW = dW[l];
    a.layer[l].db = db[l];
    if (l == 0) a.fm = nkbs_fixed_map_for(dW[0]);
  }
  a.hin = HeadIn{};
  a.x = x;
  a.y = nullptr;
  a.y0 = nullptr;
  a.selector = selector;
  a.dy = dy;
  a.dy0 = dy0;
  a.dx = dx;
  a.N = N;
  if (launch_tc<true, 0>(a, N, (cudaStream_t)stream, "nkbs_mlp_tc_bwd")) return -1;
  NKBS_CHECK_LAUNCH("nkbs_mlp_tc_bwd");
  return 0;
}

// ---- head MLPs with the input assembled in the kernel (K4 folded into K3) ----
extern "C" int nkbs_head_tc_fwd(int32_t sh_dim, int32_t geo_dim, int32_t emb_dim, const float* geo, int32_t geo_stride,
                                const float* directions, int32_t dir_per_ray, const int64_t* cam, int32_t cam_per_ray,
                                const float* emb, const float* emb_const, int32_t S, int64_t N, int32_t nl,
                                const int32_t* dims, const float* const* W, const float* const* b, int32_t relu_mask,
                                int32_t out_act, int32_t fmt, float* y, void* stream) {
  TcArgs a;
  if (build_args(a, false, nl, dims, W, b, relu_mask, out_act, fmt, "nkbs_head_tc_fwd")) return -1;
  NKBS_CHECK_ARG(N >= 0, "nkbs_head_tc_fwd: negative N");
  if (N == 0) return 0;
  int in_mode = 0;
  if (fill_head(a, "nkbs_head_tc_fwd", sh_dim, geo_dim, emb_dim, geo, geo_stride, directions, dir_per_ray, cam,
                cam_per_ray, emb, emb_const, S, N, in_mode))
    return -1;
  NKBS_CHECK_ARG(y, "nkbs_head_tc_fwd: null y");
  a.x = nullptr;
  a.y = y;
  a.y0 = nullptr;
  a.selector = nullptr;
  a.dy = nullptr;
  a.dy0 = nullptr;
  a.dx = nullptr;
  a.N = N;
  const int rc = in_mode == 1   ? launch_tc<false, 1>(a, N, (cudaStream_t)stream, "nkbs_head_tc_fwd")
                 : in_mode == 2 ? launch_tc<false, 2>(a, N, (cudaStream_t)stream, "nkbs_head_tc_fwd")
                                : launch_tc<false, 3>(a, N, (cudaStream_t)stream, "nkbs_head_tc_fwd");
  if (rc) return -1;
  NKBS_CHECK_LAUNCH("nkbs_head_tc_fwd");
  return 0;
}

extern "C" int nkbs_head_tc_bwd(int32_t sh_dim, int32_t geo_dim, int32_t emb_dim, const float* geo, int32_t geo_stride,
                                const float* directions, int32_t dir_per_ray, const int64_t* cam, int32_t cam_per_ray,
                                const float* emb, const float* emb_const, int32_t S, const float* dy, int64_t N,
                                int32_t nl, const int32_t* dims, const float* const* W, const float* const* b,
                                int32_t relu_mask, int32_t out_act, int32_t fmt, float* dgeo, int32_t dgeo_stride,
                                int32_t dgeo_lead, float* demb, float* const* dW, float* const* db, void* stream) {
  TcArgs a;
  if (build_args(a, true, nl, dims, W, b, relu_mask, out_act, fmt, "nkbs_head_tc_bwd")) return -1;
  NKBS_CHECK_ARG(N >= 0, "nkbs_head_tc_bwd: negative N");
  if (N == 0) return 0;
  int in_mode = 0;
  if (fill_head(a, "nkbs_head_tc_bwd", sh_dim, geo_dim, emb_dim, geo, geo_stride, directions, dir_per_ray, cam,
                cam_per_ray, emb, emb_const, S, N, in_mode))
    return -1;
  NKBS_CHECK_ARG(dy, "nkbs_head_tc_bwd: null dy");
  NKBS_CHECK_ARG(a.stage_ok, "nkbs_head_tc_bwd: staging does not fit for this MLP shape");
  NKBS_CHECK_ARG(!dgeo || dgeo_stride >= geo_dim, "nkbs_head_tc_bwd: bad dgeo stride");
  for (int l = 0; l < nl; ++l) {
    NKBS_CHECK_ARG(dW[l] && db[l], "nkbs_head_tc_bwd: null dW/db for layer %d", l);
    a.layer[l].dW = dW[l];
    a.layer[l].db = db[l];
    if (l == 0) a.fm = nkbs_fixed_map_for(dW[0]);
  }
  NKBS_CHECK_ARG(dgeo_lead >= 0 && (!dgeo || dgeo_lead + geo_dim <= dgeo_stride), "nkbs_head_tc_bwd: bad dgeo_lead %d", dgeo_lead);
  a.hin.dgeo = dgeo;
  a.hin.dgeo_stride = dgeo_stride;
  a.hin.dgeo_lead = dgeo ? dgeo_lead : 0;
  a.hin.demb = a.hin.cam ? demb : nullptr;
  a.fm_emb = nkbs_fixed_map_for(a.hin.demb);
  a.x = nullptr;
  a.y = nullptr;
  a.y0 = nullptr;
  a.selector = nullptr;
  a.dy = dy;
  a.dy0 = nullptr;
  a.dx = nullptr;
  a.N = N;
  const int rc = in_mode == 1   ? launch_tc<true, 1>(a, N, (cudaStream_t)stream, "nkbs_head_tc_bwd")
                 : in_mode == 2 ? launch_tc<true, 2>(a, N, (cudaStream_t)stream, "nkbs_head_tc_bwd")
                                : launch_tc<true, 3>(a, N, (cudaStream_t)stream, "nkbs_head_tc_bwd");
  if (rc) return -1;
  NKBS_CHECK_LAUNCH("nkbs_head_tc_bwd");
  return 0;
}
