// K3 (fp32 path): fused small-MLP forward / backward on CUDA cores, weights resident in shared
// memory, activations in shared memory as [feature][sample] columns.  This is the full-precision path
// (fp32 parity 1e-4 against the oracle); the tcgen05 tensor-core path lives in mlp_tc.cu.
// Replaces nerfstudio field_components.mlp.MLP (+ the field-head activations) reached from
// ref:nerfacto.py:275-311,398 and ref:semantic_nerfw.py:102-126,188.
#include "common.cuh"

namespace {

constexpr int kTile = 128;     // samples per CTA tile
constexpr int kTS = 129;       // shared-memory row stride (one pad word: conflict-free transposes)
constexpr int kThreads = 256;  // two threads per sample, each owns half of the output neurons
constexpr int kMaxWidth = 64;
constexpr int kMaxLayers = 3;

enum OutAct { ACT_NONE = 0, ACT_SIGMOID = 1, ACT_DENSITY = 2, ACT_TRANSIENT = 3 };

struct Layer {
  const float* W;  // [N,K] torch nn.Linear layout
  const float* b;  // [N]
  float* dW;
  float* db;
  int K, N, NC, Npad;
  int relu;        // ReLU after this layer
  int w_off, b_off;  // offsets (floats) into the shared weight arena
};

struct MlpArgs {
  int nl;
  Layer layer[kMaxLayers];
  const float* x;         // [N,K0]
  const float* selector;  // [N] or null (ACT_DENSITY)
  float* y;               // [N,Kn]
  float* y0;              // [N] or null: column 0 of y once more, contiguous
  const float* dy;        // [N,Kn]
  const float* dy0;       // [N] or null: added to column 0 of dy
  float* dx;              // [N,K0] or null
  FixedMap fm;            // backward, deterministic mode (common.cuh): shadow of the gradient buffer, or a null map
  int out_act;
  int64_t N;
  int w_floats;           // total floats of the weight arena
};

__device__ __forceinline__ void stage_weights(const MlpArgs& a, float* sw) {
  for (int l = 0; l < a.nl; ++l) {
    const Layer& ly = a.layer[l];
    float* wt = sw + ly.w_off;  // [K][Npad]
    for (int i = threadIdx.x; i < ly.K * ly.Npad; i += kThreads) {
      int k = i / ly.Npad, n = i - k * ly.Npad;
      wt[i] = n < ly.N ? __ldg(ly.W + (size_t)n * ly.K + k) : 0.f;
    }
    float* bs = sw + ly.b_off;
    for (int i = threadIdx.x; i < ly.Npad; i += kThreads) bs[i] = i < ly.N ? __ldg(ly.b + i) : 0.f;
  }
}

// global [rows, K] row-major  ->  shared [K][kTS]; coalesced reads, conflict-free transposed writes
__device__ __forceinline__ void load_tile(const float* g, int64_t row0, int64_t rows_total, int K, float* s) {
  int64_t rows = rows_total - row0;
  if (rows > kTile) rows = kTile;
  const float* src = g + row0 * K;
  int total = kTile * K;
  int valid = (int)rows * K;
  for (int e = threadIdx.x; e < total; e += kThreads) {
    int r = e / K, k = e - r * K;
    s[k * kTS + r] = e < valid ? __ldg(src + e) : 0.f;
  }
}

__device__ __forceinline__ void store_tile(float* g, int64_t row0, int64_t rows_total, int K, const float* s) {
  int64_t rows = rows_total - row0;
  if (rows > kTile) rows = kTile;
  float* dst = g + row0 * K;
  int valid = (int)rows * K;
  for (int e = threadIdx.x; e < valid; e += kThreads) {
    int r = e / K, k = e - r * K;
    dst[e] = s[k * kTS + r];
  }
}

// y[n] = act(b[n] + sum_k x[k] W[n,k]) for one sample column t; the two halves of the CTA take
// alternate NC-wide chunks of outputs.  Per k: one column read + NC/4 broadcast LDS.128 + NC FFMA.
template <int NC>
__device__ __forceinline__ void dense_fwd(int K, int Npad, const float* __restrict__ xin,
                                          const float* __restrict__ wt, const float* __restrict__ bs,
                                          float* __restrict__ yout, int relu, int t, int half) {
  const int nchunks = Npad / NC;
  for (int c = half; c < nchunks; c += 2) {
    float acc[NC];
#pragma unroll
    for (int j = 0; j < NC; ++j) acc[j] = bs[c * NC + j];
    const float* wrow = wt + c * NC;
#pragma unroll 2
    for (int k = 0; k < K; ++k) {
      float xv = xin[k * kTS + t];
#pragma unroll
      for (int j = 0; j < NC; j += 4) {
        float4 w = *reinterpret_cast<const float4*>(wrow + k * Npad + j);
        acc[j] = fmaf(xv, w.x, acc[j]);
        acc[j + 1] = fmaf(xv, w.y, acc[j + 1]);
        acc[j + 2] = fmaf(xv, w.z, acc[j + 2]);
        acc[j + 3] = fmaf(xv, w.w, acc[j + 3]);
      }
    }
#pragma unroll
    for (int j = 0; j < NC; ++j) {
      float v = acc[j];
      if (relu) v = fmaxf(v, 0.f);
      yout[(c * NC + j) * kTS + t] = v;
    }
  }
}

__device__ __forceinline__ void dense_fwd_dispatch(const Layer& ly, const float* xin, const float* sw, float* yout,
                                                   int t, int half) {
  const float* wt = sw + ly.w_off;
  const float* bs = sw + ly.b_off;
  switch (ly.NC) {
    case 32: dense_fwd<32>(ly.K, ly.Npad, xin, wt, bs, yout, ly.relu, t, half); break;
    case 16: dense_fwd<16>(ly.K, ly.Npad, xin, wt, bs, yout, ly.relu, t, half); break;
    case 8: dense_fwd<8>(ly.K, ly.Npad, xin, wt, bs, yout, ly.relu, t, half); break;
    default: dense_fwd<4>(ly.K, ly.Npad, xin, wt, bs, yout, ly.relu, t, half); break;
  }
}

// output activation applied in place on the last layer's raw outputs (rows n < N)
__device__ __forceinline__ float apply_out_act(int act, int n, float raw, float sel) {
  switch (act) {
    case ACT_SIGMOID: return nkbs_sigmoid(raw);
    case ACT_DENSITY: return n == 0 ? expf(raw) * sel : raw;
    case ACT_TRANSIENT: return (n >= 1 && n <= 3) ? nkbs_sigmoid(raw) : nkbs_softplus(raw);
    default: return raw;
  }
}

// d(out)/d(raw) * dy
__device__ __forceinline__ float out_act_grad(int act, int n, float raw, float sel, float dy) {
  switch (act) {
    case ACT_SIGMOID: {
      float s = nkbs_sigmoid(raw);
      return dy * s * (1.f - s);
    }
    case ACT_DENSITY:
      // trunc_exp backward: g * exp(clamp(raw, -15, 15)); then the selector product
      return n == 0 ? dy * sel * expf(fminf(fmaxf(raw, -15.f), 15.f)) : dy;
    case ACT_TRANSIENT:
      if (n >= 1 && n <= 3) {
        float s = nkbs_sigmoid(raw);
        return dy * s * (1.f - s);
      } else {
        return raw > 20.f ? dy : dy * nkbs_sigmoid(raw);
      }
    default: return dy;
  }
}

__global__ void __launch_bounds__(kThreads) mlp_fwd_kernel(const MlpArgs a) {
  nkbs_pdl_sync();
  extern __shared__ __align__(16) float smem[];
  float* sw = smem;
  float* buf0 = smem + a.w_floats;
  float* buf1 = buf0 + kMaxWidth * kTS;
  stage_weights(a, sw);
  const int t = threadIdx.x & (kTile - 1), half = threadIdx.x >> 7;
  const int64_t ntiles = (a.N + kTile - 1) / kTile;
  const int Kn = a.layer[a.nl - 1].N;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t row0 = tile * kTile;
    __syncthreads();  // weights staged / previous tile's stores done
    load_tile(a.x, row0, a.N, a.layer[0].K, buf0);
    __syncthreads();
    float* cur = buf0;
    float* nxt = buf1;
    for (int l = 0; l < a.nl; ++l) {
      dense_fwd_dispatch(a.layer[l], cur, sw, nxt, t, half);
      __syncthreads();
      float* tmp = cur;
      cur = nxt;
      nxt = tmp;
    }
    if (a.out_act != ACT_NONE) {
      float sel = 1.f;
      if (a.selector && row0 + t < a.N) sel = a.selector[row0 + t];
      for (int n = half; n < Kn; n += 2) cur[n * kTS + t] = apply_out_act(a.out_act, n, cur[n * kTS + t], sel);
      __syncthreads();
    }
    if (a.y0 && half == 0 && row0 + t < a.N) a.y0[row0 + t] = cur[t];  // feature 0 of sample t
    store_tile(a.y, row0, a.N, Kn, cur);
  }
}

// gx[k] = sum_n gy[n] W[n,k] (times the ReLU mask of the layer input).  The upstream gradient of a
// chunk of NC outputs sits in registers; the two halves of the CTA split the k range.
template <int NC>
__device__ __forceinline__ void dense_bwd_data(int K, int Npad, const float* __restrict__ gy,
                                               const float* __restrict__ wt, float* __restrict__ gx,
                                               const float* __restrict__ xin, int relu_in, int t, int half) {
  const int nchunks = Npad / NC;
  const int kh = (K + 1) >> 1;
  const int k_lo = half * kh;
  const int k_hi = (k_lo + kh < K) ? k_lo + kh : K;
  for (int c = 0; c < nchunks; ++c) {
    float d[NC];
#pragma unroll
    for (int j = 0; j < NC; ++j) d[j] = gy[(c * NC + j) * kTS + t];
    const bool last = (c == nchunks - 1);
#pragma unroll 2
    for (int k = k_lo; k < k_hi; ++k) {
      const float* wrow = wt + k * Npad + c * NC;
      float a0 = (c == 0) ? 0.f : gx[k * kTS + t], a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
      for (int j = 0; j < NC; j += 4) {
        float4 w = *reinterpret_cast<const float4*>(wrow + j);
        a0 = fmaf(d[j], w.x, a0);
        a1 = fmaf(d[j + 1], w.y, a1);
        a2 = fmaf(d[j + 2], w.z, a2);
        a3 = fmaf(d[j + 3], w.w, a3);
      }
      float g = (a0 + a1) + (a2 + a3);
      if (last && relu_in && !(xin[k * kTS + t] > 0.f)) g = 0.f;
      gx[k * kTS + t] = g;
    }
  }
}

__device__ __forceinline__ void dense_bwd_data_dispatch(const Layer& ly, const float* gy, const float* sw, float* gx,
                                                        const float* xin, int relu_in, int t, int half) {
  const float* wt = sw + ly.w_off;
  switch (ly.NC) {
    case 32: dense_bwd_data<32>(ly.K, ly.Npad, gy, wt, gx, xin, relu_in, t, half); break;
    case 16: dense_bwd_data<16>(ly.K, ly.Npad, gy, wt, gx, xin, relu_in, t, half); break;
    case 8: dense_bwd_data<8>(ly.K, ly.Npad, gy, wt, gx, xin, relu_in, t, half); break;
    default: dense_bwd_data<4>(ly.K, ly.Npad, gy, wt, gx, xin, relu_in, t, half); break;
  }
}

// dW[n,k] += sum_t x[k,t] gy[n,t]: thread (kb, nb) owns a 4x4 block of the weight gradient and keeps
// it in registers across all tiles of this persistent CTA.
__device__ __forceinline__ void dense_bwd_weight(int K, int Npad, const float* __restrict__ xin,
                                                 const float* __restrict__ gy, float accW[16], float accB[4]) {
  const int kb = threadIdx.x & 15, nb = threadIdx.x >> 4;
  const int k0 = kb * 4, n0 = nb * 4;
  if (k0 >= K || n0 >= Npad) return;
  const float* xr[4];
  bool xv[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    xv[i] = (k0 + i) < K;
    xr[i] = xin + (xv[i] ? (k0 + i) : k0) * kTS;
  }
  const float* gr = gy + n0 * kTS;
#pragma unroll 4
  for (int t = 0; t < kTile; ++t) {
    float x[4], d[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) x[i] = xv[i] ? xr[i][t] : 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) d[j] = gr[j * kTS + t];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) accW[i * 4 + j] = fmaf(x[i], d[j], accW[i * 4 + j]);
    if (kb == 0) {
#pragma unroll
      for (int j = 0; j < 4; ++j) accB[j] += d[j];
    }
  }
}

__device__ __forceinline__ void flush_weight_grad(const Layer& ly, const FixedMap& fm, const float accW[16],
                                                  const float accB[4]) {
  const int kb = threadIdx.x & 15, nb = threadIdx.x >> 4;
  const int k0 = kb * 4, n0 = nb * 4;
  if (k0 >= ly.K || n0 >= ly.N) return;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (k0 + i < ly.K && n0 + j < ly.N) nkbs_acc(ly.dW + (size_t)(n0 + j) * ly.K + k0 + i, fm, accW[i * 4 + j]);
  if (kb == 0) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (n0 + j < ly.N) nkbs_acc(ly.db + n0 + j, fm, accB[j]);
  }
}

__global__ void __launch_bounds__(kThreads) mlp_bwd_kernel(const MlpArgs a) {
  nkbs_pdl_sync();
  extern __shared__ __align__(16) float smem[];
  float* sw = smem;
  float* acts[kMaxLayers];
  float* p = smem + a.w_floats;
#pragma unroll
  for (int l = 0; l < kMaxLayers; ++l) {
    acts[l] = p;
    if (l < a.nl) p += kMaxWidth * kTS;
  }
  float* gA = p;
  float* gB = p + kMaxWidth * kTS;
  stage_weights(a, sw);
  const int t = threadIdx.x & (kTile - 1), half = threadIdx.x >> 7;
  const int64_t ntiles = (a.N + kTile - 1) / kTile;
  const Layer& last = a.layer[a.nl - 1];
  const int Kn = last.N;
  float accW[kMaxLayers][16];
  float accB[kMaxLayers][4];
#pragma unroll
  for (int l = 0; l < kMaxLayers; ++l) {
#pragma unroll
    for (int i = 0; i < 16; ++i) accW[l][i] = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) accB[l][i] = 0.f;
  }
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t row0 = tile * kTile;
    __syncthreads();
    load_tile(a.x, row0, a.N, a.layer[0].K, acts[0]);
    load_tile(a.dy, row0, a.N, Kn, gB);
    __syncthreads();
    if (a.dy0 && half == 0 && row0 + t < a.N) gB[t] += __ldg(a.dy0 + row0 + t);  // read after the barriers of the recompute
    // recompute the forward pass, keeping every layer input
#pragma unroll
    for (int l = 0; l < kMaxLayers; ++l) {
      if (l < a.nl) {
        float* out = (l == a.nl - 1) ? gA : acts[(l + 1 < kMaxLayers) ? l + 1 : l];
        dense_fwd_dispatch(a.layer[l], acts[l], sw, out, t, half);
        __syncthreads();
      }
    }
    // gradient w.r.t. the raw outputs (padding rows are exact zeros)
    {
      float sel = 1.f;
      if (a.selector && row0 + t < a.N) sel = a.selector[row0 + t];
      for (int n = half; n < last.Npad; n += 2) {
        float g = 0.f;
        if (n < Kn) g = out_act_grad(a.out_act, n, gA[n * kTS + t], sel, gB[n * kTS + t]);
        gA[n * kTS + t] = g;
      }
      __syncthreads();
    }
    float* gy = gA;
    float* gx = gB;
#pragma unroll
    for (int l = kMaxLayers - 1; l >= 0; --l) {
      if (l < a.nl) {
        const Layer& ly = a.layer[l];
        dense_bwd_weight(ly.K, ly.Npad, acts[l], gy, accW[l], accB[l]);
        if (l > 0 || a.dx) {
          int relu_in = (l > 0) ? a.layer[l > 0 ? l - 1 : 0].relu : 0;
          dense_bwd_data_dispatch(ly, gy, sw, gx, acts[l], relu_in, t, half);
          // rows K..Npad_prev-1 of the next gy must be zero for the weight-gradient pass
          if (l > 0) {
            const Layer& lp = a.layer[l - 1];
            for (int n = lp.N + half; n < lp.Npad; n += 2) gx[n * kTS + t] = 0.f;
          }
        }
        __syncthreads();
        float* tmp = gy;
        gy = gx;
        gx = tmp;
      }
    }
    if (a.dx) store_tile(a.dx, row0, a.N, a.layer[0].K, gy);
  }
#pragma unroll
  for (int l = 0; l < kMaxLayers; ++l)
    if (l < a.nl) flush_weight_grad(a.layer[l], a.fm, accW[l], accB[l]);
}

int pick_nc(int n) {
  if (n <= 4) return 4;
  if (n <= 8) return 8;
  if (n <= 16) return 16;
  return 32;
}

int build_args(MlpArgs& a, int nl, const int32_t* dims, const float* const* W, const float* const* b,
               int relu_mask, int out_act, const char* who) {
  NKBS_CHECK_ARG(nl >= 1 && nl <= kMaxLayers, "%s: 1..%d linear layers supported (got %d)", who, kMaxLayers, nl);
  NKBS_CHECK_ARG(out_act >= 0 && out_act <= 3, "%s: unknown output activation %d", who, out_act);
  int off = 0;
  for (int l = 0; l <= nl; ++l)
    NKBS_CHECK_ARG(dims[l] >= 1 && dims[l] <= kMaxWidth, "%s: layer width %d outside 1..%d", who, dims[l], kMaxWidth);
  for (int l = 0; l < nl; ++l) {
    Layer& ly = a.layer[l];
    NKBS_CHECK_ARG(W[l] && b[l], "%s: null weight/bias for layer %d", who, l);
    ly.W = W[l];
    ly.b = b[l];
    ly.dW = nullptr;
    ly.db = nullptr;
    a.fm = FixedMap{nullptr, nullptr, 0, 0.f};
    ly.K = dims[l];
    ly.N = dims[l + 1];
    ly.NC = pick_nc(ly.N);
    ly.Npad = (ly.N + ly.NC - 1) / ly.NC * ly.NC;
    ly.relu = (relu_mask >> l) & 1;
    ly.w_off = off;
    off += ly.K * ly.Npad;
    ly.b_off = off;
    off += ly.Npad;
  }
  a.nl = nl;
  a.w_floats = (off + 3) & ~3;
  a.out_act = out_act;
  return 0;
}

}  // namespace

extern "C" int nkbs_mlp_fwd(const float* x, int64_t N, int32_t nl, const int32_t* dims, const float* const* W,
                            const float* const* b, int32_t relu_mask, int32_t out_act, const float* selector,
                            float* y, float* y0, void* stream) {
  MlpArgs a;
  if (build_args(a, nl, dims, W, b, relu_mask, out_act, "nkbs_mlp_fwd")) return -1;
  NKBS_CHECK_ARG(N >= 0, "nkbs_mlp_fwd: negative N");
  if (N == 0) return 0;
  NKBS_CHECK_ARG(x && y, "nkbs_mlp_fwd: null x/y");
  a.x = x;
  a.y = y;
  a.y0 = y0;
  a.selector = selector;
  a.dy = nullptr;
  a.dy0 = nullptr;
  a.dx = nullptr;
  a.N = N;
  size_t smem = (size_t)(a.w_floats + 2 * kMaxWidth * kTS) * sizeof(float);
  static thread_local size_t configured = 0;
  if (smem > configured) {
    cudaFuncSetAttribute(mlp_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured = smem;
  }
  int64_t ntiles = (N + kTile - 1) / kTile;
  int64_t cap = (int64_t)nkbs_num_sms() * 2;
  int grid = (int)(ntiles < cap ? ntiles : cap);
  nkbs_launch(mlp_fwd_kernel, grid, kThreads, smem, (cudaStream_t)stream, a);
  NKBS_CHECK_LAUNCH("nkbs_mlp_fwd");
  return 0;
}

extern "C" int nkbs_mlp_bwd(const float* x, const float* dy, int64_t N, int32_t nl, const int32_t* dims,
                            const float* const* W, const float* const* b, int32_t relu_mask, int32_t out_act,
                            const float* selector, const float* dy0, float* dx, float* const* dW, float* const* db,
                            void* stream) {
  MlpArgs a;
  if (build_args(a, nl, dims, W, b, relu_mask, out_act, "nkbs_mlp_bwd")) return -1;
  NKBS_CHECK_ARG(N >= 0, "nkbs_mlp_bwd: negative N");
  if (N == 0) return 0;
  NKBS_CHECK_ARG(x && dy, "nkbs_mlp_bwd: null x/dy");
  for (int l = 0; l < nl; ++l) {
    NKBS_CHECK_ARG(dW[l] && db[l], "nkbs_mlp_bwd: null dW/db for layer %d", l);
    a.layer[l].dW = dW[l];
    a.layer[l].db = db[l];
    if (l == 0) a.fm = nkbs_fixed_map_for(dW[0]);
  }
  a.x = x;
  a.y = nullptr;
  a.y0 = nullptr;
  a.selector = selector;
  a.dy = dy;
  a.dy0 = dy0;
  a.dx = dx;
  a.N = N;
  size_t smem = (size_t)(a.w_floats + (nl + 2) * kMaxWidth * kTS) * sizeof(float);
  static thread_local size_t configured = 0;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(mlp_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    NKBS_CHECK_ARG(e == cudaSuccess, "nkbs_mlp_bwd: %zu bytes of shared memory refused: %s", smem, cudaGetErrorString(e));
    configured = smem;
  }
  int64_t ntiles = (N + kTile - 1) / kTile;
  int64_t cap = (int64_t)nkbs_num_sms();
  int grid = (int)(ntiles < cap ? ntiles : cap);
  nkbs_launch(mlp_bwd_kernel, grid, kThreads, smem, (cudaStream_t)stream, a);
  NKBS_CHECK_LAUNCH("nkbs_mlp_bwd");
  return 0;
}
