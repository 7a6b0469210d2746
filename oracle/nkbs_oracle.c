/* CPU oracle, C restatement of the integer / bit-exact pieces of the nerfstudio torch path:
 * hash-grid corner indices + trilinear features, the piecewise initial sampler and
 * PDF inversion.
 *
 * TEST INFRASTRUCTURE ONLY: built into oracle/liboracle.so by oracle/Makefile and loaded only by
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
 *
 * PARITY UNPINNED against third-party nerfstudio 0.3.x (not vendored, not installable here; the
 * reference ships no golden vectors).  It is pinned instead against oracle/nerfstudio_torch.py (torch
 * CPU ops) by tests/test_oracle_c.py.  One summation-order choice is fixed here that torch leaves open:
 * the PDF normaliser is the double-precision sum rounded once to float (torch.sum's vectorised float
 * order depends on the host's SIMD width); cumulative sums use a double accumulator rounded to float
 * per element, which IS what torch.cumsum does on the CPU.
 *
 * Call sites restated: ref:nerfacto.py:275-311,325-332,397-398; ref:semantic_nerfw.py:102-137,187-188.
 * Compile with -ffp-contract=off so that no a*b+c is fused.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

/* ---- SceneContraction(order=inf) + (x+2)/4 + selector ---- */
static void normalise(const float p[3], int contraction, const float aabb[6], float q[3], float* sel) {
  if (contraction) {
    float mag = fmaxf(fmaxf(fabsf(p[0]), fabsf(p[1])), fabsf(p[2]));
    for (int i = 0; i < 3; ++i) {
      float v = p[i];
      if (!(mag < 1.0f)) v = (2.0f - (1.0f / mag)) * (p[i] / mag);
      q[i] = (v + 2.0f) / 4.0f;
    }
  } else {
    for (int i = 0; i < 3; ++i) q[i] = (p[i] - aabb[i]) / (aabb[3 + i] - aabb[i]);
  }
  int in = q[0] > 0.f && q[0] < 1.f && q[1] > 0.f && q[1] < 1.f && q[2] > 0.f && q[2] < 1.f;
  *sel = in ? 1.f : 0.f;
  for (int i = 0; i < 3; ++i) q[i] = q[i] * *sel;
}

/* positions = origins + directions * (starts + ends) / 2, then normalise; mode as in nkbs_encode_fwd */
void oracle_positions(int mode, const float* x, const float* origins, const float* directions, const float* starts,
                      const float* ends, int64_t R, int32_t S, int contraction, const float* aabb, float* xn,
                      float* selector) {
  for (int64_t n = 0; n < R * (int64_t)S; ++n) {
    float p[3];
    if (mode == 2) {
      int64_t r = n / S;
      float se = starts[n] + ends[n];
      for (int i = 0; i < 3; ++i) p[i] = origins[3 * r + i] + (directions[3 * r + i] * se) / 2.0f;
    } else {
      for (int i = 0; i < 3; ++i) p[i] = x[3 * n + i];
    }
    if (mode == 0) {
      for (int i = 0; i < 3; ++i) xn[3 * n + i] = p[i];
      selector[n] = 1.f;
    } else {
      normalise(p, contraction, aabb, xn + 3 * n, selector + n);
    }
  }
}

/* ---- HashEncoding.pytorch_fwd: int64 hash, corner order h0..h7, blend order ---- */
static int64_t hash3(int64_t a, int64_t b, int64_t c, int64_t T) {
  int64_t h = (a * 1) ^ (b * 2654435761LL) ^ (c * 805459861LL);
  h %= T;
  if (h < 0) h += T; /* torch's % is a floor-mod */
  return h;
}

void oracle_hash_encode(const float* xn, int64_t N, const float* table, const float* scalings, int32_t L,
                        int32_t log2_T, float* features, int64_t* indices) {
  const int64_t T = (int64_t)1 << log2_T;
  for (int64_t n = 0; n < N; ++n) {
    for (int l = 0; l < L; ++l) {
      int64_t c[3], f[3];
      float o[3];
      for (int i = 0; i < 3; ++i) {
        float s = xn[3 * n + i] * scalings[l];
        c[i] = (int32_t)ceilf(s);
        f[i] = (int32_t)floorf(s);
        o[i] = s - (float)f[i];
      }
      int64_t h[8];
      h[0] = hash3(c[0], c[1], c[2], T);
      h[1] = hash3(c[0], f[1], c[2], T);
      h[2] = hash3(f[0], f[1], c[2], T);
      h[3] = hash3(f[0], c[1], c[2], T);
      h[4] = hash3(c[0], c[1], f[2], T);
      h[5] = hash3(c[0], f[1], f[2], T);
      h[6] = hash3(f[0], f[1], f[2], T);
      h[7] = hash3(f[0], c[1], f[2], T);
      for (int k = 0; k < 8; ++k) {
        h[k] += (int64_t)l * T;
        if (indices) indices[(n * L + l) * 8 + k] = h[k];
      }
      if (features) {
        for (int j = 0; j < 2; ++j) {
          float v[8];
          for (int k = 0; k < 8; ++k) v[k] = table[h[k] * 2 + j];
          float f03 = v[0] * o[0] + v[3] * (1 - o[0]);
          float f12 = v[1] * o[0] + v[2] * (1 - o[0]);
          float f56 = v[5] * o[0] + v[6] * (1 - o[0]);
          float f47 = v[4] * o[0] + v[7] * (1 - o[0]);
          float f0312 = f03 * o[1] + f12 * (1 - o[1]);
          float f4756 = f47 * o[1] + f56 * (1 - o[1]);
          features[n * (2 * L) + 2 * l + j] = f0312 * o[2] + f4756 * (1 - o[2]);
        }
      }
    }
  }
}

/* ---- samplers ---- */
static float spacing_fn(float x) { return x < 1.f ? x / 2.f : 1.f - 1.f / (2.f * x); }
static float spacing_inv(float x) { return x < 0.5f ? 2.f * x : 1.f / (2.f - 2.f * x); }
static float to_euclid(float s, float s_near, float s_far) { return spacing_inv(s * s_far + (1.f - s) * s_near); }

/* lin = torch.linspace(0,1,S+1) supplied by the caller (its rounding is torch's) */
void oracle_initial_sampler(int64_t R, int32_t S, const float* lin, const float* t_rand, int rand_per_bin,
                            const float* nears, const float* fars, float* sbins, float* ebins) {
  for (int64_t r = 0; r < R; ++r) {
    float s_near = spacing_fn(nears[r]), s_far = spacing_fn(fars[r]);
    for (int i = 0; i <= S; ++i) {
      float b = lin[i];
      if (t_rand) {
        float tr = rand_per_bin ? t_rand[r * (S + 1) + i] : t_rand[r];
        float upper = i < S ? (lin[i + 1] + lin[i]) / 2.f : lin[i];
        float lower = i > 0 ? (lin[i] + lin[i - 1]) / 2.f : lin[i];
        b = lower + (upper - lower) * tr;
      }
      sbins[r * (S + 1) + i] = b;
      ebins[r * (S + 1) + i] = to_euclid(b, s_near, s_far);
    }
  }
}

static float nan_to_num0(float t) {
  if (t != t) return 0.f;
  if (t > FLT_MAX) return FLT_MAX;
  if (t < -FLT_MAX) return -FLT_MAX;
  return t;
}

/* PDFSampler with the explicit cdf array and a binary searchsorted(side="right"), i.e. the textbook form
 * (the CUDA kernel uses a merge walk instead; results must agree bit for bit). */
void oracle_pdf_sample(int64_t R, int32_t S, int32_t nb, const float* weights, float anneal, float hist_pad,
                       float eps, const float* sbins, const float* u_base, const float* rand, int rand_per_bin,
                       const float* nears, const float* fars, float* out_sbins, float* out_ebins,
                       int64_t* out_inds) {
  float* w = (float*)malloc(sizeof(float) * (size_t)S);
  float* cdf = (float*)malloc(sizeof(float) * (size_t)(S + 1));
  for (int64_t r = 0; r < R; ++r) {
    double dsum = 0.0;
    for (int s = 0; s < S; ++s) {
      float v = weights[r * S + s];
      if (anneal != 1.0f) v = powf(v, anneal);
      w[s] = v + hist_pad;
      dsum += (double)w[s];
    }
    float sum = (float)dsum;
    float padding = eps - sum > 0.f ? eps - sum : 0.f;
    float pad_each = padding / (float)S;
    sum = sum + padding;
    double run = 0.0;
    cdf[0] = 0.f;
    for (int s = 0; s < S; ++s) {
      float pdf = (w[s] + pad_each) / sum;
      run += (double)pdf;
      float c = (float)run;
      cdf[s + 1] = c < 1.f ? c : 1.f;
    }
    float s_near = spacing_fn(nears[r]), s_far = spacing_fn(fars[r]);
    const float* eb = sbins + r * (S + 1);
    for (int i = 0; i < nb; ++i) {
      float u = u_base[i];
      if (rand)
        u = u + (rand_per_bin ? rand[r * nb + i] : rand[r]) / (float)nb;
      else
        u = u + (float)(1.0 / (2.0 * (double)nb));
      /* searchsorted right: first index with cdf[idx] > u */
      int lo = 0, hi = S + 1;
      while (lo < hi) {
        int mid = (lo + hi) / 2;
        if (cdf[mid] <= u) lo = mid + 1; else hi = mid;
      }
      int idx = lo;
      int below = idx - 1 < 0 ? 0 : (idx - 1 > S ? S : idx - 1);
      int above = idx < 0 ? 0 : (idx > S ? S : idx);
      float t = nan_to_num0((u - cdf[below]) / (cdf[above] - cdf[below]));
      t = t < 0.f ? 0.f : (t > 1.f ? 1.f : t);
      float b = eb[below] + t * (eb[above] - eb[below]);
      out_sbins[r * nb + i] = b;
      out_ebins[r * nb + i] = to_euclid(b, s_near, s_far);
      if (out_inds) out_inds[r * nb + i] = idx;
    }
  }
  free(w);
  free(cdf);
}
