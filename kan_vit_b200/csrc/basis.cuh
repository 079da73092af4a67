// On-the-fly basis expansions Phi_m(x) and their derivatives for the five KAN families
// (+ identity for nn.Linear).  Every GEMM kernel in this library calls these from its
// prologue / epilogue; the [T, in*M] basis tensor is never materialised in HBM.
//
// A "sink" receives (m, value) pairs: sink.put(m, v).  Dense sinks (kDense = true) are
// given every m in [0, M) exactly once (zeros included); sparse sinks only non-zeros.
#pragma once
#include "common.cuh"

namespace kv {

struct BasisCtx {
  const float* fp;  // this group's prologue parameters (see kv_kan_fparam_count)
  int in;           // in_features
  int G;            // grid_size | degree | num_grids
  int k;            // spline order (effkan)
  int M;            // basis count incl. base branch
  float rbf_denom;  // fastkan denominator
  uint32_t flags;
  // effkan, tensor-core kernels only (make_ctx_tc): every feature shares one uniform knot vector t0 + j h, so the span
  // index and the de Boor triangle need no knot loads.  The exact fp32 kernels keep uniform = 0 (stored knots).
  int uniform;
  float t0, inv_h;
};

__device__ __forceinline__ BasisCtx make_ctx(const kv_kan_desc_t& d, const float* fparams, int g) {
  BasisCtx c;
  c.fp = fparams ? fparams + (size_t)g * d.fparam_stride : nullptr;
  c.in = d.in_features;
  c.G = d.grid_size;
  c.k = d.spline_order;
  c.M = d.M;
  c.rbf_denom = d.rbf_denom;
  c.flags = d.flags;
  c.uniform = 0;
  c.t0 = 0.0f;
  c.inv_h = 0.0f;
  return c;
}

// tensor-core kernels: as make_ctx, plus the uniform-grid probe kv_kan_pack wrote behind the knots of an effkan group
__device__ __forceinline__ BasisCtx make_ctx_tc(const kv_kan_desc_t& d, const float* fparams, int g) {
  BasisCtx c = make_ctx(d, fparams, g);
  if (d.family == KV_EFFKAN && c.fp != nullptr) {
    const float* probe = c.fp + (size_t)d.in_features * (d.grid_size + 2 * d.spline_order + 1);
    c.uniform = probe[0] == 1.0f ? 1 : 0;
    c.t0 = probe[1];
    c.inv_h = probe[2];
  }
  return c;
}

__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + expf(-x)); }
__device__ __forceinline__ float silu_grad_f(float x) {
  float s = 1.0f / (1.0f + expf(-x));
  return s * (1.0f + x * (1.0f - s));
}

// FastKAN: value fed to the RBFs = LayerNorm(x)_i (models/fastkan.py:57,68) or raw x.
__device__ __forceinline__ float fastkan_xn(const BasisCtx& c, int i, float x, float mean, float rstd) {
  if (c.flags & KV_FLAG_NO_LAYERNORM) return x;
  return (x - mean) * rstd * c.fp[i] + c.fp[c.in + i];
}

// ---------------------------------------------------------------------------------
// B-splines: local Cox-de Boor triangle on the knot span that contains x.
// Same arithmetic as the reference's dense recursion for the (at most k+1) non-zero
// functions (models/effkan.py:115-125): terms it multiplies by a zero lower-order
// basis are simply not formed.  Returns j (index of the order-0 indicator that is 1)
// or -1 when x lies outside [t_0, t_last) (all bases zero, effkan.py:115).
// N[r], r = 0..level, holds B_{j-level+r, level}.
// ---------------------------------------------------------------------------------
__device__ __forceinline__ int bspline_find_span(const float* __restrict__ t, int nknots, float x) {
  const float t0 = t[0], tl = t[nknots - 1];
  if (!(x >= t0) || !(x < tl)) return -1;
  // the largest j <= nknots - 2 with t[j] <= x: a guess from the grid ends (exact on a uniform grid up to rounding),
  // then a walk -- a few steps on the quantile grids of update_grid, instead of one comparison per knot
  int j = (int)((x - t0) * (float)(nknots - 1) / (tl - t0));
  j = j < 0 ? 0 : (j > nknots - 2 ? nknots - 2 : j);
  while (j > 0 && x < t[j]) --j;
  while (j < nknots - 2 && x >= t[j + 1]) ++j;
  return j;
}

// advance the triangle from level p-1 to level p (in place, high r to low r)
__device__ __forceinline__ void bspline_level(const float* __restrict__ t, int nknots, int j, int p,
                                              float x, float (&N)[KV_MAX_SPLINE_ORDER + 1]) {
#pragma unroll
  for (int r = KV_MAX_SPLINE_ORDER; r >= 0; --r) {
    if (r <= p) {
      int a = j - p + r;  // function index at level p
      float v = 0.0f;
      if (a >= 0 && a + p + 1 <= nknots - 1) {
        float lo = (r >= 1) ? N[r >= 1 ? r - 1 : 0] : 0.0f;  // B_{a,p-1}
        float hi = (r <= p - 1) ? N[r] : 0.0f;               // B_{a+1,p-1}
        float left = (x - t[a]) / (t[a + p] - t[a]);
        float right = (t[a + p + 1] - x) / (t[a + p + 1] - t[a + 1]);
        // a zero lower-order basis contributes exactly 0 in the reference (finite
        // coefficient times 0); skip the term so that 0/0 knot gaps cannot leak in.
        v = (lo != 0.0f ? left * lo : 0.0f) + (hi != 0.0f ? right * hi : 0.0f);
      }
      N[r] = v;
    }
  }
}

template <class Sink>
__device__ __forceinline__ void effkan_eval(const BasisCtx& c, int i, float x, Sink& s) {
  const int nknots = c.G + 2 * c.k + 1;
  const float* t = c.fp + (size_t)i * nknots;
  const int nb = c.G + c.k;
  if (Sink::kDense) {
    for (int m = 0; m < nb; ++m) s.put(m, 0.0f);
  }
  int j = bspline_find_span(t, nknots, x);
  if (j >= 0) {
    float N[KV_MAX_SPLINE_ORDER + 1];
#pragma unroll
    for (int r = 0; r <= KV_MAX_SPLINE_ORDER; ++r) N[r] = 0.0f;
    N[0] = 1.0f;
#pragma unroll
    for (int p = 1; p <= KV_MAX_SPLINE_ORDER; ++p)
      if (p <= c.k) bspline_level(t, nknots, j, p, x, N);
#pragma unroll
    for (int r = 0; r <= KV_MAX_SPLINE_ORDER; ++r) {
      int a = j - c.k + r;
      if (r <= c.k && a >= 0 && a < nb) s.put(a, N[r]);
    }
  }
  s.put(nb, silu_f(x));
}

// d/dx: B'_{a,k} = k/(t_{a+k}-t_a) B_{a,k-1} - k/(t_{a+k+1}-t_{a+1}) B_{a+1,k-1}
template <class Sink>
__device__ __forceinline__ void effkan_deriv(const BasisCtx& c, int i, float x, Sink& s) {
  const int nknots = c.G + 2 * c.k + 1;
  const float* t = c.fp + (size_t)i * nknots;
  const int nb = c.G + c.k;
  if (Sink::kDense) {
    for (int m = 0; m < nb; ++m) s.put(m, 0.0f);
  }
  if (c.uniform && c.k >= 1) {
    // uniform knots (tensor-core kernels): span and triangle in units of h, no knot loads;
    // B'_{a,k} = (B_{a,k-1} - B_{a+1,k-1}) / h
    const float sx = (x - c.t0) * c.inv_h;
    if (sx >= 0.0f && sx < (float)(nknots - 1)) {
      const int j = min((int)sx, nknots - 2);
      const float u = sx - (float)j;
      float N[KV_MAX_SPLINE_ORDER + 1];
#pragma unroll
      for (int r = 0; r <= KV_MAX_SPLINE_ORDER; ++r) N[r] = 0.0f;
      N[0] = 1.0f;
#pragma unroll
      for (int p = 1; p <= KV_MAX_SPLINE_ORDER; ++p) {
        if (p <= c.k - 1) {
          float saved = 0.0f;
          const float ip = 1.0f / (float)p;
#pragma unroll
          for (int r = 0; r < p; ++r) {
            const float temp = N[r] * ip;
            N[r] = saved + ((float)(r + 1) - u) * temp;
            saved = (u + (float)(p - r - 1)) * temp;
          }
          N[p] = saved;
        }
      }
#pragma unroll
      for (int r = 0; r <= KV_MAX_SPLINE_ORDER; ++r) {
        const int a = j - c.k + r;
        if (r <= c.k && a >= 0 && a < nb) {
          const float lo = (r >= 1) ? N[r >= 1 ? r - 1 : 0] : 0.0f;
          const float hi = (r <= c.k - 1) ? N[r] : 0.0f;
          s.put(a, (lo - hi) * c.inv_h);
        }
      }
    }
    s.put(nb, silu_grad_f(x));
    return;
  }
  int j = bspline_find_span(t, nknots, x);
  if (j >= 0 && c.k >= 1) {
    float N[KV_MAX_SPLINE_ORDER + 1];
#pragma unroll
    for (int r = 0; r <= KV_MAX_SPLINE_ORDER; ++r) N[r] = 0.0f;
    N[0] = 1.0f;
#pragma unroll
    for (int p = 1; p <= KV_MAX_SPLINE_ORDER; ++p)
      if (p <= c.k - 1) bspline_level(t, nknots, j, p, x, N);
    // N[r'] = B_{j-(k-1)+r', k-1}, r' = 0..k-1
    const float kf = (float)c.k;
#pragma unroll
    for (int r = 0; r <= KV_MAX_SPLINE_ORDER; ++r) {
      int a = j - c.k + r;
      if (r <= c.k && a >= 0 && a < nb) {
        float lo = (r >= 1) ? N[r >= 1 ? r - 1 : 0] : 0.0f;
        float hi = (r <= c.k - 1) ? N[r] : 0.0f;
        float v = 0.0f;
        if (lo != 0.0f) v += kf / (t[a + c.k] - t[a]) * lo;
        if (hi != 0.0f) v -= kf / (t[a + c.k + 1] - t[a + 1]) * hi;
        s.put(a, v);
      }
    }
  }
  s.put(nb, silu_grad_f(x));
}

// ---------------------------------------------------------------------------------
// family dispatch
// ---------------------------------------------------------------------------------
template <int F, class Sink>
__device__ __forceinline__ void basis_eval(const BasisCtx& c, int i, float x, float xn, Sink& s) {
  if (F == KV_LINEAR) {
    s.put(0, x);
  } else if (F == KV_EFFKAN) {
    effkan_eval(c, i, x, s);
  } else if (F == KV_CHEBY) {
    // T_n(tanh x) by the three-term recurrence (ChebyKAN/ChebyKANLayer_unoptimized.py:22-27);
    // the reference evaluates cos(n acos(tanh x)) (models/cheby.py:37-44).
    float t = tanhf(x);
    float t0 = 1.0f, t1 = t;
    s.put(0, 1.0f);
    if (c.M > 1) s.put(1, t);
    for (int n = 2; n < c.M; ++n) {
      float t2 = 2.0f * t * t1 - t0;
      s.put(n, t2);
      t0 = t1;
      t1 = t2;
    }
  } else if (F == KV_FASTKAN) {
    const float* g = c.fp + 2 * c.in;
    for (int j = 0; j < c.G; ++j) {
      float z = (xn - g[j]) / c.rbf_denom;
      s.put(j, expf(-(z * z)));
    }
    if (!(c.flags & KV_FLAG_NO_BASE)) s.put(c.G, silu_f(x));
  } else if (F == KV_SINE) {
    const float* freq = c.fp;
    const float* ph = c.fp + c.G + (size_t)i * c.G;
    for (int g = 0; g < c.G; ++g) s.put(g, sinf(x * freq[g] + ph[g]));
  } else if (F == KV_FOURIER) {
    for (int k = 1; k <= c.G; ++k) {
      float sn, cs;
      sincosf((float)k * x, &sn, &cs);
      s.put(k - 1, cs);
      s.put(c.G + k - 1, sn);
    }
  }
}

// derivative of each basis function w.r.t. its argument: x, except fastkan's RBFs
// (m < G) which are differentiated w.r.t. xn = LayerNorm(x)_i.
template <int F, class Sink>
__device__ __forceinline__ void basis_deriv(const BasisCtx& c, int i, float x, float xn, Sink& s) {
  if (F == KV_LINEAR) {
    s.put(0, 1.0f);
  } else if (F == KV_EFFKAN) {
    effkan_deriv(c, i, x, s);
  } else if (F == KV_CHEBY) {
    // dT_n/dx = n U_{n-1}(t) (1 - t^2), t = tanh x  (finite where the reference's
    // autograd through acos yields NaN at |t| == 1)
    float t = tanhf(x);
    float w = 1.0f - t * t;
    float u0 = 1.0f, u1 = 2.0f * t;
    s.put(0, 0.0f);
    if (c.M > 1) s.put(1, w);
    for (int n = 2; n < c.M; ++n) {
      s.put(n, (float)n * u1 * w);
      float u2 = 2.0f * t * u1 - u0;
      u0 = u1;
      u1 = u2;
    }
  } else if (F == KV_FASTKAN) {
    const float* g = c.fp + 2 * c.in;
    for (int j = 0; j < c.G; ++j) {
      float z = (xn - g[j]) / c.rbf_denom;
      s.put(j, -2.0f * z / c.rbf_denom * expf(-(z * z)));
    }
    if (!(c.flags & KV_FLAG_NO_BASE)) s.put(c.G, silu_grad_f(x));
  } else if (F == KV_SINE) {
    const float* freq = c.fp;
    const float* ph = c.fp + c.G + (size_t)i * c.G;
    for (int g = 0; g < c.G; ++g) s.put(g, freq[g] * cosf(x * freq[g] + ph[g]));
  } else if (F == KV_FOURIER) {
    for (int k = 1; k <= c.G; ++k) {
      float sn, cs;
      sincosf((float)k * x, &sn, &cs);
      s.put(k - 1, -(float)k * sn);
      s.put(c.G + k - 1, (float)k * cs);
    }
  }
}

}  // namespace kv
