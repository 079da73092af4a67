// HBM-bound kernels of the path: LayerNorm fwd/bwd (+ residual gradient), FastKAN's
// inner LayerNorm backward, patchify and its adjoint, class-token rows, row gather /
// scatter.  All coalesced along the feature dimension; one warp per row.
#include "internal.cuh"

namespace kv {

constexpr int LN_THREADS = 256;
constexpr int LN_WARPS = LN_THREADS / 32;

// ------------------------------------------------------------------------------
// nn.LayerNorm forward (model.py:22,24,107)
// ------------------------------------------------------------------------------
__global__ void __launch_bounds__(LN_THREADS)
layernorm_fwd_kernel(const float* __restrict__ X, const float* __restrict__ gamma,
                     const float* __restrict__ beta, float* __restrict__ Y, float* __restrict__ mean,
                     float* __restrict__ rstd, int T, int d, int ldx, float eps) {
  const int lane = threadIdx.x & 31;
  for (int t = blockIdx.x * LN_WARPS + (threadIdx.x >> 5); t < T; t += gridDim.x * LN_WARPS) {
    const float* x = X + (size_t)t * ldx;
    float s = 0.0f;
    for (int c = lane; c < d; c += 32) s += x[c];
    const float mu = warp_sum(s) / (float)d;
    float sq = 0.0f;
    for (int c = lane; c < d; c += 32) {
      const float dl = x[c] - mu;
      sq = fmaf(dl, dl, sq);
    }
    const float rs = rsqrtf(warp_sum(sq) / (float)d + eps);
    float* y = Y + (size_t)t * d;
    for (int c = lane; c < d; c += 32) {
      float v = (x[c] - mu) * rs;
      if (gamma) v *= gamma[c];
      if (beta) v += beta[c];
      y[c] = v;
    }
    if (lane == 0) {
      if (mean) mean[t] = mu;
      if (rstd) rstd[t] = rs;
    }
  }
}

// ------------------------------------------------------------------------------
// LayerNorm backward over column slices.  Work item = (row t, slice s); for each of the
// `share` parameter sets on that slice:
//   gy = dY * gamma;  dX (+)= rstd * (gy - mean(gy) - xhat * mean(gy * xhat))
//   dgamma += dY * xhat;  dbeta += dY          (block-level smem partials, then atomics)
// nn.LayerNorm: n_slices = share = 1, dY row stride = width, accumulate = false.
// FastKAN:      dY = dXn [T][n_slices*share*width], accumulate = true.
// ------------------------------------------------------------------------------
__global__ void __launch_bounds__(LN_THREADS)
layernorm_bwd_kernel(const float* __restrict__ X, int ldx, const float* __restrict__ dY, int ldy,
                     const float* __restrict__ gamma, int gamma_stride,
                     const float* __restrict__ mean, const float* __restrict__ rstd,
                     const float* __restrict__ dresidual, float* __restrict__ dX, int lddx,
                     float* __restrict__ dgamma, float* __restrict__ dbeta, int dparam_stride,
                     int T, int n_slices, int share, int width, int accumulate) {
  extern __shared__ float acc_s[];  // [n_slices*share][2][width]
  const int n_sets = n_slices * share;
  for (int p = threadIdx.x; p < n_sets * 2 * width; p += LN_THREADS) acc_s[p] = 0.0f;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const long long total = (long long)T * n_slices;
  for (long long w = (long long)blockIdx.x * LN_WARPS + (threadIdx.x >> 5); w < total;
       w += (long long)gridDim.x * LN_WARPS) {
    const int s = (int)(w % n_slices);
    const long long t = w / n_slices;
    const float* x = X + t * ldx + (size_t)s * width;
    float* dx = dX + t * lddx + (size_t)s * width;
    const float mu = mean[w], rs = rstd[w];
    for (int q = 0; q < share; ++q) {
      const int set = s * share + q;
      const float* gy = dY + t * ldy + (size_t)set * width;
      const float* gm = gamma ? gamma + (size_t)set * gamma_stride : nullptr;
      float* accg = acc_s + (size_t)set * 2 * width;
      float* accb = accg + width;
      float s1 = 0.0f, s2 = 0.0f;
      for (int c = lane; c < width; c += 32) {
        const float xh = (x[c] - mu) * rs;
        const float g = gy[c];
        const float gg = gm ? g * gm[c] : g;
        s1 += gg;
        s2 = fmaf(gg, xh, s2);
        if (dgamma) atomicAdd(&accg[c], g * xh);
        if (dbeta) atomicAdd(&accb[c], g);
      }
      s1 = warp_sum(s1) / (float)width;
      s2 = warp_sum(s2) / (float)width;
      for (int c = lane; c < width; c += 32) {
        const float xh = (x[c] - mu) * rs;
        const float g = gy[c];
        const float gg = gm ? g * gm[c] : g;
        float v = rs * (gg - s1 - xh * s2);
        if (accumulate || q > 0) v += dx[c];
        else if (dresidual) v += dresidual[t * lddx + (size_t)s * width + c];
        dx[c] = v;
      }
    }
  }
  __syncthreads();
  for (int p = threadIdx.x; p < n_sets * width; p += LN_THREADS) {
    const int set = p / width, c = p % width;
    if (dgamma) atomicAdd(&dgamma[(size_t)set * dparam_stride + c], acc_s[(size_t)set * 2 * width + c]);
    if (dbeta) atomicAdd(&dbeta[(size_t)set * dparam_stride + c], acc_s[(size_t)set * 2 * width + width + c]);
  }
}

// ------------------------------------------------------------------------------
// Register-resident variants for the plain nn.LayerNorm case with d % 4 == 0, d <= 1024 and 16-byte
// aligned rows: a lane owns the float4 columns {lane + 32 k}, the row is read ONCE with 16-byte loads,
// and (backward) the d gamma / d beta partial sums stay in registers across all rows a warp visits;
// they meet in shared memory once per block and leave with one atomic per column per block.
// ------------------------------------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(LN_THREADS)
layernorm_fwd_vec_kernel(const float* __restrict__ X, const float* __restrict__ gamma,
                         const float* __restrict__ beta, float* __restrict__ Y, float* __restrict__ mean,
                         float* __restrict__ rstd, int T, int d, int ldx, float eps) {
  const int lane = threadIdx.x & 31;
  const int nv = d >> 2;
  float4 gm[NV], bt[NV];
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const int c = lane + 32 * k;
    gm[k] = (gamma && c < nv) ? reinterpret_cast<const float4*>(gamma)[c] : make_float4(1.f, 1.f, 1.f, 1.f);
    bt[k] = (beta && c < nv) ? reinterpret_cast<const float4*>(beta)[c] : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const float inv_d = 1.0f / (float)d;
  for (int t = blockIdx.x * LN_WARPS + (threadIdx.x >> 5); t < T; t += gridDim.x * LN_WARPS) {
    const float4* x4 = reinterpret_cast<const float4*>(X + (size_t)t * ldx);
    float4 x[NV];
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = lane + 32 * k;
      x[k] = c < nv ? __ldcs(x4 + c) : make_float4(0.f, 0.f, 0.f, 0.f);
      s += (x[k].x + x[k].y) + (x[k].z + x[k].w);
    }
    const float mu = warp_sum(s) * inv_d;
    float sq = 0.0f;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      if (lane + 32 * k < nv) {
        const float a = x[k].x - mu, b = x[k].y - mu, c2 = x[k].z - mu, e = x[k].w - mu;
        sq += (a * a + b * b) + (c2 * c2 + e * e);
      }
    }
    const float rs = rsqrtf(warp_sum(sq) * inv_d + eps);
    float4* y4 = reinterpret_cast<float4*>(Y + (size_t)t * d);
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = lane + 32 * k;
      if (c < nv) {
        float4 o;
        o.x = (x[k].x - mu) * rs * gm[k].x + bt[k].x;
        o.y = (x[k].y - mu) * rs * gm[k].y + bt[k].y;
        o.z = (x[k].z - mu) * rs * gm[k].z + bt[k].z;
        o.w = (x[k].w - mu) * rs * gm[k].w + bt[k].w;
        y4[c] = o;
      }
    }
    if (lane == 0) {
      if (mean) mean[t] = mu;
      if (rstd) rstd[t] = rs;
    }
  }
}

template <int NV>
__global__ void __launch_bounds__(LN_THREADS, 2)
layernorm_bwd_vec_kernel(const float* __restrict__ X, int ldx, const float* __restrict__ dY,
                         const float* __restrict__ gamma, const float* __restrict__ mean,
                         const float* __restrict__ rstd, const float* __restrict__ dresidual,
                         float* __restrict__ dX, float* __restrict__ dgamma, float* __restrict__ dbeta, int T,
                         int d) {
  // smem: gamma [d], then per-warp private partial sums [LN_WARPS][2][d] (plain loads/stores: a lane owns its
  // columns), which keeps the register count low enough for two blocks per SM
  extern __shared__ float4 ln_smem4[];
  float* gamma_s = reinterpret_cast<float*>(ln_smem4);
  float* acc_s = gamma_s + d;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nv = d >> 2;
  for (int p = threadIdx.x; p < d; p += LN_THREADS) gamma_s[p] = gamma ? gamma[p] : 1.0f;
  for (int p = threadIdx.x; p < LN_WARPS * 2 * d; p += LN_THREADS) acc_s[p] = 0.0f;
  __syncthreads();
  const float4* gm4 = reinterpret_cast<const float4*>(gamma_s);
  float4* ag4 = reinterpret_cast<float4*>(acc_s + (size_t)warp * 2 * d);
  float4* ab4 = ag4 + nv;
  const bool want_params = dgamma != nullptr || dbeta != nullptr;
  const float inv_d = 1.0f / (float)d;
  for (int t = blockIdx.x * LN_WARPS + warp; t < T; t += gridDim.x * LN_WARPS) {
    const float4* x4 = reinterpret_cast<const float4*>(X + (size_t)t * ldx);
    const float4* g4 = reinterpret_cast<const float4*>(dY + (size_t)t * d);
    const float4* r4 = dresidual ? reinterpret_cast<const float4*>(dresidual + (size_t)t * d) : nullptr;
    float4 xh[NV], gg[NV], rr[NV];
    // every load of the row is issued before the first use
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = lane + 32 * k;
      const bool in = c < nv;
      xh[k] = in ? __ldcs(x4 + c) : make_float4(0.f, 0.f, 0.f, 0.f);
      gg[k] = in ? __ldcs(g4 + c) : make_float4(0.f, 0.f, 0.f, 0.f);
      rr[k] = (in && r4) ? __ldcs(r4 + c) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    const float mu = mean[t], rs = rstd[t];
    float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = lane + 32 * k;
      if (c < nv) {
        const float4 g = gg[k];
        xh[k] = make_float4((xh[k].x - mu) * rs, (xh[k].y - mu) * rs, (xh[k].z - mu) * rs, (xh[k].w - mu) * rs);
        if (want_params) {
          float4 a = ag4[c], b = ab4[c];
          a.x = fmaf(g.x, xh[k].x, a.x); a.y = fmaf(g.y, xh[k].y, a.y);
          a.z = fmaf(g.z, xh[k].z, a.z); a.w = fmaf(g.w, xh[k].w, a.w);
          b.x += g.x; b.y += g.y; b.z += g.z; b.w += g.w;
          ag4[c] = a;
          ab4[c] = b;
        }
        const float4 gm = gm4[c];
        gg[k] = make_float4(g.x * gm.x, g.y * gm.y, g.z * gm.z, g.w * gm.w);
        s1 += (gg[k].x + gg[k].y) + (gg[k].z + gg[k].w);
        s2 += (gg[k].x * xh[k].x + gg[k].y * xh[k].y) + (gg[k].z * xh[k].z + gg[k].w * xh[k].w);
      }
    }
    s1 = warp_sum(s1) * inv_d;
    s2 = warp_sum(s2) * inv_d;
    float4* dx4 = reinterpret_cast<float4*>(dX + (size_t)t * d);
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = lane + 32 * k;
      if (c < nv) {
        float4 o;
        o.x = rs * (gg[k].x - s1 - xh[k].x * s2) + rr[k].x;
        o.y = rs * (gg[k].y - s1 - xh[k].y * s2) + rr[k].y;
        o.z = rs * (gg[k].z - s1 - xh[k].z * s2) + rr[k].z;
        o.w = rs * (gg[k].w - s1 - xh[k].w * s2) + rr[k].w;
        dx4[c] = o;
      }
    }
  }
  __syncthreads();
  if (want_params) {
    for (int c = threadIdx.x; c < 2 * d; c += LN_THREADS) {
      float v = 0.0f;
#pragma unroll
      for (int w = 0; w < LN_WARPS; ++w) v += acc_s[(size_t)w * 2 * d + c];
      if (c < d) {
        if (dgamma) atomicAdd(&dgamma[c], v);
      } else if (dbeta) {
        atomicAdd(&dbeta[c - d], v);
      }
    }
  }
}

static bool ln_vec_ok(const void* a, const void* b, const void* c, const void* e, int d, int ldx) {
  const uintptr_t bits = (uintptr_t)a | (uintptr_t)b | (uintptr_t)c | (uintptr_t)e;
  return (bits & 15) == 0 && d % 4 == 0 && ldx % 4 == 0 && d <= 1024;
}

// ------------------------------------------------------------------------------
// patchify (model.py:111-126): patches[b, i*np+j, c*ph*pw + y*pw + x] = img[b,c,i*ph+y,j*pw+x]
// One thread per image element, indexed in IMAGE order so reads coalesce (rows of W
// floats); writes land in runs of pw floats.
// ------------------------------------------------------------------------------
__global__ void patchify_kernel(const float* __restrict__ src, float* __restrict__ dst, size_t total,
                                int C, int H, int W, int np, int adjoint) {
  const int ph = H / np, pw = W / np;
  for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < total;
       p += (size_t)gridDim.x * blockDim.x) {
    const int xw = (int)(p % W);
    size_t r = p / W;
    const int yh = (int)(r % H);
    r /= H;
    const int c = (int)(r % C);
    const size_t b = r / C;
    const int i = yh / ph, y = yh % ph, j = xw / pw, x = xw % pw;
    const size_t q = ((b * np * np + (size_t)i * np + j) * C + c) * ph * pw + (size_t)y * pw + x;
    if (adjoint) dst[p] = src[q];  // d images <- d patches
    else dst[q] = src[p];          // patches <- images
  }
}

__global__ void cls_rows_fwd_kernel(const float* __restrict__ v_class, const float* __restrict__ pos,
                                    float* __restrict__ tokens, int B, int N, int d) {
  for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < (size_t)B * d;
       p += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(p % d);
    const size_t b = p / d;
    tokens[b * N * d + c] = v_class[c] + (pos ? pos[c] : 0.0f);
  }
}

// dv_class[c] = sum_b dtokens[b, 0, c]; one thread per column, fixed order (deterministic)
__global__ void cls_rows_bwd_kernel(const float* __restrict__ dtokens, float* __restrict__ dv_class, int B,
                                    int N, int d) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= d) return;
  float s = 0.0f;
  for (int b = 0; b < B; ++b) s += dtokens[(size_t)b * N * d + c];
  dv_class[c] = s;
}

__global__ void gather_rows_kernel(const float* __restrict__ X, float* __restrict__ Y, int B,
                                   int row_stride, int d) {
  for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < (size_t)B * d;
       p += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(p % d);
    const size_t b = p / d;
    Y[p] = X[b * row_stride * d + c];
  }
}
__global__ void scatter_rows_kernel(const float* __restrict__ dY, float* __restrict__ dX, int B,
                                    int row_stride, int d) {
  for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < (size_t)B * d;
       p += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(p % d);
    const size_t b = p / d;
    dX[b * row_stride * d + c] += dY[p];
  }
}

static int grid_for(size_t n, int threads) {
  size_t b = (n + threads - 1) / threads;
  if (b > 148 * 16) b = 148 * 16;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace kv

using namespace kv;

extern "C" int kv_layernorm_fwd(const float* X, const float* gamma, const float* beta, float* Y,
                                float* mean, float* rstd, int T, int d, int ldx, float eps, void* stream) {
  if (T == 0) return KV_OK;
  if (X == nullptr || Y == nullptr || T < 0 || d <= 0 || ldx < d) return KV_ERR_BAD_ARG;
  int blocks = ceil_div(T, LN_WARPS);
  if (ln_vec_ok(X, Y, gamma, beta, d, ldx)) {
    if (blocks > 148 * 4) blocks = 148 * 4;
    const int nvl = ceil_div(d / 4, 32);
#define KV_LN_FWD(NV) layernorm_fwd_vec_kernel<NV><<<blocks, LN_THREADS, 0, (cudaStream_t)stream>>>(X, gamma, beta, Y, mean, rstd, T, d, ldx, eps)
    switch (nvl) {
      case 1: KV_LN_FWD(1); break;
      case 2: KV_LN_FWD(2); break;
      case 3: KV_LN_FWD(3); break;
      case 4: KV_LN_FWD(4); break;
      case 5: case 6: KV_LN_FWD(6); break;
      default: KV_LN_FWD(8); break;
    }
#undef KV_LN_FWD
    KV_CHECK_LAUNCH("layernorm_fwd_vec");
    return KV_OK;
  }
  if (blocks > 148 * 8) blocks = 148 * 8;
  layernorm_fwd_kernel<<<blocks, LN_THREADS, 0, (cudaStream_t)stream>>>(X, gamma, beta, Y, mean, rstd, T, d, ldx, eps);
  KV_CHECK_LAUNCH("layernorm_fwd");
  return KV_OK;
}

extern "C" int kv_layernorm_bwd(const float* X, const float* dY, const float* gamma, const float* mean,
                                const float* rstd, const float* dresidual, float* dX, float* dgamma,
                                float* dbeta, int T, int d, int ldx, float eps, void* stream) {
  (void)eps;
  if (T == 0) return KV_OK;
  if (X == nullptr || dY == nullptr || mean == nullptr || rstd == nullptr || dX == nullptr || T < 0 ||
      d <= 0 || ldx < d)
    return KV_ERR_BAD_ARG;
  if (T == 0) return KV_OK;
  const size_t smem = sizeof(float) * 2 * (size_t)d;
  if (smem > 200 * 1024) return KV_ERR_BAD_SHAPE;
  if (ln_vec_ok(X, dY, dX, dresidual, d, ldx) && ln_vec_ok(gamma, dgamma, dbeta, nullptr, d, ldx)) {
    int vb = ceil_div(T, LN_WARPS * 8);
    if (vb > 148 * 2) vb = 148 * 2;
    const int nvl = ceil_div(d / 4, 32);
    const size_t vsmem = sizeof(float) * (size_t)d * (1 + 2 * LN_WARPS);
#define KV_LN_BWD(NV)                                                                                              \
  do {                                                                                                             \
    if (vsmem > 48 * 1024)                                                                                         \
      KV_CUDA(cudaFuncSetAttribute(layernorm_bwd_vec_kernel<NV>, cudaFuncAttributeMaxDynamicSharedMemorySize,      \
                                   (int)vsmem), "layernorm_bwd_vec smem");                                         \
    layernorm_bwd_vec_kernel<NV><<<vb, LN_THREADS, vsmem, (cudaStream_t)stream>>>(X, ldx, dY, gamma, mean, rstd,   \
                                                                                 dresidual, dX, dgamma, dbeta, T, d); \
  } while (0)
    switch (nvl) {
      case 1: KV_LN_BWD(1); break;
      case 2: KV_LN_BWD(2); break;
      case 3: KV_LN_BWD(3); break;
      case 4: KV_LN_BWD(4); break;
      case 5: case 6: KV_LN_BWD(6); break;
      default: KV_LN_BWD(8); break;
    }
#undef KV_LN_BWD
    KV_CHECK_LAUNCH("layernorm_bwd_vec");
    return KV_OK;
  }
  int blocks = ceil_div(T, LN_WARPS * 4);
  if (blocks > 148 * 4) blocks = 148 * 4;
  if (smem > 48 * 1024)
    KV_CUDA(cudaFuncSetAttribute(layernorm_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
            "layernorm_bwd smem");
  layernorm_bwd_kernel<<<blocks, LN_THREADS, smem, (cudaStream_t)stream>>>(
      X, ldx, dY, d, gamma, d, mean, rstd, dresidual, dX, d, dgamma, dbeta, d, T, 1, 1, d, 0);
  KV_CHECK_LAUNCH("layernorm_bwd");
  return KV_OK;
}

extern "C" int kv_fastkan_ln_bwd(const kv_kan_desc_t* d, const float* X, const float* dXn,
                                 const float* fparams, const float* mean, const float* rstd, float* dX,
                                 float* dfparams, void* stream) {
  int rc = check_desc(d);
  if (rc != KV_OK) return rc;
  if (d->family != KV_FASTKAN) return KV_ERR_BAD_FAMILY;
  if (X == nullptr || dXn == nullptr || fparams == nullptr || mean == nullptr || rstd == nullptr ||
      dX == nullptr)
    return KV_ERR_BAD_ARG;
  if (d->T == 0) return KV_OK;
  const int in = d->in_features;
  const int n_slices = d->n_groups / d->x_share;
  const size_t smem = sizeof(float) * 2 * (size_t)in * d->n_groups;
  if (smem > 200 * 1024) return KV_ERR_BAD_SHAPE;
  const long long total = (long long)d->T * n_slices;
  long long blocks = (total + LN_WARPS * 4 - 1) / (LN_WARPS * 4);
  if (blocks > 148 * 2) blocks = 148 * 2;
  if (smem > 48 * 1024)
    KV_CUDA(cudaFuncSetAttribute(layernorm_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
            "fastkan_ln_bwd smem");
  // gamma of set g = fparams[g*stride + 0..in); dgamma -> dfparams[g*stride + 0..in), dbeta -> +in
  layernorm_bwd_kernel<<<(int)blocks, LN_THREADS, smem, (cudaStream_t)stream>>>(
      X, d->ldx, dXn, d->n_groups * in, fparams, d->fparam_stride, mean, rstd, nullptr, dX, d->ldx,
      dfparams, dfparams ? dfparams + in : nullptr, d->fparam_stride, d->T, n_slices, d->x_share, in, 1);
  KV_CHECK_LAUNCH("fastkan_ln_bwd");
  return KV_OK;
}

static int patchify_common(const float* a, float* b, int B, int C, int H, int W, int np, int adjoint,
                           void* stream) {
  if (a == nullptr || b == nullptr || B < 0 || C <= 0 || H <= 0 || W <= 0 || np <= 0) return KV_ERR_BAD_ARG;
  if (H % np != 0 || W % np != 0) return KV_ERR_BAD_SHAPE;  // model.py:58-59
  const size_t total = (size_t)B * C * H * W;
  if (total == 0) return KV_OK;
  patchify_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(a, b, total, C, H, W, np, adjoint);
  KV_CHECK_LAUNCH("patchify");
  return KV_OK;
}

extern "C" int kv_patchify(const float* images, float* patches, int B, int C, int H, int W,
                           int n_patches, void* stream) {
  return patchify_common(images, patches, B, C, H, W, n_patches, 0, stream);
}
extern "C" int kv_patchify_bwd(const float* dpatches, float* dimages, int B, int C, int H, int W,
                               int n_patches, void* stream) {
  return patchify_common(dpatches, dimages, B, C, H, W, n_patches, 1, stream);
}

extern "C" int kv_cls_rows_fwd(const float* v_class, const float* pos, float* tokens, int B, int N, int d,
                               void* stream) {
  if (v_class == nullptr || tokens == nullptr || B < 0 || N <= 0 || d <= 0) return KV_ERR_BAD_ARG;
  if (B == 0) return KV_OK;
  cls_rows_fwd_kernel<<<grid_for((size_t)B * d, 256), 256, 0, (cudaStream_t)stream>>>(v_class, pos, tokens, B, N, d);
  KV_CHECK_LAUNCH("cls_rows_fwd");
  return KV_OK;
}
extern "C" int kv_cls_rows_bwd(const float* dtokens, float* dv_class, int B, int N, int d, void* stream) {
  if (dtokens == nullptr || dv_class == nullptr || B < 0 || N <= 0 || d <= 0) return KV_ERR_BAD_ARG;
  cls_rows_bwd_kernel<<<ceil_div(d, 128), 128, 0, (cudaStream_t)stream>>>(dtokens, dv_class, B, N, d);
  KV_CHECK_LAUNCH("cls_rows_bwd");
  return KV_OK;
}
extern "C" int kv_gather_rows(const float* X, float* Y, int B, int row_stride, int d, void* stream) {
  if (X == nullptr || Y == nullptr || B < 0 || row_stride <= 0 || d <= 0) return KV_ERR_BAD_ARG;
  if (B == 0) return KV_OK;
  gather_rows_kernel<<<grid_for((size_t)B * d, 256), 256, 0, (cudaStream_t)stream>>>(X, Y, B, row_stride, d);
  KV_CHECK_LAUNCH("gather_rows");
  return KV_OK;
}
extern "C" int kv_scatter_rows(const float* dY, float* dX, int B, int row_stride, int d, void* stream) {
  if (dY == nullptr || dX == nullptr || B < 0 || row_stride <= 0 || d <= 0) return KV_ERR_BAD_ARG;
  if (B == 0) return KV_OK;
  scatter_rows_kernel<<<grid_for((size_t)B * d, 256), 256, 0, (cudaStream_t)stream>>>(dY, dX, B, row_stride, d);
  KV_CHECK_LAUNCH("scatter_rows");
  return KV_OK;
}
