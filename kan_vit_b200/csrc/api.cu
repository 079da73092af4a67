// C-ABI surface of libkanvit_b200: validation, parameter packing / gradient unpacking,
// and dispatch of the KAN entry points to the SIMT (fp32) or tcgen05 (bf16) kernels.
#include <cstdio>
#include <cstring>
#include "internal.cuh"

namespace kv {

thread_local char g_cuda_err[256] = "";
unsigned long long g_launches = 0;

int record_cuda_error(cudaError_t e, const char* where) {
  snprintf(g_cuda_err, sizeof(g_cuda_err), "%s: %s", where, cudaGetErrorString(e));
  return KV_ERR_CUDA;
}

static int family_M(const kv_kan_desc_t& d) {
  switch (d.family) {
    case KV_LINEAR: return 1;
    case KV_EFFKAN: return d.grid_size + d.spline_order + 1;
    case KV_CHEBY: return d.grid_size + 1;
    case KV_FASTKAN: return d.grid_size + ((d.flags & KV_FLAG_NO_BASE) ? 0 : 1);
    case KV_SINE: return d.grid_size;
    case KV_FOURIER: return 2 * d.grid_size;
    default: return -1;
  }
}

static int fparam_count(const kv_kan_desc_t& d) {
  switch (d.family) {
    case KV_EFFKAN: return d.in_features * (d.grid_size + 2 * d.spline_order + 1) + 4;  // + grid probe
    case KV_FASTKAN: return 2 * d.in_features + d.grid_size;
    case KV_SINE: return d.grid_size + d.in_features * d.grid_size;
    default: return 0;
  }
}

int check_desc(const kv_kan_desc_t* d) {
  if (d == nullptr) return KV_ERR_BAD_ARG;
  if (d->family < KV_LINEAR || d->family > KV_FOURIER) return KV_ERR_BAD_FAMILY;
  if (d->T < 0 || d->in_features <= 0 || d->out_features <= 0 || d->n_groups <= 0 || d->x_share <= 0)
    return KV_ERR_BAD_ARG;
  if (d->n_groups % d->x_share != 0) return KV_ERR_BAD_SHAPE;
  if (d->family != KV_LINEAR && d->grid_size <= 0 && d->family != KV_CHEBY) return KV_ERR_BAD_SHAPE;
  if (d->family == KV_EFFKAN && (d->spline_order < 0 || d->spline_order > KV_MAX_SPLINE_ORDER))
    return KV_ERR_BAD_SHAPE;
  const int M = family_M(*d);
  if (M <= 0 || M > KV_MAX_M || M != d->M) return KV_ERR_BAD_SHAPE;
  if (d->fparam_stride < fparam_count(*d)) return KV_ERR_BAD_SHAPE;
  const int n_slices = d->n_groups / d->x_share;
  if (d->ldx < n_slices * d->in_features || d->ldy < d->n_groups * d->out_features) return KV_ERR_BAD_SHAPE;
  if (d->rows_in < 0 || (d->rows_in > 0 && (d->rows_out < d->rows_in + d->row_shift || d->row_shift < 0)))
    return KV_ERR_BAD_SHAPE;
  if (d->family == KV_FASTKAN && !(d->rbf_denom != 0.0f)) return KV_ERR_BAD_ARG;
  return KV_OK;
}

// ------------------------------------------------------------------------------
// pack: reference parameter layouts -> [g][out][in*M] (K index i*M+m)
// ------------------------------------------------------------------------------
__global__ void pack_weights_kernel(const kv_kan_desc_t d, const uint64_t* __restrict__ tbl,
                                    float* __restrict__ wp_f32) {
  const int g = blockIdx.y;
  const int M = d.M, in = d.in_features, out = d.out_features;
  const size_t per_group = (size_t)out * in * M;
  for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < per_group;
       p += (size_t)gridDim.x * blockDim.x) {
    const int m = (int)(p % M);
    const size_t r = p / M;
    const int i = (int)(r % in);
    const int o = (int)(r / in);
    wp_f32[(size_t)g * per_group + p] = load_packed_weight(d, tbl, g, o, i, m);
  }
}

__global__ void pack_fparams_kernel(const kv_kan_desc_t d, const uint64_t* __restrict__ tbl,
                                    float* __restrict__ fparams, float* __restrict__ bias) {
  const int g = blockIdx.y;
  const int in = d.in_features, out = d.out_features, G = d.grid_size;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const int nth = gridDim.x * blockDim.x;
  if (fparams != nullptr) {
    float* fp = fparams + (size_t)g * d.fparam_stride;
    if (d.family == KV_EFFKAN) {
      const int n = in * (G + 2 * d.spline_order + 1);
      const float* grid = slot(tbl, g, 3);
      for (int p = tid; p < n; p += nth) fp[p] = grid[p];
    } else if (d.family == KV_FASTKAN) {
      const float* w = slot(tbl, g, 0);
      const float* b = slot(tbl, g, 1);
      const float* gr = slot(tbl, g, 2);
      for (int p = tid; p < in; p += nth) {
        fp[p] = w ? w[p] : 1.0f;
        fp[in + p] = b ? b[p] : 0.0f;
      }
      for (int p = tid; p < G; p += nth) fp[2 * in + p] = gr[p];
    } else if (d.family == KV_SINE) {
      const float* fr = slot(tbl, g, 1);
      const float* ph = slot(tbl, g, 2);
      for (int p = tid; p < G; p += nth) fp[p] = fr[p];
      for (int p = tid; p < in * G; p += nth) fp[G + p] = ph[p];
    }
  }
  if (bias != nullptr) {
    int s = -1;
    if (d.family == KV_LINEAR) s = 1;
    else if (d.family == KV_FASTKAN) s = 5;
    else if (d.family == KV_SINE) s = 3;
    else if (d.family == KV_FOURIER) s = 1;
    const float* b = s >= 0 ? slot(tbl, g, s) : nullptr;
    for (int p = tid; p < out; p += nth) bias[(size_t)g * out + p] = b ? b[p] : 0.0f;
  }
}

// effkan: is every feature's knot vector the same uniform grid t0 + j*h?  Written behind the knots as
// {uniform ? 1 : 0, t0, 1/h, 0}; the tensor-core kernels pick their closed-form prologue from it.
__global__ void effkan_probe_kernel(const kv_kan_desc_t d, const uint64_t* __restrict__ tbl,
                                    float* __restrict__ fparams) {
  const int g = blockIdx.x;
  const int nk = d.grid_size + 2 * d.spline_order + 1;
  const int n = d.in_features * nk;
  const float* grid = slot(tbl, g, 3);
  const float t0 = grid[0], span = grid[nk - 1] - grid[0];
  const float h = span / (float)(nk - 1);
  int ok = (span > 0.0f) ? 1 : 0;
  for (int p = threadIdx.x; p < n; p += blockDim.x) {
    const int j = p % nk;
    if (!(fabsf(grid[p] - (t0 + (float)j * h)) <= 1e-5f * span)) ok = 0;
  }
  ok = __syncthreads_and(ok);
  if (threadIdx.x == 0) {
    float* fp = fparams + (size_t)g * d.fparam_stride + n;
    fp[0] = ok ? 1.0f : 0.0f;
    fp[1] = t0;
    fp[2] = ok ? 1.0f / h : 0.0f;
    fp[3] = 0.0f;
  }
}

// ------------------------------------------------------------------------------
// unpack: packed gradients -> reference parameter layouts
// ------------------------------------------------------------------------------
__global__ void unpack_grads_kernel(const kv_kan_desc_t d, const float* __restrict__ dWp,
                                    const float* __restrict__ dfparams, const float* __restrict__ dbias,
                                    const uint64_t* __restrict__ tbl, const uint64_t* __restrict__ gtbl) {
  const int g = blockIdx.y;
  const int M = d.M, in = d.in_features, out = d.out_features, G = d.grid_size;
  const bool accum = (d.flags & KV_FLAG_ACCUM_GRADS) != 0;  // += into the caller's gradient buffers
  auto put = [accum](float* dst, float v) { *dst = accum ? *dst + v : v; };
  const size_t n_oi = (size_t)out * in;
  const float* dW = dWp ? dWp + (size_t)g * n_oi * M : nullptr;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t nth = (size_t)gridDim.x * blockDim.x;
  if (dW != nullptr) {
    for (size_t p = tid; p < n_oi; p += nth) {  // one (o,i) pair per iteration
      const int i = (int)(p % in), o = (int)(p / in);
      const float* src = dW + p * M;
      switch (d.family) {
        case KV_LINEAR: {
          float* gw = slot_out(gtbl, g, 0);
          if (gw) put(gw + p, src[0]);
          break;
        }
        case KV_EFFKAN: {
          const int nb = G + d.spline_order;
          float* gb = slot_out(gtbl, g, 0);
          float* gs = slot_out(gtbl, g, 1);
          float* gsc = slot_out(gtbl, g, 2);
          const float* sw = slot(tbl, g, 1);
          const float* sc = slot(tbl, g, 2);
          if (gb) put(gb + p, src[nb]);
          const float scale = sc ? sc[p] : 1.0f;
          float acc = 0.0f;
          if (nb == 8 && gs && ((((uintptr_t)gs) | ((uintptr_t)sw)) & 15) == 0) {
            // default effkan (8 coefficients per (o,i) = 32 bytes): 16-byte loads / stores of the row
            const float4* sw4 = reinterpret_cast<const float4*>(sw + p * 8);
            float4* gs4 = reinterpret_cast<float4*>(gs + p * 8);
            const float4 w0 = __ldcs(sw4), w1 = __ldcs(sw4 + 1);
            float4 o0 = make_float4(src[0] * scale, src[1] * scale, src[2] * scale, src[3] * scale);
            float4 o1 = make_float4(src[4] * scale, src[5] * scale, src[6] * scale, src[7] * scale);
            acc = (src[0] * w0.x + src[1] * w0.y) + (src[2] * w0.z + src[3] * w0.w) +
                  (src[4] * w1.x + src[5] * w1.y) + (src[6] * w1.z + src[7] * w1.w);
            if (accum) {
              const float4 a0 = gs4[0], a1 = gs4[1];
              o0.x += a0.x; o0.y += a0.y; o0.z += a0.z; o0.w += a0.w;
              o1.x += a1.x; o1.y += a1.y; o1.z += a1.z; o1.w += a1.w;
            }
            gs4[0] = o0;
            gs4[1] = o1;
            if (gsc) put(gsc + p, acc);
            break;
          }
          for (int m = 0; m < nb; ++m) {
            const float gv = src[m];
            if (gs) put(gs + p * nb + m, gv * scale);
            acc += gv * sw[p * nb + m];
          }
          if (gsc) put(gsc + p, acc);
          break;
        }
        case KV_CHEBY: {
          float* gc = slot_out(gtbl, g, 0);
          if (gc)
            for (int m = 0; m < M; ++m) put(gc + ((size_t)i * out + o) * M + m, src[m]);
          break;
        }
        case KV_FASTKAN: {
          float* gs = slot_out(gtbl, g, 3);
          float* gb = slot_out(gtbl, g, 4);
          if (gs)
            for (int m = 0; m < G; ++m) put(gs + (size_t)o * in * G + (size_t)i * G + m, src[m]);
          if (gb && M > G) put(gb + p, src[G]);
          break;
        }
        case KV_SINE: {
          float* ga = slot_out(gtbl, g, 0);
          if (ga)
            for (int m = 0; m < G; ++m) put(ga + p * G + m, src[m]);
          break;
        }
        case KV_FOURIER: {
          float* gf = slot_out(gtbl, g, 0);
          if (gf)
            for (int m = 0; m < M; ++m) {
              const int part = m / G, k = m % G;
              put(gf + (((size_t)part * out + o) * in + i) * G + k, src[m]);
            }
          break;
        }
      }
    }
  }
  if (dbias != nullptr) {
    int s = -1;
    if (d.family == KV_LINEAR) s = 1;
    else if (d.family == KV_FASTKAN) s = 5;
    else if (d.family == KV_SINE) s = 3;
    else if (d.family == KV_FOURIER) s = 1;
    float* gb = s >= 0 ? slot_out(gtbl, g, s) : nullptr;
    if (gb)
      for (size_t p = tid; p < (size_t)out; p += nth) put(gb + p, dbias[(size_t)g * out + p]);
  }
  if (dfparams != nullptr) {
    const float* dfp = dfparams + (size_t)g * d.fparam_stride;
    if (d.family == KV_FASTKAN) {
      float* gw = slot_out(gtbl, g, 0);
      float* gbb = slot_out(gtbl, g, 1);
      for (size_t p = tid; p < (size_t)in; p += nth) {
        if (gw) put(gw + p, dfp[p]);
        if (gbb) put(gbb + p, dfp[in + p]);
      }
    } else if (d.family == KV_SINE) {
      float* gf = slot_out(gtbl, g, 1);
      if (gf)
        for (size_t p = tid; p < (size_t)G; p += nth) put(gf + p, dfp[p]);
    }
  }
}

// ------------------------------------------------------------------------------
// per-(row, slice) LayerNorm statistics
// ------------------------------------------------------------------------------
__global__ void row_stats_kernel(const float* __restrict__ X, int T, int ldx, int n_slices, int width,
                                 float eps, float* __restrict__ mean, float* __restrict__ rstd) {
  const int warps_per_block = blockDim.x / 32;
  const int lane = threadIdx.x & 31;
  const long long total = (long long)T * n_slices;
  for (long long w = (long long)blockIdx.x * warps_per_block + threadIdx.x / 32; w < total;
       w += (long long)gridDim.x * warps_per_block) {
    const int s = (int)(w % n_slices);
    const long long t = w / n_slices;
    const float* row = X + t * ldx + (size_t)s * width;
    float sum = 0.0f;
    for (int c = lane; c < width; c += 32) sum += row[c];
    const float mu = warp_sum(sum) / (float)width;
    float sq = 0.0f;
    for (int c = lane; c < width; c += 32) {
      const float dlt = row[c] - mu;
      sq += dlt * dlt;
    }
    const float var = warp_sum(sq) / (float)width;
    if (lane == 0) {
      mean[w] = mu;
      rstd[w] = rsqrtf(var + eps);
    }
  }
}

}  // namespace kv

// ================================================================================
// extern "C"
// ================================================================================
using namespace kv;

extern "C" {

int kv_version(void) { return KV_ABI_VERSION; }

const char* kv_status_string(int status) {
  switch (status) {
    case KV_OK: return "ok";
    case KV_ERR_BAD_ARG: return "bad argument (null pointer or negative size)";
    case KV_ERR_BAD_SHAPE: return "shape not supported";
    case KV_ERR_BAD_FAMILY: return "unknown basis family";
    case KV_ERR_ALIGNMENT: return "pointer or leading dimension misaligned";
    case KV_ERR_UNSUPPORTED: return "combination not supported in this precision mode";
    case KV_ERR_CUDA: return "CUDA error";
    default: return "unknown status";
  }
}

const char* kv_last_cuda_error(void) { return g_cuda_err; }

unsigned long long kv_launch_count(void) { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

// Tensor-core view of a descriptor.  Groups that share an input slice (x_share > 1: the q/k/v maps of one
// head, attention.py:196-197) and one uniform knot vector are ONE layer with x_share*out outputs: the
// packed weights [g][out][K], bias [g][out], dY/Y columns g*out+o and dWp are the same memory either way,
// so the merge is a reinterpretation.  The basis is then evaluated once per slice instead of once per
// group, and the gradient kernels (one group per slice) apply.  Prologue parameters are read from the
// first group of each slice (fparam_stride scaled), which the uniform-grid flag makes equal to the others.
static kv_kan_desc_t tc_view(const kv_kan_desc_t& d) {
  kv_kan_desc_t v = d;
  // effkan needs one shared knot vector; Linear / ChebyKAN / Fourier have no per-module prologue parameters
  const bool mergeable = (d.family == KV_EFFKAN && (d.flags & KV_FLAG_UNIFORM_GRID)) || d.family == KV_LINEAR ||
                         d.family == KV_CHEBY || d.family == KV_FOURIER;
  if ((d.flags & KV_FLAG_BF16_TC) && mergeable && d.x_share > 1 &&
      d.n_groups % d.x_share == 0 && d.out_features * d.x_share <= 512) {
    v.n_groups = d.n_groups / d.x_share;
    v.out_features = d.out_features * d.x_share;
    v.fparam_stride = d.fparam_stride * d.x_share;
    v.x_share = 1;
  }
  return v;
}

int kv_kan_fparam_count(const kv_kan_desc_t* d) {
  if (d == nullptr) return KV_ERR_BAD_ARG;
  if (d->family < KV_LINEAR || d->family > KV_FOURIER) return KV_ERR_BAD_FAMILY;
  return fparam_count(*d);
}

int kv_kan_pack(const kv_kan_desc_t* d, const uint64_t* ptr_table, float* wp_f32, void* wtiles_bf16,
                void* wtiles_t_bf16, float* fparams, float* bias, void* stream) {
  int rc = check_desc(d);
  if (rc != KV_OK) return rc;
  if (ptr_table == nullptr) return KV_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if (wp_f32 != nullptr) {
    const size_t per_group = (size_t)d->out_features * d->in_features * d->M;
    int bx = (int)((per_group + 255) / 256);
    if (bx > 2048) bx = 2048;
    pack_weights_kernel<<<dim3(bx, d->n_groups), 256, 0, st>>>(*d, ptr_table, wp_f32);
    KV_CHECK_LAUNCH("pack_weights");
  }
  if (fparams != nullptr || bias != nullptr) {
    pack_fparams_kernel<<<dim3(8, d->n_groups), 256, 0, st>>>(*d, ptr_table, fparams, bias);
    KV_CHECK_LAUNCH("pack_fparams");
    if (fparams != nullptr && d->family == KV_EFFKAN) {
      effkan_probe_kernel<<<d->n_groups, 1024, 0, st>>>(*d, ptr_table, fparams);
      KV_CHECK_LAUNCH("effkan_probe");
    }
  }
  if (wtiles_bf16 != nullptr || wtiles_t_bf16 != nullptr) {
    rc = tc_supported(*d);
    if (rc != KV_OK) return rc;
    if (fparam_count(*d) > 0 && fparams == nullptr) return KV_ERR_BAD_ARG;  // layout choice reads the grid probe
    rc = tc_pack(tc_view(*d), *d, ptr_table, fparams, wtiles_bf16, wtiles_t_bf16, st);
    if (rc != KV_OK) return rc;
  }
  return KV_OK;
}

int kv_kan_tc_bytes(const kv_kan_desc_t* d, size_t* wtiles_bytes, size_t* wtiles_t_bytes, size_t* dy_image_bytes) {
  int rc = check_desc(d);
  if (rc != KV_OK) return rc;
  rc = tc_supported(*d);
  if (rc != KV_OK) return rc;
  const kv_kan_desc_t v = tc_view(*d);
  if (wtiles_bytes) *wtiles_bytes = tc_wtiles_bytes(v);
  if (wtiles_t_bytes) *wtiles_t_bytes = tc_wtiles_t_bytes(v);
  if (dy_image_bytes) *dy_image_bytes = tc_dy_image_bytes(v);
  return KV_OK;
}

int kv_kan_dy_image_geom(const kv_kan_desc_t* d, int* t8_groups, int* o8_groups) {
  int rc = check_desc(d);
  if (rc != KV_OK) return rc;
  if (t8_groups == nullptr || o8_groups == nullptr) return KV_ERR_BAD_ARG;
  const kv_kan_desc_t v = tc_view(*d);
  if (tc_dy_image_bytes(v) == 0) return KV_ERR_UNSUPPORTED;
  tc_dy_image_geom(v, t8_groups, o8_groups);
  return KV_OK;
}

int kv_kan_dy_image(const kv_kan_desc_t* d, const float* dY, const float* act_mask, void* dy_image, void* stream) {
  int rc = check_desc(d);
  if (rc != KV_OK) return rc;
  if (d->T == 0) return KV_OK;
  if (dY == nullptr || dy_image == nullptr) return KV_ERR_BAD_ARG;
  if ((d->flags & KV_FLAG_RELU) && act_mask == nullptr) return KV_ERR_BAD_ARG;
  return tc_dy_image(tc_view(*d), dY, act_mask, dy_image, (cudaStream_t)stream);
}

int kv_row_stats(const float* X, int T, int ldx, int n_slices, int width, float eps, float* mean,
                 float* rstd, void* stream) {
  if (T == 0) return KV_OK;
  if (X == nullptr || mean == nullptr || rstd == nullptr || T < 0 || n_slices <= 0 || width <= 0)
    return KV_ERR_BAD_ARG;
  const long long total = (long long)T * n_slices;
  int blocks = (int)((total + 7) / 8);
  if (blocks > 148 * 8) blocks = 148 * 8;
  row_stats_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(X, T, ldx, n_slices, width, eps, mean, rstd);
  KV_CHECK_LAUNCH("row_stats");
  return KV_OK;
}

int kv_kan_fwd(const kv_kan_desc_t* d, const float* X, const float* wp_f32, const void* wtiles_bf16,
               const float* fparams, const float* bias, const float* mean, const float* rstd,
               const float* residual, const float* pos, float* Y, void* stream) {
  int rc = check_desc(d);
  if (rc != KV_OK) return rc;
  if (d->T == 0) return KV_OK;
  if (X == nullptr || Y == nullptr) return KV_ERR_BAD_ARG;
  if (fparam_count(*d) > 0 && fparams == nullptr) return KV_ERR_BAD_ARG;
  if (d->family == KV_FASTKAN && !(d->flags & KV_FLAG_NO_LAYERNORM) && (mean == nullptr || rstd == nullptr))
    return KV_ERR_BAD_ARG;
  if (d->T == 0) return KV_OK;
  if (d->flags & KV_FLAG_BF16_TC) {
    if (wtiles_bf16 == nullptr) return KV_ERR_BAD_ARG;
    rc = tc_supported(*d);
    if (rc != KV_OK) return rc;
    return tc_fwd(tc_view(*d), X, wtiles_bf16, fparams, bias, mean, rstd, residual, pos, Y, (cudaStream_t)stream);
  }
  if (wp_f32 == nullptr) return KV_ERR_BAD_ARG;
  return simt_fwd(*d, X, wp_f32, fparams, bias, mean, rstd, residual, pos, Y, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------
// kv_kan_fwd_noalloc: pack + exact fp32 forward of one (small) KAN layer without any device allocation.
// The packed weights, prologue parameters and bias -- the scratch that kv_kan_pack normally writes into buffers the
// caller allocated -- live in ONE file-scope device array.  Every GPU that loads the library holds the array until
// the process exits; every call shares it (calls that overlap on the GPU race), and it keeps the last call's
// contents: each call rewrites every element it reads (pack_weights / pack_fparams cover the whole packed layer,
// bias included), so nothing has to start at zero.
// ------------------------------------------------------------------------------
static __device__ __align__(256) float g_noalloc_scratch[KV_NOALLOC_SCRATCH_FLOATS];

// The array's name is a host-side stand-in in host code, not a device address: the address comes from
// cudaGetSymbolAddress, once per device, with that device current.
static int noalloc_scratch_ptr(float** out) {
  static float* ptrs[64] = {nullptr};   // per device ordinal; every writer stores the same value
  int dev = 0;
  KV_CUDA(cudaGetDevice(&dev), "kv_kan_fwd_noalloc: cudaGetDevice");
  if (dev < 0 || dev >= 64) return KV_ERR_UNSUPPORTED;
  float* p = __atomic_load_n(&ptrs[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    KV_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch),
            "kv_kan_fwd_noalloc: cudaGetSymbolAddress");
    if (p == nullptr) return KV_ERR_CUDA;
    __atomic_store_n(&ptrs[dev], p, __ATOMIC_RELEASE);
  }
  *out = p;
  return KV_OK;
}

int kv_kan_fwd_noalloc(const kv_kan_desc_t* d, const uint64_t* ptr_table, const float* X, const float* mean,
                       const float* rstd, const float* residual, const float* pos, float* Y, void* stream) {
  int rc = check_desc(d);
  if (rc != KV_OK) return rc;
  if (d->flags & KV_FLAG_BF16_TC) return KV_ERR_UNSUPPORTED;   // exact fp32 kernels only: no operand images
  if (ptr_table == nullptr) return KV_ERR_BAD_ARG;
  // scratch layout (floats, every part 256-byte aligned): wp [g][out][in*M] | fparams [g][stride] | bias [g][out]
  auto up64 = [](size_t n) { return (n + 63) / 64 * 64; };
  const size_t n_wp = up64((size_t)d->n_groups * d->out_features * d->in_features * d->M);
  const size_t n_fp = fparam_count(*d) > 0 ? up64((size_t)d->n_groups * (size_t)d->fparam_stride) : 0;
  const size_t n_bias = up64((size_t)d->n_groups * d->out_features);
  if (fparam_count(*d) > 0 && d->fparam_stride < fparam_count(*d)) return KV_ERR_BAD_ARG;
  if (n_wp + n_fp + n_bias > (size_t)KV_NOALLOC_SCRATCH_FLOATS) {
    snprintf(g_cuda_err, sizeof(g_cuda_err),
             "kv_kan_fwd_noalloc: the packed layer needs %zu scratch floats, the static scratch holds %d "
             "(KV_NOALLOC_SCRATCH_FLOATS); use kv_kan_pack + kv_kan_fwd with caller-allocated buffers",
             n_wp + n_fp + n_bias, (int)KV_NOALLOC_SCRATCH_FLOATS);
    return KV_ERR_BAD_SHAPE;
  }
  float* scratch = nullptr;
  rc = noalloc_scratch_ptr(&scratch);
  if (rc != KV_OK) return rc;
  float* wp = scratch;
  float* fp = n_fp ? scratch + n_wp : nullptr;
  float* bias = scratch + n_wp + n_fp;
  rc = kv_kan_pack(d, ptr_table, wp, nullptr, nullptr, fp, bias, stream);
  if (rc != KV_OK) return rc;
  return kv_kan_fwd(d, X, wp, nullptr, fp, bias, mean, rstd, residual, pos, Y, stream);
}

int kv_kan_fwd_attn_image(const kv_kan_desc_t* d, const float* X, const void* wtiles_bf16, const float* fparams,
                          const float* bias, const float* mean, const float* rstd, void* attn_image,
                          int tokens_per_sample, void* stream) {
  int rc = check_desc(d);
  if (rc != KV_OK) return rc;
  if (!(d->flags & KV_FLAG_BF16_TC)) return KV_ERR_UNSUPPORTED;
  if (d->T == 0) return KV_OK;
  if (X == nullptr || wtiles_bf16 == nullptr || attn_image == nullptr) return KV_ERR_BAD_ARG;
  if (fparam_count(*d) > 0 && fparams == nullptr) return KV_ERR_BAD_ARG;
  rc = tc_supported(*d);
  if (rc != KV_OK) return rc;
  const kv_kan_desc_t v = tc_view(*d);
  if (v.n_groups == d->n_groups && d->x_share != 1) return KV_ERR_UNSUPPORTED;   // the q/k/v maps did not merge
  return tc_fwd(v, X, wtiles_bf16, fparams, bias, mean, rstd, nullptr, nullptr, nullptr, (cudaStream_t)stream,
                attn_image, tokens_per_sample);
}

int kv_kan_bwd_input(const kv_kan_desc_t* d, const float* X, const float* dY, const float* act_mask,
                     const float* wp_f32, const void* wtiles_t_bf16, const void* dy_image,
                     const float* fparams, const float* mean, const float* rstd, float* dX, float* dXn,
                     float* dfparams, void* stream) {
  int rc = check_desc(d);
  if (rc != KV_OK) return rc;
  if (d->T == 0) return KV_OK;
  const bool tc_path = (d->flags & KV_FLAG_BF16_TC) && wtiles_t_bf16 != nullptr && dy_image != nullptr;
  // dY may be NULL on the tensor-core path when the caller produced the bf16 dY image itself (attention backward)
  if (X == nullptr || dX == nullptr || (dY == nullptr && !(tc_path && d->family != KV_LINEAR))) return KV_ERR_BAD_ARG;
  if ((d->flags & KV_FLAG_RELU) && act_mask == nullptr) return KV_ERR_BAD_ARG;
  if (fparam_count(*d) > 0 && fparams == nullptr) return KV_ERR_BAD_ARG;
  if (d->family == KV_FASTKAN && !(d->flags & KV_FLAG_NO_LAYERNORM) &&
      (mean == nullptr || rstd == nullptr || dXn == nullptr))
    return KV_ERR_BAD_ARG;
  if (d->T == 0) return KV_OK;
  if (tc_path)
    return tc_bwd_input(tc_view(*d), X, dY, act_mask, dy_image, wtiles_t_bf16, fparams, mean, rstd, dX, dXn, dfparams,
                        (cudaStream_t)stream);
  if (wp_f32 == nullptr) return KV_ERR_BAD_ARG;
  return simt_bwd_input(*d, X, dY, act_mask, wp_f32, fparams, mean, rstd, dX, dXn, dfparams,
                        (cudaStream_t)stream);
}

int kv_kan_bwd_weight(const kv_kan_desc_t* d, const float* X, const float* dY, const float* act_mask,
                      const void* dy_image, const float* fparams, const float* mean, const float* rstd, float* dWp,
                      float* dbias, void* stream) {
  int rc = check_desc(d);
  if (rc != KV_OK) return rc;
  if (d->T == 0) return KV_OK;
  const bool tc_w = (d->flags & KV_FLAG_BF16_TC) && dy_image != nullptr && tc_bwd_weight_available(tc_view(*d));
  // dY may be NULL on the tensor-core path (the caller made the image) as long as no bias gradient is wanted
  if (X == nullptr || dWp == nullptr || (dY == nullptr && !(tc_w && dbias == nullptr))) return KV_ERR_BAD_ARG;
  if ((d->flags & KV_FLAG_RELU) && act_mask == nullptr) return KV_ERR_BAD_ARG;
  if (fparam_count(*d) > 0 && fparams == nullptr) return KV_ERR_BAD_ARG;
  if (d->family == KV_FASTKAN && !(d->flags & KV_FLAG_NO_LAYERNORM) && (mean == nullptr || rstd == nullptr))
    return KV_ERR_BAD_ARG;
  if (d->T == 0) return KV_OK;
  if (tc_w) {
    rc = tc_bwd_weight(tc_view(*d), X, dy_image, fparams, mean, rstd, dWp, (cudaStream_t)stream);
    if (rc != KV_OK) return rc;
    if (dbias == nullptr) return KV_OK;
    return simt_col_sums(*d, dY, act_mask, dbias, (cudaStream_t)stream);
  }
  return simt_bwd_weight(*d, X, dY, act_mask, fparams, mean, rstd, dWp, dbias, (cudaStream_t)stream);
}

int kv_kan_unpack_grads(const kv_kan_desc_t* d, const float* dWp, const float* dfparams,
                        const float* dbias, const uint64_t* ptr_table, const uint64_t* grad_table,
                        void* stream) {
  int rc = check_desc(d);
  if (rc != KV_OK) return rc;
  if (ptr_table == nullptr || grad_table == nullptr) return KV_ERR_BAD_ARG;
  const size_t n_oi = (size_t)d->out_features * d->in_features;
  int bx = (int)((n_oi + 255) / 256);
  if (bx > 1024) bx = 1024;
  unpack_grads_kernel<<<dim3(bx, d->n_groups), 256, 0, (cudaStream_t)stream>>>(*d, dWp, dfparams, dbias,
                                                                               ptr_table, grad_table);
  KV_CHECK_LAUNCH("unpack_grads");
  return KV_OK;
}

}  // extern "C"
