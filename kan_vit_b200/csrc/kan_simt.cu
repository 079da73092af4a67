// fp32 SIMT "basis-prologue GEMM" kernels: forward, input gradient, weight gradient.
// This is the exact-arithmetic path (parity target 1e-5 against the reference's fp32
// modules); the tcgen05 bf16 path lives in kan_tc.cu and shares basis.cuh.
//
// Layouts (see include/kanvit_b200.h):  X [T][ldx] fp32, Wp [g][out][in*M] fp32 with
// K index i*M+m, Y [rows][ldy] fp32.  The basis tile is built in shared memory only.
#include "basis.cuh"

namespace kv {

constexpr int BM = 64;       // rows per CTA tile
constexpr int BN = 64;       // outputs per CTA tile
constexpr int KC = 64;       // max basis columns per K chunk ( >= KV_MAX_M )
constexpr int NTHREADS = 256;
constexpr int LDA = BM + 4;  // As[k][row]
constexpr int LDB = BN + 4;  // Bs[k][o]

struct ColSink {  // writes Phi_m into a k-major shared tile: base[m * stride]
  static constexpr bool kDense = true;
  float* base;
  int stride;
  __device__ __forceinline__ void put(int m, float v) { base[m * stride] = v; }
};
struct RowSink {  // writes Phi_m into a row-major shared tile: base[m]
  static constexpr bool kDense = true;
  float* base;
  __device__ __forceinline__ void put(int m, float v) { base[m] = v; }
};
struct DotSink {  // accumulates sum_m v_m * s[m]
  static constexpr bool kDense = false;
  const float* s;
  float acc;
  __device__ __forceinline__ void put(int m, float v) { acc += v * s[m]; }
};
struct FastkanDotSink {  // RBF part (m < G) and base part (m == G) kept apart
  static constexpr bool kDense = false;
  const float* s;
  int G;
  float acc_rbf, acc_base;
  __device__ __forceinline__ void put(int m, float v) {
    if (m < G) acc_rbf += v * s[m]; else acc_base += v * s[m];
  }
};

// ------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------
template <int F>
__global__ void __launch_bounds__(NTHREADS)
kan_fwd_f32_kernel(const kv_kan_desc_t d, const float* __restrict__ X, const float* __restrict__ Wp,
                   const float* __restrict__ fparams, const float* __restrict__ bias,
                   const float* __restrict__ mean, const float* __restrict__ rstd,
                   const float* __restrict__ residual, const float* __restrict__ pos,
                   float* __restrict__ Y) {
  __shared__ __align__(16) float As[KC * LDA];
  __shared__ __align__(16) float Bs[KC * LDB];
  const int tid = threadIdx.x;
  const int g = blockIdx.z;
  const int t0 = blockIdx.x * BM;
  const int o0 = blockIdx.y * BN;
  const int in = d.in_features, out = d.out_features, M = d.M;
  const int K = in * M;
  const int slice = g / d.x_share;
  const int n_slices = d.n_groups / d.x_share;
  const int xoff = slice * in;
  const BasisCtx c = make_ctx(d, fparams, g);
  const float* Wg = Wp + (size_t)g * out * K;
  const int BI = KC / M;  // features per chunk (M <= KC checked on the host)
  const int tx = tid % 16, ty = tid / 16;

  float acc[4][4];
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int q = 0; q < 4; ++q) acc[r][q] = 0.0f;

  for (int i0 = 0; i0 < in; i0 += BI) {
    const int nf = min(BI, in - i0);
    const int cw = nf * M;
    // ---- prologue: basis tile As[il*M+m][row]
    for (int p = tid; p < BM * nf; p += NTHREADS) {
      const int row = p % BM, il = p / BM;
      const int t = t0 + row;
      ColSink s{&As[(il * M) * LDA + row], LDA};
      if (t < d.T) {
        const float x = X[(size_t)t * d.ldx + xoff + i0 + il];
        float xn = x;
        if (F == KV_FASTKAN) {
          const float mu = (d.flags & KV_FLAG_NO_LAYERNORM) ? 0.f : mean[(size_t)t * n_slices + slice];
          const float rs = (d.flags & KV_FLAG_NO_LAYERNORM) ? 1.f : rstd[(size_t)t * n_slices + slice];
          xn = fastkan_xn(c, i0 + il, x, mu, rs);
        }
        basis_eval<F>(c, i0 + il, x, xn, s);
      } else {
        for (int m = 0; m < M; ++m) s.put(m, 0.0f);
      }
    }
    // ---- weight tile Bs[kk][o]
    for (int p = tid; p < BN * cw; p += NTHREADS) {
      const int kk = p % cw, o = p / cw;
      float w = 0.0f;
      if (o0 + o < out) w = Wg[(size_t)(o0 + o) * K + (size_t)i0 * M + kk];
      Bs[kk * LDB + o] = w;
    }
    __syncthreads();
    // two-level summation: the K chunk is summed on its own and then added to the running total, so the rounding
    // of a K = in*M <= 27 648 contraction grows like sqrt(chunks), not sqrt(K) (the 1e-5 fp32 parity bar)
    float part[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int q = 0; q < 4; ++q) part[r][q] = 0.0f;
    for (int kk = 0; kk < cw; ++kk) {
      const float4 a = *reinterpret_cast<const float4*>(&As[kk * LDA + ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[kk * LDB + tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w};
      const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int q = 0; q < 4; ++q) part[r][q] = fmaf(av[r], bv[q], part[r][q]);
    }
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[r][q] += part[r][q];
    __syncthreads();
  }
  // ---- epilogue
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int t = t0 + ty * 4 + r;
    if (t >= d.T) continue;
    const int prow = map_row(t, d.rows_in, d.rows_out, d.row_shift);
    const int posrow = d.rows_in ? (d.row_shift + t % d.rows_in) : t;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int o = o0 + tx * 4 + q;
      if (o >= out) continue;
      const int col = g * out + o;
      float v = acc[r][q];
      if (bias) v += bias[col];
      if (pos) v += pos[(size_t)posrow * (d.n_groups * out) + col];
      if (d.flags & KV_FLAG_RELU) v = fmaxf(v, 0.0f);
      if (residual) v += residual[(size_t)prow * d.ldy + col];
      Y[(size_t)prow * d.ldy + col] = v;
    }
  }
}

// ------------------------------------------------------------------------------
// input gradient
// ------------------------------------------------------------------------------
constexpr int BO = 32;  // outputs per reduction step

template <int F>
__global__ void __launch_bounds__(NTHREADS)
kan_bwd_input_f32_kernel(const kv_kan_desc_t d, const float* __restrict__ X,
                         const float* __restrict__ dY, const float* __restrict__ act_mask,
                         const float* __restrict__ Wp, const float* __restrict__ fparams,
                         const float* __restrict__ mean, const float* __restrict__ rstd,
                         float* __restrict__ dX, float* __restrict__ dXn,
                         float* __restrict__ dfparams) {
  constexpr int LDS = KC + 1;
  constexpr int SMEM_FLOATS = (BM * LDS > BO * (LDA + LDB)) ? BM * LDS : BO * (LDA + LDB);
  __shared__ __align__(16) float smem[SMEM_FLOATS];  // As+Bs during the GEMM, S afterwards
  __shared__ float DX[BM * (KC + 1)];
  __shared__ float dfreq_s[KV_MAX_M];
  float* As = smem;             // [BO][LDA]
  float* Bs = smem + BO * LDA;  // [BO][LDB]
  float* S = smem;              // [BM][LDS]
  const int tid = threadIdx.x;
  const int slice = blockIdx.z;
  const int n_slices = d.n_groups / d.x_share;
  const int t0 = blockIdx.x * BM;
  const int in = d.in_features, out = d.out_features, M = d.M;
  const int K = in * M;
  const int BI = KC / M;
  const int i0 = blockIdx.y * BI;
  const int nf = min(BI, in - i0);
  const int cw = nf * M;
  const int xoff = slice * in;
  const int tx = tid % 16, ty = tid / 16;

  for (int p = tid; p < BM * (KC + 1); p += NTHREADS) DX[p] = 0.0f;

  for (int gs = 0; gs < d.x_share; ++gs) {
    const int g = slice * d.x_share + gs;
    const BasisCtx c = make_ctx(d, fparams, g);
    const float* Wg = Wp + (size_t)g * out * K + (size_t)i0 * M;
    if (F == KV_SINE) {
      if (tid < KV_MAX_M) dfreq_s[tid] = 0.0f;
    }
    float acc[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[r][q] = 0.0f;
    __syncthreads();  // S of the previous group fully consumed; DX zeroed

    for (int ob = 0; ob < out; ob += BO) {
      // As[oo][row] = dY[row, g*out + ob + oo]
      for (int p = tid; p < BM * BO; p += NTHREADS) {
        const int oo = p % BO, row = p / BO;
        const int t = t0 + row;
        float v = 0.0f;
        if (t < d.T && ob + oo < out) {
          const size_t idx = (size_t)map_row(t, d.rows_in, d.rows_out, d.row_shift) * d.ldy + g * out + ob + oo;
          v = dY[idx];
          if ((d.flags & KV_FLAG_RELU) && !(act_mask[idx] > 0.0f)) v = 0.0f;
        }
        As[oo * LDA + row] = v;
      }
      // Bs[oo][kk] = Wp[g][ob+oo][i0*M + kk]
      for (int p = tid; p < BO * cw; p += NTHREADS) {
        const int kk = p % cw, oo = p / cw;
        Bs[oo * LDB + kk] = (ob + oo < out) ? Wg[(size_t)(ob + oo) * K + kk] : 0.0f;
      }
      if (cw < KC) {  // keep the unused columns finite
        for (int p = tid; p < BO * (KC - cw); p += NTHREADS) {
          const int kk = cw + p % (KC - cw), oo = p / (KC - cw);
          Bs[oo * LDB + kk] = 0.0f;
        }
      }
      __syncthreads();
      float part[4][4];   // two-level summation over the outputs (see the forward kernel)
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int q = 0; q < 4; ++q) part[r][q] = 0.0f;
#pragma unroll 8
      for (int oo = 0; oo < BO; ++oo) {
        const float4 a = *reinterpret_cast<const float4*>(&As[oo * LDA + ty * 4]);
        const float4 b = *reinterpret_cast<const float4*>(&Bs[oo * LDB + tx * 4]);
        const float av[4] = {a.x, a.y, a.z, a.w};
        const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
          for (int q = 0; q < 4; ++q) part[r][q] = fmaf(av[r], bv[q], part[r][q]);
      }
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[r][q] += part[r][q];
      __syncthreads();
    }
    // S[row][kk] = (dY . Wp)[row, i0*M + kk]
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int q = 0; q < 4; ++q) S[(ty * 4 + r) * LDS + tx * 4 + q] = acc[r][q];
    __syncthreads();
    // contraction with Phi'
    for (int p0 = 0; p0 < BM * nf; p0 += NTHREADS) {
      const int p = p0 + tid;
      const bool active = p < BM * nf;
      const int row = active ? p % BM : 0, il = active ? p / BM : 0;
      const int t = t0 + row;
      const bool live = active && t < d.T;
      float x = 0.0f;
      if (live) x = X[(size_t)t * d.ldx + xoff + i0 + il];
      const float* srow = &S[row * LDS + il * M];
      if (F == KV_SINE) {
        const float* freq = c.fp;
        const float* ph = c.fp + c.G + (size_t)(i0 + il) * c.G;
        float dx = 0.0f;
        for (int q = 0; q < c.G; ++q) {
          float contrib = 0.0f;
          if (live) {
            const float cs = cosf(x * freq[q] + ph[q]);
            const float sv = srow[q];
            dx += freq[q] * cs * sv;
            contrib = x * cs * sv;
          }
          contrib = warp_sum(contrib);
          if ((tid & 31) == 0) atomicAdd(&dfreq_s[q], contrib);
        }
        if (live) DX[row * (KC + 1) + il] += dx;
      } else if (F == KV_FASTKAN) {
        if (live) {
          float mu = 0.f, rs = 1.f;
          if (!(d.flags & KV_FLAG_NO_LAYERNORM)) {
            mu = mean[(size_t)t * n_slices + slice];
            rs = rstd[(size_t)t * n_slices + slice];
          }
          const float xn = fastkan_xn(c, i0 + il, x, mu, rs);
          FastkanDotSink s{srow, c.G, 0.0f, 0.0f};
          basis_deriv<F>(c, i0 + il, x, xn, s);
          float dx = s.acc_base;
          if (d.flags & KV_FLAG_NO_LAYERNORM) dx += s.acc_rbf;
          else dXn[(size_t)t * (d.n_groups * in) + (size_t)g * in + i0 + il] = s.acc_rbf;
          DX[row * (KC + 1) + il] += dx;
        }
      } else {
        if (live) {
          DotSink s{srow, 0.0f};
          basis_deriv<F>(c, i0 + il, x, x, s);
          DX[row * (KC + 1) + il] += s.acc;
        }
      }
    }
    __syncthreads();
    if (F == KV_SINE) {
      if (tid < c.G && dfparams) atomicAdd(&dfparams[(size_t)g * d.fparam_stride + tid], dfreq_s[tid]);
    }
  }
  // dX tile store (feature index fastest)
  for (int p = tid; p < BM * nf; p += NTHREADS) {
    const int il = p % nf, row = p / nf;
    const int t = t0 + row;
    if (t < d.T) dX[(size_t)t * d.ldx + xoff + i0 + il] = DX[row * (KC + 1) + il];
  }
}

// ------------------------------------------------------------------------------
// weight gradient (packed layout), split over row ranges with fp32 atomics
// ------------------------------------------------------------------------------
constexpr int BT = 32;  // rows per reduction step

template <int F>
__global__ void __launch_bounds__(NTHREADS)
kan_bwd_weight_f32_kernel(const kv_kan_desc_t d, int tsplit, const float* __restrict__ X,
                          const float* __restrict__ dY, const float* __restrict__ act_mask,
                          const float* __restrict__ fparams, const float* __restrict__ mean,
                          const float* __restrict__ rstd, float* __restrict__ dWp,
                          float* __restrict__ dbias) {
  constexpr int LDP = KC + 1;
  __shared__ __align__(16) float As[BT * LDA];  // dY tile [tt][o]
  __shared__ float Ps[BT * LDP];                // Phi tile [tt][kk]
  const int tid = threadIdx.x;
  const int g = blockIdx.z / tsplit;
  const int ts = blockIdx.z % tsplit;
  const int in = d.in_features, out = d.out_features, M = d.M;
  const int K = in * M;
  const int BI = KC / M;
  const int i0 = blockIdx.x * BI;
  const int nf = min(BI, in - i0);
  const int cw = nf * M;
  const int o0 = blockIdx.y * BN;
  const int slice = g / d.x_share;
  const int n_slices = d.n_groups / d.x_share;
  const int xoff = slice * in;
  const BasisCtx c = make_ctx(d, fparams, g);
  const int tx = tid % 16, ty = tid / 16;  // ty -> outputs, tx -> basis columns

  const int n_tiles = ceil_div(d.T, BT);
  const int tiles_per = ceil_div(n_tiles, tsplit);
  const int tile_lo = ts * tiles_per;
  const int tile_hi = min(n_tiles, tile_lo + tiles_per);

  float acc[4][4];
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int q = 0; q < 4; ++q) acc[r][q] = 0.0f;
  float bsum = 0.0f;

  // zero the unused basis columns once
  for (int p = tid; p < BT * LDP; p += NTHREADS) Ps[p] = 0.0f;
  __syncthreads();

  for (int tile = tile_lo; tile < tile_hi; ++tile) {
    const int tb = tile * BT;
    for (int p = tid; p < BT * BN; p += NTHREADS) {
      const int o = p % BN, tt = p / BN;
      const int t = tb + tt;
      float v = 0.0f;
      if (t < d.T && o0 + o < out) {
        const size_t idx = (size_t)map_row(t, d.rows_in, d.rows_out, d.row_shift) * d.ldy + g * out + o0 + o;
        v = dY[idx];
        if ((d.flags & KV_FLAG_RELU) && !(act_mask[idx] > 0.0f)) v = 0.0f;
      }
      As[tt * LDA + o] = v;
    }
    for (int p = tid; p < BT * nf; p += NTHREADS) {
      const int tt = p % BT, il = p / BT;
      const int t = tb + tt;
      RowSink s{&Ps[tt * LDP + il * M]};
      if (t < d.T) {
        const float x = X[(size_t)t * d.ldx + xoff + i0 + il];
        float xn = x;
        if (F == KV_FASTKAN) {
          const float mu = (d.flags & KV_FLAG_NO_LAYERNORM) ? 0.f : mean[(size_t)t * n_slices + slice];
          const float rs = (d.flags & KV_FLAG_NO_LAYERNORM) ? 1.f : rstd[(size_t)t * n_slices + slice];
          xn = fastkan_xn(c, i0 + il, x, mu, rs);
        }
        basis_eval<F>(c, i0 + il, x, xn, s);
      } else {
        for (int m = 0; m < M; ++m) s.put(m, 0.0f);
      }
    }
    __syncthreads();
    float part[4][4];   // two-level summation over the rows (see the forward kernel)
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int q = 0; q < 4; ++q) part[r][q] = 0.0f;
#pragma unroll 4
    for (int tt = 0; tt < BT; ++tt) {
      const float4 a = *reinterpret_cast<const float4*>(&As[tt * LDA + ty * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w};
      float bv[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) bv[q] = Ps[tt * LDP + tx * 4 + q];
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int q = 0; q < 4; ++q) part[r][q] = fmaf(av[r], bv[q], part[r][q]);
    }
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[r][q] += part[r][q];
    if (dbias != nullptr && blockIdx.x == 0 && tid < BN) {
      for (int tt = 0; tt < BT; ++tt) bsum += As[tt * LDA + tid];
    }
    __syncthreads();
  }
  float* dWg = dWp + (size_t)g * out * K + (size_t)i0 * M;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int o = o0 + ty * 4 + r;
    if (o >= out) continue;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int kk = tx * 4 + q;
      if (kk < cw) atomicAdd(&dWg[(size_t)o * K + kk], acc[r][q]);
    }
  }
  if (dbias != nullptr && blockIdx.x == 0 && tid < BN && o0 + tid < out)
    atomicAdd(&dbias[(size_t)g * out + o0 + tid], bsum);
}

// ------------------------------------------------------------------------------
// host dispatch
// ------------------------------------------------------------------------------
template <int F>
static int launch_fwd(const kv_kan_desc_t& d, const float* X, const float* Wp, const float* fparams,
                      const float* bias, const float* mean, const float* rstd, const float* residual,
                      const float* pos, float* Y, cudaStream_t st) {
  dim3 grid(ceil_div(d.T, BM), ceil_div(d.out_features, BN), d.n_groups);
  kan_fwd_f32_kernel<F><<<grid, NTHREADS, 0, st>>>(d, X, Wp, fparams, bias, mean, rstd, residual, pos, Y);
  KV_CHECK_LAUNCH("kan_fwd_f32");
  return KV_OK;
}
template <int F>
static int launch_bwd_input(const kv_kan_desc_t& d, const float* X, const float* dY, const float* mask,
                            const float* Wp, const float* fparams, const float* mean, const float* rstd,
                            float* dX, float* dXn, float* dfparams, cudaStream_t st) {
  const int BI = KC / d.M;
  dim3 grid(ceil_div(d.T, BM), ceil_div(d.in_features, BI), d.n_groups / d.x_share);
  kan_bwd_input_f32_kernel<F><<<grid, NTHREADS, 0, st>>>(d, X, dY, mask, Wp, fparams, mean, rstd, dX, dXn, dfparams);
  KV_CHECK_LAUNCH("kan_bwd_input_f32");
  return KV_OK;
}
template <int F>
static int launch_bwd_weight(const kv_kan_desc_t& d, const float* X, const float* dY, const float* mask,
                             const float* fparams, const float* mean, const float* rstd, float* dWp,
                             float* dbias, cudaStream_t st) {
  const int BI = KC / d.M;
  const int kchunks = ceil_div(d.in_features, BI);
  const int otiles = ceil_div(d.out_features, BN);
  const int base = kchunks * otiles * d.n_groups;
  int tsplit = ceil_div(2 * 148, base);
  const int n_tiles = ceil_div(d.T, BT);
  if (tsplit > n_tiles) tsplit = n_tiles;
  if (tsplit < 1) tsplit = 1;
  if ((long long)d.n_groups * tsplit > 65535) tsplit = 65535 / d.n_groups;
  dim3 grid(kchunks, otiles, d.n_groups * tsplit);
  kan_bwd_weight_f32_kernel<F><<<grid, NTHREADS, 0, st>>>(d, tsplit, X, dY, mask, fparams, mean, rstd, dWp, dbias);
  KV_CHECK_LAUNCH("kan_bwd_weight_f32");
  return KV_OK;
}

#define KV_DISPATCH_FAMILY(fam, CALL)                          \
  switch (fam) {                                               \
    case KV_LINEAR: return CALL(KV_LINEAR);                    \
    case KV_EFFKAN: return CALL(KV_EFFKAN);                    \
    case KV_CHEBY: return CALL(KV_CHEBY);                      \
    case KV_FASTKAN: return CALL(KV_FASTKAN);                  \
    case KV_SINE: return CALL(KV_SINE);                        \
    case KV_FOURIER: return CALL(KV_FOURIER);                  \
    default: return KV_ERR_BAD_FAMILY;                         \
  }

int simt_fwd(const kv_kan_desc_t& d, const float* X, const float* Wp, const float* fparams,
             const float* bias, const float* mean, const float* rstd, const float* residual,
             const float* pos, float* Y, cudaStream_t st) {
#define CALL(F) launch_fwd<F>(d, X, Wp, fparams, bias, mean, rstd, residual, pos, Y, st)
  KV_DISPATCH_FAMILY(d.family, CALL)
#undef CALL
}
int simt_bwd_input(const kv_kan_desc_t& d, const float* X, const float* dY, const float* mask,
                   const float* Wp, const float* fparams, const float* mean, const float* rstd,
                   float* dX, float* dXn, float* dfparams, cudaStream_t st) {
#define CALL(F) launch_bwd_input<F>(d, X, dY, mask, Wp, fparams, mean, rstd, dX, dXn, dfparams, st)
  KV_DISPATCH_FAMILY(d.family, CALL)
#undef CALL
}
// dbias[g][o] += sum_t dY[t, g*out + o] (x ReLU mask): companion of the tensor-core weight gradient
__global__ void col_sums_kernel(const kv_kan_desc_t d, int rows_per_block, const float* __restrict__ dY,
                                const float* __restrict__ act_mask, float* __restrict__ dbias) {
  __shared__ float part[8][33];
  const int ncols = d.n_groups * d.out_features;
  const int col = blockIdx.x * 32 + (threadIdx.x & 31);
  const int sub = threadIdx.x >> 5;
  const int r0 = blockIdx.y * rows_per_block;
  const int r1 = min(d.T, r0 + rows_per_block);
  float acc = 0.0f;
  if (col < ncols) {
    for (int t = r0 + sub; t < r1; t += 8) {
      const size_t idx = (size_t)map_row(t, d.rows_in, d.rows_out, d.row_shift) * d.ldy + col;
      float v = dY[idx];
      if ((d.flags & KV_FLAG_RELU) && !(act_mask[idx] > 0.0f)) v = 0.0f;
      acc += v;
    }
  }
  part[sub][threadIdx.x & 31] = acc;
  __syncthreads();
  if (sub == 0 && col < ncols) {
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += part[k][threadIdx.x];
    atomicAdd(&dbias[col], s);
  }
}

int simt_col_sums(const kv_kan_desc_t& d, const float* dY, const float* mask, float* dbias, cudaStream_t st) {
  const int ncols = d.n_groups * d.out_features;
  int ysplit = ceil_div(d.T, 2048);
  if (ysplit > 64) ysplit = 64;
  const int rows_per_block = ceil_div(d.T, ysplit);
  col_sums_kernel<<<dim3(ceil_div(ncols, 32), ysplit), 256, 0, st>>>(d, rows_per_block, dY, mask, dbias);
  KV_CHECK_LAUNCH("col_sums_kernel");
  return KV_OK;
}

int simt_bwd_weight(const kv_kan_desc_t& d, const float* X, const float* dY, const float* mask,
                    const float* fparams, const float* mean, const float* rstd, float* dWp,
                    float* dbias, cudaStream_t st) {
#define CALL(F) launch_bwd_weight<F>(d, X, dY, mask, fparams, mean, rstd, dWp, dbias, st)
  KV_DISPATCH_FAMILY(d.family, CALL)
#undef CALL
}

}  // namespace kv
