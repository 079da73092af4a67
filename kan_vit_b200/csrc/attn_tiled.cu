// Tiled exact fp32 attention for sequences whose K / V (and Q / dO in the backward) do not fit one CTA's shared
// memory: the fallback of kv_attn_{fwd,bwd} (attn.cu) and kv_attn_general_{fwd,bwd} (attn_general.cu), which keep a whole
// head resident and used to return KV_ERR_BAD_SHAPE beyond ~215 / ~430 tokens.  The reference accepts any length: MSA
// materialises the full score matrix (attention.py:199-200), FlashAttentionFunction walks q / k buckets
// (utils.py:137-295).  Same arithmetic and special cases as attn_general.cu (hidden scores -FLT_MAX, mask-false weights
// zeroed after the exponential, row sum clamped at 1e-10), in the flash form of the reference's own bucket loops:
//   forward   one CTA = 32 query rows of one (sample, head); keys / values stream through shared memory in tiles of
//             64 with a running row max and row sum (utils.py:197-222);
//   backward  column pass (one CTA = 32 keys, queries stream in tiles of 64: dK, dV) and row pass (one CTA = 32
//             queries, keys stream: dQ), both from p = exp(s - lse) and D_i (utils.py:270-289).  D_i = sum(dO_i * O_i)
//             when the caller has O (general entry point); the MSA entry point has no O, so a first row pass forms
//             D_i = sum_j p_ij dp_ij and parks it in element 0 of the row's dQ slot, which the row pass that writes
//             dQ reads before it overwrites the row (three launches on one stream, no scratch allocation).
// One warp per query (or key) row, lanes over the streamed tile, then lanes over the head dimension: no atomics,
// deterministic.  Heads are addressed as column h * hs of a row-major matrix with row stride ld, which covers both
// layouts: 'b n (h d)' (hs = d_head) and the grouped q/k/v layer's [T][H][3][d_head] (hs = 3 d_head).
#include <cfloat>
#include "internal.cuh"

namespace kv {

constexpr int TT_THREADS = 256;
constexpr int TT_WARPS = TT_THREADS / 32;
constexpr int TT_RPW = 4;                     // rows per warp
constexpr int TT_ROWS = TT_WARPS * TT_RPW;    // rows (queries or keys) per CTA
constexpr int TT_TILE = 64;                   // streamed rows per tile
constexpr int TT_MAX_DH = 128;
constexpr int TT_NC = TT_MAX_DH / 32;         // head-dimension elements per lane

__device__ __forceinline__ void tt_stage(float* dst, const float* __restrict__ src, int rows, int dh, int lds,
                                         size_t row_stride, int tid) {
  for (int p = tid; p < rows * dh; p += TT_THREADS) {
    const int c = p % dh, n = p / dh;
    dst[n * lds + c] = src[(size_t)n * row_stride + c];
  }
}

__device__ __forceinline__ bool tt_hidden(const AttnTiled& a, int i, int j, int diff) { return a.causal && j > i - diff; }

// ------------------------------------------------------------------------------------------------
// forward: out_i = softmax_j(s_ij) v_j (+ residual), lse_i
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TT_THREADS)
attn_tiled_fwd_kernel(AttnTiled a, const float* __restrict__ residual, float* __restrict__ out, int ldo, int hso,
                      float* __restrict__ lse) {
  extern __shared__ float sm[];
  const int dh = a.dh, Nq = a.Nq, Nk = a.Nk, lds = dh + 1;
  float* Ks = sm;
  float* Vs = Ks + TT_TILE * lds;
  float* Qs = Vs + TT_TILE * lds;            // [TT_ROWS][dh]
  float* Ps = Qs + TT_ROWS * dh;             // [TT_WARPS][TT_TILE]
  const int b = blockIdx.y / a.H, h = blockIdx.y % a.H;
  const int q0 = blockIdx.x * TT_ROWS;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float* qb = a.q + (size_t)b * Nq * a.ldq + (size_t)h * a.hs;
  const float* kb = a.k + (size_t)b * Nk * a.ldk + (size_t)h * a.hs;
  const float* vb = a.v + (size_t)b * Nk * a.ldv + (size_t)h * a.hs;
  const int nq = min(TT_ROWS, Nq - q0);
  tt_stage(Qs, qb + (size_t)q0 * a.ldq, nq, dh, dh, a.ldq, tid);
  const float scale = rsqrtf((float)dh);
  const int diff = Nk > Nq ? Nk - Nq : 0;
  const unsigned char* mb = a.mask ? a.mask + (size_t)b * a.msb + (size_t)h * a.msh : nullptr;
  float* prow = Ps + warp * TT_TILE;
  float m[TT_RPW], l[TT_RPW], acc[TT_RPW][TT_NC];
#pragma unroll
  for (int r = 0; r < TT_RPW; ++r) {
    m[r] = -FLT_MAX;
    l[r] = 0.0f;
#pragma unroll
    for (int cc = 0; cc < TT_NC; ++cc) acc[r][cc] = 0.0f;
  }
  for (int k0 = 0; k0 < Nk; k0 += TT_TILE) {
    const int tk = min(TT_TILE, Nk - k0);
    __syncthreads();   // the previous tile is consumed (and Qs is staged)
    tt_stage(Ks, kb + (size_t)k0 * a.ldk, tk, dh, lds, a.ldk, tid);
    tt_stage(Vs, vb + (size_t)k0 * a.ldv, tk, dh, lds, a.ldv, tid);
    __syncthreads();
#pragma unroll
    for (int r = 0; r < TT_RPW; ++r) {
      const int i = q0 + warp * TT_RPW + r;
      if (i >= Nq) continue;                 // warp-uniform
      const float* q = Qs + (warp * TT_RPW + r) * dh;
      const unsigned char* mrow = mb ? mb + (size_t)i * a.msq + k0 : nullptr;
      float mx = -FLT_MAX;
      for (int jj = lane; jj < tk; jj += 32) {
        const float* k = Ks + jj * lds;
        float s = 0.0f;
        for (int c = 0; c < dh; ++c) s = fmaf(q[c], k[c], s);
        s *= scale;
        if ((mrow && !mrow[jj]) || tt_hidden(a, i, k0 + jj, diff)) s = -FLT_MAX;
        prow[jj] = s;
        mx = fmaxf(mx, s);
      }
      const float m_new = fmaxf(m[r], warp_max(mx));
      const float corr = expf(m[r] - m_new);
      float sum = 0.0f;
      for (int jj = lane; jj < tk; jj += 32) {
        float e = expf(prow[jj] - m_new);
        if (mrow && !mrow[jj]) e = 0.0f;
        prow[jj] = e;
        sum += e;
      }
      l[r] = l[r] * corr + warp_sum(sum);
      m[r] = m_new;
      __syncwarp();
#pragma unroll
      for (int cc = 0; cc < TT_NC; ++cc) {
        const int c = lane + 32 * cc;
        if (c < dh) {
          float s2 = 0.0f;
          for (int jj = 0; jj < tk; ++jj) s2 = fmaf(prow[jj], Vs[jj * lds + c], s2);
          acc[r][cc] = acc[r][cc] * corr + s2;
        }
      }
      __syncwarp();
    }
  }
#pragma unroll
  for (int r = 0; r < TT_RPW; ++r) {
    const int i = q0 + warp * TT_RPW + r;
    if (i >= Nq) continue;
    const float sum = fmaxf(l[r], 1e-10f);
    const float inv = 1.0f / sum;
    const size_t o = ((size_t)b * Nq + i) * ldo + (size_t)h * hso;
#pragma unroll
    for (int cc = 0; cc < TT_NC; ++cc) {
      const int c = lane + 32 * cc;
      if (c < dh) out[o + c] = acc[r][cc] * inv + (residual ? residual[o + c] : 0.0f);
    }
    if (lane == 0) lse[((size_t)b * a.H + h) * Nq + i] = logf(sum) + m[r];
  }
}

// ------------------------------------------------------------------------------------------------
// backward, row pass.  WRITE_D: D_i = sum_j p_ij dp_ij -> dq[i][0] (MSA entry point: no O);
// otherwise dQ_i = sum_j p_ij scale (dp_ij - D_i) k_j with D_i = dO_i . O_i (o != nullptr) or read from dq[i][0].
// ------------------------------------------------------------------------------------------------
template <bool WRITE_D>
__global__ void __launch_bounds__(TT_THREADS)
attn_tiled_bwd_rows_kernel(AttnTiled a, const float* __restrict__ o, const float* __restrict__ dout, int ldo, int hso,
                           const float* __restrict__ lse, float* __restrict__ dq) {
  extern __shared__ float sm[];
  const int dh = a.dh, Nq = a.Nq, Nk = a.Nk, lds = dh + 1;
  float* Ks = sm;
  float* Vs = Ks + TT_TILE * lds;
  float* Qs = Vs + TT_TILE * lds;            // [TT_ROWS][dh]
  float* Gs = Qs + TT_ROWS * dh;             // dO [TT_ROWS][dh]
  float* Ps = Gs + TT_ROWS * dh;             // [TT_WARPS][TT_TILE]
  const int b = blockIdx.y / a.H, h = blockIdx.y % a.H;
  const int q0 = blockIdx.x * TT_ROWS;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float* qb = a.q + (size_t)b * Nq * a.ldq + (size_t)h * a.hs;
  const float* kb = a.k + (size_t)b * Nk * a.ldk + (size_t)h * a.hs;
  const float* vb = a.v + (size_t)b * Nk * a.ldv + (size_t)h * a.hs;
  const float* gb = dout + (size_t)b * Nq * ldo + (size_t)h * hso;
  float* dqb = dq + (size_t)b * Nq * a.ldq + (size_t)h * a.hs;
  const int nq = min(TT_ROWS, Nq - q0);
  tt_stage(Qs, qb + (size_t)q0 * a.ldq, nq, dh, dh, a.ldq, tid);
  tt_stage(Gs, gb + (size_t)q0 * ldo, nq, dh, dh, ldo, tid);
  __syncthreads();
  const float scale = rsqrtf((float)dh);
  const int diff = Nk > Nq ? Nk - Nq : 0;
  const unsigned char* mb = a.mask ? a.mask + (size_t)b * a.msb + (size_t)h * a.msh : nullptr;
  float* grow = Ps + warp * TT_TILE;
  float li[TT_RPW], di[TT_RPW], acc[TT_RPW][TT_NC];
#pragma unroll
  for (int r = 0; r < TT_RPW; ++r) {
    const int i = q0 + warp * TT_RPW + r;
    li[r] = 0.0f;
    di[r] = 0.0f;
#pragma unroll
    for (int cc = 0; cc < TT_NC; ++cc) acc[r][cc] = 0.0f;
    if (i >= Nq) continue;
    li[r] = lse[((size_t)b * a.H + h) * Nq + i];
    if (!WRITE_D) {
      if (o != nullptr) {
        const float* og = o + ((size_t)b * Nq + i) * ldo + (size_t)h * hso;
        const float* g = Gs + (warp * TT_RPW + r) * dh;
        float d = 0.0f;
        for (int c = lane; c < dh; c += 32) d = fmaf(g[c], og[c], d);
        di[r] = warp_sum(d);
      } else {
        di[r] = dqb[(size_t)i * a.ldq];        // parked by the WRITE_D pass; this warp overwrites the row at the end
      }
    }
  }
  for (int k0 = 0; k0 < Nk; k0 += TT_TILE) {
    const int tk = min(TT_TILE, Nk - k0);
    __syncthreads();
    tt_stage(Ks, kb + (size_t)k0 * a.ldk, tk, dh, lds, a.ldk, tid);
    tt_stage(Vs, vb + (size_t)k0 * a.ldv, tk, dh, lds, a.ldv, tid);
    __syncthreads();
#pragma unroll
    for (int r = 0; r < TT_RPW; ++r) {
      const int i = q0 + warp * TT_RPW + r;
      if (i >= Nq) continue;
      const float* q = Qs + (warp * TT_RPW + r) * dh;
      const float* g = Gs + (warp * TT_RPW + r) * dh;
      const unsigned char* mrow = mb ? mb + (size_t)i * a.msq + k0 : nullptr;
      float dsum = 0.0f;
      for (int jj = lane; jj < tk; jj += 32) {
        const float* k = Ks + jj * lds;
        const float* v = Vs + jj * lds;
        float s = 0.0f, dp = 0.0f;
        for (int c = 0; c < dh; ++c) {
          s = fmaf(q[c], k[c], s);
          dp = fmaf(g[c], v[c], dp);
        }
        s *= scale;
        if (tt_hidden(a, i, k0 + jj, diff)) s = -FLT_MAX;
        const float p = (mrow && !mrow[jj]) ? 0.0f : expf(s - li[r]);
        if (WRITE_D) dsum = fmaf(p, dp, dsum);
        else grow[jj] = p * scale * (dp - di[r]);
      }
      if (WRITE_D) {
        di[r] += warp_sum(dsum);
      } else {
        __syncwarp();
#pragma unroll
        for (int cc = 0; cc < TT_NC; ++cc) {
          const int c = lane + 32 * cc;
          if (c < dh) {
            float s2 = 0.0f;
            for (int jj = 0; jj < tk; ++jj) s2 = fmaf(grow[jj], Ks[jj * lds + c], s2);
            acc[r][cc] += s2;
          }
        }
        __syncwarp();
      }
    }
  }
#pragma unroll
  for (int r = 0; r < TT_RPW; ++r) {
    const int i = q0 + warp * TT_RPW + r;
    if (i >= Nq) continue;
    if (WRITE_D) {
      if (lane == 0) dqb[(size_t)i * a.ldq] = di[r];
    } else {
#pragma unroll
      for (int cc = 0; cc < TT_NC; ++cc) {
        const int c = lane + 32 * cc;
        if (c < dh) dqb[(size_t)i * a.ldq + c] = acc[r][cc];
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// backward, column pass: dK_j = sum_i p_ij scale (dp_ij - D_i) q_i, dV_j = sum_i p_ij dO_i.
// One CTA = 32 keys (resident), queries / dO / lse / D stream in tiles of 64.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TT_THREADS)
attn_tiled_bwd_cols_kernel(AttnTiled a, const float* __restrict__ o, const float* __restrict__ dout, int ldo, int hso,
                           const float* __restrict__ lse, const float* __restrict__ dq_parked, float* __restrict__ dk,
                           float* __restrict__ dv) {
  extern __shared__ float sm[];
  const int dh = a.dh, Nq = a.Nq, Nk = a.Nk, lds = dh + 1;
  float* Qs = sm;                            // [TT_TILE][lds]
  float* Gs = Qs + TT_TILE * lds;            // dO
  float* Kb = Gs + TT_TILE * lds;            // [TT_ROWS][dh]
  float* Vb = Kb + TT_ROWS * dh;
  float* Ls = Vb + TT_ROWS * dh;             // [TT_TILE]
  float* Ds = Ls + TT_TILE;
  float* Ps = Ds + TT_TILE;                  // [TT_WARPS][2][TT_TILE]
  const int b = blockIdx.y / a.H, h = blockIdx.y % a.H;
  const int j0 = blockIdx.x * TT_ROWS;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float* qb = a.q + (size_t)b * Nq * a.ldq + (size_t)h * a.hs;
  const float* kb = a.k + (size_t)b * Nk * a.ldk + (size_t)h * a.hs;
  const float* vb = a.v + (size_t)b * Nk * a.ldv + (size_t)h * a.hs;
  const float* gb = dout + (size_t)b * Nq * ldo + (size_t)h * hso;
  const int nk = min(TT_ROWS, Nk - j0);
  tt_stage(Kb, kb + (size_t)j0 * a.ldk, nk, dh, dh, a.ldk, tid);
  tt_stage(Vb, vb + (size_t)j0 * a.ldv, nk, dh, dh, a.ldv, tid);
  const float scale = rsqrtf((float)dh);
  const int diff = Nk > Nq ? Nk - Nq : 0;
  const unsigned char* mb = a.mask ? a.mask + (size_t)b * a.msb + (size_t)h * a.msh : nullptr;
  float* p_row = Ps + warp * 2 * TT_TILE;
  float* g_row = p_row + TT_TILE;
  float ak[TT_RPW][TT_NC], av[TT_RPW][TT_NC];
#pragma unroll
  for (int r = 0; r < TT_RPW; ++r)
#pragma unroll
    for (int cc = 0; cc < TT_NC; ++cc) ak[r][cc] = av[r][cc] = 0.0f;
  for (int i0 = 0; i0 < Nq; i0 += TT_TILE) {
    const int tq = min(TT_TILE, Nq - i0);
    __syncthreads();
    tt_stage(Qs, qb + (size_t)i0 * a.ldq, tq, dh, lds, a.ldq, tid);
    tt_stage(Gs, gb + (size_t)i0 * ldo, tq, dh, lds, ldo, tid);
    for (int ii = tid; ii < tq; ii += TT_THREADS) Ls[ii] = lse[((size_t)b * a.H + h) * Nq + i0 + ii];
    __syncthreads();
    // D of the tile's rows: dO_i . O_i (general entry point) or the value parked by the first row pass
    for (int ii = warp; ii < tq; ii += TT_WARPS) {
      if (o != nullptr) {
        const float* og = o + ((size_t)b * Nq + i0 + ii) * ldo + (size_t)h * hso;
        float d = 0.0f;
        for (int c = lane; c < dh; c += 32) d = fmaf(Gs[ii * lds + c], og[c], d);
        d = warp_sum(d);
        if (lane == 0) Ds[ii] = d;
      } else if (lane == 0) {
        Ds[ii] = dq_parked[((size_t)b * Nq + i0 + ii) * a.ldq + (size_t)h * a.hs];
      }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < TT_RPW; ++r) {
      const int j = j0 + warp * TT_RPW + r;
      if (j >= Nk) continue;                 // warp-uniform
      const float* k = Kb + (warp * TT_RPW + r) * dh;
      const float* v = Vb + (warp * TT_RPW + r) * dh;
      for (int ii = lane; ii < tq; ii += 32) {
        const float* q = Qs + ii * lds;
        const float* g = Gs + ii * lds;
        float s = 0.0f, dp = 0.0f;
        for (int c = 0; c < dh; ++c) {
          s = fmaf(q[c], k[c], s);
          dp = fmaf(g[c], v[c], dp);
        }
        s *= scale;
        if (tt_hidden(a, i0 + ii, j, diff)) s = -FLT_MAX;
        const bool off = mb && !mb[(size_t)(i0 + ii) * a.msq + j];
        const float p = off ? 0.0f : expf(s - Ls[ii]);
        p_row[ii] = p;
        g_row[ii] = p * scale * (dp - Ds[ii]);
      }
      __syncwarp();
#pragma unroll
      for (int cc = 0; cc < TT_NC; ++cc) {
        const int c = lane + 32 * cc;
        if (c < dh) {
          float sk = 0.0f, sv = 0.0f;
          for (int ii = 0; ii < tq; ++ii) {
            sk = fmaf(g_row[ii], Qs[ii * lds + c], sk);
            sv = fmaf(p_row[ii], Gs[ii * lds + c], sv);
          }
          ak[r][cc] += sk;
          av[r][cc] += sv;
        }
      }
      __syncwarp();
    }
  }
  float* dkb = dk + (size_t)b * Nk * a.ldk + (size_t)h * a.hs;
  float* dvb = dv + (size_t)b * Nk * a.ldv + (size_t)h * a.hs;
#pragma unroll
  for (int r = 0; r < TT_RPW; ++r) {
    const int j = j0 + warp * TT_RPW + r;
    if (j >= Nk) continue;
#pragma unroll
    for (int cc = 0; cc < TT_NC; ++cc) {
      const int c = lane + 32 * cc;
      if (c < dh) {
        dkb[(size_t)j * a.ldk + c] = ak[r][cc];
        dvb[(size_t)j * a.ldv + c] = av[r][cc];
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static size_t tt_rows_smem(int dh, bool bwd) {
  return sizeof(float) * ((size_t)2 * TT_TILE * (dh + 1) + (size_t)(bwd ? 2 : 1) * TT_ROWS * dh + TT_WARPS * TT_TILE);
}
static size_t tt_cols_smem(int dh) {
  return sizeof(float) * ((size_t)2 * TT_TILE * (dh + 1) + (size_t)2 * TT_ROWS * dh + 2 * TT_TILE +
                          (size_t)TT_WARPS * 2 * TT_TILE);
}

static int tt_check(const AttnTiled& a) {
  if (a.dh <= 0 || a.dh > TT_MAX_DH) return KV_ERR_BAD_SHAPE;
  if ((long long)a.B * a.H > 65535) return KV_ERR_BAD_SHAPE;   // grid.y
  return KV_OK;
}

int attn_tiled_fwd(const AttnTiled& a, const float* residual, float* out, int ldo, int hso, float* lse,
                   cudaStream_t st) {
  const int rc = tt_check(a);
  if (rc != KV_OK) return rc;
  const size_t smem = tt_rows_smem(a.dh, false);
  KV_SMEM_OPT_IN(attn_tiled_fwd_kernel, (int)tt_rows_smem(TT_MAX_DH, false), "attn_tiled_fwd smem");
  dim3 grid(ceil_div(a.Nq, TT_ROWS), a.B * a.H);
  attn_tiled_fwd_kernel<<<grid, TT_THREADS, smem, st>>>(a, residual, out, ldo, hso, lse);
  KV_CHECK_LAUNCH("attn_tiled_fwd");
  return KV_OK;
}

int attn_tiled_bwd(const AttnTiled& a, const float* o, const float* dout, int ldo, int hso, const float* lse,
                   float* dq, float* dk, float* dv, cudaStream_t st) {
  const int rc = tt_check(a);
  if (rc != KV_OK) return rc;
  const size_t smem_r = tt_rows_smem(a.dh, true), smem_c = tt_cols_smem(a.dh);
  KV_SMEM_OPT_IN(attn_tiled_bwd_rows_kernel<true>, (int)tt_rows_smem(TT_MAX_DH, true), "attn_tiled_bwd smem");
  KV_SMEM_OPT_IN(attn_tiled_bwd_rows_kernel<false>, (int)tt_rows_smem(TT_MAX_DH, true), "attn_tiled_bwd smem");
  KV_SMEM_OPT_IN(attn_tiled_bwd_cols_kernel, (int)tt_cols_smem(TT_MAX_DH), "attn_tiled_bwd smem");
  dim3 grid_r(ceil_div(a.Nq, TT_ROWS), a.B * a.H), grid_c(ceil_div(a.Nk, TT_ROWS), a.B * a.H);
  if (o == nullptr) {   // no forward output: D_i from a row pass of its own, parked in dq[i][0]
    attn_tiled_bwd_rows_kernel<true><<<grid_r, TT_THREADS, smem_r, st>>>(a, nullptr, dout, ldo, hso, lse, dq);
    KV_CHECK_LAUNCH("attn_tiled_bwd D pass");
  }
  attn_tiled_bwd_cols_kernel<<<grid_c, TT_THREADS, smem_c, st>>>(a, o, dout, ldo, hso, lse, dq, dk, dv);
  KV_CHECK_LAUNCH("attn_tiled_bwd column pass");
  attn_tiled_bwd_rows_kernel<false><<<grid_r, TT_THREADS, smem_r, st>>>(a, o, dout, ldo, hso, lse, dq);
  KV_CHECK_LAUNCH("attn_tiled_bwd row pass");
  return KV_OK;
}

}  // namespace kv
