// Attention of the reference's FlashAttention module (attention.py:61-109 -> utils.py:137-295): separate
// q / k / v matrices in 'b n (h d)' layout (what the bias-free to_q / to_kv projections write), a key
// length that may differ from the query length (context), an optional boolean mask and the reference's
// causal rule.  Exact fp32, one CTA per (sample, head), K/V (and Q/dO in the backward) resident in
// shared memory, one warp per query row (forward, backward row pass) or key row (backward column pass):
// no atomics, deterministic.
//
// The arithmetic follows the reference tile loops with a single tile (the bucket sizes only change the
// summation order):
//   s_ij = q_i.k_j * d_h^-0.5;   masked (mask false or causal) -> -FLT_MAX              utils.py:186-195
//   causal: key j is visible to query i iff j <= i - max(Nk - Nq, 0)                    utils.py:142,175,192-194
//   p_ij = exp(s_ij - max_j s_ij), mask-false entries set to 0; row sum clamped at 1e-10  utils.py:197-206
//   o_i = sum_j p_ij v_j / rowsum;  lse_i = log(rowsum) + max                            utils.py:208-222
// so a row whose keys are all causally hidden attends uniformly (every s = -FLT_MAX), and a row whose keys
// are all mask-false gives o = 0 -- both as in the reference.  Backward (utils.py:270-289):
//   p_ij = exp(s_ij - lse_i) with only the causal fill applied to s, then mask-false entries set to 0;
//   D_i = sum_c dO_ic O_ic;  ds = p * scale * (dO.v - D);  dq = ds k;  dk = ds^T q;  dv = p^T dO.
#include <cfloat>
#include "internal.cuh"

namespace kv {

constexpr int AG_THREADS = 256;
constexpr int AG_WARPS = AG_THREADS / 32;
constexpr int AG_MAX_DH = 128;

struct AttnGeneral {
  const float *q, *k, *v;
  int ldq, ldk, ldv;          // row strides (floats); head h starts at column h * dh
  const unsigned char* mask;  // nullable; element (b, h, i, j) at b*msb + h*msh + i*msq + j (0 stride = broadcast)
  long long msb, msh, msq;
  int causal;
  int B, H, Nq, Nk, dh;
};

__device__ __forceinline__ void load_rows(float* dst, const float* __restrict__ src, int N, int dh, int lds,
                                          size_t row_stride, int tid) {
  for (int p = tid; p < N * dh; p += AG_THREADS) {
    const int c = p % dh, n = p / dh;
    dst[n * lds + c] = src[(size_t)n * row_stride + c];
  }
}

__global__ void __launch_bounds__(AG_THREADS)
attn_general_fwd_kernel(AttnGeneral a, float* __restrict__ out, int ldo, float* __restrict__ lse) {
  extern __shared__ float sm[];
  const int dh = a.dh, Nq = a.Nq, Nk = a.Nk, lds = dh + 1;
  float* Ks = sm;
  float* Vs = Ks + Nk * lds;
  float* Ps = Vs + Nk * lds;          // [AG_WARPS][Nk]
  float* Qw = Ps + AG_WARPS * Nk;     // [AG_WARPS][dh]
  const int b = blockIdx.x / a.H, h = blockIdx.x % a.H;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  load_rows(Ks, a.k + (size_t)b * Nk * a.ldk + (size_t)h * dh, Nk, dh, lds, a.ldk, tid);
  load_rows(Vs, a.v + (size_t)b * Nk * a.ldv + (size_t)h * dh, Nk, dh, lds, a.ldv, tid);
  __syncthreads();
  const float scale = rsqrtf((float)dh);
  const int diff = Nk > Nq ? Nk - Nq : 0;
  float* prow = Ps + warp * Nk;
  float* q = Qw + warp * dh;
  const unsigned char* mb = a.mask ? a.mask + (size_t)b * a.msb + (size_t)h * a.msh : nullptr;
  for (int i = warp; i < Nq; i += AG_WARPS) {
    const float* qg = a.q + ((size_t)b * Nq + i) * a.ldq + (size_t)h * dh;
    for (int c = lane; c < dh; c += 32) q[c] = qg[c];
    __syncwarp();
    const unsigned char* mrow = mb ? mb + (size_t)i * a.msq : nullptr;
    float mx = -FLT_MAX;
    for (int j = lane; j < Nk; j += 32) {
      const float* k = Ks + j * lds;
      float s = 0.0f;
      for (int c = 0; c < dh; ++c) s = fmaf(q[c], k[c], s);
      s *= scale;
      if ((mrow && !mrow[j]) || (a.causal && j > i - diff)) s = -FLT_MAX;
      prow[j] = s;
      mx = fmaxf(mx, s);
    }
    mx = warp_max(mx);
    float sum = 0.0f;
    for (int j = lane; j < Nk; j += 32) {
      float e = expf(prow[j] - mx);
      if (mrow && !mrow[j]) e = 0.0f;
      prow[j] = e;
      sum += e;
    }
    sum = fmaxf(warp_sum(sum), 1e-10f);
    __syncwarp();
    const float inv = 1.0f / sum;
    float* o = out + ((size_t)b * Nq + i) * ldo + (size_t)h * dh;
    for (int c = lane; c < dh; c += 32) {
      float acc = 0.0f;
      for (int j = 0; j < Nk; ++j) acc = fmaf(prow[j], Vs[j * lds + c], acc);
      o[c] = acc * inv;
    }
    if (lane == 0) lse[((size_t)b * a.H + h) * Nq + i] = logf(sum) + mx;
    __syncwarp();
  }
}

__global__ void __launch_bounds__(AG_THREADS)
attn_general_bwd_kernel(AttnGeneral a, const float* __restrict__ o, const float* __restrict__ dout, int ldo,
                        const float* __restrict__ lse, float* __restrict__ dq, float* __restrict__ dk,
                        float* __restrict__ dv) {
  extern __shared__ float sm[];
  const int dh = a.dh, Nq = a.Nq, Nk = a.Nk, lds = dh + 1;
  const int Nm = Nq > Nk ? Nq : Nk;
  float* Qs = sm;
  float* Gs = Qs + Nq * lds;  // dO
  float* Ks = Gs + Nq * lds;
  float* Vs = Ks + Nk * lds;
  float* Ls = Vs + Nk * lds;  // lse[Nq]
  float* Ds = Ls + Nq;        // D[Nq]
  float* Ps = Ds + Nq;        // [AG_WARPS][2][Nm]
  const int b = blockIdx.x / a.H, h = blockIdx.x % a.H;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  load_rows(Qs, a.q + (size_t)b * Nq * a.ldq + (size_t)h * dh, Nq, dh, lds, a.ldq, tid);
  load_rows(Gs, dout + (size_t)b * Nq * ldo + (size_t)h * dh, Nq, dh, lds, ldo, tid);
  load_rows(Ks, a.k + (size_t)b * Nk * a.ldk + (size_t)h * dh, Nk, dh, lds, a.ldk, tid);
  load_rows(Vs, a.v + (size_t)b * Nk * a.ldv + (size_t)h * dh, Nk, dh, lds, a.ldv, tid);
  for (int i = tid; i < Nq; i += AG_THREADS) Ls[i] = lse[((size_t)b * a.H + h) * Nq + i];
  __syncthreads();
  // D_i = sum_c dO_ic O_ic (utils.py:277), one warp per row
  for (int i = warp; i < Nq; i += AG_WARPS) {
    const float* og = o + ((size_t)b * Nq + i) * ldo + (size_t)h * dh;
    float d = 0.0f;
    for (int c = lane; c < dh; c += 32) d = fmaf(Gs[i * lds + c], og[c], d);
    d = warp_sum(d);
    if (lane == 0) Ds[i] = d;
  }
  __syncthreads();
  const float scale = rsqrtf((float)dh);
  const int diff = Nk > Nq ? Nk - Nq : 0;
  const unsigned char* mb = a.mask ? a.mask + (size_t)b * a.msb + (size_t)h * a.msh : nullptr;
  float* p_row = Ps + warp * 2 * Nm;
  float* g_row = p_row + Nm;

  // ---- row pass: dQ_i = sum_j ds_ij k_j
  for (int i = warp; i < Nq; i += AG_WARPS) {
    const float* q = Qs + i * lds;
    const float* go = Gs + i * lds;
    const unsigned char* mrow = mb ? mb + (size_t)i * a.msq : nullptr;
    const float li = Ls[i], di = Ds[i];
    for (int j = lane; j < Nk; j += 32) {
      const float* k = Ks + j * lds;
      const float* v = Vs + j * lds;
      float s = 0.0f, dp = 0.0f;
      for (int c = 0; c < dh; ++c) {
        s = fmaf(q[c], k[c], s);
        dp = fmaf(go[c], v[c], dp);
      }
      s *= scale;
      if (a.causal && j > i - diff) s = -FLT_MAX;
      const float p = (mrow && !mrow[j]) ? 0.0f : expf(s - li);
      g_row[j] = p * scale * (dp - di);
    }
    __syncwarp();
    float* dqg = dq + ((size_t)b * Nq + i) * a.ldq + (size_t)h * dh;
    for (int c = lane; c < dh; c += 32) {
      float acc = 0.0f;
      for (int j = 0; j < Nk; ++j) acc = fmaf(g_row[j], Ks[j * lds + c], acc);
      dqg[c] = acc;
    }
    __syncwarp();
  }
  // ---- column pass: dK_j = sum_i ds_ij q_i, dV_j = sum_i p_ij dO_i
  for (int j = warp; j < Nk; j += AG_WARPS) {
    const float* k = Ks + j * lds;
    const float* v = Vs + j * lds;
    for (int i = lane; i < Nq; i += 32) {
      const float* q = Qs + i * lds;
      const float* go = Gs + i * lds;
      float s = 0.0f, dp = 0.0f;
      for (int c = 0; c < dh; ++c) {
        s = fmaf(q[c], k[c], s);
        dp = fmaf(go[c], v[c], dp);
      }
      s *= scale;
      if (a.causal && j > i - diff) s = -FLT_MAX;
      const bool off = mb && !mb[(size_t)i * a.msq + j];
      const float p = off ? 0.0f : expf(s - Ls[i]);
      p_row[i] = p;
      g_row[i] = p * scale * (dp - Ds[i]);
    }
    __syncwarp();
    float* dkg = dk + ((size_t)b * Nk + j) * a.ldk + (size_t)h * dh;
    float* dvg = dv + ((size_t)b * Nk + j) * a.ldv + (size_t)h * dh;
    for (int c = lane; c < dh; c += 32) {
      float sk = 0.0f, sv = 0.0f;
      for (int i = 0; i < Nq; ++i) {
        sk = fmaf(g_row[i], Qs[i * lds + c], sk);
        sv = fmaf(p_row[i], Gs[i * lds + c], sv);
      }
      dkg[c] = sk;
      dvg[c] = sv;
    }
    __syncwarp();
  }
}

static size_t ag_fwd_smem(int Nk, int dh) {
  return sizeof(float) * ((size_t)2 * Nk * (dh + 1) + (size_t)AG_WARPS * (Nk + dh));
}
static size_t ag_bwd_smem(int Nq, int Nk, int dh) {
  const int Nm = Nq > Nk ? Nq : Nk;
  return sizeof(float) * ((size_t)2 * (Nq + Nk) * (dh + 1) + 2 * (size_t)Nq + (size_t)AG_WARPS * 2 * Nm);
}

static int ag_check(const kv_attn_general_t* a) {
  if (a == nullptr) return KV_ERR_BAD_ARG;
  if (a->B < 0 || a->H <= 0 || a->Nq <= 0 || a->Nk <= 0 || a->d_head <= 0) return KV_ERR_BAD_ARG;
  if (a->d_head > AG_MAX_DH) return KV_ERR_BAD_SHAPE;
  if (a->ldq < a->H * a->d_head || a->ldk < a->H * a->d_head || a->ldv < a->H * a->d_head || a->ldo < a->H * a->d_head)
    return KV_ERR_BAD_ARG;
  return KV_OK;
}

static AttnGeneral ag_params(const kv_attn_general_t* a, const float* q, const float* k, const float* v,
                             const unsigned char* mask) {
  AttnGeneral p;
  p.q = q; p.k = k; p.v = v;
  p.ldq = a->ldq; p.ldk = a->ldk; p.ldv = a->ldv;
  p.mask = mask;
  p.msb = a->mask_stride_b; p.msh = a->mask_stride_h; p.msq = a->mask_stride_q;
  p.causal = a->causal;
  p.B = a->B; p.H = a->H; p.Nq = a->Nq; p.Nk = a->Nk; p.dh = a->d_head;
  return p;
}

static AttnTiled ag_tiled(const kv_attn_general_t* a, const float* q, const float* k, const float* v,
                          const unsigned char* mask) {
  AttnTiled t;
  t.q = q; t.k = k; t.v = v;
  t.ldq = a->ldq; t.ldk = a->ldk; t.ldv = a->ldv;
  t.hs = a->d_head;
  t.mask = mask;
  t.msb = a->mask_stride_b; t.msh = a->mask_stride_h; t.msq = a->mask_stride_q;
  t.causal = a->causal;
  t.B = a->B; t.H = a->H; t.Nq = a->Nq; t.Nk = a->Nk; t.dh = a->d_head;
  return t;
}

}  // namespace kv

using namespace kv;

extern "C" int kv_attn_general_fwd(const kv_attn_general_t* a, const float* q, const float* k, const float* v,
                                   const unsigned char* mask, float* out, float* lse, void* stream) {
  const int rc = ag_check(a);
  if (rc != KV_OK) return rc;
  const size_t smem = ag_fwd_smem(a->Nk, a->d_head);
  if (a->B == 0) return KV_OK;
  if (!q || !k || !v || !out || !lse) return KV_ERR_BAD_ARG;
  if (smem > 227 * 1024)   // K / V do not fit one CTA: tiled kernels (attn_tiled.cu)
    return attn_tiled_fwd(ag_tiled(a, q, k, v, mask), nullptr, out, a->ldo, a->d_head, lse, (cudaStream_t)stream);
  KV_CUDA(cudaFuncSetAttribute(attn_general_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
          "attn_general_fwd smem");
  attn_general_fwd_kernel<<<a->B * a->H, AG_THREADS, smem, (cudaStream_t)stream>>>(ag_params(a, q, k, v, mask), out,
                                                                                   a->ldo, lse);
  KV_CHECK_LAUNCH("attn_general_fwd");
  return KV_OK;
}

extern "C" int kv_attn_general_bwd(const kv_attn_general_t* a, const float* q, const float* k, const float* v,
                                   const unsigned char* mask, const float* out, const float* dout,
                                   const float* lse, float* dq, float* dk, float* dv, void* stream) {
  const int rc = ag_check(a);
  if (rc != KV_OK) return rc;
  const size_t smem = ag_bwd_smem(a->Nq, a->Nk, a->d_head);
  if (a->B == 0) return KV_OK;
  if (!q || !k || !v || !out || !dout || !lse || !dq || !dk || !dv) return KV_ERR_BAD_ARG;
  if (smem > 227 * 1024)
    return attn_tiled_bwd(ag_tiled(a, q, k, v, mask), out, dout, a->ldo, a->d_head, lse, dq, dk, dv,
                          (cudaStream_t)stream);
  KV_CUDA(cudaFuncSetAttribute(attn_general_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
          "attn_general_bwd smem");
  attn_general_bwd_kernel<<<a->B * a->H, AG_THREADS, smem, (cudaStream_t)stream>>>(
      ag_params(a, q, k, v, mask), out, dout, a->ldo, lse, dq, dk, dv);
  KV_CHECK_LAUNCH("attn_general_bwd");
  return KV_OK;
}
