// Fused softmax attention per (sample, head): forward with saved log-sum-exp, and the
// recompute backward (algorithm of utils.py:137-295 of the reference with a single
// key tile, since N <= 256 here).  fp32 SIMT: one CTA per (b, h), Q/K/V (and dO in
// the backward) resident in shared memory, one warp per query row (forward, backward
// row pass) or per key row (backward column pass).  No atomics: deterministic.
//
// qkv layout [B*N][H][3][d_h] (written by the grouped KAN layer), out [B*N][H*d_h].
#include <cfloat>
#include "internal.cuh"

namespace kv {

constexpr int ATT_THREADS = 256;
constexpr int ATT_WARPS = ATT_THREADS / 32;
constexpr int ATT_MAX_DH = 128;

__device__ __forceinline__ void load_head_tile(float* dst, const float* __restrict__ src, int N, int dh,
                                               int lds, size_t row_stride, int tid, int nth) {
  // dst[n*lds + c] = src[n*row_stride + c]
  for (int p = tid; p < N * dh; p += nth) {
    const int c = p % dh, n = p / dh;
    dst[n * lds + c] = src[(size_t)n * row_stride + c];
  }
}

__global__ void __launch_bounds__(ATT_THREADS)
attn_fwd_kernel(const float* __restrict__ qkv, const float* __restrict__ residual,
                float* __restrict__ out, float* __restrict__ lse, int N, int H, int dh) {
  extern __shared__ float sm[];
  const int lds = dh + 1;
  float* Qs = sm;
  float* Ks = Qs + N * lds;
  float* Vs = Ks + N * lds;
  float* Ps = Vs + N * lds;  // [ATT_WARPS][N]
  const int b = blockIdx.x / H, h = blockIdx.x % H;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const size_t rs = (size_t)H * 3 * dh;
  const float* base = qkv + (size_t)b * N * rs + (size_t)h * 3 * dh;
  load_head_tile(Qs, base, N, dh, lds, rs, tid, ATT_THREADS);
  load_head_tile(Ks, base + dh, N, dh, lds, rs, tid, ATT_THREADS);
  load_head_tile(Vs, base + 2 * dh, N, dh, lds, rs, tid, ATT_THREADS);
  __syncthreads();
  const float scale = rsqrtf((float)dh);
  float* prow = Ps + warp * N;
  for (int i = warp; i < N; i += ATT_WARPS) {
    const float* q = Qs + i * lds;
    float mx = -FLT_MAX;
    for (int j = lane; j < N; j += 32) {
      const float* k = Ks + j * lds;
      float s = 0.0f;
      for (int c = 0; c < dh; ++c) s = fmaf(q[c], k[c], s);
      s *= scale;
      prow[j] = s;
      mx = fmaxf(mx, s);
    }
    mx = warp_max(mx);
    float sum = 0.0f;
    for (int j = lane; j < N; j += 32) {
      const float e = expf(prow[j] - mx);
      prow[j] = e;
      sum += e;
    }
    sum = warp_sum(sum);
    __syncwarp();
    const float inv = 1.0f / sum;
    const size_t orow = ((size_t)b * N + i) * (size_t)(H * dh) + (size_t)h * dh;
    for (int c = lane; c < dh; c += 32) {
      float acc = 0.0f;
      for (int j = 0; j < N; ++j) acc = fmaf(prow[j], Vs[j * lds + c], acc);
      acc *= inv;
      if (residual) acc += residual[orow + c];
      out[orow + c] = acc;
    }
    if (lane == 0 && lse != nullptr) lse[((size_t)b * H + h) * N + i] = logf(sum) + mx;
    __syncwarp();
  }
}

__global__ void __launch_bounds__(ATT_THREADS)
attn_bwd_kernel(const float* __restrict__ qkv, const float* __restrict__ dout,
                const float* __restrict__ lse, float* __restrict__ dqkv, int N, int H, int dh) {
  extern __shared__ float sm[];
  const int lds = dh + 1;
  float* Qs = sm;
  float* Ks = Qs + N * lds;
  float* Vs = Ks + N * lds;
  float* Os = Vs + N * lds;  // dO
  float* Ls = Os + N * lds;  // lse[N]
  float* Ds = Ls + N;        // D[N]
  float* Ps = Ds + N;        // [ATT_WARPS][2][N]
  const int b = blockIdx.x / H, h = blockIdx.x % H;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const size_t rs = (size_t)H * 3 * dh;
  const size_t head_off = (size_t)b * N * rs + (size_t)h * 3 * dh;
  const float* base = qkv + head_off;
  float* dbase = dqkv + head_off;
  load_head_tile(Qs, base, N, dh, lds, rs, tid, ATT_THREADS);
  load_head_tile(Ks, base + dh, N, dh, lds, rs, tid, ATT_THREADS);
  load_head_tile(Vs, base + 2 * dh, N, dh, lds, rs, tid, ATT_THREADS);
  load_head_tile(Os, dout + (size_t)b * N * (size_t)(H * dh) + (size_t)h * dh, N, dh, lds,
                 (size_t)H * dh, tid, ATT_THREADS);
  for (int i = tid; i < N; i += ATT_THREADS) Ls[i] = lse[((size_t)b * H + h) * N + i];
  __syncthreads();
  const float scale = rsqrtf((float)dh);
  float* p_row = Ps + warp * 2 * N;  // p_ij
  float* g_row = p_row + N;          // dp_ij, then ds_ij

  // ---- row pass: D_i and dQ_i
  for (int i = warp; i < N; i += ATT_WARPS) {
    const float* q = Qs + i * lds;
    const float* go = Os + i * lds;
    const float li = Ls[i];
    float dsum = 0.0f;
    for (int j = lane; j < N; j += 32) {
      const float* k = Ks + j * lds;
      const float* v = Vs + j * lds;
      float s = 0.0f, dp = 0.0f;
      for (int c = 0; c < dh; ++c) {
        s = fmaf(q[c], k[c], s);
        dp = fmaf(go[c], v[c], dp);
      }
      const float p = expf(s * scale - li);
      p_row[j] = p;
      g_row[j] = dp;
      dsum = fmaf(p, dp, dsum);
    }
    dsum = warp_sum(dsum);  // D_i = sum_j p_ij dp_ij  ( = sum_c dO_ic O_ic, utils.py:277 )
    for (int j = lane; j < N; j += 32) g_row[j] = p_row[j] * scale * (g_row[j] - dsum);
    if (lane == 0) Ds[i] = dsum;
    __syncwarp();
    for (int c = lane; c < dh; c += 32) {
      float acc = 0.0f;
      for (int j = 0; j < N; ++j) acc = fmaf(g_row[j], Ks[j * lds + c], acc);
      dbase[(size_t)i * rs + c] = acc;
    }
    __syncwarp();
  }
  __syncthreads();
  // ---- column pass: dK_j and dV_j
  for (int j = warp; j < N; j += ATT_WARPS) {
    const float* k = Ks + j * lds;
    const float* v = Vs + j * lds;
    for (int i = lane; i < N; i += 32) {
      const float* q = Qs + i * lds;
      const float* go = Os + i * lds;
      float s = 0.0f, dp = 0.0f;
      for (int c = 0; c < dh; ++c) {
        s = fmaf(q[c], k[c], s);
        dp = fmaf(go[c], v[c], dp);
      }
      const float p = expf(s * scale - Ls[i]);
      p_row[i] = p;
      g_row[i] = p * scale * (dp - Ds[i]);
    }
    __syncwarp();
    for (int c = lane; c < dh; c += 32) {
      float dk = 0.0f, dv = 0.0f;
      for (int i = 0; i < N; ++i) {
        dk = fmaf(g_row[i], Qs[i * lds + c], dk);
        dv = fmaf(p_row[i], Os[i * lds + c], dv);
      }
      dbase[(size_t)j * rs + dh + c] = dk;
      dbase[(size_t)j * rs + 2 * dh + c] = dv;
    }
    __syncwarp();
  }
}

// the grouped q/k/v layout [B*N][H][3][d_h] as the tiled kernels address it: q, k, v are column offsets 0, d_h, 2 d_h
// of head column h * 3 d_h
static AttnTiled msa_tiled(const float* qkv, int B, int N, int H, int dh) {
  AttnTiled a;
  a.q = qkv; a.k = qkv + dh; a.v = qkv + 2 * dh;
  a.ldq = a.ldk = a.ldv = 3 * H * dh;
  a.hs = 3 * dh;
  a.mask = nullptr;
  a.msb = a.msh = a.msq = 0;
  a.causal = 0;
  a.B = B; a.H = H; a.Nq = N; a.Nk = N; a.dh = dh;
  return a;
}

static size_t attn_fwd_smem(int N, int dh) { return sizeof(float) * ((size_t)3 * N * (dh + 1) + ATT_WARPS * N); }
static size_t attn_bwd_smem(int N, int dh) {
  return sizeof(float) * ((size_t)4 * N * (dh + 1) + 2 * N + ATT_WARPS * 2 * N);
}

}  // namespace kv

using namespace kv;

extern "C" int kv_attn_fwd(const float* qkv, const float* residual, float* out, float* lse, int B, int N,
                           int H, int d_head, void* stream) {
  if (qkv == nullptr || out == nullptr || B < 0 || N <= 0 || H <= 0 || d_head <= 0) return KV_ERR_BAD_ARG;
  if (d_head > ATT_MAX_DH) return KV_ERR_BAD_SHAPE;
  const size_t smem = attn_fwd_smem(N, d_head);
  if (smem > 227 * 1024) {   // the head does not fit one CTA: tiled kernels (attn_tiled.cu)
    if (B == 0) return KV_OK;
    return attn_tiled_fwd(msa_tiled(qkv, B, N, H, d_head), residual, out, H * d_head, d_head, lse, (cudaStream_t)stream);
  }
  if (B == 0) return KV_OK;
  KV_CUDA(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
          "attn_fwd smem");
  attn_fwd_kernel<<<B * H, ATT_THREADS, smem, (cudaStream_t)stream>>>(qkv, residual, out, lse, N, H, d_head);
  KV_CHECK_LAUNCH("attn_fwd");
  return KV_OK;
}

extern "C" int kv_attn_bwd(const float* qkv, const float* dout, const float* lse, float* dqkv, int B,
                           int N, int H, int d_head, void* stream) {
  if (qkv == nullptr || dout == nullptr || lse == nullptr || dqkv == nullptr || B < 0 || N <= 0 ||
      H <= 0 || d_head <= 0)
    return KV_ERR_BAD_ARG;
  if (d_head > ATT_MAX_DH) return KV_ERR_BAD_SHAPE;
  const size_t smem = attn_bwd_smem(N, d_head);
  if (smem > 227 * 1024) {
    if (B == 0) return KV_OK;
    return attn_tiled_bwd(msa_tiled(qkv, B, N, H, d_head), nullptr, dout, H * d_head, d_head, lse, dqkv, dqkv + d_head,
                          dqkv + 2 * d_head, (cudaStream_t)stream);
  }
  if (B == 0) return KV_OK;
  KV_CUDA(cudaFuncSetAttribute(attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
          "attn_bwd smem");
  attn_bwd_kernel<<<B * H, ATT_THREADS, smem, (cudaStream_t)stream>>>(qkv, dout, lse, dqkv, N, H, d_head);
  KV_CHECK_LAUNCH("attn_bwd");
  return KV_OK;
}
