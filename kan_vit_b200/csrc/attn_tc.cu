// Tensor-core fused softmax attention per (sample, head), bf16 operands / fp32 accumulation:
// forward = flash-style online softmax over key chunks with the log-sum-exp saved, backward = the
// recompute scheme of utils.py:229-295 of the reference (p = exp(s - lse), D_i = sum_j p_ij dp_ij,
// ds = p * scale * (dp - D)).  One CTA per (b, h): Q/K/V (and dO) of the head are converted to bf16
// once and stay in shared memory; every warp owns 16-query (or 16-key) row blocks and keeps the
// score tiles in registers, so neither S nor P ever exists in memory.  No atomics: deterministic.
//
// These problems are tiny (N <= 197 tokens, d_h <= 64: < 1 % of the step's FLOPs), far below the
// 128-row tile a tcgen05 instruction wants per CTA, so the kernel uses warp-level mma.sync
// m16n8k16 with ldmatrix-fed fragments; the point is latency and shared-memory residency.
//
// qkv layout [B*N][H][3][d_h] fp32 (written by the grouped KAN layer), out [B*N][H*d_h] fp32.
#include <cfloat>
#include "internal.cuh"

namespace kv {
namespace atc {

constexpr int KC = 32;  // keys (queries in the column pass) per register chunk
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
// D(16x8) += A(16x16, row) * B(16x8, col), bf16 operands, fp32 accumulators
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// A fragments (16 rows x DHP) of row block `r0` of a row-major [rows][LD] bf16 tile
template <int DHP>
__device__ __forceinline__ void load_a_frags(uint32_t (&f)[DHP / 16][4], const __nv_bfloat16* tile, int r0, int lane) {
  constexpr int LD = DHP + 8;
  const int row = r0 + (lane & 7) + ((lane >> 3) & 1) * 8;
  const uint32_t base = smem_u32(tile + row * LD + (lane >> 4) * 8);
#pragma unroll
  for (int ks = 0; ks < DHP / 16; ++ks) ldsm_x4(f[ks], base + ks * 32);
}

// C[KC/8 tiles](16 x KC) = A(16 x DHP) . B^T where B is a row-major [n][LD] tile, rows n0 .. n0+KC
template <int DHP>
__device__ __forceinline__ void mm_abt(float (&c)[KC / 8][4], const uint32_t (&a)[DHP / 16][4],
                                       const __nv_bfloat16* tile, int n0, int lane) {
  constexpr int LD = DHP + 8;
#pragma unroll
  for (int nt = 0; nt < KC / 8; ++nt)
#pragma unroll
    for (int e = 0; e < 4; ++e) c[nt][e] = 0.0f;
  const int row = n0 + (lane & 7) + (lane >> 4) * 8;
  const uint32_t base = smem_u32(tile + row * LD + ((lane >> 3) & 1) * 8);
#pragma unroll
  for (int ks = 0; ks < DHP / 16; ++ks) {
#pragma unroll
    for (int np = 0; np < KC / 16; ++np) {
      uint32_t b[4];
      ldsm_x4(b, base + (np * 16 * LD + ks * 16) * 2);
      mma16816(c[2 * np], a[ks], b[0], b[1]);
      mma16816(c[2 * np + 1], a[ks], b[2], b[3]);
    }
  }
}

// acc[DHP/8 tiles](16 x DHP) += P(16 x KC, from registers) . B where B is a row-major [k][LD] tile, rows k0 .. k0+KC
template <int DHP>
__device__ __forceinline__ void mm_pb(float (&acc)[DHP / 8][4], const uint32_t (&p)[KC / 16][4],
                                      const __nv_bfloat16* tile, int k0, int lane) {
  constexpr int LD = DHP + 8;
  const int row = k0 + (lane & 7) + ((lane >> 3) & 1) * 8;
  const uint32_t base = smem_u32(tile + row * LD + (lane >> 4) * 8);
#pragma unroll
  for (int js = 0; js < KC / 16; ++js) {
#pragma unroll
    for (int cp = 0; cp < DHP / 16; ++cp) {
      uint32_t b[4];
      ldsm_x4_t(b, base + (js * 16 * LD + cp * 16) * 2);
      mma16816(acc[2 * cp], p[js], b[0], b[1]);
      mma16816(acc[2 * cp + 1], p[js], b[2], b[3]);
    }
  }
}

// accumulator tiles (16 x KC, fp32) -> A fragments of the next product (bf16)
__device__ __forceinline__ void c_to_a(uint32_t (&a)[KC / 16][4], const float (&c)[KC / 8][4]) {
#pragma unroll
  for (int js = 0; js < KC / 16; ++js) {
    a[js][0] = pack2(c[2 * js][0], c[2 * js][1]);
    a[js][1] = pack2(c[2 * js][2], c[2 * js][3]);
    a[js][2] = pack2(c[2 * js + 1][0], c[2 * js + 1][1]);
    a[js][3] = pack2(c[2 * js + 1][2], c[2 * js + 1][3]);
  }
}

// fp32 global rows -> bf16 shared tiles.  src row n holds `n_parts` consecutive d_h-wide parts (q,k,v or dO).
template <int DHP>
__device__ __forceinline__ void stage_rows(__nv_bfloat16* const* tiles, int n_parts, const float* __restrict__ src,
                                           size_t row_stride, int N, int Np, int dh, int tid, int nth) {
  constexpr int LD = DHP + 8;
  constexpr int U = 8;  // 16-byte loads in flight per thread: the staging is latency-bound otherwise
  const int f_per_row = n_parts * dh / 4;
  const int total = Np * f_per_row;
  for (int base = tid; base < total; base += U * nth) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int idx = base + u * nth;
      const int n = idx / f_per_row, f = idx - n * f_per_row;
      v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (idx < total && n < N) v[u] = __ldcs(reinterpret_cast<const float4*>(src + (size_t)n * row_stride + f * 4));
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int idx = base + u * nth;
      if (idx < total) {
        const int n = idx / f_per_row, f = idx - n * f_per_row;
        const int col = f * 4, part = col / dh, c = col - part * dh;
        uint2 w;
        w.x = pack2(v[u].x, v[u].y);
        w.y = pack2(v[u].z, v[u].w);
        *reinterpret_cast<uint2*>(tiles[part] + n * LD + c) = w;
      }
    }
  }
  if (dh < DHP) {
    const int padw = (DHP - dh) / 4;  // dh % 4 == 0
    for (int idx = tid; idx < Np * padw * n_parts; idx += nth) {
      const int part = idx / (Np * padw), r = idx - part * (Np * padw);
      const int n = r / padw, c = dh + (r - n * padw) * 4;
      *reinterpret_cast<uint2*>(tiles[part] + n * LD + c) = make_uint2(0u, 0u);
    }
  }
}

// ------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------
template <int DHP>
__global__ void __launch_bounds__(256)
attn_tc_fwd_kernel(const float* __restrict__ qkv, const float* __restrict__ residual, float* __restrict__ out,
                   float* __restrict__ lse, int N, int H, int dh) {
  constexpr int LD = DHP + 8;
  extern __shared__ __align__(16) unsigned char smraw[];
  const int Np = (N + KC - 1) / KC * KC;
  __nv_bfloat16* Qs = reinterpret_cast<__nv_bfloat16*>(smraw);
  __nv_bfloat16* Ks = Qs + Np * LD;
  __nv_bfloat16* Vs = Ks + Np * LD;
  const int b = blockIdx.x / H, h = blockIdx.x % H;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const size_t rs = (size_t)H * 3 * dh;
  {
    __nv_bfloat16* tiles[3] = {Qs, Ks, Vs};
    stage_rows<DHP>(tiles, 3, qkv + (size_t)b * N * rs + (size_t)h * 3 * dh, rs, N, Np, dh, tid, blockDim.x);
  }
  __syncthreads();
  const float sl2 = rsqrtf((float)dh) * LOG2E;
  const int nrb = (N + 15) / 16;
  for (int rb = warp; rb < nrb; rb += nwarps) {
    if (residual) {
      // the epilogue adds the residual rows of this block: start pulling them towards L2 now
      // (16 rows x d_h floats; a lane takes one 128-byte line: row lane/2, half lane%2)
      const int pr = rb * 16 + (lane >> 1), pc = (lane & 1) * 32;
      if (pr < N && pc < dh) {
        const float* pa = residual + ((size_t)b * N + pr) * (size_t)(H * dh) + (size_t)h * dh + pc;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(pa));
      }
    }
    uint32_t qf[DHP / 16][4];
    load_a_frags<DHP>(qf, Qs, rb * 16, lane);
    float o[DHP / 8][4];
#pragma unroll
    for (int ct = 0; ct < DHP / 8; ++ct)
#pragma unroll
      for (int e = 0; e < 4; ++e) o[ct][e] = 0.0f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.0f, l1 = 0.0f;
    for (int k0 = 0; k0 < Np; k0 += KC) {
      float s[KC / 8][4];
      mm_abt<DHP>(s, qf, Ks, k0, lane);
      float c0 = -INFINITY, c1 = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < KC / 8; ++nt) {
        const int j = k0 + nt * 8 + 2 * t;
        s[nt][0] = (j < N) ? s[nt][0] * sl2 : -INFINITY;
        s[nt][1] = (j + 1 < N) ? s[nt][1] * sl2 : -INFINITY;
        s[nt][2] = (j < N) ? s[nt][2] * sl2 : -INFINITY;
        s[nt][3] = (j + 1 < N) ? s[nt][3] * sl2 : -INFINITY;
        c0 = fmaxf(c0, fmaxf(s[nt][0], s[nt][1]));
        c1 = fmaxf(c1, fmaxf(s[nt][2], s[nt][3]));
      }
      c0 = fmaxf(c0, __shfl_xor_sync(0xffffffffu, c0, 1));
      c0 = fmaxf(c0, __shfl_xor_sync(0xffffffffu, c0, 2));
      c1 = fmaxf(c1, __shfl_xor_sync(0xffffffffu, c1, 1));
      c1 = fmaxf(c1, __shfl_xor_sync(0xffffffffu, c1, 2));
      const float n0 = fmaxf(m0, c0), n1 = fmaxf(m1, c1);  // finite: every chunk holds a valid key
      const float a0 = exp2f(m0 - n0), a1 = exp2f(m1 - n1);
      m0 = n0;
      m1 = n1;
      float r0 = 0.0f, r1 = 0.0f;
#pragma unroll
      for (int nt = 0; nt < KC / 8; ++nt) {
        s[nt][0] = exp2f(s[nt][0] - n0);
        s[nt][1] = exp2f(s[nt][1] - n0);
        s[nt][2] = exp2f(s[nt][2] - n1);
        s[nt][3] = exp2f(s[nt][3] - n1);
        r0 += s[nt][0] + s[nt][1];
        r1 += s[nt][2] + s[nt][3];
      }
      l0 = l0 * a0 + r0;
      l1 = l1 * a1 + r1;
#pragma unroll
      for (int ct = 0; ct < DHP / 8; ++ct) {
        o[ct][0] *= a0;
        o[ct][1] *= a0;
        o[ct][2] *= a1;
        o[ct][3] *= a1;
      }
      uint32_t pf[KC / 16][4];
      c_to_a(pf, s);
      mm_pb<DHP>(o, pf, Vs, k0, lane);
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = 1.0f / l0, i1 = 1.0f / l1;
    const int r0 = rb * 16 + g, r1 = r0 + 8;
#pragma unroll
    for (int ct = 0; ct < DHP / 8; ++ct) {
      const int c = ct * 8 + 2 * t;
      if (c < dh) {
        if (r0 < N) {
          const size_t off = ((size_t)b * N + r0) * (size_t)(H * dh) + (size_t)h * dh + c;
          float2 v = make_float2(o[ct][0] * i0, o[ct][1] * i0);
          if (residual) {
            const float2 r = *reinterpret_cast<const float2*>(residual + off);
            v.x += r.x;
            v.y += r.y;
          }
          *reinterpret_cast<float2*>(out + off) = v;
        }
        if (r1 < N) {
          const size_t off = ((size_t)b * N + r1) * (size_t)(H * dh) + (size_t)h * dh + c;
          float2 v = make_float2(o[ct][2] * i1, o[ct][3] * i1);
          if (residual) {
            const float2 r = *reinterpret_cast<const float2*>(residual + off);
            v.x += r.x;
            v.y += r.y;
          }
          *reinterpret_cast<float2*>(out + off) = v;
        }
      }
    }
    if (t == 0 && lse != nullptr) {
      if (r0 < N) lse[((size_t)b * H + h) * N + r0] = (m0 + log2f(l0)) * LN2;
      if (r1 < N) lse[((size_t)b * H + h) * N + r1] = (m1 + log2f(l1)) * LN2;
    }
  }
}

// ------------------------------------------------------------------------------------------
// backward: row pass (warp = 16 queries: D_i, dQ_i) then column pass (warp = 16 keys: dK_j, dV_j)
// ------------------------------------------------------------------------------------------
template <int DHP>
__global__ void __launch_bounds__(448)
attn_tc_bwd_kernel(const float* __restrict__ qkv, const float* __restrict__ dout, const float* __restrict__ lse,
                   float* __restrict__ dqkv, int N, int H, int dh) {
  constexpr int LD = DHP + 8;
  extern __shared__ __align__(16) unsigned char smraw[];
  const int Np = (N + KC - 1) / KC * KC;
  __nv_bfloat16* Qs = reinterpret_cast<__nv_bfloat16*>(smraw);
  __nv_bfloat16* Ks = Qs + Np * LD;
  __nv_bfloat16* Vs = Ks + Np * LD;
  __nv_bfloat16* Os = Vs + Np * LD;                   // dO
  float* Ls = reinterpret_cast<float*>(Os + Np * LD);  // lse in log2 units, +inf on the padding rows
  float* Ds = Ls + Np;
  const int b = blockIdx.x / H, h = blockIdx.x % H;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const size_t rs = (size_t)H * 3 * dh;
  const size_t head_off = (size_t)b * N * rs + (size_t)h * 3 * dh;
  {
    __nv_bfloat16* tiles[3] = {Qs, Ks, Vs};
    stage_rows<DHP>(tiles, 3, qkv + head_off, rs, N, Np, dh, tid, blockDim.x);
    __nv_bfloat16* otile[1] = {Os};
    stage_rows<DHP>(otile, 1, dout + (size_t)b * N * (size_t)(H * dh) + (size_t)h * dh, (size_t)H * dh, N, Np, dh,
                    tid, blockDim.x);
  }
  for (int i = tid; i < Np; i += blockDim.x) {
    Ls[i] = (i < N) ? lse[((size_t)b * H + h) * N + i] * LOG2E : INFINITY;
    Ds[i] = 0.0f;  // rows beyond the last 16-row block are never written by the row pass (0 * garbage must stay 0)
  }
  __syncthreads();
  const float scale = rsqrtf((float)dh);
  const float sl2 = scale * LOG2E;
  float* dbase = dqkv + head_off;
  const int nrb = (N + 15) / 16;

  // ---- row pass
  for (int rb = warp; rb < nrb; rb += nwarps) {
    uint32_t qf[DHP / 16][4], of[DHP / 16][4];
    load_a_frags<DHP>(qf, Qs, rb * 16, lane);
    load_a_frags<DHP>(of, Os, rb * 16, lane);
    const int r0 = rb * 16 + g, r1 = r0 + 8;
    const float L0 = Ls[r0], L1 = Ls[r1];
    float d0 = 0.0f, d1 = 0.0f;
    for (int k0 = 0; k0 < Np; k0 += KC) {
      float s[KC / 8][4], dp[KC / 8][4];
      mm_abt<DHP>(s, qf, Ks, k0, lane);
      mm_abt<DHP>(dp, of, Vs, k0, lane);
#pragma unroll
      for (int nt = 0; nt < KC / 8; ++nt) {
        const int j = k0 + nt * 8 + 2 * t;
        const float p0 = (j < N) ? exp2f(s[nt][0] * sl2 - L0) : 0.0f;
        const float p1 = (j + 1 < N) ? exp2f(s[nt][1] * sl2 - L0) : 0.0f;
        const float p2 = (j < N) ? exp2f(s[nt][2] * sl2 - L1) : 0.0f;
        const float p3 = (j + 1 < N) ? exp2f(s[nt][3] * sl2 - L1) : 0.0f;
        d0 = fmaf(p0, dp[nt][0], fmaf(p1, dp[nt][1], d0));
        d1 = fmaf(p2, dp[nt][2], fmaf(p3, dp[nt][3], d1));
      }
    }
    d0 += __shfl_xor_sync(0xffffffffu, d0, 1);
    d0 += __shfl_xor_sync(0xffffffffu, d0, 2);
    d1 += __shfl_xor_sync(0xffffffffu, d1, 1);
    d1 += __shfl_xor_sync(0xffffffffu, d1, 2);
    if (t == 0) {
      Ds[r0] = d0;  // D_i = sum_j p_ij dp_ij ( = sum_c dO_ic O_ic, utils.py:277 )
      Ds[r1] = d1;
    }
    float dq[DHP / 8][4];
#pragma unroll
    for (int ct = 0; ct < DHP / 8; ++ct)
#pragma unroll
      for (int e = 0; e < 4; ++e) dq[ct][e] = 0.0f;
    for (int k0 = 0; k0 < Np; k0 += KC) {
      float s[KC / 8][4], dp[KC / 8][4];
      mm_abt<DHP>(s, qf, Ks, k0, lane);
      mm_abt<DHP>(dp, of, Vs, k0, lane);
#pragma unroll
      for (int nt = 0; nt < KC / 8; ++nt) {
        const int j = k0 + nt * 8 + 2 * t;
        const float p0 = (j < N) ? exp2f(s[nt][0] * sl2 - L0) : 0.0f;
        const float p1 = (j + 1 < N) ? exp2f(s[nt][1] * sl2 - L0) : 0.0f;
        const float p2 = (j < N) ? exp2f(s[nt][2] * sl2 - L1) : 0.0f;
        const float p3 = (j + 1 < N) ? exp2f(s[nt][3] * sl2 - L1) : 0.0f;
        s[nt][0] = p0 * scale * (dp[nt][0] - d0);
        s[nt][1] = p1 * scale * (dp[nt][1] - d0);
        s[nt][2] = p2 * scale * (dp[nt][2] - d1);
        s[nt][3] = p3 * scale * (dp[nt][3] - d1);
      }
      uint32_t dsf[KC / 16][4];
      c_to_a(dsf, s);
      mm_pb<DHP>(dq, dsf, Ks, k0, lane);
    }
#pragma unroll
    for (int ct = 0; ct < DHP / 8; ++ct) {
      const int c = ct * 8 + 2 * t;
      if (c < dh) {
        if (r0 < N) *reinterpret_cast<float2*>(dbase + (size_t)r0 * rs + c) = make_float2(dq[ct][0], dq[ct][1]);
        if (r1 < N) *reinterpret_cast<float2*>(dbase + (size_t)r1 * rs + c) = make_float2(dq[ct][2], dq[ct][3]);
      }
    }
  }
  __syncthreads();

  // ---- column pass: S^T = K Q^T, dP^T = V dO^T; dV += P^T dO, dK += dS^T Q
  for (int cb = warp; cb < nrb; cb += nwarps) {
    uint32_t kf[DHP / 16][4], vf[DHP / 16][4];
    load_a_frags<DHP>(kf, Ks, cb * 16, lane);
    load_a_frags<DHP>(vf, Vs, cb * 16, lane);
    float dk[DHP / 8][4], dv[DHP / 8][4];
#pragma unroll
    for (int ct = 0; ct < DHP / 8; ++ct)
#pragma unroll
      for (int e = 0; e < 4; ++e) dk[ct][e] = dv[ct][e] = 0.0f;
    for (int i0 = 0; i0 < Np; i0 += KC) {
      float st[KC / 8][4], dpt[KC / 8][4];
      mm_abt<DHP>(st, kf, Qs, i0, lane);
      mm_abt<DHP>(dpt, vf, Os, i0, lane);
#pragma unroll
      for (int nt = 0; nt < KC / 8; ++nt) {
        const int i = i0 + nt * 8 + 2 * t;
        const float2 Li = *reinterpret_cast<const float2*>(Ls + i);
        const float2 Di = *reinterpret_cast<const float2*>(Ds + i);
        const float p0 = exp2f(st[nt][0] * sl2 - Li.x);  // 0 on padding queries (Ls = +inf)
        const float p1 = exp2f(st[nt][1] * sl2 - Li.y);
        const float p2 = exp2f(st[nt][2] * sl2 - Li.x);
        const float p3 = exp2f(st[nt][3] * sl2 - Li.y);
        st[nt][0] = p0;
        st[nt][1] = p1;
        st[nt][2] = p2;
        st[nt][3] = p3;
        dpt[nt][0] = p0 * scale * (dpt[nt][0] - Di.x);
        dpt[nt][1] = p1 * scale * (dpt[nt][1] - Di.y);
        dpt[nt][2] = p2 * scale * (dpt[nt][2] - Di.x);
        dpt[nt][3] = p3 * scale * (dpt[nt][3] - Di.y);
      }
      uint32_t pf[KC / 16][4], dsf[KC / 16][4];
      c_to_a(pf, st);
      c_to_a(dsf, dpt);
      mm_pb<DHP>(dv, pf, Os, i0, lane);
      mm_pb<DHP>(dk, dsf, Qs, i0, lane);
    }
    const int j0 = cb * 16 + g, j1 = j0 + 8;
#pragma unroll
    for (int ct = 0; ct < DHP / 8; ++ct) {
      const int c = ct * 8 + 2 * t;
      if (c < dh) {
        if (j0 < N) {
          *reinterpret_cast<float2*>(dbase + (size_t)j0 * rs + dh + c) = make_float2(dk[ct][0], dk[ct][1]);
          *reinterpret_cast<float2*>(dbase + (size_t)j0 * rs + 2 * dh + c) = make_float2(dv[ct][0], dv[ct][1]);
        }
        if (j1 < N) {
          *reinterpret_cast<float2*>(dbase + (size_t)j1 * rs + dh + c) = make_float2(dk[ct][2], dk[ct][3]);
          *reinterpret_cast<float2*>(dbase + (size_t)j1 * rs + 2 * dh + c) = make_float2(dv[ct][2], dv[ct][3]);
        }
      }
    }
  }
}

static size_t fwd_smem(int N, int dhp) { return (size_t)3 * ((N + KC - 1) / KC * KC) * (dhp + 8) * 2; }
static size_t bwd_smem(int N, int dhp) {
  const size_t np = (N + KC - 1) / KC * KC;
  return (size_t)4 * np * (dhp + 8) * 2 + 2 * np * sizeof(float);
}
// warps per CTA: the 16-row blocks spread evenly over at most `max_warps` warps
static int warps_for(int N, int max_warps = 8) {
  const int nrb = (N + 15) / 16;
  const int rounds = (nrb + max_warps - 1) / max_warps;
  return (nrb + rounds - 1) / rounds;
}
static int dhp_for(int dh) { return dh <= 32 ? 32 : 64; }

}  // namespace atc
}  // namespace kv

using namespace kv;

static int attn_tc_check(int B, int N, int H, int dh, size_t smem) {
  if (B < 0 || N <= 0 || H <= 0 || dh <= 0) return KV_ERR_BAD_ARG;
  if (dh > 64 || (dh % 4) != 0 || smem > 227 * 1024) return KV_ERR_UNSUPPORTED;
  return KV_OK;
}

extern "C" int kv_attn_fwd_bf16(const float* qkv, const float* residual, float* out, float* lse, int B, int N,
                                int H, int d_head, void* stream) {
  const int dhp = atc::dhp_for(d_head);
  const size_t smem = atc::fwd_smem(N > 0 ? N : 1, dhp);
  const int rc = attn_tc_check(B, N, H, d_head, smem);  // shape verdict first: callers probe with B = 0
  if (rc != KV_OK) return rc;
  if (B == 0) return KV_OK;
  if (qkv == nullptr || out == nullptr) return KV_ERR_BAD_ARG;
  if (B == 0) return KV_OK;
  const int threads = 32 * atc::warps_for(N);
  if (dhp == 32) {
    KV_CUDA(cudaFuncSetAttribute(atc::attn_tc_fwd_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
            "attn_tc_fwd smem");
    atc::attn_tc_fwd_kernel<32><<<B * H, threads, smem, (cudaStream_t)stream>>>(qkv, residual, out, lse, N, H, d_head);
  } else {
    KV_CUDA(cudaFuncSetAttribute(atc::attn_tc_fwd_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
            "attn_tc_fwd smem");
    atc::attn_tc_fwd_kernel<64><<<B * H, threads, smem, (cudaStream_t)stream>>>(qkv, residual, out, lse, N, H, d_head);
  }
  KV_CHECK_LAUNCH("attn_tc_fwd");
  return KV_OK;
}

extern "C" int kv_attn_bwd_bf16(const float* qkv, const float* dout, const float* lse, float* dqkv, int B, int N,
                                int H, int d_head, void* stream) {
  const int dhp = atc::dhp_for(d_head);
  const size_t smem = atc::bwd_smem(N > 0 ? N : 1, dhp);
  const int rc = attn_tc_check(B, N, H, d_head, smem);
  if (rc != KV_OK) return rc;
  if (B == 0) return KV_OK;
  if (qkv == nullptr || dout == nullptr || lse == nullptr || dqkv == nullptr) return KV_ERR_BAD_ARG;
  if (B == 0) return KV_OK;
  const int threads = 32 * atc::warps_for(N, 14);  // one CTA per SM (four staged tiles): more warps, one round
  if (dhp == 32) {
    KV_CUDA(cudaFuncSetAttribute(atc::attn_tc_bwd_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
            "attn_tc_bwd smem");
    atc::attn_tc_bwd_kernel<32><<<B * H, threads, smem, (cudaStream_t)stream>>>(qkv, dout, lse, dqkv, N, H, d_head);
  } else {
    KV_CUDA(cudaFuncSetAttribute(atc::attn_tc_bwd_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
            "attn_tc_bwd smem");
    atc::attn_tc_bwd_kernel<64><<<B * H, threads, smem, (cudaStream_t)stream>>>(qkv, dout, lse, dqkv, N, H, d_head);
  }
  KV_CHECK_LAUNCH("attn_tc_bwd");
  return KV_OK;
}
