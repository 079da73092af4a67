// Blackwell-native fused softmax attention per (sample, head): S = Q K^T and O = P V on tcgen05 with the
// score / output accumulators in TMEM, operands bulk-copied (TMA engine) from a bf16 "attention image" that
// the grouped q/k/v GEMM writes in the tensor-core operand layout, several heads in flight per CTA.
//
// Spec: MSA.forward, attention.py:199-200 of the reference (softmax(q k^T / sqrt(d_h)) v per sample and head);
// the log-sum-exp / recompute-backward scheme is FlashAttentionFunction's (utils.py:137-227, 229-295):
// p = exp(s - lse), D_i = sum_j p_ij dp_ij, ds = p * scale * (dp - D).
//
// Attention image (written by kan_tc_fwd_kernel's image epilogue, kv_kan_fwd with KV_FLAG_OUT_ATTN_IMAGE):
//   img[b][h][part (q, k, v)][n8][d8][n % 8][d % 8]  bf16, n < NPAD = roundup(N, 16), d < DHP = roundup(d_h, 16),
//   zero padded.  One part is one contiguous block = one bulk copy, and its 8 x 8 blocks are exactly the
//   no-swizzle core matrices of tcgen05's shared-memory descriptors, so the same bytes serve as a K-major operand
//   (contraction over d: Q, K in S = Q K^T; dO, V in dP = dO V^T) and as an MN-major operand (contraction over the
//   token index: V in O = P V, dO / Q in dV = P^T dO / dK = dS^T Q, K in dQ = dS K): only LBO / SBO differ.
//
// Forward, one CTA per SM, persistent over heads (cta_group::1, M = 128 query rows per tile, <= 2 tiles):
//   warp 8 (loader)   bulk copies of Q,K and of V, one head ahead of the compute
//   warp 9 (MMA)      S_t = Q_t K^T  -> TMEM columns [256 t, 256 t + NPAD);  O_t = P_t V -> TMEM columns [256 t, +DHP)
//   warps 0-3 / 4-7   tile 0 / 1: thread = query row.  Pass 1 reads S for the row maximum, pass 2 re-reads it,
//                     exponentiates (exp2, scale folded in), sums, and stores bf16 P as the A operand of O = P V;
//                     then the epilogue reads O, divides by the row sum, adds the residual (model.py:34) and
//                     writes the output row and the log-sum-exp.
#include <cfloat>
#include "internal.cuh"
#include "tc_ptx.cuh"

namespace kv {
namespace at5 {

constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;
constexpr int FWD_THREADS = 320;

__host__ __device__ inline int npad_of(int N) { return (N + 15) / 16 * 16; }
__host__ __device__ inline int dhp_of(int dh) { return (dh + 15) / 16 * 16; }

struct FwdSmem {   // byte offsets into dynamic shared memory
  int q, k, v, p0, p1, bars, total;
};
__host__ __device__ inline FwdSmem fwd_smem(int N, int dh) {
  const int NPAD = npad_of(N), DHP = dhp_of(dh);
  const int ntile = (N + 127) / 128;
  FwdSmem s;
  s.q = 0;
  s.k = s.q + ntile * 128 * DHP * 2;
  s.v = s.k + NPAD * DHP * 2;
  s.p0 = s.v + NPAD * DHP * 2;
  s.p1 = s.p0 + 128 * NPAD * 2;
  s.bars = s.p1 + (ntile > 1 ? 128 * NPAD * 2 : 0);
  s.total = s.bars + 256;
  return s;
}

__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

struct FwdParams {
  const __nv_bfloat16* img;   // attention image
  const float* residual;      // [B*N][H*dh] or null
  float* out;                 // [B*N][H*dh]
  float* lse;                 // [B][H][N]
  int B, N, H, dh;
  int n_heads_total;          // B * H
};

// barrier indices
enum { B_QK_FULL = 0, B_V_FULL, B_QK_EMPTY, B_V_EMPTY, B_S_FULL, B_P_FULL = B_S_FULL + 2, B_O_FULL = B_P_FULL + 2,
       B_O_EMPTY = B_O_FULL + 2, B_COUNT = B_O_EMPTY + 2 };

__global__ void __launch_bounds__(FWD_THREADS, 1) attn5_fwd_kernel(const FwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int N = p.N, dh = p.dh, H = p.H;
  const int NPAD = npad_of(N), DHP = dhp_of(dh);
  const int ntile = (N + 127) / 128;
  const FwdSmem L = fwd_smem(N, dh);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bars);
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto bar = [&](int i) { return bar0 + 8u * i; };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + B_COUNT);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t part_bytes = (uint32_t)NPAD * DHP * 2;

  if (warp == 9 && lane == 0) {
    for (int i = 0; i < B_COUNT; ++i) {
      const bool warp_count = (i >= B_P_FULL && i < B_O_FULL) || i >= B_O_EMPTY;   // arrivals: one lane per warp
      ptx::mbar_init(bar(i), warp_count ? 4 : 1);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 9) {
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 512);
    ptx::tmem_relinquish();
  }
  // Q rows beyond NPAD (up to the end of the last 128-row tile) are read by the MMA but never loaded: zero them
  // once so that the rows of S / O they produce stay finite
  {
    const int q_bytes = ntile * 128 * DHP * 2;
    for (int i = (int)part_bytes + threadIdx.x * 16; i < q_bytes; i += FWD_THREADS * 16)
      *reinterpret_cast<uint4*>(smem + L.q + i) = make_uint4(0, 0, 0, 0);
    ptx::fence_proxy_async_smem();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int n_iter = (p.n_heads_total - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 8) {
    // ================= loader =================
    if (lane == 0) {
      for (int it = 0; it < n_iter; ++it) {
        const size_t head = (size_t)blockIdx.x + (size_t)it * gridDim.x;
        const uint8_t* src = reinterpret_cast<const uint8_t*>(p.img) + head * 3 * (size_t)part_bytes;
        const uint32_t ph = (uint32_t)(it & 1) ^ 1u;
        ptx::mbar_wait(bar(B_QK_EMPTY), ph);
        ptx::mbar_arrive_expect_tx(bar(B_QK_FULL), 2 * part_bytes);
        ptx::bulk_g2s(ptx::smem_u32(smem + L.q), src, part_bytes, bar(B_QK_FULL));
        ptx::bulk_g2s(ptx::smem_u32(smem + L.k), src + part_bytes, part_bytes, bar(B_QK_FULL));
        ptx::mbar_wait(bar(B_V_EMPTY), ph);
        ptx::mbar_arrive_expect_tx(bar(B_V_FULL), part_bytes);
        ptx::bulk_g2s(ptx::smem_u32(smem + L.v), src + 2 * (size_t)part_bytes, part_bytes, bar(B_V_FULL));
      }
    }
    __syncwarp();
  } else if (warp == 9) {
    // ================= MMA issuer =================
    const uint32_t idesc_s = ptx::make_idesc_bf16(128, NPAD);            // S_t[128, NPAD] = Q_t . K^T
    const uint32_t idesc_o = ptx::make_idesc_bf16(128, DHP, 0, 1);       // O_t[128, DHP] = P_t . V  (V MN-major)
    const uint32_t rg = (uint32_t)DHP * 16;                              // bytes per 8-row group of an image part
    const uint64_t dq0 = ptx::make_smem_desc(ptx::smem_u32(smem + L.q), 128, rg);
    const uint64_t dk0 = ptx::make_smem_desc(ptx::smem_u32(smem + L.k), 128, rg);
    const uint64_t dv0 = ptx::make_smem_desc(ptx::smem_u32(smem + L.v), rg, 128);
    const uint64_t dp0[2] = {ptx::make_smem_desc(ptx::smem_u32(smem + L.p0), 2048, 128),
                             ptx::make_smem_desc(ptx::smem_u32(smem + L.p1), 2048, 128)};
    for (int it = 0; it < n_iter; ++it) {
      const uint32_t ph = (uint32_t)(it & 1);
      ptx::mbar_wait(bar(B_QK_FULL), ph);
      for (int t = 0; t < ntile; ++t) {
        ptx::mbar_wait(bar(B_O_EMPTY + t), ph ^ 1u);   // the epilogue of the previous head has read O_t
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
          const uint64_t dq = dq0 + (uint64_t)((t * 16 * rg) >> 4);
          for (int ks = 0; ks < DHP / 16; ++ks)
            ptx::mma_bf16_ss(tmem_base + t * 256, dq + (uint64_t)(ks * 16), dk0 + (uint64_t)(ks * 16), idesc_s, ks > 0);
          ptx::tc_commit(bar(B_S_FULL + t));
        }
        __syncwarp();
      }
      if (ptx::elect_one()) ptx::tc_commit(bar(B_QK_EMPTY));
      __syncwarp();
      ptx::mbar_wait(bar(B_V_FULL), ph);
      for (int t = 0; t < ntile; ++t) {
        ptx::mbar_wait(bar(B_P_FULL + t), ph);
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
          for (int ks = 0; ks < NPAD / 16; ++ks)
            ptx::mma_bf16_ss(tmem_base + t * 256, dp0[t] + (uint64_t)(ks * 256), dv0 + (uint64_t)((ks * 2 * rg) >> 4),
                             idesc_o, ks > 0);
          ptx::tc_commit(bar(B_O_FULL + t));
        }
        __syncwarp();
      }
      if (ptx::elect_one()) ptx::tc_commit(bar(B_V_EMPTY));
      __syncwarp();
    }
  } else if ((warp >> 2) < ntile) {
    // ================= softmax + epilogue: thread = query row =================
    const int t = warp >> 2;
    const int rl = (warp & 3) * 32 + lane;           // row inside the tile = TMEM lane
    const int n = t * 128 + rl;                      // token index inside the sample
    const bool row_ok = n < N;
    const uint32_t trow = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + t * 256;
    uint8_t* sp = smem + (t ? L.p1 : L.p0) + (rl >> 3) * 128 + (rl & 7) * 16;   // + key8 * 2048
    const float cs = rsqrtf((float)dh) * LOG2E;      // exp(s * scale) = exp2(s * cs)
    for (int it = 0; it < n_iter; ++it) {
      const size_t head = (size_t)blockIdx.x + (size_t)it * gridDim.x;
      const int b = (int)(head / H), h = (int)(head % H);
      const uint32_t ph = (uint32_t)(it & 1);
      ptx::mbar_wait(bar(B_S_FULL + t), ph);
      ptx::tc_fence_after();
      // pass 1: row maximum over the valid keys
      float mx = -FLT_MAX;
      for (int c0 = 0; c0 < NPAD; c0 += 32) {
        uint32_t r[32];
        if (c0 + 32 <= NPAD) {
          ptx::tmem_ld32(trow + c0, r);
        } else {
          uint32_t r16[16];
          tmem_ld16(trow + c0, r16);
#pragma unroll
          for (int e = 0; e < 16; ++e) { r[e] = r16[e]; r[16 + e] = 0xff7fffffu; }   // -FLT_MAX
        }
        ptx::tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; ++e) mx = fmaxf(mx, (c0 + e < N) ? __uint_as_float(r[e]) : -FLT_MAX);
      }
      // pass 2: p = exp2((s - mx) * cs), row sum, bf16 P into the A operand of O = P V
      const float mxc = mx * cs;
      float sum = 0.0f;
      for (int c0 = 0; c0 < NPAD; c0 += 32) {
        uint32_t r[32];
        if (c0 + 32 <= NPAD) {
          ptx::tmem_ld32(trow + c0, r);
        } else {
          uint32_t r16[16];
          tmem_ld16(trow + c0, r16);
#pragma unroll
          for (int e = 0; e < 16; ++e) { r[e] = r16[e]; r[16 + e] = 0u; }
        }
        ptx::tmem_ld_wait();
        float pv[32];
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const float v = exp2f(fmaf(__uint_as_float(r[e]), cs, -mxc));
          pv[e] = (row_ok && c0 + e < N) ? v : 0.0f;
          sum += pv[e];
        }
#pragma unroll
        for (int k8 = 0; k8 < 4; ++k8) {
          if (c0 + 8 * k8 < NPAD)
            *reinterpret_cast<uint4*>(sp + (size_t)((c0 >> 3) + k8) * 2048) =
                make_uint4(pack2(pv[8 * k8], pv[8 * k8 + 1]), pack2(pv[8 * k8 + 2], pv[8 * k8 + 3]),
                           pack2(pv[8 * k8 + 4], pv[8 * k8 + 5]), pack2(pv[8 * k8 + 6], pv[8 * k8 + 7]));
        }
      }
      ptx::tc_fence_before();
      ptx::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(bar(B_P_FULL + t));
      // epilogue: O / sum (+ residual), log-sum-exp
      ptx::mbar_wait(bar(B_O_FULL + t), ph);
      ptx::tc_fence_after();
      const float inv = row_ok ? 1.0f / sum : 0.0f;
      const size_t orow = ((size_t)b * N + (row_ok ? n : 0)) * ((size_t)H * dh) + (size_t)h * dh;
      for (int c0 = 0; c0 < DHP; c0 += 16) {
        uint32_t r16[16];
        tmem_ld16(trow + c0, r16);
        ptx::tmem_ld_wait();
        if (row_ok) {
#pragma unroll
          for (int e4 = 0; e4 < 4; ++e4) {
            const int d0 = c0 + 4 * e4;
            if (d0 < dh) {     // dh % 4 == 0 (checked by the host)
              float4 o = make_float4(__uint_as_float(r16[4 * e4]) * inv, __uint_as_float(r16[4 * e4 + 1]) * inv,
                                     __uint_as_float(r16[4 * e4 + 2]) * inv, __uint_as_float(r16[4 * e4 + 3]) * inv);
              if (p.residual != nullptr) {
                const float4 rv = *reinterpret_cast<const float4*>(p.residual + orow + d0);
                o.x += rv.x; o.y += rv.y; o.z += rv.z; o.w += rv.w;
              }
              *reinterpret_cast<float4*>(p.out + orow + d0) = o;
            }
          }
        }
      }
      if (row_ok && p.lse != nullptr) p.lse[((size_t)b * H + h) * N + n] = (mxc + log2f(sum)) * LN2;
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(bar(B_O_EMPTY + t));
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace at5

// shape verdict shared by the launchers (callers probe with B = 0)
static int attn5_check(int B, int N, int H, int dh) {
  if (B < 0 || N <= 0 || H <= 0 || dh <= 0) return KV_ERR_BAD_ARG;
  if (N > 256 || dh > 64 || (dh % 8) != 0) return KV_ERR_UNSUPPORTED;   // two 128-row tiles; 16-byte d chunks
  return KV_OK;
}

}  // namespace kv

using namespace kv;

extern "C" size_t kv_attn_image_bytes(int B, int N, int H, int d_head) {
  if (B <= 0 || N <= 0 || H <= 0 || d_head <= 0) return 0;
  return (size_t)B * H * 3 * at5::npad_of(N) * at5::dhp_of(d_head) * 2;
}

extern "C" int kv_attn_image_fwd(const void* image, const float* residual, float* out, float* lse, int B, int N,
                                 int H, int d_head, void* stream) {
  const int rc = attn5_check(B, N, H, d_head);
  if (rc != KV_OK) return rc;
  if (B == 0) return KV_OK;
  if (image == nullptr || out == nullptr) return KV_ERR_BAD_ARG;
  if ((reinterpret_cast<uintptr_t>(image) & 15) || (reinterpret_cast<uintptr_t>(out) & 15) ||
      (residual && (reinterpret_cast<uintptr_t>(residual) & 15)))
    return KV_ERR_ALIGNMENT;
  const at5::FwdSmem L = at5::fwd_smem(N, d_head);
  KV_SMEM_OPT_IN(at5::attn5_fwd_kernel, L.total, "attn5_fwd smem attribute");
  at5::FwdParams p;
  p.img = reinterpret_cast<const __nv_bfloat16*>(image);
  p.residual = residual;
  p.out = out;
  p.lse = lse;
  p.B = B; p.N = N; p.H = H; p.dh = d_head;
  p.n_heads_total = B * H;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = p.n_heads_total < sms ? p.n_heads_total : sms;
  at5::attn5_fwd_kernel<<<grid, at5::FWD_THREADS, L.total, (cudaStream_t)stream>>>(p);
  KV_CHECK_LAUNCH("attn5_fwd_kernel");
  return KV_OK;
}
