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
  if (L.total > 227 * 1024) return KV_ERR_BAD_SHAPE;
  KV_SMEM_OPT_IN(at5::attn5_fwd_kernel, 227 * 1024, "attn5_fwd smem attribute");
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

// =====================================================================================================
// Backward (recompute scheme of utils.py:229-295): per head
//   S = Q K^T, dP = dO V^T            (tcgen05, blocks of 128 queries x <= 128 keys, accumulators in TMEM)
//   p = exp(s * scale - lse),  D_i = sum_j p_ij dp_ij,  ds = p * (dp - D_i) * scale      (threads, thread = query row)
//   dQ += dS K,  dV += P^T dO,  dK += dS^T Q                                             (tcgen05, P / dS via shared memory)
// Two passes over the (query tile, key block) grid: pass A only accumulates D_i (it needs every key of the row),
// pass B forms P and dS and feeds the three gradient GEMMs.  dQ (both query tiles), dV and dK (current key block)
// accumulate in TMEM; the gradients leave the kernel as the bf16 dY image of the q/k/v layer's gradient kernels
// (kan_tc.cu: [head][t8][o8][t % 8][o % 8], o = part * d_h + d), so neither an fp32 dqkv nor a dy_image pass exists.
//
// TMEM columns: S [0,128)  dP [128,256)  dQ tile 0 / 1 [256,320) / [320,384)  dV [384,448)  dK [448,512).
// Warps 0-7: compute (warps w and w+4 share TMEM lane quadrant w % 4 and split a block's key columns), warp 8:
// bulk-copy loader (Q, K, V), warp 9: MMA issuer (one thread).
// =====================================================================================================
namespace kv {
namespace at5 {

constexpr int BWD_THREADS = 320;

struct BwdSmem {
  int q, k, v, dO, p, ds, lse, dpart, bars, total;
};
__host__ __device__ inline BwdSmem bwd_smem(int N, int dh) {
  const int NPAD = npad_of(N), DHP = dhp_of(dh);
  const int ntile = (N + 127) / 128;
  BwdSmem s;
  s.p = 0;                                   // P and dS first: the MN-major reads of a short last key block run past
  s.ds = s.p + 128 * 128 * 2;                // their valid key8 blocks into what follows (finite data, unused rows)
  s.q = s.ds + 128 * 128 * 2;
  s.dO = s.q + ntile * 128 * DHP * 2;
  s.k = s.dO + ntile * 128 * DHP * 2;
  s.v = s.k + NPAD * DHP * 2;
  s.lse = s.v + NPAD * DHP * 2;
  s.dpart = s.lse + 256 * 4;
  s.bars = s.dpart + 2 * 256 * 4;
  s.total = s.bars + 256;
  return s;
}

struct BwdParams {
  const __nv_bfloat16* img;   // attention image (q, k, v)
  const float* dout;          // [B*N][H*dh] gradient of the attention output
  const float* lse;           // [B][H][N]
  __nv_bfloat16* dyimg;       // dY image of the q/k/v layer: [H][T8][O8][8][8]
  int B, N, H, dh;
  int T8, O8;
  int n_heads_total;
};

enum { C_QKV_FULL = 0, C_QKV_EMPTY, C_DO_FULL, C_SD_FULL, C_CD_DONE, C_PDS_EMPTY, C_KV_FULL, C_KV_EMPTY, C_DQ_FULL,
       C_DQ_EMPTY, C_COUNT };

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(BWD_THREADS, 1) attn5_bwd_kernel(const BwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int N = p.N, dh = p.dh, H = p.H;
  const int NPAD = npad_of(N), DHP = dhp_of(dh);
  const int ntile = (N + 127) / 128;          // query tiles
  const int nkb = (NPAD + 127) / 128;         // key blocks
  const BwdSmem L = bwd_smem(N, dh);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bars);
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto bar = [&](int i) { return bar0 + 8u * i; };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + C_COUNT);
  float* sL = reinterpret_cast<float*>(smem + L.lse);       // lse in log2 units, +inf on the padding rows
  float* sDp = reinterpret_cast<float*>(smem + L.dpart);    // [2 column halves][256 rows] partial D_i
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t part_bytes = (uint32_t)NPAD * DHP * 2;
  const uint32_t rg = (uint32_t)DHP * 16;                   // bytes per 8-row group of an image part
  const int n_steps = ntile * nkb;                          // (query tile, key block) pairs per pass

  if (warp == 9 && lane == 0) {
    for (int i = 0; i < C_COUNT; ++i) {
      const bool by_warps = (i == C_DO_FULL || i == C_CD_DONE || i == C_KV_EMPTY || i == C_DQ_EMPTY);
      ptx::mbar_init(bar(i), by_warps ? 8 : 1);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 9) {
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 512);
    ptx::tmem_relinquish();
  }
  {  // rows of the Q / dO tiles beyond NPAD are MMA operands but never written: zero once (dO: 0 * garbage = NaN)
    const int tile_bytes = ntile * 128 * DHP * 2;
    for (int i = (int)part_bytes + threadIdx.x * 16; i < tile_bytes; i += BWD_THREADS * 16) {
      *reinterpret_cast<uint4*>(smem + L.q + i) = make_uint4(0, 0, 0, 0);
      *reinterpret_cast<uint4*>(smem + L.dO + i) = make_uint4(0, 0, 0, 0);
    }
    for (int i = threadIdx.x * 16; i < 2 * 128 * 128 * 2; i += BWD_THREADS * 16)   // P, dS: finite from the start
      *reinterpret_cast<uint4*>(smem + L.p + i) = make_uint4(0, 0, 0, 0);
    ptx::fence_proxy_async_smem();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int n_iter = (p.n_heads_total - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  constexpr uint32_t COL_S = 0, COL_DP = 128, COL_DQ = 256, COL_DV = 384, COL_DK = 448;

  if (warp == 8) {
    // ================= loader: Q, K, V of the next head as soon as this head's MMAs are done =================
    if (lane == 0) {
      for (int it = 0; it < n_iter; ++it) {
        const size_t head = (size_t)blockIdx.x + (size_t)it * gridDim.x;
        const uint8_t* src = reinterpret_cast<const uint8_t*>(p.img) + head * 3 * (size_t)part_bytes;
        ptx::mbar_wait(bar(C_QKV_EMPTY), (uint32_t)(it & 1) ^ 1u);
        ptx::mbar_arrive_expect_tx(bar(C_QKV_FULL), 3 * part_bytes);
        ptx::bulk_g2s(ptx::smem_u32(smem + L.q), src, part_bytes, bar(C_QKV_FULL));
        ptx::bulk_g2s(ptx::smem_u32(smem + L.k), src + part_bytes, part_bytes, bar(C_QKV_FULL));
        ptx::bulk_g2s(ptx::smem_u32(smem + L.v), src + 2 * (size_t)part_bytes, part_bytes, bar(C_QKV_FULL));
      }
    }
    __syncwarp();
  } else if (warp == 9) {
    // ================= MMA issuer (one thread) =================
    if (lane == 0) {
      const uint32_t a_q = ptx::smem_u32(smem + L.q), a_k = ptx::smem_u32(smem + L.k), a_v = ptx::smem_u32(smem + L.v);
      const uint32_t a_do = ptx::smem_u32(smem + L.dO), a_p = ptx::smem_u32(smem + L.p), a_ds = ptx::smem_u32(smem + L.ds);
      const uint32_t id_o = ptx::make_idesc_bf16(128, DHP, 0, 1);   // dQ: A K-major (dS), B MN-major (K)
      const uint32_t id_t = ptx::make_idesc_bf16(128, DHP, 1, 1);   // dV / dK: A MN-major (P^T, dS^T), B MN-major
      // S and dP of step (t, kb)
      auto issue_sd = [&](int t, int kb) {
        const int KW = min(128, NPAD - 128 * kb);
        const uint32_t id_s = ptx::make_idesc_bf16(128, KW);
        const uint64_t dq = ptx::make_smem_desc(a_q + t * 16 * rg, 128, rg);
        const uint64_t dk = ptx::make_smem_desc(a_k + kb * 16 * rg, 128, rg);
        const uint64_t ddo = ptx::make_smem_desc(a_do + t * 16 * rg, 128, rg);
        const uint64_t dv = ptx::make_smem_desc(a_v + kb * 16 * rg, 128, rg);
        for (int ks = 0; ks < DHP / 16; ++ks)
          ptx::mma_bf16_ss(tmem_base + COL_S, dq + (uint64_t)(ks * 16), dk + (uint64_t)(ks * 16), id_s, ks > 0);
        for (int ks = 0; ks < DHP / 16; ++ks)
          ptx::mma_bf16_ss(tmem_base + COL_DP, ddo + (uint64_t)(ks * 16), dv + (uint64_t)(ks * 16), id_s, ks > 0);
        ptx::tc_commit(bar(C_SD_FULL));
      };
      auto step_tk = [&](int j, int& t, int& kb) {   // pass A: tile-major, pass B: key-block-major
        if (j < n_steps) { t = j / nkb; kb = j - t * nkb; }
        else { const int jb = j - n_steps; kb = jb / ntile; t = jb - kb * ntile; }
      };
      uint32_t n_cd = 0, n_blk = 0;
      for (int it = 0; it < n_iter; ++it) {
        const uint32_t ph = (uint32_t)(it & 1);
        ptx::mbar_wait(bar(C_QKV_FULL), ph);
        ptx::mbar_wait(bar(C_DO_FULL), ph);
        ptx::tc_fence_after();
        int t, kb;
        step_tk(0, t, kb);
        issue_sd(t, kb);
        for (int j = 0; j < 2 * n_steps; ++j) {
          ptx::mbar_wait(bar(C_CD_DONE), n_cd & 1u);   // step j: S / dP consumed (pass B: P / dS written)
          ++n_cd;
          ptx::tc_fence_after();
          if (j + 1 < 2 * n_steps) {
            step_tk(j + 1, t, kb);
            issue_sd(t, kb);
          }
          if (j >= n_steps) {
            step_tk(j, t, kb);
            const int KW = min(128, NPAD - 128 * kb);
            if (j == n_steps) ptx::mbar_wait(bar(C_DQ_EMPTY), ph ^ 1u);          // dQ of the previous head drained
            if (t == 0) ptx::mbar_wait(bar(C_KV_EMPTY), (n_blk & 1u) ^ 1u);      // dV / dK of the previous block drained
            ptx::tc_fence_after();
            // dQ_t += dS . K_kb
            {
              const uint64_t da = ptx::make_smem_desc(a_ds, 2048, 128);
              const uint64_t db = ptx::make_smem_desc(a_k + kb * 16 * rg, rg, 128);
              for (int ks = 0; ks < KW / 16; ++ks)
                ptx::mma_bf16_ss(tmem_base + COL_DQ + 64 * t, da + (uint64_t)(ks * 256), db + (uint64_t)((ks * 2 * rg) >> 4),
                                 id_o, (kb > 0) || (ks > 0));
            }
            // dV_kb += P^T . dO_t ;  dK_kb += dS^T . Q_t   (contraction over the 128 query rows of the tile)
            {
              const uint64_t dap = ptx::make_smem_desc(a_p, 128, 2048);
              const uint64_t das = ptx::make_smem_desc(a_ds, 128, 2048);
              const uint64_t dbo = ptx::make_smem_desc(a_do + t * 16 * rg, rg, 128);
              const uint64_t dbq = ptx::make_smem_desc(a_q + t * 16 * rg, rg, 128);
              for (int ks = 0; ks < 8; ++ks)
                ptx::mma_bf16_ss(tmem_base + COL_DV, dap + (uint64_t)(ks * 16), dbo + (uint64_t)((ks * 2 * rg) >> 4), id_t,
                                 (t > 0) || (ks > 0));
              for (int ks = 0; ks < 8; ++ks)
                ptx::mma_bf16_ss(tmem_base + COL_DK, das + (uint64_t)(ks * 16), dbq + (uint64_t)((ks * 2 * rg) >> 4), id_t,
                                 (t > 0) || (ks > 0));
            }
            ptx::tc_commit(bar(C_PDS_EMPTY));
            if (t == ntile - 1) {
              ptx::tc_commit(bar(C_KV_FULL));
              ++n_blk;
            }
          }
        }
        ptx::tc_commit(bar(C_DQ_FULL));
        ptx::tc_commit(bar(C_QKV_EMPTY));
      }
    }
    __syncwarp();
  } else {
    // ================= compute warps: thread = query row (softmax gradient) / key row (dV, dK drain) =================
    const int q4 = warp & 3, hf = warp >> 2;       // TMEM lane quadrant, column half of a key block
    const int rl = q4 * 32 + lane;                 // row inside a 128-row tile = TMEM lane
    const uint32_t trow = tmem_base + ((uint32_t)(q4 * 32) << 16);
    const int ctid = threadIdx.x;                  // 0..255
    const float scale = rsqrtf((float)dh);
    const float cs = scale * LOG2E;
    const int O8 = p.O8;
    uint32_t n_sd = 0, n_b = 0, n_kv = 0;
    // 16-byte chunk (8 consecutive d of part `part`) of token row tg in the dY image
    auto dy_chunk = [&](int h, size_t tg, int part, int d0) {
      const int o = part * dh + d0;
      return reinterpret_cast<uint4*>(p.dyimg + ((((size_t)h * p.T8 + (tg >> 3)) * O8 + (o >> 3)) * 8 + (tg & 7)) * 8);
    };
    // rows [row0, row0 + 128) x DHP columns at TMEM column `col` -> part `part` of the dY image (thread = row rl)
    auto drain = [&](uint32_t col, int b, int h, int row0, int part) {
      const int n = row0 + rl;
      for (int c0 = 0; c0 < DHP; c0 += 16) {
        uint32_t r[16];
        tmem_ld16(trow + col + c0, r);
        ptx::tmem_ld_wait();
        if (n < N) {
#pragma unroll
          for (int k8 = 0; k8 < 2; ++k8) {
            const int d0 = c0 + 8 * k8;
            if (d0 < dh)
              *dy_chunk(h, (size_t)b * N + n, part, d0) =
                  make_uint4(pack2(__uint_as_float(r[8 * k8]), __uint_as_float(r[8 * k8 + 1])),
                             pack2(__uint_as_float(r[8 * k8 + 2]), __uint_as_float(r[8 * k8 + 3])),
                             pack2(__uint_as_float(r[8 * k8 + 4]), __uint_as_float(r[8 * k8 + 5])),
                             pack2(__uint_as_float(r[8 * k8 + 6]), __uint_as_float(r[8 * k8 + 7])));
          }
        }
      }
    };
    for (int it = 0; it < n_iter; ++it) {
      const size_t head = (size_t)blockIdx.x + (size_t)it * gridDim.x;
      const int b = (int)(head / H), h = (int)(head % H);
      const uint32_t ph = (uint32_t)(it & 1);
      // ---- stage dO (fp32 -> bf16 image layout) and the log-sum-exp of the head.  The previous head's MMAs have
      //      all completed (its DQ_FULL was waited for below), so the buffers are free.
      {
        const int f4_per_row = dh / 4;
        const int total = NPAD * f4_per_row;
        const float* src = p.dout + (size_t)b * N * ((size_t)H * dh) + (size_t)h * dh;
        for (int base = ctid; base < total; base += 256 * 4) {
          float4 v[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int idx = base + u * 256;
            v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (idx < total) {
              const int n = idx / f4_per_row, f = idx - n * f4_per_row;
              if (n < N) v[u] = __ldg(reinterpret_cast<const float4*>(src + (size_t)n * ((size_t)H * dh)) + f);
            }
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int idx = base + u * 256;
            if (idx < total) {
              const int n = idx / f4_per_row, f = idx - n * f4_per_row;
              const int d0 = f * 4;
              uint2* dst = reinterpret_cast<uint2*>(smem + L.dO + (size_t)(n >> 3) * rg + (d0 >> 3) * 128 + (n & 7) * 16 +
                                                    (d0 & 7) * 2);
              *dst = make_uint2(pack2(v[u].x, v[u].y), pack2(v[u].z, v[u].w));
            }
          }
        }
        if (dh < DHP) {   // zero the d padding of dO (d_h = 24 -> 32): the K-major dP contraction runs over DHP
          for (int idx = ctid; idx < NPAD * ((DHP - dh) / 8); idx += 256) {
            const int n = idx / ((DHP - dh) / 8), d0 = dh + (idx % ((DHP - dh) / 8)) * 8;
            *reinterpret_cast<uint4*>(smem + L.dO + (size_t)(n >> 3) * rg + (d0 >> 3) * 128 + (n & 7) * 16) =
                make_uint4(0, 0, 0, 0);
          }
        }
        for (int n = ctid; n < 256; n += 256)
          sL[n] = (n < N) ? p.lse[((size_t)b * H + h) * N + n] * LOG2E : INFINITY;
        ptx::fence_proxy_async_smem();
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (lane == 0) ptx::mbar_arrive(bar(C_DO_FULL));
      }
      float l2[2], Dacc[2] = {0.f, 0.f};
      l2[0] = sL[rl];
      l2[1] = sL[128 + rl];
      // ---- pass A: D_i = sum_j p_ij dp_ij over this warp's column half of every key block
      for (int t = 0; t < ntile; ++t) {
        for (int kb = 0; kb < nkb; ++kb) {
          const int KW = min(128, NPAD - 128 * kb);
          ptx::mbar_wait(bar(C_SD_FULL), n_sd & 1u);
          ++n_sd;
          ptx::tc_fence_after();
          float acc = 0.f;
          for (int c0 = 64 * hf; c0 < min(KW, 64 * hf + 64); c0 += 16) {
            uint32_t s16[16], d16[16];
            tmem_ld16(trow + COL_S + c0, s16);
            tmem_ld16(trow + COL_DP + c0, d16);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 16; ++e) {
              const float pe = ex2(fmaf(__uint_as_float(s16[e]), cs, -l2[t]));
              acc = (128 * kb + c0 + e < N) ? fmaf(pe, __uint_as_float(d16[e]), acc) : acc;
            }
          }
          Dacc[t] += acc;
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(bar(C_CD_DONE));
        }
      }
      sDp[hf * 256 + rl] = Dacc[0];
      sDp[hf * 256 + 128 + rl] = Dacc[1];
      asm volatile("bar.sync 1, 256;" ::: "memory");
      float Di[2];
      Di[0] = sDp[rl] + sDp[256 + rl];
      Di[1] = sDp[128 + rl] + sDp[256 + 128 + rl];
      // ---- pass B: P and dS blocks -> shared memory (A operands of dQ / dV / dK), then the gradient drains
      for (int kb = 0; kb < nkb; ++kb) {
        const int KW = min(128, NPAD - 128 * kb);
        for (int t = 0; t < ntile; ++t) {
          ptx::mbar_wait(bar(C_SD_FULL), n_sd & 1u);
          ++n_sd;
          ptx::mbar_wait(bar(C_PDS_EMPTY), (n_b & 1u) ^ 1u);   // the MMAs of the previous step have read P / dS
          ++n_b;
          ptx::tc_fence_after();
          const bool row_ok = t * 128 + rl < N;
          uint8_t* prow = smem + L.p + (rl >> 3) * 128 + (rl & 7) * 16;
          uint8_t* srow = smem + L.ds + (rl >> 3) * 128 + (rl & 7) * 16;
          for (int c0 = 64 * hf; c0 < min(KW, 64 * hf + 64); c0 += 16) {
            uint32_t s16[16], d16[16];
            tmem_ld16(trow + COL_S + c0, s16);
            tmem_ld16(trow + COL_DP + c0, d16);
            ptx::tmem_ld_wait();
            float pv[16], dv[16];
#pragma unroll
            for (int e = 0; e < 16; ++e) {
              const bool ok = row_ok && (128 * kb + c0 + e < N);
              const float pe = ok ? ex2(fmaf(__uint_as_float(s16[e]), cs, -l2[t])) : 0.0f;
              pv[e] = pe;
              dv[e] = pe * (__uint_as_float(d16[e]) - Di[t]) * scale;
            }
#pragma unroll
            for (int k8 = 0; k8 < 2; ++k8) {
              const int blk = (c0 >> 3) + k8;
              *reinterpret_cast<uint4*>(prow + blk * 2048) =
                  make_uint4(pack2(pv[8 * k8], pv[8 * k8 + 1]), pack2(pv[8 * k8 + 2], pv[8 * k8 + 3]),
                             pack2(pv[8 * k8 + 4], pv[8 * k8 + 5]), pack2(pv[8 * k8 + 6], pv[8 * k8 + 7]));
              *reinterpret_cast<uint4*>(srow + blk * 2048) =
                  make_uint4(pack2(dv[8 * k8], dv[8 * k8 + 1]), pack2(dv[8 * k8 + 2], dv[8 * k8 + 3]),
                             pack2(dv[8 * k8 + 4], dv[8 * k8 + 5]), pack2(dv[8 * k8 + 6], dv[8 * k8 + 7]));
            }
          }
          ptx::tc_fence_before();
          ptx::fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(bar(C_CD_DONE));
        }
        // dV (warps 0-3) and dK (warps 4-7) of this key block: thread = key row
        ptx::mbar_wait(bar(C_KV_FULL), n_kv & 1u);
        ++n_kv;
        ptx::tc_fence_after();
        drain(hf == 0 ? COL_DV : COL_DK, b, h, 128 * kb, hf == 0 ? 2 : 1);
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(bar(C_KV_EMPTY));
      }
      // dQ: warps 0-3 query tile 0, warps 4-7 query tile 1
      ptx::mbar_wait(bar(C_DQ_FULL), ph);
      ptx::tc_fence_after();
      if (hf < ntile) drain(COL_DQ + 64 * hf, b, h, 128 * hf, 0);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(bar(C_DQ_EMPTY));
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
}  // namespace kv

extern "C" int kv_attn_image_bwd(const void* image, const float* dout, const float* lse, void* dy_image, int B, int N,
                                 int H, int d_head, int T8, int O8, void* stream) {
  const int rc = attn5_check(B, N, H, d_head);
  if (rc != KV_OK) return rc;
  if (B == 0) return KV_OK;
  if (image == nullptr || dout == nullptr || lse == nullptr || dy_image == nullptr) return KV_ERR_BAD_ARG;
  if ((reinterpret_cast<uintptr_t>(image) & 15) || (reinterpret_cast<uintptr_t>(dout) & 15) ||
      (reinterpret_cast<uintptr_t>(dy_image) & 15))
    return KV_ERR_ALIGNMENT;
  if ((long long)T8 * 8 < (long long)B * N || O8 * 8 < 3 * d_head) return KV_ERR_BAD_SHAPE;
  cudaStream_t st = (cudaStream_t)stream;
  // zero padding of the dY image: token rows beyond B*N (the last, partially filled 8-row group included: the
  // kernel then writes its valid rows) and, when 3 * d_head < 8 * O8, the unused output columns
  const size_t group_bytes = (size_t)O8 * 128;   // one 8-row group of one head
  if (O8 * 8 != 3 * d_head) {
    KV_CUDA(cudaMemsetAsync(dy_image, 0, (size_t)H * T8 * group_bytes, st), "attn5_bwd zero dY image");
  } else {
    const size_t t8_0 = (size_t)B * N / 8;
    if (t8_0 < (size_t)T8)
      KV_CUDA(cudaMemset2DAsync(reinterpret_cast<uint8_t*>(dy_image) + t8_0 * group_bytes, (size_t)T8 * group_bytes, 0,
                                ((size_t)T8 - t8_0) * group_bytes, (size_t)H, st),
              "attn5_bwd zero dY image tail");
  }
  const at5::BwdSmem L = at5::bwd_smem(N, d_head);
  if (L.total > 227 * 1024) return KV_ERR_BAD_SHAPE;
  KV_SMEM_OPT_IN(at5::attn5_bwd_kernel, 227 * 1024, "attn5_bwd smem attribute");
  at5::BwdParams p;
  p.img = reinterpret_cast<const __nv_bfloat16*>(image);
  p.dout = dout;
  p.lse = lse;
  p.dyimg = reinterpret_cast<__nv_bfloat16*>(dy_image);
  p.B = B; p.N = N; p.H = H; p.dh = d_head;
  p.T8 = T8; p.O8 = O8;
  p.n_heads_total = B * H;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = p.n_heads_total < sms ? p.n_heads_total : sms;
  at5::attn5_bwd_kernel<<<grid, at5::BWD_THREADS, L.total, st>>>(p);
  KV_CHECK_LAUNCH("attn5_bwd_kernel");
  return KV_OK;
}
