// tcgen05 / TMEM bf16 "basis-prologue GEMM" (forward):  Y[T,out] = Phi(X)[T, in*M] . Wp[out, in*M]^T
//
// The A operand does not exist in memory: producer warps read X, evaluate the basis expansion
// (basis.cuh) and write bf16 values straight into shared memory in the tcgen05 operand layout.
// The B operand (packed weights) is pre-tiled by tc_pack into exactly the shared-memory image
// one MMA needs, so a tile arrives with ONE bulk async copy (TMA engine) per K=16 unit.
// Accumulators live in TMEM; the epilogue reads them with tcgen05.ld and fuses bias / positional
// embedding / ReLU / residual.
//
// Work decomposition
//   K is walked in "units" of 16: unit u = (chunk c = u / M, basis m = u % M) covers input features
//   [16c, 16c+16) for ONE basis function m.  (A permutation of the K index i*M+m -- the sum is the
//   same.)  One unit = one tcgen05.mma K step: A unit 128 rows x 16 (4 KB), B unit BN rows x 16.
//   CTA tile: 128 rows x BN outputs, BN <= 512 = the whole TMEM (two N<=256 MMAs per unit).
//
// Warp roles (320 threads): warp 0 = TMEM allocator + B loader, warp 1 = barrier init + MMA issuer,
//   warps 2..9 = A producers (basis prologue), then epilogue.
#include "basis.cuh"
#include "internal.cuh"
#include "tc_ptx.cuh"

namespace kv {
namespace tc {

constexpr int BM = 128;              // rows per CTA tile (UMMA M)
constexpr int KU = 16;               // K per unit (UMMA K for bf16)
constexpr int A_UNIT_BYTES = BM * KU * 2;
constexpr int NA = 24;               // A ring slots (>= max M on this path)
constexpr int A_RING_BYTES = NA * A_UNIT_BYTES;
constexpr int B_RING_BYTES = 96 * 1024;
constexpr int NB_MAX = 16;
constexpr int N_PRODUCER_WARPS = 8;
constexpr int NTHREADS = 32 * (2 + N_PRODUCER_WARPS);
constexpr int BAR_BYTES = 1024;
constexpr int SMEM_BYTES = A_RING_BYTES + B_RING_BYTES + BAR_BYTES;
constexpr int TMEM_COLS = 512;
constexpr int MAX_M_TC = NA;

struct Geom {
  int n_chunks;  // ceil(in / 16)
  int n_units;   // n_chunks * M
  int n_tiles;   // output tiles per group
  int BN;        // outputs per tile (multiple of 16, <= 512)
  int NB;        // B ring slots
  __host__ __device__ size_t tile_elems() const { return (size_t)n_units * BN * KU; }
};

__host__ __device__ inline Geom make_geom(int in, int out, int M) {
  Geom g;
  g.n_chunks = (in + KU - 1) / KU;
  g.n_units = g.n_chunks * M;
  const int out16 = (out + 15) / 16 * 16;
  g.n_tiles = (out16 + 511) / 512;
  g.BN = ((out16 + g.n_tiles - 1) / g.n_tiles + 15) / 16 * 16;
  int nb = B_RING_BYTES / (g.BN * KU * 2);
  g.NB = nb > NB_MAX ? NB_MAX : nb;
  return g;
}

// ----------------------------------------------------------------------------------------
// weight pre-tiling: wp_f32 [g][out][in*M] (K index i*M+m)  ->  wtiles bf16
//   [g][n_tile][unit][k-half h][row group rg = BN/8][8 rows][8 elements]
// = for every unit the no-swizzle K-major operand image (core matrix = 8 rows x 16 B).
// ----------------------------------------------------------------------------------------
__global__ void tc_pack_kernel(const float* __restrict__ wp, __nv_bfloat16* __restrict__ wt, int n_groups,
                               int in, int out, int M, Geom gm) {
  const size_t per_tile = gm.tile_elems();
  const size_t total = (size_t)n_groups * gm.n_tiles * per_tile;
  const int K = in * M;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (size_t)gridDim.x * blockDim.x) {
    size_t r = idx;
    const int e = r % 8; r /= 8;
    const int r8 = r % 8; r /= 8;
    const int rg = r % (gm.BN / 8); r /= (gm.BN / 8);
    const int h = r % 2; r /= 2;
    const int u = r % gm.n_units; r /= gm.n_units;
    const int nt = r % gm.n_tiles; r /= gm.n_tiles;
    const int g = (int)r;
    const int o = nt * gm.BN + rg * 8 + r8;
    const int c = u / M, m = u - c * M;
    const int i = c * KU + h * 8 + e;
    float v = 0.0f;
    if (o < out && i < in) v = wp[((size_t)g * out + o) * K + (size_t)i * M + m];
    wt[idx] = __float2bfloat16_rn(v);
  }
}

// writes Phi_m(x) as bf16 into the A ring: unit (u0 + m) -> slot (s0 + m) mod NA
struct UnitSink {
  static constexpr bool kDense = true;
  uint8_t* sA;
  int s0;   // slot of the chunk's unit m = 0
  int off;  // byte offset of (row, feature) inside a unit
  __device__ __forceinline__ void put(int m, float v) {
    int s = s0 + m;
    if (s >= NA) s -= NA;
    *reinterpret_cast<__nv_bfloat16*>(sA + s * A_UNIT_BYTES + off) = __float2bfloat16_rn(v);
  }
};

struct FwdParams {
  kv_kan_desc_t d;
  const float* X;
  const __nv_bfloat16* wt;
  const float* fparams;
  const float* bias;
  const float* mean;
  const float* rstd;
  const float* residual;
  const float* pos;
  float* Y;
  Geom gm;
};

template <int F>
__global__ void __launch_bounds__(NTHREADS, 1) kan_tc_fwd_kernel(const FwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;
  uint8_t* sB = smem + A_RING_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + A_RING_BYTES + B_RING_BYTES);
  // barrier map: [0,NA) fullA, [NA,2NA) emptyA, [2NA, 2NA+NB_MAX) fullB, [.., +NB_MAX) emptyB, then tmem_full
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto fullA = [&](int s) { return bar0 + 8u * s; };
  auto emptyA = [&](int s) { return bar0 + 8u * (NA + s); };
  auto fullB = [&](int s) { return bar0 + 8u * (2 * NA + s); };
  auto emptyB = [&](int s) { return bar0 + 8u * (2 * NA + NB_MAX + s); };
  const uint32_t tmem_full = bar0 + 8u * (2 * NA + 2 * NB_MAX);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * NA + 2 * NB_MAX + 1);

  const kv_kan_desc_t& d = p.d;
  const Geom gm = p.gm;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_tile = blockIdx.x, n_tile = blockIdx.y, g = blockIdx.z;
  const int M = d.M;
  const int n_units = gm.n_units;
  const int BN = gm.BN, NB = gm.NB;
  const uint32_t b_unit_bytes = (uint32_t)BN * KU * 2;

  if (warp == 1 && lane == 0) {
    for (int s = 0; s < NA; ++s) {
      ptx::mbar_init(fullA(s), N_PRODUCER_WARPS);
      ptx::mbar_init(emptyA(s), 1);
    }
    for (int s = 0; s < NB_MAX; ++s) {
      ptx::mbar_init(fullB(s), 1);
      ptx::mbar_init(emptyB(s), 1);
    }
    ptx::mbar_init(tmem_full, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 0) {
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================= B loader: one bulk copy per unit =================
    if (lane == 0) {
      const uint8_t* src = reinterpret_cast<const uint8_t*>(p.wt) +
                           ((size_t)g * gm.n_tiles + n_tile) * gm.tile_elems() * 2;
      for (int u = 0; u < n_units; ++u) {
        const int s = u % NB;
        const uint32_t it = u / NB;
        ptx::mbar_wait(emptyB(s), (it & 1) ^ 1);
        ptx::mbar_arrive_expect_tx(fullB(s), b_unit_bytes);
        ptx::bulk_g2s(ptx::smem_u32(sB) + s * b_unit_bytes, src + (size_t)u * b_unit_bytes, b_unit_bytes,
                      fullB(s));
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      const int n1 = BN > 256 ? 256 : BN;
      const int n2 = BN - n1;
      const uint32_t idesc1 = ptx::make_idesc_bf16(BM, n1);
      const uint32_t idesc2 = n2 ? ptx::make_idesc_bf16(BM, n2) : 0u;
      const uint32_t a_lbo = (BM / 8) * 128, b_lbo = (uint32_t)(BN / 8) * 128, sbo = 128;
      for (int u = 0; u < n_units; ++u) {
        const int sa = u % NA, sb = u % NB;
        ptx::mbar_wait(fullA(sa), (u / NA) & 1);
        ptx::mbar_wait(fullB(sb), (u / NB) & 1);
        ptx::tc_fence_after();
        const uint64_t da = ptx::make_smem_desc(ptx::smem_u32(sA) + sa * A_UNIT_BYTES, a_lbo, sbo);
        const uint32_t b_addr = ptx::smem_u32(sB) + sb * b_unit_bytes;
        ptx::mma_bf16_ss(tmem_base, da, ptx::make_smem_desc(b_addr, b_lbo, sbo), idesc1, u > 0);
        if (n2)
          ptx::mma_bf16_ss(tmem_base + 256, da, ptx::make_smem_desc(b_addr + 256 * 16, b_lbo, sbo), idesc2,
                           u > 0);
        ptx::tc_commit(emptyA(sa));
        ptx::tc_commit(emptyB(sb));
      }
      ptx::tc_commit(tmem_full);
    }
    __syncwarp();
  } else {
    // ================= A producers: basis prologue =================
    const int pw = warp - 2;
    const int f8 = lane & 7, r4 = lane >> 3;
    const int n_slices = d.n_groups / d.x_share;
    const int slice = g / d.x_share;
    const int xoff = slice * d.in_features;
    const int t0 = m_tile * BM;
    const BasisCtx c = make_ctx(d, p.fparams, g);
    for (int ch = 0; ch < gm.n_chunks; ++ch) {
      const int u0 = ch * M;
      const int s0 = u0 % NA;
      // wait until the MMAs that read these slots one ring pass ago have completed
      for (int m = 0; m < M; ++m) {
        const int u = u0 + m;
        ptx::mbar_wait(emptyA(u % NA), ((u / NA) & 1) ^ 1);
      }
#pragma unroll 1
      for (int step = 0; step < 8; ++step) {
        const int row = pw * 16 + (step >> 1) * 4 + r4;
        const int h = step & 1;
        const int i = ch * KU + h * 8 + f8;
        const int t = t0 + row;
        UnitSink sink{sA, s0, h * 2048 + (row >> 3) * 128 + (row & 7) * 16 + f8 * 2};
        if (i < d.in_features) {
          const float x = (t < d.T) ? p.X[(size_t)t * d.ldx + xoff + i] : 0.0f;
          float xn = x;
          if (F == KV_FASTKAN) {
            float mu = 0.f, rs = 1.f;
            if (!(d.flags & KV_FLAG_NO_LAYERNORM) && t < d.T) {
              mu = p.mean[(size_t)t * n_slices + slice];
              rs = p.rstd[(size_t)t * n_slices + slice];
            }
            xn = fastkan_xn(c, i, x, mu, rs);
          }
          basis_eval<F>(c, i, x, xn, sink);
        } else {
          for (int m = 0; m < M; ++m) sink.put(m, 0.0f);
        }
      }
      ptx::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        for (int m = 0; m < M; ++m) ptx::mbar_arrive(fullA((u0 + m) % NA));
      }
    }
    // ================= epilogue =================
    ptx::mbar_wait(tmem_full, 0);
    ptx::tc_fence_after();
    const int q = warp & 3;             // TMEM lane quadrant this warp may read
    const int half = pw >> 2;           // two warps per quadrant split the column blocks
    const int row = q * 32 + lane;
    const int t = t0 + row;
    const int out = d.out_features;
    const int ncols_total = d.n_groups * out;
    const int prow = map_row(t < d.T ? t : 0, d.rows_in, d.rows_out, d.row_shift);
    const int posrow = d.rows_in ? (d.row_shift + (t < d.T ? t : 0) % d.rows_in) : t;
    for (int cb = half; cb * 32 < BN; cb += 2) {
      uint32_t r[32];
      ptx::tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + cb * 32, r);
      ptx::tmem_ld_wait();
      if (t < d.T) {
        const int o0 = n_tile * BN + cb * 32;
        const int col0 = g * out + o0;
        float* yrow = p.Y + (size_t)prow * d.ldy + col0;
        const float* rrow = p.residual ? p.residual + (size_t)prow * d.ldy + col0 : nullptr;
        const float* prow_p = p.pos ? p.pos + (size_t)posrow * ncols_total + col0 : nullptr;
        const float* brow = p.bias ? p.bias + col0 : nullptr;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          if (o0 + j < out) {
            float v = __uint_as_float(r[j]);
            if (brow) v += brow[j];
            if (prow_p) v += prow_p[j];
            if (d.flags & KV_FLAG_RELU) v = fmaxf(v, 0.0f);
            if (rrow) v += rrow[j];
            yrow[j] = v;
          }
        }
      }
    }
    ptx::tc_fence_before();
  }
  __syncthreads();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

template <int F>
static int launch_fwd(const FwdParams& p, cudaStream_t st) {
  static bool attr_set = false;  // benign race: the attribute value is always the same
  if (!attr_set) {
    KV_CUDA(cudaFuncSetAttribute(kan_tc_fwd_kernel<F>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES),
            "tc_fwd smem attribute");
    attr_set = true;
  }
  dim3 grid(ceil_div(p.d.T, BM), p.gm.n_tiles, p.d.n_groups);
  kan_tc_fwd_kernel<F><<<grid, NTHREADS, SMEM_BYTES, st>>>(p);
  KV_CHECK_LAUNCH("kan_tc_fwd_kernel");
  return KV_OK;
}

}  // namespace tc

int tc_supported(const kv_kan_desc_t& d) {
  if (d.M > tc::MAX_M_TC) return KV_ERR_UNSUPPORTED;
  return KV_OK;
}

size_t tc_wtiles_bytes(const kv_kan_desc_t& d) {
  const tc::Geom gm = tc::make_geom(d.in_features, d.out_features, d.M);
  return (size_t)d.n_groups * gm.n_tiles * gm.tile_elems() * 2;
}

size_t tc_wtiles_t_bytes(const kv_kan_desc_t&) { return 0; }
bool tc_bwd_weight_available(const kv_kan_desc_t&) { return false; }

int tc_pack(const kv_kan_desc_t& d, const float* wp_f32, void* wtiles, void* wtiles_t, cudaStream_t st) {
  (void)wtiles_t;
  if (wtiles == nullptr) return KV_OK;
  const tc::Geom gm = tc::make_geom(d.in_features, d.out_features, d.M);
  const size_t total = (size_t)d.n_groups * gm.n_tiles * gm.tile_elems();
  const int threads = 256;
  const int blocks = (int)((total + threads - 1) / threads > 148 * 16 ? 148 * 16 : (total + threads - 1) / threads);
  tc::tc_pack_kernel<<<blocks, threads, 0, st>>>(wp_f32, reinterpret_cast<__nv_bfloat16*>(wtiles), d.n_groups,
                                                 d.in_features, d.out_features, d.M, gm);
  KV_CHECK_LAUNCH("tc_pack_kernel");
  return KV_OK;
}

int tc_fwd(const kv_kan_desc_t& d, const float* X, const void* wtiles, const float* fparams,
           const float* bias, const float* mean, const float* rstd, const float* residual,
           const float* pos, float* Y, cudaStream_t st) {
  if (tc_supported(d) != KV_OK) return KV_ERR_UNSUPPORTED;
  tc::FwdParams p;
  p.d = d;
  p.X = X;
  p.wt = reinterpret_cast<const __nv_bfloat16*>(wtiles);
  p.fparams = fparams;
  p.bias = bias;
  p.mean = mean;
  p.rstd = rstd;
  p.residual = residual;
  p.pos = pos;
  p.Y = Y;
  p.gm = tc::make_geom(d.in_features, d.out_features, d.M);
  switch (d.family) {
    case KV_LINEAR: return tc::launch_fwd<KV_LINEAR>(p, st);
    case KV_EFFKAN: return tc::launch_fwd<KV_EFFKAN>(p, st);
    case KV_CHEBY: return tc::launch_fwd<KV_CHEBY>(p, st);
    case KV_FASTKAN: return tc::launch_fwd<KV_FASTKAN>(p, st);
    case KV_SINE: return tc::launch_fwd<KV_SINE>(p, st);
    case KV_FOURIER: return tc::launch_fwd<KV_FOURIER>(p, st);
  }
  return KV_ERR_BAD_FAMILY;
}

// backward tensor-core kernels: not built yet -- the bf16 path uses the fp32 SIMT gradients
int tc_bwd_input(const kv_kan_desc_t&, const float*, const float*, const float*, const void*, const float*,
                 const float*, const float*, float*, float*, float*, cudaStream_t) { return KV_ERR_UNSUPPORTED; }
int tc_bwd_weight(const kv_kan_desc_t&, const float*, const float*, const float*, const float*, const float*,
                  const float*, float*, float*, cudaStream_t) { return KV_ERR_UNSUPPORTED; }

}  // namespace kv
