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
#include <cstdio>
#include <cstdlib>
#include <type_traits>
#include <cuda.h>   // CUtensorMap and the encode entry point's types (the function itself is fetched at run time)
#include "basis.cuh"
#include "internal.cuh"
#include "tc_ptx.cuh"

namespace kv {
namespace tc {

constexpr int BM = 128;              // rows per CTA tile (UMMA M)
constexpr int KU = 16;               // K per unit (UMMA K for bf16)
constexpr int A_UNIT_BYTES = BM * KU * 2;
constexpr int A_UNITS = 18;          // A ring capacity in units: two default-effkan chunks (the producers need ~1/3 of a
                                     // chunk's MMA time, so two slots keep the tensor pipe fed); the memory goes to B
constexpr int A_RING_BYTES = A_UNITS * A_UNIT_BYTES;
constexpr int B_RING_BYTES = 144 * 1024;   // 6 stages of 3 units at BN = 512: the weight stream's round trip
                                           // (commit -> loader -> L2 -> peer relay) needs more than 4 stages in flight
constexpr int NB_MAX = 16;
constexpr int NCH_MAX = A_UNITS;
constexpr int N_PRODUCER_WARPS = 8;
constexpr int NTHREADS = 32 * (2 + N_PRODUCER_WARPS);
constexpr int BAR_BYTES = 1024;
constexpr int DBG_BYTES = 4 * 1024;   // clock stamps of timing experiments (d.reserved & 32): 4 roles x 64 entries x 4 words
constexpr int SMEM_BYTES = A_RING_BYTES + B_RING_BYTES + BAR_BYTES + DBG_BYTES;
constexpr int TMEM_COLS = 512;
constexpr int MAX_M_TC = KV_MAX_M;  // forward / weight gradient: any M (chunks are walked in parts of <= 9 bases)
constexpr int MAX_M_TC_DX = 32;    // input gradient: one 16-feature chunk = 16 M <= 512 TMEM columns

// K-order of the units inside a 16-feature chunk
enum { LAYOUT_BASIS_MAJOR = 0,  // unit m = 16 features x basis m            (every family)
       LAYOUT_EFFKAN8 = 1 };    // units 0..7 = feature pair (2q, 2q+1) x the 8 B-splines, unit 8 = 16 x SiLU

struct Geom {
  int n_chunks;  // ceil(in / 16)
  int n_units;   // n_chunks * M
  int n_tiles;   // output tiles per group
  int BN;        // outputs per tile (multiple of 16, <= 512)
  int S;         // units per B stage (one bulk copy, one barrier round trip)
  int NB;        // B ring slots (stages)
  int NCH;       // A ring slots (each holds one PART of a chunk: Mp units)
  int npart;     // parts per chunk: the bases of a chunk are produced / consumed in npart slices of <= Mp bases,
  int Mp;        //   so that the ring always holds >= 2 slots (M > 13) and layers with M > 27 fit at all
  int effkan8;   // host-side candidate for LAYOUT_EFFKAN8 (the device-side grid probe decides)
  int loopC;     // the K walk of the forward kernels: loopC chunk groups of loopM units, cut into npart parts of Mp.
  int loopM;     //   KAN families: (n_chunks, M).  nn.Linear (M = 1): ONE group of n_chunks units in parts of 6
                 //   consecutive 16-feature chunks, so a ring slot / barrier round trip carries 6 MMAs instead of 1
  __host__ __device__ size_t tile_elems() const { return (size_t)n_units * BN * KU; }
};

// max_bn: 512 for the one-tile-per-CTA-pair kernels (forward KV_FWD_BN=512, weight gradient), 256 for the
// persistent forward kernel (two accumulators in TMEM)
__host__ __device__ inline Geom make_geom(const kv_kan_desc_t& d, int max_bn = 512) {
  Geom g;
  g.n_chunks = (d.in_features + KU - 1) / KU;
  g.n_units = g.n_chunks * d.M;
  const int out16 = (d.out_features + 15) / 16 * 16;
  g.n_tiles = (out16 + max_bn - 1) / max_bn;
  g.BN = ((out16 + g.n_tiles - 1) / g.n_tiles + 15) / 16 * 16;
  const int half_unit = g.BN * KU;  // bytes; each CTA of the pair stages half of a unit: BN/2 rows
  if (d.family == KV_LINEAR && d.M == 1) {
    g.loopC = 1;
    g.loopM = g.n_chunks;
    g.Mp = g.n_chunks < 6 ? g.n_chunks : 6;
    g.npart = (g.n_chunks + g.Mp - 1) / g.Mp;
  } else {
    g.loopC = g.n_chunks;
    g.loopM = d.M;
    g.npart = (d.M + 8) / 9;   // parts of <= 9 bases: at least two ring slots
    g.Mp = (d.M + g.npart - 1) / g.npart;
  }
  g.S = 24576 / half_unit;
  g.S = g.S < 1 ? 1 : (g.S > 8 ? 8 : g.S);
  // parts of equal size that split into whole stages of 3 units: the MMA thread's unrolled issue loop applies
  if (g.loopM % g.Mp == 0 && g.Mp % 3 == 0 && g.S >= 3) g.S = 3;
  int nb = B_RING_BYTES / (g.S * half_unit);
  g.NB = nb > NB_MAX ? NB_MAX : nb;
  g.NCH = A_UNITS / g.Mp;
  g.effkan8 = (d.family == KV_EFFKAN && d.spline_order == 3 && d.grid_size == 5 &&
               (d.flags & KV_FLAG_UNIFORM_GRID)) ? 1 : 0;
  return g;
}

// The layout the kernels of one group use: LAYOUT_EFFKAN8 iff the layer is the default-shaped
// effkan AND the grid probe (written by kv_kan_pack behind the knots) found one uniform grid.
__device__ __forceinline__ int group_layout(const kv_kan_desc_t& d, const Geom& gm, const float* fparams, int g) {
  if (!gm.effkan8 || fparams == nullptr) return LAYOUT_BASIS_MAJOR;
  const float* probe = fparams + (size_t)g * d.fparam_stride +
                       (size_t)d.in_features * (d.grid_size + 2 * d.spline_order + 1);
  return probe[0] == 1.0f ? LAYOUT_EFFKAN8 : LAYOUT_BASIS_MAJOR;
}

// (feature i, basis m) of K element (unit q of chunk c, k-half h, element e)
__device__ __forceinline__ void unit_to_im(int layout, int M, int c, int q, int h, int e, int& i, int& m) {
  if (layout == LAYOUT_EFFKAN8 && q < 8) {
    i = c * KU + 2 * q + h;
    m = e;
  } else {
    i = c * KU + h * 8 + e;
    m = (layout == LAYOUT_EFFKAN8) ? 8 : q;
  }
}

// ----------------------------------------------------------------------------------------
// weight pre-tiling: module tensors (pointer table of kv_kan_pack)  ->  wtiles bf16
//   [g][n_tile][CTA rank][unit][sub-MMA j][k-half h][row group][8 rows][8 elements]
// = for every unit and each CTA of the pair the no-swizzle K-major operand image of its half of the
//   rows of each N<=256 sub-MMA (core matrix = 8 rows x 16 B).  One thread = one core-matrix row (16 B).
// d is the tensor-core view of the layer, d0 the caller's descriptor: view group g / output o is
// group g * xs + o / d0.out, output o % d0.out of the pointer table (xs = merged q/k/v maps, api.cu tc_view).
// ----------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t pack_bf16_pair(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// the 8 K elements e = 0..7 of (chunk c, unit q, k-half h) for output o0 of group g0 of the pointer table
__device__ __forceinline__ void load_w8(const kv_kan_desc_t& d0, const uint64_t* __restrict__ tbl, int g0, int o0,
                                        int layout, int c, int q, int h, float (&v)[8]) {
  const int in = d0.in_features;
#pragma unroll
  for (int e = 0; e < 8; ++e) v[e] = 0.0f;
  if (o0 >= d0.out_features) return;
  if (layout == LAYOUT_EFFKAN8) {
    if (q < 8) {  // the 8 B-spline coefficients of one feature: 32 contiguous bytes of spline_weight
      const int i = c * KU + 2 * q + h;
      if (i >= in) return;
      const float* w = slot(tbl, g0, 1) + ((size_t)o0 * in + i) * 8;
      const float* sc = slot(tbl, g0, 2);
      const float sv = sc ? sc[(size_t)o0 * in + i] : 1.0f;  // scaled_spline_weight, models/effkan.py:166-172
      if ((reinterpret_cast<uintptr_t>(w) & 15) == 0) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(w)), b = __ldg(reinterpret_cast<const float4*>(w) + 1);
        v[0] = a.x * sv; v[1] = a.y * sv; v[2] = a.z * sv; v[3] = a.w * sv;
        v[4] = b.x * sv; v[5] = b.y * sv; v[6] = b.z * sv; v[7] = b.w * sv;
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = w[e] * sv;
      }
    } else {      // SiLU branch: 8 consecutive features of base_weight
      const int i0 = c * KU + h * 8;
      const float* w = slot(tbl, g0, 0) + (size_t)o0 * in + i0;
      if (i0 + 8 <= in && (reinterpret_cast<uintptr_t>(w) & 15) == 0) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(w)), b = __ldg(reinterpret_cast<const float4*>(w) + 1);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e)
          if (i0 + e < in) v[e] = w[e];
      }
    }
    return;
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    int i, m;
    unit_to_im(layout, d0.M, c, q, h, e, i, m);
    if (i < in) v[e] = load_packed_weight(d0, tbl, g0, o0, i, m);
  }
}

__global__ void tc_pack_kernel(const kv_kan_desc_t d, const kv_kan_desc_t d0, const uint64_t* __restrict__ tbl,
                               const float* __restrict__ fparams, __nv_bfloat16* __restrict__ wt, Geom gm) {
  const size_t total8 = (size_t)d.n_groups * gm.n_tiles * gm.tile_elems() / 8;
  const int M = d.M, xs = d.n_groups == d0.n_groups ? 1 : d0.x_share;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total8;
       idx += (size_t)gridDim.x * blockDim.x) {
    // [g][n_tile][CTA rank][unit][sub-MMA j][k-half h][row group][8 rows]  (x 8 elements)
    size_t r = idx;
    const int r8 = r % 8; r /= 8;
    int pr = r % (gm.BN / 8); r /= (gm.BN / 8);   // core-matrix position inside this rank's half unit
    const int u = r % gm.n_units; r /= gm.n_units;
    const int rank = r % 2; r /= 2;
    const int nt = r % gm.n_tiles; r /= gm.n_tiles;
    const int g = (int)r;
    const int n1 = gm.BN > 256 ? 256 : gm.BN, n2 = gm.BN - n1;
    int j = 0, nj = n1;
    if (pr >= n1 / 8) { pr -= n1 / 8; j = 1; nj = n2; }
    const int h = pr / (nj / 16), rgl = pr % (nj / 16);
    const int o = nt * gm.BN + (j ? n1 : 0) + rank * (nj / 2) + rgl * 8 + r8;
    const int c = u / M, q = u - c * M;
    float v[8];
    if (o < d.out_features) {
      load_w8(d0, tbl, g * xs + o / d0.out_features, o % d0.out_features, group_layout(d, gm, fparams, g), c, q, h, v);
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] = 0.0f;
    }
    *reinterpret_cast<uint4*>(wt + idx * 8) = make_uint4(pack_bf16_pair(v[0], v[1]), pack_bf16_pair(v[2], v[3]),
                                                         pack_bf16_pair(v[4], v[5]), pack_bf16_pair(v[6], v[7]));
  }
}

// Linear input gradient: dX = dY . W is the forward of the TRANSPOSED layer (in' = out, out' = in), so its
// operand image is the forward image of W^T.  dt = descriptor of that transposed layer (geometry gmt), d / d0 as
// above.  Element (output row o' = input feature of the layer, K element i' = output of the layer).
__global__ void tc_pack_linear_t_kernel(const kv_kan_desc_t dt, const kv_kan_desc_t d, const kv_kan_desc_t d0,
                                        const uint64_t* __restrict__ tbl, __nv_bfloat16* __restrict__ wt, Geom gmt) {
  const size_t total8 = (size_t)dt.n_groups * gmt.n_tiles * gmt.tile_elems() / 8;
  const int xs = d.n_groups == d0.n_groups ? 1 : d0.x_share;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total8;
       idx += (size_t)gridDim.x * blockDim.x) {
    size_t r = idx;
    const int r8 = r % 8; r /= 8;
    int pr = r % (gmt.BN / 8); r /= (gmt.BN / 8);
    const int u = r % gmt.n_units; r /= gmt.n_units;
    const int rank = r % 2; r /= 2;
    const int nt = r % gmt.n_tiles; r /= gmt.n_tiles;
    const int g = (int)r;
    const int n1 = gmt.BN > 256 ? 256 : gmt.BN, n2 = gmt.BN - n1;
    int j = 0, nj = n1;
    if (pr >= n1 / 8) { pr -= n1 / 8; j = 1; nj = n2; }
    const int h = pr / (nj / 16), rgl = pr % (nj / 16);
    const int op = nt * gmt.BN + (j ? n1 : 0) + rank * (nj / 2) + rgl * 8 + r8;  // o' : input feature of the layer
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int ip = u * KU + h * 8 + e;  // i' : output of the (merged) layer
      v[e] = 0.0f;
      if (op < d.in_features && ip < d.out_features)
        v[e] = load_packed_weight(d0, tbl, g * xs + ip / d0.out_features, ip % d0.out_features, op, 0);
    }
    *reinterpret_cast<uint4*>(wt + idx * 8) = make_uint4(pack_bf16_pair(v[0], v[1]), pack_bf16_pair(v[2], v[3]),
                                                         pack_bf16_pair(v[4], v[5]), pack_bf16_pair(v[6], v[7]));
  }
}

// scattered bf16 stores of Phi_m(x) into a chunk slot (basis-major layout): unit m at m * A_UNIT_BYTES
struct UnitSink {
  static constexpr bool kDense = true;
  uint8_t* base;  // chunk slot + byte offset of (row, feature) inside a unit
  __device__ __forceinline__ void put(int m, float v) {
    *reinterpret_cast<__nv_bfloat16*>(base + m * A_UNIT_BYTES) = __float2bfloat16_rn(v);
  }
};

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float fast_silu(float x) { return __fdividef(x, 1.0f + __expf(-x)); }
// (a == b) ? x : y as setp + selp (the ternary chains of the gradient epilogue otherwise compile to branches)
__device__ __forceinline__ float sel_eq(int a, int b, float x, float y) {
  float r;
  asm("{ .reg .pred p; setp.eq.s32 p, %1, %2; selp.f32 %0, %3, %4, p; }" : "=f"(r) : "r"(a), "r"(b), "f"(x), "f"(y));
  return r;
}
__device__ __forceinline__ uint64_t shl64(uint64_t v, int n) {  // PTX shifts clamp: n >= 64 (or "negative") -> 0
  uint64_t r;
  asm("shl.b64 %0, %1, %2;" : "=l"(r) : "l"(v), "r"(n));
  return r;
}
__device__ __forceinline__ uint64_t shr64(uint64_t v, int n) {
  uint64_t r;
  asm("shr.u64 %0, %1, %2;" : "=l"(r) : "l"(v), "r"(n));
  return r;
}
__device__ __forceinline__ void sts128(uint8_t* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  *reinterpret_cast<uint4*>(p) = make_uint4(a, b, c, d);
}

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
  const float* in_mask;  // Linear only, nullable: the input is X * (in_mask > 0), same addressing as X
  // q/k/v layer of MSA (attention.py:196-197), merged view (group = head, out = 3 * d_head): instead of fp32 Y the
  // epilogue writes the bf16 attention image of attn_tc5.cu, img[b][h][part][n8][d8][n%8][d%8], zero padding included
  __nv_bfloat16* attn_img;
  int attn_N;            // tokens per sample (rows t = b * attn_N + n)
};

// 8 consecutive features of one row (zero beyond `in` / beyond T)
__device__ __forceinline__ void load_x8(const float* __restrict__ xrow, int i0, int in, bool row_ok, bool vec,
                                        float (&x)[8]) {
  if (row_ok && vec && i0 + 8 <= in) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(xrow + i0));
    const float4 b = __ldg(reinterpret_cast<const float4*>(xrow + i0 + 4));
    x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w;
    x[4] = b.x; x[5] = b.y; x[6] = b.z; x[7] = b.w;
  } else {
#pragma unroll
    for (int q = 0; q < 8; ++q) x[q] = (row_ok && i0 + q < in) ? __ldg(xrow + i0 + q) : 0.0f;
  }
}

// ---- A producers -------------------------------------------------------------------------
// Every producer thread owns one row r of the tile and 8 consecutive features (k-half wg) of the
// chunk; `base` = chunk slot + the row's 16-byte lane inside a core-matrix column.

// Default effkan (G=5, k=3) on a uniform grid, LAYOUT_EFFKAN8: the 8 B-spline values of ONE input are
// one 16-byte store.  The 4 non-zero cubic pieces are packed to bf16 and shifted into place.
__device__ __forceinline__ void effkan8_chunk(uint8_t* base, int wg, const float (&x)[8], float t0, float inv_h) {
  float sv[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const float xv = x[q];
    const float u = (xv - t0) * inv_h;
    const float jf = floorf(u);
    const float s = u - jf;
    const int j = (int)jf;
    const bool ok = (u >= 0.0f) && (u < 11.0f);  // x in [t_0, t_11): models/effkan.py:115
    const float s2 = s * s, s3 = s2 * s, om = 1.0f - s;
    const float w0 = om * om * om * (1.0f / 6.0f);
    const float w1 = fmaf(0.5f, s3, fmaf(-1.0f, s2, 2.0f / 3.0f));
    const float w2 = fmaf(-0.5f, s3, fmaf(0.5f, s2, fmaf(0.5f, s, 1.0f / 6.0f)));
    const float w3 = s3 * (1.0f / 6.0f);
    uint64_t W = (uint64_t)pack_bf16(w0, w1) | ((uint64_t)pack_bf16(w2, w3) << 32);
    if (!ok) W = 0;
    const int b = (j - 3) * 16;  // basis a = j-3+r sits at 16-bit lane a of the 128-bit vector
    const uint64_t lo = shl64(W, b) | shr64(W, -b);
    const uint64_t hi = shl64(W, b - 64) | shr64(W, 64 - b);
    sts128(base + (wg * 4 + (q >> 1)) * A_UNIT_BYTES + (q & 1) * 2048, (uint32_t)lo, (uint32_t)(lo >> 32),
           (uint32_t)hi, (uint32_t)(hi >> 32));
    sv[q] = fast_silu(xv);
  }
  sts128(base + 8 * A_UNIT_BYTES + wg * 2048, pack_bf16(sv[0], sv[1]), pack_bf16(sv[2], sv[3]),
         pack_bf16(sv[4], sv[5]), pack_bf16(sv[6], sv[7]));
}

// The same for a thread that owns 4 consecutive features 4*qg .. 4*qg+3 of the chunk (16 producer warps: twice the
// warps per scheduler to hide the latency of the dependent cubic / shift / pack chains).
__device__ __forceinline__ void effkan8_chunk4(uint8_t* base, int qg, const float (&x)[4], float t0, float inv_h) {
  float sv[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float xv = x[q];
    const float u = (xv - t0) * inv_h;
    const float jf = floorf(u);
    const float s = u - jf;
    const int j = (int)jf;
    const bool ok = (u >= 0.0f) && (u < 11.0f);  // x in [t_0, t_11): models/effkan.py:115
    const float s2 = s * s, s3 = s2 * s, om = 1.0f - s;
    const float w0 = om * om * om * (1.0f / 6.0f);
    const float w1 = fmaf(0.5f, s3, fmaf(-1.0f, s2, 2.0f / 3.0f));
    const float w2 = fmaf(-0.5f, s3, fmaf(0.5f, s2, fmaf(0.5f, s, 1.0f / 6.0f)));
    const float w3 = s3 * (1.0f / 6.0f);
    uint64_t W = (uint64_t)pack_bf16(w0, w1) | ((uint64_t)pack_bf16(w2, w3) << 32);
    if (!ok) W = 0;
    const int b = (j - 3) * 16;
    const uint64_t lo = shl64(W, b) | shr64(W, -b);
    const uint64_t hi = shl64(W, b - 64) | shr64(W, 64 - b);
    // feature f = 4*qg + q: unit f / 2, k-half f % 2
    sts128(base + (qg * 2 + (q >> 1)) * A_UNIT_BYTES + (q & 1) * 2048, (uint32_t)lo, (uint32_t)(lo >> 32),
           (uint32_t)hi, (uint32_t)(hi >> 32));
    sv[q] = fast_silu(xv);
  }
  // SiLU unit: k-half (4*qg) / 8, elements (4*qg) % 8 .. +3 of the row's 16-byte vector
  *reinterpret_cast<uint2*>(base + 8 * A_UNIT_BYTES + (qg >> 1) * 2048 + (qg & 1) * 8) =
      make_uint2(pack_bf16(sv[0], sv[1]), pack_bf16(sv[2], sv[3]));
}

// Cubic splines (spline_order 3, the only order the reference's models build): the 4 B-splines that are non-zero at x
// for feature i, N[r] = B_{jb + r, 3}(x), jb = j - 3 for the knot span j that holds x (far negative, N = 0, when x lies
// outside [t_0, t_last) or i is beyond the layer's features).
//   c.uniform: one uniform knot vector t0 + j h (any grid_size) -- the four cubic pieces in closed form, no knot loads,
//   no divisions;
//   otherwise: span search on the feature's own knots (a guess from the grid ends, corrected by a short walk), the 6
//   knots around the span and the triangular de Boor scheme (6 divisions) -- every term the reference's full recursion
//   (models/effkan.py:99-132) forms for a non-zero basis, none of the zero ones.
// (An earlier version unrolled the triangle up to KV_MAX_SPLINE_ORDER = 7 with run-time order checks: ~560 instructions
// per feature, a producer loop that no longer fitted the instruction cache -- 22 % of the stall samples were no_inst.)
__device__ __forceinline__ void effkan_local_basis3(const BasisCtx& c, int i, float xv, float (&N)[4], int& jb) {
  const int nk = c.G + 7;
  N[0] = N[1] = N[2] = N[3] = 0.0f;
  jb = -(1 << 20);
  if (i >= c.in) return;
  if (c.uniform) {
    const float sx = (xv - c.t0) * c.inv_h;
    if (!(sx >= 0.0f) || !(sx < (float)(nk - 1))) return;
    const int j = min((int)sx, nk - 2);
    const float u = sx - (float)j;
    const float s2 = u * u, s3 = s2 * u, om = 1.0f - u;
    N[0] = om * om * om * (1.0f / 6.0f);
    N[1] = fmaf(0.5f, s3, fmaf(-1.0f, s2, 2.0f / 3.0f));
    N[2] = fmaf(-0.5f, s3, fmaf(0.5f, s2, fmaf(0.5f, u, 1.0f / 6.0f)));
    N[3] = s3 * (1.0f / 6.0f);
    jb = j - 3;
    return;
  }
  const float* t = c.fp + (size_t)i * nk;
  const int j = bspline_find_span(t, nk, xv);
  if (j < 0) return;
  float left[4], right[4];
#pragma unroll
  for (int pp = 1; pp <= 3; ++pp) {
    left[pp] = xv - t[max(j + 1 - pp, 0)];
    right[pp] = t[min(j + pp, nk - 1)] - xv;
  }
  N[0] = 1.0f;
#pragma unroll
  for (int pp = 1; pp <= 3; ++pp) {
    float saved = 0.0f;
#pragma unroll
    for (int r = 0; r < pp; ++r) {
      const float den = right[r + 1] + left[pp - r];
      const float temp = den != 0.0f ? __fdividef(N[r], den) : 0.0f;
      N[r] = saved + right[r + 1] * temp;
      saved = left[pp - r] * temp;
    }
    N[pp] = saved;
  }
  jb = j - 3;
}

// Basis-major layout, basis index outermost: 8 features x one m = one 16-byte store.
// basis_octet: the values of 8 consecutive features (one 16-byte vector per basis function m) for the bases
// m0 <= m < m0 + Mt, written at dst + (m - m0) * stride.
struct StrideSink {  // per-feature scattered bf16 stores (generic Cox-de Boor path)
  static constexpr bool kDense = true;
  uint8_t* base;
  int stride, m0, Mt;
  __device__ __forceinline__ void put(int m, float v) {
    if (m >= m0 && m < m0 + Mt) *reinterpret_cast<__nv_bfloat16*>(base + (m - m0) * stride) = __float2bfloat16_rn(v);
  }
};

// effkan on a grid other than the default one: the non-zero splines of the 8 features of an octet (once per chunk) ...
struct EffkanOctet {
  float N[8][4];
  int jb[8];   // index of the first non-zero basis (span - 3); far negative when x lies outside [t_0, t_last)
};
__device__ __forceinline__ void effkan_octet_basis(const BasisCtx& c, int i0, const float (&x)[8], EffkanOctet& e) {
  if (c.k != 3) return;   // other orders: effkan_scatter_part takes the per-feature Cox-de Boor route
#pragma unroll
  for (int q = 0; q < 8; ++q) effkan_local_basis3(c, i0 + q, x[q], e.N[q], e.jb[q]);
}
// ... and the rows m0 <= m < m0 + Mt of its operand image (row m = 8 features x basis m = 16 bytes, at dst + (m - m0) *
// stride): M 16-byte stores of zeros, then the 4 values of each feature as 2-byte stores into the rows of its span's
// window (same thread, same addresses: program order), then the SiLU row.  ~50 stores per octet and part instead of a
// select chain per (basis, feature): the producers of grid 10 / 20 layers were issue-bound on those chains.
__device__ __forceinline__ void effkan_scatter_part(uint8_t* dst, int stride, int m0, int Mt, const BasisCtx& c, int i0,
                                                    const float (&x)[8], const EffkanOctet& e) {
  const int nb = c.G + c.k;
  if (c.k != 3) {
    // any other spline order: the full recursion per feature, dense 2-byte stores (slow, never built by the reference)
#pragma unroll 1
    for (int q = 0; q < 8; ++q) {
      StrideSink sink{dst + q * 2, stride, m0, Mt};
      if (i0 + q < c.in) {
        basis_eval<KV_EFFKAN>(c, i0 + q, x[q], x[q], sink);
      } else {
        for (int m = m0; m < m0 + Mt; ++m) sink.put(m, 0.0f);
      }
    }
    return;
  }
  const int mlo = max(m0, 0), mhi = min(nb, m0 + Mt);
  for (int m = mlo; m < mhi; ++m) sts128(dst + (m - m0) * stride, 0u, 0u, 0u, 0u);
#pragma unroll
  for (int q = 0; q < 8; ++q) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int m = e.jb[q] + r;
      if (m >= mlo && m < mhi)
        *reinterpret_cast<__nv_bfloat16*>(dst + (m - m0) * stride + q * 2) = __float2bfloat16_rn(e.N[q][r]);
    }
  }
  if (nb >= m0 && nb < m0 + Mt) {
    float v[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) v[q] = (i0 + q < c.in) ? fast_silu(x[q]) : 0.0f;
    sts128(dst + (nb - m0) * stride, pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]),
           pack_bf16(v[6], v[7]));
  }
}

template <int F>
__device__ __forceinline__ void basis_octet(uint8_t* dst, int stride, int m0, int Mt, const float (&x)[8],
                                            const BasisCtx& c, int i0, float mu, float rs) {
  const int M = c.M;
  float v[8];
  auto emit = [&](int m) {
    if (m >= m0 && m < m0 + Mt)
      sts128(dst + (m - m0) * stride, pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]),
             pack_bf16(v[6], v[7]));
  };
  if (F == KV_EFFKAN) {
    // B-splines on arbitrary per-feature knots (grid_size != 5, or any grid after update_grid), models/effkan.py:99-132.
    // At most k + 1 bases are non-zero at x: those of the knot span j that holds it.  Per feature: find the span (a
    // guess from the grid ends, corrected by a short walk), load the 2k knots around it and run the triangular
    // de Boor scheme (k (k + 1) / 2 divisions) -- every term the reference's full recursion forms for a non-zero
    // basis, none of the zero ones.  The rows of the part are then written as zeros (one 16-byte store per basis index)
    // and the k + 1 values of each feature scattered over them (effkan_scatter_part).
    EffkanOctet e;
    effkan_octet_basis(c, i0, x, e);
    effkan_scatter_part(dst, stride, m0, Mt, c, i0, x, e);
    return;
  }
  if (F == KV_LINEAR) {
#pragma unroll
    for (int q = 0; q < 8; ++q) v[q] = x[q];
    emit(0);
  } else if (F == KV_CHEBY) {
    float t[8], p0[8], p1[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      t[q] = 1.0f - __fdividef(2.0f, 1.0f + __expf(2.0f * x[q]));  // tanh
      p0[q] = 1.0f;
      p1[q] = t[q];
      v[q] = 1.0f;
    }
    emit(0);
    if (M > 1) {
#pragma unroll
      for (int q = 0; q < 8; ++q) v[q] = t[q];
      emit(1);
    }
    for (int m = 2; m < M; ++m) {
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        v[q] = 2.0f * t[q] * p1[q] - p0[q];
        p0[q] = p1[q];
        p1[q] = v[q];
      }
      emit(m);
    }
  } else if (F == KV_FASTKAN) {
    float xn[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int i = min(i0 + q, c.in - 1);
      xn[q] = fastkan_xn(c, i, x[q], mu, rs);
    }
    const float* gr = c.fp + 2 * c.in;
    const float inv_den = 1.0f / c.rbf_denom;
    for (int m = max(m0, 0); m < min(c.G, m0 + Mt); ++m) {   // only the bases of this part
      const float gm = gr[m];
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float z = (xn[q] - gm) * inv_den;
        v[q] = __expf(-(z * z));
      }
      emit(m);
    }
    if (!(c.flags & KV_FLAG_NO_BASE) && c.G >= m0 && c.G < m0 + Mt) {
#pragma unroll
      for (int q = 0; q < 8; ++q) v[q] = fast_silu(x[q]);
      emit(c.G);
    }
  } else if (F == KV_SINE) {
    const float* freq = c.fp;
    const float* ph = c.fp + c.G;
    for (int m = max(m0, 0); m < min(c.G, m0 + Mt); ++m) {
      const float fr = freq[m];
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const int i = min(i0 + q, c.in - 1);
        v[q] = __sinf(fmaf(x[q], fr, ph[(size_t)i * c.G + m]));
      }
      emit(m);
    }
  } else if (F == KV_FOURIER) {
    // bases m < G: cos((m+1) x), m >= G: sin((m-G+1) x).  Inside a run of consecutive harmonics the angle-addition
    // recurrence steps by x; each run (a part may start anywhere, and crosses m = G at most once) starts from a
    // direct evaluation.
    float c1[8], s1[8], ck[8], sk[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) __sincosf(x[q], &s1[q], &c1[q]);
    int m = max(m0, 0);
    const int mend = min(2 * c.G, m0 + Mt);
    while (m < mend) {
      const bool is_sin = m >= c.G;
      const int run_end = min(mend, is_sin ? 2 * c.G : c.G);
      const float kf = (float)((is_sin ? m - c.G : m) + 1);
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        if (kf == 1.0f) { ck[q] = c1[q]; sk[q] = s1[q]; }
        else __sincosf(kf * x[q], &sk[q], &ck[q]);
      }
      for (; m < run_end; ++m) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          v[q] = is_sin ? sk[q] : ck[q];
          const float nc = ck[q] * c1[q] - sk[q] * s1[q];
          const float ns = sk[q] * c1[q] + ck[q] * s1[q];
          ck[q] = nc;
          sk[q] = ns;
        }
        emit(m);
      }
    }
  }
}

// out-of-line copy for kernels whose hot path must not inherit the register pressure of the arbitrary-knot branch
// (the persistent forward kernel runs 448 threads: 128 registers per thread)
template <int F>
__device__ __noinline__ void basis_octet_call(uint8_t* dst, int stride, int m0, int Mt, const float (&x)[8],
                                              const BasisCtx& c, int i0, float mu, float rs) {
  basis_octet<F>(dst, stride, m0, Mt, x, c, i0, mu, rs);
}

template <int F>
__device__ __forceinline__ void basis_major_chunk(uint8_t* base, int wg, const float (&x)[8], const BasisCtx& c,
                                                  int i0, float mu, float rs) {
  basis_octet<F>(base + wg * 2048, A_UNIT_BYTES, 0, c.M, x, c, i0, mu, rs);
}

// NPW = producer / epilogue warps: 8 (a thread owns 8 features of a chunk) or 16 (4 features; effkan only)
// nn.Linear A producer of one tile (both forward kernels, 256 producer threads = 8 warps): the operand is the bf16
// image of X itself.  A ring slot carries Mp (<= 6) consecutive 16-feature chunks of the tile's 128 rows = 16 row
// groups x Mp chunks of (8 rows x 16 features); a warp instruction moves one such block: lane = (row of the group,
// float4 of the chunk), i.e. 8 rows x 64 contiguous bytes on the load side and the 8-row core matrices of the operand
// layout on the store side (one 8-byte store per lane, two shared-memory wavefronts).  The first version gave each
// thread a row (32 cache lines, 16 bytes each, per warp instruction): clock stamps showed the producers at 35 k cycles
// per 256 x 256 x 768 tile (7 k of MMA work) and the epilogue's stores queueing behind them in the load / store unit.
// (Loads are not prefetched across slots: three ring slots and eight warps cover the latency, and the registers go to
// the eight epilogue warps of the persistent kernel.)  mtile != nullptr: the ReLU gate of a
// transposed layer (input gradient of Linear + ReLU), applied while converting.  xtile / mtile point at (row t0 of
// the tile, first feature of the slice).
__device__ __forceinline__ void linear_produce_tile(const Geom& gm, const float* __restrict__ xtile,
                                                    const float* __restrict__ mtile, int rows_valid, int ldx, bool vec,
                                                    bool mvec, int in, uint8_t* sA, uint32_t bar0, int NCH, int ptid,
                                                    int lane, int& slot_i) {
  constexpr int LC = 6;          // >= Geom::Mp of a Linear layer
  constexpr int NV = 2 * LC;     // blocks per warp and slot: 16 row groups x 6 chunks / 8 warps
  const int Mp = gm.Mp, n_chunks = gm.n_chunks;
  const int warp = ptid >> 5;
  const int rl = lane >> 2, f4 = lane & 3;               // row of the 8-row group, float4 of the 16-feature chunk
  // block j of this warp: chunk u = j % LC of the slot, row group 2 * warp + j / LC
  auto load_slot = [&](int c0, float4 (&v)[NV]) {
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int u = j % LC, row = (2 * warp + j / LC) * 8 + rl;
      const int i0 = (c0 + u) * KU + f4 * 4;               // first feature of the float4
      v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (u >= Mp || c0 + u >= n_chunks || row >= rows_valid || i0 >= in) continue;
      const float* src = xtile + (size_t)row * ldx + i0;
      if (vec && i0 + 4 <= in) {
        v[j] = __ldg(reinterpret_cast<const float4*>(src));
      } else {
        v[j].x = __ldg(src);
        if (i0 + 1 < in) v[j].y = __ldg(src + 1);
        if (i0 + 2 < in) v[j].z = __ldg(src + 2);
        if (i0 + 3 < in) v[j].w = __ldg(src + 3);
      }
      if (mtile != nullptr) {
        const float* ms = mtile + (size_t)row * ldx + i0;
        float4 mk = make_float4(0.f, 0.f, 0.f, 0.f);
        if (mvec && i0 + 4 <= in) {
          mk = __ldg(reinterpret_cast<const float4*>(ms));
        } else {
          mk.x = __ldg(ms);
          if (i0 + 1 < in) mk.y = __ldg(ms + 1);
          if (i0 + 2 < in) mk.z = __ldg(ms + 2);
          if (i0 + 3 < in) mk.w = __ldg(ms + 3);
        }
        v[j].x = mk.x > 0.0f ? v[j].x : 0.0f;
        v[j].y = mk.y > 0.0f ? v[j].y : 0.0f;
        v[j].z = mk.z > 0.0f ? v[j].z : 0.0f;
        v[j].w = mk.w > 0.0f ? v[j].w : 0.0f;
      }
    }
  };
  float4 cur[NV];
  // byte offset of this lane inside a unit: k-half f4 / 2, row group, row, half of the 16-byte row
  const int lane_off = (f4 >> 1) * 2048 + rl * 16 + (f4 & 1) * 8;
  for (int pt = 0; pt < gm.npart; ++pt, ++slot_i) {
    load_slot(pt * Mp, cur);
    const int sc = slot_i % NCH;
    const int Mt = min(Mp, n_chunks - pt * Mp);
    ptx::mbar_wait(bar0 + 8u * (NCH_MAX + sc), ((slot_i / NCH) & 1) ^ 1);   // emptyA(sc)
    uint8_t* base = sA + sc * Mp * A_UNIT_BYTES + lane_off;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int u = j % LC, rg = 2 * warp + j / LC;
      if (u < Mt)
        *reinterpret_cast<uint2*>(base + u * A_UNIT_BYTES + rg * 128) =
            make_uint2(pack_bf16(cur[j].x, cur[j].y), pack_bf16(cur[j].z, cur[j].w));
    }
    ptx::fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) ptx::mbar_arrive_cluster(ptx::mapa(bar0 + 8u * sc, 0));  // the leader's fullA(sc)
  }
}

template <int F, int NPW>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(32 * (2 + NPW), 1) kan_tc_fwd_kernel(const FwdParams p) {
  constexpr int NTH = 32 * (2 + NPW);
  constexpr int FPT = 64 / NPW;  // features per producer thread
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;
  uint8_t* sB = smem + A_RING_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + A_RING_BYTES + B_RING_BYTES);
  // barrier map: fullA[NCH_MAX], emptyA[NCH_MAX], fullB[NB_MAX], emptyB[NB_MAX], tmem_full
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto fullA = [&](int s) { return bar0 + 8u * s; };
  auto emptyA = [&](int s) { return bar0 + 8u * (NCH_MAX + s); };
  auto fullB = [&](int s) { return bar0 + 8u * (2 * NCH_MAX + s); };
  auto emptyB = [&](int s) { return bar0 + 8u * (2 * NCH_MAX + NB_MAX + s); };
  const uint32_t tmem_full = bar0 + 8u * (2 * NCH_MAX + 2 * NB_MAX);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * NCH_MAX + 2 * NB_MAX + 1);

  const kv_kan_desc_t& d = p.d;
  const Geom gm = p.gm;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();  // 0 = leader of the CTA pair
  const int n_tile = (blockIdx.x >> 1) % gm.n_tiles;
  const int m_tile = ((blockIdx.x >> 1) / gm.n_tiles) * 2 + (int)rank;
  const int g = blockIdx.z;
  const int M = d.M;
  const int n_units = gm.n_units;
  const int BN = gm.BN, NB = gm.NB, NCH = gm.NCH;
  const uint32_t b_unit_bytes = (uint32_t)BN * KU;  // this CTA's half of the B unit: BN/2 rows x 16 x 2 B
  const int t0 = m_tile * BM;

  if (warp == 1 && lane == 0) {
    for (int s = 0; s < NCH_MAX; ++s) {
      ptx::mbar_init(fullA(s), 2 * NPW);  // (leader) local + peer producer warps
      ptx::mbar_init(emptyA(s), 1);
    }
    for (int s = 0; s < NB_MAX; ++s) {
      ptx::mbar_init(fullB(s), rank == 0 ? 2 : 1);  // leader: own bulk copy + the peer's "my half landed"
      ptx::mbar_init(emptyB(s), 1);
    }
    ptx::mbar_init(tmem_full, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 0) {
    ptx::tmem_alloc2(ptx::smem_u32(tmem_slot), TMEM_COLS);
    ptx::tmem_relinquish2();
  }
  ptx::tc_fence_before();
  ptx::cluster_sync_all();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // timing experiments only (d.reserved & 32): first tile stamps clock64 into the head of Y
  const bool dbg = (d.reserved & 32) && (blockIdx.x >> 1) == 0 && blockIdx.z == 0;
  uint32_t* dbuf = reinterpret_cast<uint32_t*>(smem + A_RING_BYTES + B_RING_BYTES + BAR_BYTES);
  auto stamp = [&](int role, int u, int k) {
    if (dbg && u < 64) dbuf[(role * 64 + u) * 4 + k] = (uint32_t)clock64();
  };

  if (threadIdx.x == 64) stamp(2, 61, 1);
  const int S = gm.S;
  const int n_stages = (n_units + S - 1) / S;
  if (warp == 0) {
    // ================= B loader: this CTA's half of S units per bulk copy =================
    if (lane == 0) {
      const uint8_t* src = reinterpret_cast<const uint8_t*>(p.wt) +
                           (((size_t)g * gm.n_tiles + n_tile) * 2 + rank) * (size_t)n_units * b_unit_bytes;
      int s = 0;
      uint32_t ph = 1;  // a fresh barrier passes a wait on parity 1
      for (int st = 0; st < n_stages; ++st) {
        const int nu = min(S, n_units - st * S);
        ptx::mbar_wait(emptyB(s), ph);
        stamp(1, st, 0);
        if (d.reserved & 16) {           // timing experiments: no weight traffic
          ptx::mbar_arrive(fullB(s));
        } else {
          ptx::mbar_arrive_expect_tx(fullB(s), nu * b_unit_bytes);
          ptx::bulk_g2s(ptx::smem_u32(sB) + s * S * b_unit_bytes, src + (size_t)st * S * b_unit_bytes,
                        nu * b_unit_bytes, fullB(s));
        }
        stamp(1, st, 1);
        if (++s == NB) { s = 0; ph ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (rank == 0) {
      // ================= MMA issuer (leader CTA) =================
      // Measured (scripts/mma_pace.cu, profiles/r02_mma_pace.txt, and clock stamps inside this kernel): the tensor
      // pipe retires a 256 x 512 x 16 unit in 256 cycles, but an issue loop that spends ~60 instructions per unit in
      // ONE thread (election, warp sync, fence, 64-bit descriptor arithmetic, per-unit branches) cannot issue faster
      // than ~450 cycles per unit: the issuing thread, not the tensor pipe, paced the round-1 kernels.  The fast path
      // below is walked by the whole converged warp with warp-uniform operands and elects inside the asm blocks, so
      // it compiles to straight-line uniform-datapath code: ~7 instructions per MMA, one wait and one commit per
      // stage of 3 units.
      const int n1 = BN > 256 ? 256 : BN;
      const int n2 = BN - n1;
      const uint32_t idesc1 = ptx::make_idesc_bf16(2 * BM, n1);
      const uint32_t idesc2 = n2 ? ptx::make_idesc_bf16(2 * BM, n2) : 0u;
      const uint32_t sbo = 128;
      // descriptors differ only in the 14-bit start-address field: add (bytes >> 4) per step
      const uint64_t da0 = ptx::make_smem_desc(ptx::smem_u32(sA), (BM / 8) * 128, sbo);
      const uint64_t db1_0 = ptx::make_smem_desc(ptx::smem_u32(sB), (uint32_t)n1 * 8, sbo);
      const uint64_t db2_0 = ptx::make_smem_desc(ptx::smem_u32(sB) + n1 * 16, (uint32_t)n2 * 8, sbo);
      const uint32_t a_slot_step = (uint32_t)(gm.Mp * A_UNIT_BYTES) >> 4, b_unit_step = b_unit_bytes >> 4;
      const int Mp = gm.Mp;
      const int LM = gm.loopM;   // units per chunk group (M; nn.Linear: all chunks of the layer)
      const bool fast = !(d.reserved & 2048) && S == 3 && (LM % Mp) == 0 && (Mp % 3) == 0 && !(d.reserved & 4);
      if (fast) {
        const uint32_t a_hi = (uint32_t)(da0 >> 32), b1_hi = (uint32_t)(db1_0 >> 32), b2_hi = (uint32_t)(db2_0 >> 32);
        const uint32_t a_lo0 = (uint32_t)da0, b1_lo0 = (uint32_t)db1_0, b2_lo0 = (uint32_t)db2_0;
        const uint32_t a_unit = A_UNIT_BYTES >> 4, stage_step = 3u * b_unit_step;
        const int stages_per_part = Mp / 3;
        const int n_parts_total = gm.loopC * (LM / Mp);
        int sc = 0, sb = 0;
        uint32_t phA = 0, phB = 0, acc = 0;
        int gst = 0;
        for (int part = 0; part < n_parts_total; ++part) {
          if (dbg && lane == 0) stamp(0, gst, 2);
          ptx::mbar_wait(fullA(sc), phA);
          ptx::tc_fence_after();
          uint32_t a_lo = a_lo0 + (uint32_t)sc * a_slot_step;
          for (int st = 0; st < stages_per_part; ++st, ++gst) {
            if (dbg && lane == 0) stamp(0, gst, 0);
            ptx::mbar_wait(fullB(sb), phB);
            ptx::tc_fence_after();
            if (dbg && lane == 0) stamp(0, gst, 1);
            const uint32_t b1 = b1_lo0 + (uint32_t)sb * stage_step, b2 = b2_lo0 + (uint32_t)sb * stage_step;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
              ptx::mma2_bf16_ss_lohi_elect(tmem_base, a_lo + k * a_unit, a_hi, b1 + k * b_unit_step, b1_hi, idesc1, acc);
              if (n2)
                ptx::mma2_bf16_ss_lohi_elect(tmem_base + 256, a_lo + k * a_unit, a_hi, b2 + k * b_unit_step, b2_hi, idesc2,
                                             acc);
              acc = 1;
            }
            ptx::tc_commit2_elect(emptyB(sb));
            if (++sb == NB) { sb = 0; phB ^= 1; }
            a_lo += 3 * a_unit;
          }
          ptx::tc_commit2_elect(emptyA(sc));
          if (++sc == NCH) { sc = 0; phA ^= 1; }
        }
        ptx::tc_commit2_elect(tmem_full);
      } else if (!(d.reserved & (4 | 32 | 2048))) {
        // parts that do not split into whole stages of 3 (M = 14: two parts of 7, ...): the same warp-wide issue with
        // the stage bookkeeping per unit
        const uint32_t a_hi = (uint32_t)(da0 >> 32), b1_hi = (uint32_t)(db1_0 >> 32), b2_hi = (uint32_t)(db2_0 >> 32);
        const uint32_t a_lo0 = (uint32_t)da0, b1_lo0 = (uint32_t)db1_0, b2_lo0 = (uint32_t)db2_0;
        const uint32_t a_unit = A_UNIT_BYTES >> 4;
        int sc = 0, sb = 0, ub = 0, u = 0;
        uint32_t phA = 0, phB = 0, acc = 0;
        for (int ch = 0; ch < gm.loopC; ++ch) {
          for (int m0 = 0; m0 < LM; m0 += Mp) {
            const int m1 = min(LM, m0 + Mp);
            ptx::mbar_wait(fullA(sc), phA);
            ptx::tc_fence_after();
            uint32_t a_lo = a_lo0 + (uint32_t)sc * a_slot_step;
            for (int q = m0; q < m1; ++q, ++u) {
              if (ub == 0) {
                ptx::mbar_wait(fullB(sb), phB);
                ptx::tc_fence_after();
              }
              const uint32_t boff = (uint32_t)(sb * S + ub) * b_unit_step;
              ptx::mma2_bf16_ss_lohi_elect(tmem_base, a_lo, a_hi, b1_lo0 + boff, b1_hi, idesc1, acc);
              if (n2) ptx::mma2_bf16_ss_lohi_elect(tmem_base + 256, a_lo, a_hi, b2_lo0 + boff, b2_hi, idesc2, acc);
              acc = 1;
              a_lo += a_unit;
              if (ub + 1 == S || u + 1 == n_units) {
                ptx::tc_commit2_elect(emptyB(sb));
                ub = 0;
                if (++sb == NB) { sb = 0; phB ^= 1; }
              } else {
                ++ub;
              }
            }
            ptx::tc_commit2_elect(emptyA(sc));
            if (++sc == NCH) { sc = 0; phA ^= 1; }
          }
        }
        ptx::tc_commit2_elect(tmem_full);
      } else if (lane == 0) {
        // timing experiments (stamps, no-MMA runs): one thread
        int sc = 0, sb = 0, ub = 0, u = 0;
        uint32_t phA = 0, phB = 0;
        uint64_t db1 = db1_0, db2 = db2_0;   // B descriptors of the current unit
        for (int ch = 0; ch < gm.loopC; ++ch) {
          for (int m0 = 0; m0 < LM; m0 += Mp) {     // one ring slot per part of the chunk
            const int m1 = min(LM, m0 + Mp);
            if (dbg) stamp(0, u, 2);
            ptx::mbar_wait(fullA(sc), phA);
            ptx::tc_fence_after();
            uint64_t da = da0 + (uint64_t)(sc * a_slot_step);
            for (int q = m0; q < m1; ++q, ++u) {
              if (ub == 0) {
                if (dbg) stamp(0, u, 0);
                ptx::mbar_wait(fullB(sb), phB);
                ptx::tc_fence_after();
                if (dbg) stamp(0, u, 1);
              }
              if (!(d.reserved & 4)) {
                ptx::mma2_bf16_ss(tmem_base, da, db1, idesc1, u > 0);
                if (n2) ptx::mma2_bf16_ss(tmem_base + 256, da, db2, idesc2, u > 0);
              }
              if (dbg) stamp(0, u, 3);
              da += A_UNIT_BYTES >> 4;
              const bool last_of_stage = (ub + 1 == S) || (u + 1 == n_units);
              if (last_of_stage) {
                ptx::tc_commit2(emptyB(sb));
                ub = 0;
                if (++sb == NB) { sb = 0; phB ^= 1; }
                db1 = db1_0 + (uint64_t)(sb * S) * b_unit_step;
                db2 = db2_0 + (uint64_t)(sb * S) * b_unit_step;
              } else {
                ++ub;
                db1 += b_unit_step;
                db2 += b_unit_step;
              }
            }
            ptx::tc_commit2(emptyA(sc));
            if (++sc == NCH) { sc = 0; phA ^= 1; }
          }
        }
        ptx::tc_commit2(tmem_full);
      }
    } else if (lane == 0) {
      // ================= peer CTA: tell the leader when our half of each B stage has landed ===========
      int s = 0;
      uint32_t ph = 0;
      for (int st = 0; st < n_stages; ++st) {
        ptx::mbar_wait(fullB(s), ph);
        stamp(3, st, 0);
        ptx::mbar_arrive_cluster(ptx::mapa(fullB(s), 0));
        stamp(3, st, 1);
        if (++s == NB) { s = 0; ph ^= 1; }
      }
    }
    __syncwarp();
  } else {
    // ================= A producers: basis prologue =================
    const int ptid = threadIdx.x - 64;
    const int r = ptid & (BM - 1), wg = ptid >> 7;  // wg: which FPT-feature slice of the 16-feature chunk
    const int n_slices = d.n_groups / d.x_share;
    const int slice = g / d.x_share;
    const int xoff = slice * d.in_features;
    const int in = d.in_features;
    const int t = t0 + r;
    const bool row_ok = t < d.T;
    const float* xrow = p.X + (size_t)(row_ok ? t : 0) * d.ldx + xoff;
    const bool vec = ((d.ldx & 3) == 0) && ((xoff & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.X) & 15) == 0);
    const BasisCtx c = make_ctx_tc(d, p.fparams, g);
    const int layout = group_layout(d, gm, p.fparams, g);
    float gt0 = 0.f, ginv_h = 0.f;
    if (layout == LAYOUT_EFFKAN8) {
      const float* probe = c.fp + (size_t)in * (d.grid_size + 2 * d.spline_order + 1);
      gt0 = probe[1];
      ginv_h = probe[2];
    }
    float mu = 0.f, rs = 1.f;
    if (F == KV_FASTKAN && !(d.flags & KV_FLAG_NO_LAYERNORM) && row_ok) {
      mu = p.mean[(size_t)t * n_slices + slice];
      rs = p.rstd[(size_t)t * n_slices + slice];
    }
    const int row_off = (r >> 3) * 128 + (r & 7) * 16;
    // Linear as the input gradient of a ReLU layer: dY is gated by the layer's own forward output
    const float* mrow = (F == KV_LINEAR && p.in_mask != nullptr)
                            ? p.in_mask + (size_t)(row_ok ? t : 0) * d.ldx + xoff : nullptr;
    const bool mvec = vec && ((reinterpret_cast<uintptr_t>(p.in_mask) & 15) == 0);
    // FPT consecutive features of this row, starting at i0g (zero beyond `in` / beyond T)
    auto load_xf = [&](const float* row, bool v16, int i0g, float (&v)[FPT]) {
      if constexpr (FPT == 8) {
        load_x8(row, i0g, in, row_ok, v16, v);
      } else {
        if (row_ok && v16 && i0g + 4 <= in) {
          const float4 a = __ldg(reinterpret_cast<const float4*>(row + i0g));
          v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
        } else {
#pragma unroll
          for (int q = 0; q < 4; ++q) v[q] = (row_ok && i0g + q < in) ? __ldg(row + i0g + q) : 0.0f;
        }
      }
    };
    auto gate = [&](int i0g, float (&v)[FPT]) {
      if (F == KV_LINEAR && mrow != nullptr) {
        float mk[FPT];
        load_xf(mrow, mvec, i0g, mk);
#pragma unroll
        for (int q = 0; q < FPT; ++q) v[q] = mk[q] > 0.0f ? v[q] : 0.0f;
      }
    };
    if (ptid == 0) stamp(2, 62, 0);
    if constexpr (F == KV_LINEAR) {
      static_assert(F != KV_LINEAR || FPT == 8, "the Linear producers own 8 features per thread");
      int slot_i = 0;
      const size_t toff = (size_t)t0 * d.ldx + xoff;   // (row t0 of the tile, first feature of the slice)
      linear_produce_tile(gm, p.X + toff, p.in_mask != nullptr ? p.in_mask + toff : nullptr, min(BM, d.T - t0), d.ldx, vec,
                          mvec, in, sA, bar0, NCH, ptid, lane, slot_i);
    } else {
    // two copies of the chunk loop (default-grid closed form / everything else): the registers of the other-grid
    // route (EffkanOctet: 40 values per thread) must not push spills into the default route
    auto run_chunks = [&](auto e8_tag) {
    constexpr bool E8 = decltype(e8_tag)::value;
    float cur[FPT];
    load_xf(xrow, vec, wg * FPT, cur);
    gate(wg * FPT, cur);
    for (int ch = 0; ch < gm.n_chunks; ++ch) {
      float nxt[FPT];
      if (ch + 1 < gm.n_chunks) {
        load_xf(xrow, vec, (ch + 1) * KU + wg * FPT, nxt);
        gate((ch + 1) * KU + wg * FPT, nxt);
      } else {
#pragma unroll
        for (int q = 0; q < FPT; ++q) nxt[q] = 0.0f;
      }
      const int i0 = ch * KU + wg * FPT;
      // effkan on a non-default grid: the non-zero splines of the thread's 8 features, once per chunk (all parts)
      EffkanOctet eo;
      if constexpr (F == KV_EFFKAN && FPT == 8 && !E8) {
        if (!(d.reserved & 8)) effkan_octet_basis(c, i0, cur, eo);
      }
      for (int pt = 0; pt < gm.npart; ++pt) {      // one ring slot per part: bases [m0, m0 + Mt) of the chunk
      const int slot_i = ch * gm.npart + pt;
      const int sc = slot_i % NCH;
      const int m0 = pt * gm.Mp, Mt = min(gm.Mp, M - m0);
      ptx::mbar_wait(emptyA(sc), ((slot_i / NCH) & 1) ^ 1);
      if (ptid == 0) stamp(2, ch, 0);
      uint8_t* base = sA + sc * gm.Mp * A_UNIT_BYTES + row_off;
      if (d.reserved & 8) {              // timing experiments: no basis evaluation
      } else
      if constexpr (F == KV_EFFKAN) {
        if constexpr (E8) {                         // default grid, M = 9: one part
          if constexpr (FPT == 8) effkan8_chunk(base, wg, cur, gt0, ginv_h);
          else effkan8_chunk4(base, wg, cur, gt0, ginv_h);
        } else {
          // any other grid: zero rows + the k + 1 values of each feature scattered (effkan_scatter_part)
          if constexpr (FPT == 8) {
            effkan_scatter_part(base + wg * 2048, A_UNIT_BYTES, m0, Mt, c, i0, cur, eo);
          } else {
#pragma unroll 1
            for (int q = 0; q < FPT; ++q) {   // feature f = wg*FPT + q of the chunk sits in k-half f / 8 at element f % 8
              const int f = wg * FPT + q;
              StrideSink sink{base + (f >> 3) * 2048 + (f & 7) * 2, A_UNIT_BYTES, m0, Mt};
              if (i0 + q < in) {
                basis_eval<F>(c, i0 + q, cur[q], cur[q], sink);
              } else {
                for (int m = m0; m < m0 + Mt; ++m) sink.put(m, 0.0f);
              }
            }
          }
        }
      } else {
        static_assert(F == KV_EFFKAN || FPT == 8, "the octet producers own 8 features per thread");
        if constexpr (FPT == 8) basis_octet<F>(base + wg * 2048, A_UNIT_BYTES, m0, Mt, cur, c, i0, mu, rs);
      }
      ptx::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(ptx::mapa(fullA(sc), 0));  // the leader's barrier
      if (ptid == 0) stamp(2, ch, 1);
      }
#pragma unroll
      for (int q = 0; q < FPT; ++q) cur[q] = nxt[q];
    }
    };
    if (F == KV_EFFKAN && layout == LAYOUT_EFFKAN8) {
      if constexpr (F == KV_EFFKAN) run_chunks(std::true_type{});
    } else {
      run_chunks(std::false_type{});
    }
    }
    // ================= epilogue: TMEM -> registers -> smem transpose -> coalesced global stores =========
    if (ptid == 0) stamp(2, 62, 1);
    ptx::mbar_wait(tmem_full, 0);
    ptx::tc_fence_after();
    if (ptid == 0) stamp(2, 62, 2);
    const int pw = warp - 2;
    const int q4 = warp & 3;            // TMEM lane quadrant this warp may read
    const int half = pw >> 2;           // NPW / 4 warps per quadrant split the column blocks
    const int out = d.out_features;
    const int ncols_total = d.n_groups * out;
    constexpr int LDT = 36;             // staging row stride in floats (conflict-free float4 access)
    static_assert(NPW * 32 * LDT * 4 <= A_RING_BYTES, "epilogue staging lives in the idle A ring");
    float* stg = reinterpret_cast<float*>(sA) + pw * (32 * LDT);  // the operand rings are idle by now
    const int c4 = (lane & 7) * 4, rsub = lane >> 3;
    // the 8 rows this lane stores in the row-wise phase
    long long yoff[8];
    int posrow[8];
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int te = t0 + q4 * 32 + it * 4 + rsub;
      if (te < d.T) {
        yoff[it] = (long long)map_row(te, d.rows_in, d.rows_out, d.row_shift) * d.ldy;
        posrow[it] = d.rows_in ? (d.row_shift + te % d.rows_in) : te;
      } else {
        yoff[it] = -1;
        posrow[it] = 0;
      }
    }
    const int colbase = g * out + n_tile * BN;
    const bool vec_y = ((d.ldy & 3) == 0) && ((colbase & 3) == 0) && ((ncols_total & 3) == 0) &&
                       ((reinterpret_cast<uintptr_t>(p.Y) & 15) == 0) &&
                       (!p.residual || (reinterpret_cast<uintptr_t>(p.residual) & 15) == 0) &&
                       (!p.pos || (reinterpret_cast<uintptr_t>(p.pos) & 15) == 0) &&
                       (!p.bias || (reinterpret_cast<uintptr_t>(p.bias) & 15) == 0);
    if (p.attn_img != nullptr) {
      // ---- attention-image epilogue: thread = token row (TMEM lane), 16-byte chunks of 8 consecutive d
      const int te = t0 + q4 * 32 + lane;
      const bool ok = te < d.T;
      const int Ntok = p.attn_N, dh = out / 3;
      const int NPAD = (Ntok + 15) / 16 * 16, DHP = (dh + 15) / 16 * 16;
      const int bs = ok ? te / Ntok : 0, n = ok ? te - bs * Ntok : 0;
      const size_t part_elems = (size_t)NPAD * DHP;
      __nv_bfloat16* hb = p.attn_img + ((size_t)bs * d.n_groups + g) * 3 * part_elems;
      auto chunk_at = [&](int part, int nn, int dd) {
        return reinterpret_cast<uint4*>(hb + part * part_elems + ((size_t)(nn >> 3) * (DHP >> 3) + (dd >> 3)) * 64 + (nn & 7) * 8);
      };
      const uint4 zero4 = make_uint4(0u, 0u, 0u, 0u);
      for (int cb = half; cb * 32 < BN; cb += NPW / 4) {
        uint32_t rr[32];
        ptx::tmem_ld32(tmem_base + ((uint32_t)(q4 * 32) << 16) + cb * 32, rr);
        ptx::tmem_ld_wait();
        if (!ok) continue;
#pragma unroll
        for (int k8 = 0; k8 < 4; ++k8) {
          const int o = n_tile * BN + cb * 32 + 8 * k8;   // first output of the chunk (out % 8 == 0, d_head % 8 == 0)
          if (cb * 32 + 8 * k8 >= BN || o >= out) continue;
          float v[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(rr[8 * k8 + e]) + (p.bias ? p.bias[g * out + o + e] : 0.0f);
          const int part = o / dh, dd = o - part * dh;
          *chunk_at(part, n, dd) = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]),
                                              pack_bf16(v[6], v[7]));
          const bool last_d = (dd + 8 == dh);
          if (last_d)
            for (int dp = dh; dp < DHP; dp += 8) *chunk_at(part, n, dp) = zero4;          // d padding of this row
          if (n == Ntok - 1) {                                                            // row padding of the sample
            for (int nn = Ntok; nn < NPAD; ++nn) {
              *chunk_at(part, nn, dd) = zero4;
              if (last_d)
                for (int dp = dh; dp < DHP; dp += 8) *chunk_at(part, nn, dp) = zero4;
            }
          }
        }
      }
    } else
    for (int cb = half; cb * 32 < BN; cb += NPW / 4) {
      const int sidx = 60 - (cb - half) / (NPW / 4);     // timing experiments: the first two column blocks of warp 2
      const bool st_on = ptid == 0 && sidx >= 59;
      if (st_on) stamp(2, sidx, 0);
      uint32_t rr[32];
      ptx::tmem_ld32(tmem_base + ((uint32_t)(q4 * 32) << 16) + cb * 32, rr);
      ptx::tmem_ld_wait();
      if (st_on) stamp(2, sidx, 1);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        *reinterpret_cast<uint4*>(stg + lane * LDT + 4 * k) = make_uint4(rr[4 * k], rr[4 * k + 1], rr[4 * k + 2], rr[4 * k + 3]);
      __syncwarp();
      if (st_on) stamp(2, sidx, 2);
      const int cl = cb * 32 + c4;           // column inside the tile
      const int o = n_tile * BN + cl;        // output feature of the group
      const int col = colbase + cl;          // column of Y
      float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
      const bool full4 = vec_y && (cl + 3 < BN) && (o + 3 < out);
      if (p.bias && full4) bv = *reinterpret_cast<const float4*>(p.bias + col);
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const float4 a = *reinterpret_cast<const float4*>(stg + (it * 4 + rsub) * LDT + c4);
        if (yoff[it] < 0) continue;
        float v[4] = {a.x, a.y, a.z, a.w};
        if (full4) {
          v[0] += bv.x; v[1] += bv.y; v[2] += bv.z; v[3] += bv.w;
          if (p.pos) {
            const float4 pv = *reinterpret_cast<const float4*>(p.pos + (size_t)posrow[it] * ncols_total + col);
            v[0] += pv.x; v[1] += pv.y; v[2] += pv.z; v[3] += pv.w;
          }
          if (d.flags & KV_FLAG_RELU) {
#pragma unroll
            for (int e = 0; e < 4; ++e) v[e] = fmaxf(v[e], 0.0f);
          }
          if (p.residual) {
            const float4 rv = *reinterpret_cast<const float4*>(p.residual + yoff[it] + col);
            v[0] += rv.x; v[1] += rv.y; v[2] += rv.z; v[3] += rv.w;
          }
          *reinterpret_cast<float4*>(p.Y + yoff[it] + col) = make_float4(v[0], v[1], v[2], v[3]);
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            if (cl + e < BN && o + e < out) {
              float w = v[e];
              if (p.bias) w += p.bias[col + e];
              if (p.pos) w += p.pos[(size_t)posrow[it] * ncols_total + col + e];
              if (d.flags & KV_FLAG_RELU) w = fmaxf(w, 0.0f);
              if (p.residual) w += p.residual[yoff[it] + col + e];
              p.Y[yoff[it] + col + e] = w;
            }
          }
        }
      }
      __syncwarp();
      if (st_on) stamp(2, sidx, 3);
    }
    ptx::tc_fence_before();
    if (ptid == 0) stamp(2, 62, 3);
  }
  ptx::cluster_sync_all();  // the peer's smem / TMEM stay alive until every MMA and epilogue is done
  if (dbg) {
    if (threadIdx.x == 64) stamp(2, 61, 2);
    __syncthreads();
    for (int i = threadIdx.x; i < 1024; i += NTH)
      reinterpret_cast<uint32_t*>(p.Y)[(size_t)(rank * 128 + i / 512) * d.ldy + (i % 512)] = dbuf[i];
  }
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc2(tmem_base, TMEM_COLS);
  }
}


// =========================================================================================
// Persistent forward kernel (BN <= 256): one CTA pair per SM pair walks the (group, row block, output tile) list.
//   * the accumulator is double-buffered in TMEM (2 x 256 columns): four dedicated epilogue warps drain tile i while
//     the tensor pipe works on tile i + 1, so the epilogue (TMEM -> registers -> coalesced stores, the HBM write
//     burst of every CTA at once) and the per-tile set-up (TMEM allocation, barrier init, cluster sync, first
//     operand fetch: ~6000 cycles) no longer sit between two main loops -- in the one-tile-per-CTA kernel they took
//     ~30 % of a 768 -> 3072 tile and most of a 64 -> 192 (q/k/v) tile;
//   * the operand rings simply keep running across tiles (loader and producers run ahead into the next tile);
//   * WEIGHT-RESIDENT mode: when the weight image of a whole output tile fits the B ring (the per-head q/k/v maps:
//     36 units x 6 KB = 108 KB per CTA), each pair walks a CONTIGUOUS range of row blocks of one (group, output tile)
//     and the image is fetched once per range instead of once per tile -- the q/k/v layer re-read 108 KB per CTA
//     and tile from L2 for a 3 500-cycle main loop;
//   * the producers fetch the first chunk of the next tile's X while they evaluate the last chunk of the current one,
//     and keep the per-group prologue constants while the group does not change.
// Warps: 0 loader, 1 MMA issuer (leader) / B relay (peer), 2-9 basis producers, 10-17 epilogue.
// =========================================================================================
constexpr int P_NPW = 8;                       // producer warps
// epilogue warps: two per TMEM lane quadrant, alternating column blocks (q/k/v forward 0.52 -> 0.44 ms against one per
// quadrant: a warp's stores are latency-bound).  576 threads leave 96 registers per thread; the nn.Linear producers
// therefore hold ONE slot of X in registers (with the next slot prefetched as well they spilled: 356 -> 217 TFLOP/s;
// single-buffered with eight epilogue warps: 385).
__host__ __device__ constexpr int p_new(int F) { return 8; }
__host__ __device__ constexpr int p_threads(int F) { return 32 * (2 + P_NPW + p_new(F)); }
constexpr int P_B_RING = 112 * 1024;
constexpr int P_STG_BYTES = 8 * 32 * 36 * 4;   // epilogue staging: 32 rows x 36 floats per warp, up to 8 warps
constexpr int P_SMEM = A_RING_BYTES + P_B_RING + P_STG_BYTES + BAR_BYTES;

template <int F>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(p_threads(F), 1) kan_tc_fwd_persistent_kernel(const FwdParams p) {
  constexpr int P_NEW = p_new(F);
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;
  uint8_t* sB = smem + A_RING_BYTES;
  float* sStg = reinterpret_cast<float*>(smem + A_RING_BYTES + P_B_RING);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + A_RING_BYTES + P_B_RING + P_STG_BYTES);
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto fullA = [&](int s) { return bar0 + 8u * s; };
  auto emptyA = [&](int s) { return bar0 + 8u * (NCH_MAX + s); };
  auto fullB = [&](int s) { return bar0 + 8u * (2 * NCH_MAX + s); };
  auto emptyB = [&](int s) { return bar0 + 8u * (2 * NCH_MAX + NB_MAX + s); };
  auto tmemFull = [&](int b) { return bar0 + 8u * (2 * NCH_MAX + 2 * NB_MAX + b); };
  auto tmemEmpty = [&](int b) { return bar0 + 8u * (2 * NCH_MAX + 2 * NB_MAX + 2 + b); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * NCH_MAX + 2 * NB_MAX + 4);

  const kv_kan_desc_t& d = p.d;
  const Geom gm = p.gm;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();  // 0 = leader of the CTA pair
  const int M = d.M, Mp = gm.Mp;
  const int n_units = gm.n_units;
  const int BN = gm.BN, NCH = gm.NCH, S = gm.S;
  const uint32_t b_unit_bytes = (uint32_t)BN * KU;  // this CTA's half of the B unit: BN/2 rows x 16 x 2 B
  int NB = P_B_RING / (S * (int)b_unit_bytes);
  NB = NB > NB_MAX ? NB_MAX : NB;
  const int n_stages = (n_units + S - 1) / S;
  // tile list: output tiles of one row block adjacent (their X tile is shared through L2 at the same time)
  const int n_pairs = ceil_div(d.T, 2 * BM);
  const int total_tiles = d.n_groups * n_pairs * gm.n_tiles;
  // weight-resident mode: the B ring holds the whole weight image of a tile (stage st lives in ring slot st)
  const bool resident = n_stages <= NB;
  const int G = (int)gridDim.y, y = (int)blockIdx.y;
  const int per = total_tiles / G, rem = total_tiles % G;
  const int n_my = resident ? per + (y < rem ? 1 : 0) : (total_tiles - y + G - 1) / G;
  const int first = y * per + (y < rem ? y : rem);
  auto tile_of = [&](int it, int& g, int& mp, int& n_tile) {
    if (resident) {            // contiguous range, row block fastest: (g, n_tile) changes at most a few times
      const int tile = first + it;
      mp = tile % n_pairs;
      const int r = tile / n_pairs;
      n_tile = r % gm.n_tiles;
      g = r / gm.n_tiles;
    } else {                   // interleaved, output tile fastest: the pairs running together share the X rows
      const int tile = y + it * G;
      n_tile = tile % gm.n_tiles;
      const int r = tile / gm.n_tiles;
      mp = r % n_pairs;
      g = r / n_pairs;
    }
  };
  // resident mode: tile `it` needs a fresh weight image iff its (g, n_tile) differs from the previous tile's
  auto reload = [&](int it) { return it == 0 || (first + it) / n_pairs != (first + it - 1) / n_pairs; };

  if (warp == 1 && lane == 0) {
    for (int s = 0; s < NCH_MAX; ++s) {
      ptx::mbar_init(fullA(s), 2 * P_NPW);  // (leader) local + peer producer warps
      ptx::mbar_init(emptyA(s), 1);
    }
    for (int s = 0; s < NB_MAX; ++s) {
      ptx::mbar_init(fullB(s), rank == 0 ? 2 : 1);  // leader: own bulk copy + the peer's "my half landed"
      ptx::mbar_init(emptyB(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(tmemFull(b), 1);
      ptx::mbar_init(tmemEmpty(b), 2 * P_NEW);   // (leader) the epilogue warps of both CTAs
    }
    ptx::fence_mbar_init();
  }
  if (warp == 0) {
    ptx::tmem_alloc2(ptx::smem_u32(tmem_slot), TMEM_COLS);
    ptx::tmem_relinquish2();
  }
  ptx::tc_fence_before();
  ptx::cluster_sync_all();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================= B loader: this CTA's half of S units per bulk copy, tile after tile =================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 1;  // a fresh barrier passes a wait on parity 1
      for (int it = 0; it < n_my; ++it) {
        int g, mp, n_tile;
        tile_of(it, g, mp, n_tile);
        const uint8_t* src = reinterpret_cast<const uint8_t*>(p.wt) +
                             (((size_t)g * gm.n_tiles + n_tile) * 2 + rank) * (size_t)n_units * b_unit_bytes;
        if (resident) {
          if (!reload(it)) continue;
          s = 0;               // slot = stage; ph flips once per reload (set below)
        }
        for (int st = 0; st < n_stages; ++st) {
          const int nu = min(S, n_units - st * S);
          ptx::mbar_wait(emptyB(s), ph);
          ptx::mbar_arrive_expect_tx(fullB(s), nu * b_unit_bytes);
          ptx::bulk_g2s(ptx::smem_u32(sB) + s * S * b_unit_bytes, src + (size_t)st * S * b_unit_bytes,
                        nu * b_unit_bytes, fullB(s));
          if (++s == NB && !resident) { s = 0; ph ^= 1; }
        }
        if (resident) ph ^= 1;
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (rank == 0) {
      // ================= MMA issuer (leader CTA): the whole converged warp, election inside the asm blocks ==========
      const uint32_t idesc = ptx::make_idesc_bf16(2 * BM, BN);
      const uint64_t da0 = ptx::make_smem_desc(ptx::smem_u32(sA), (BM / 8) * 128, 128);
      const uint64_t db0 = ptx::make_smem_desc(ptx::smem_u32(sB), (uint32_t)BN * 8, 128);
      const uint32_t a_hi = (uint32_t)(da0 >> 32), b_hi = (uint32_t)(db0 >> 32);
      const uint32_t a_lo0 = (uint32_t)da0, b_lo0 = (uint32_t)db0;
      const uint32_t a_unit = A_UNIT_BYTES >> 4, b_unit_step = b_unit_bytes >> 4;
      const uint32_t a_slot_step = (uint32_t)(Mp * A_UNIT_BYTES) >> 4;
      const int LM = gm.loopM;   // units per chunk group (M; nn.Linear: all chunks of the layer)
      const bool fast = S == 3 && (LM % Mp) == 0 && (Mp % 3) == 0;
      const int n_parts_total = gm.loopC * gm.npart;
      int sc = 0, sb = 0;
      uint32_t phA = 0, phB = 0, n_reloads = 0;
      for (int it = 0; it < n_my; ++it) {
        const int buf = it & 1;
        // the epilogue warps of both CTAs have drained this accumulator (its use two tiles ago)
        ptx::mbar_wait_cluster(tmemEmpty(buf), (uint32_t)((it >> 1) & 1) ^ 1u);
        ptx::tc_fence_after();
        const uint32_t tacc = tmem_base + buf * 256;
        uint32_t acc = 0;
        // resident weights: wait for the image only when it was (re)loaded for this tile, hand the slots back to the
        // loader only when the next tile needs another image
        const bool wait_b = !resident || reload(it);
        const bool free_b = !resident || (it + 1 < n_my && reload(it + 1));
        if (resident) {
          sb = 0;
          if (wait_b) phB = n_reloads++ & 1u;
        }
        if (fast) {
          const int stages_per_part = Mp / 3;
          for (int part = 0; part < n_parts_total; ++part) {
            ptx::mbar_wait(fullA(sc), phA);
            ptx::tc_fence_after();
            uint32_t a_lo = a_lo0 + (uint32_t)sc * a_slot_step;
            for (int st = 0; st < stages_per_part; ++st) {
              if (wait_b) {
                ptx::mbar_wait(fullB(sb), phB);
                ptx::tc_fence_after();
              }
              const uint32_t b_lo = b_lo0 + (uint32_t)sb * 3u * b_unit_step;
#pragma unroll
              for (int k = 0; k < 3; ++k) {
                ptx::mma2_bf16_ss_lohi_elect(tacc, a_lo + k * a_unit, a_hi, b_lo + k * b_unit_step, b_hi, idesc, acc);
                acc = 1;
              }
              if (free_b) ptx::tc_commit2_elect(emptyB(sb));
              if (++sb == NB) { sb = 0; phB ^= 1; }
              a_lo += 3 * a_unit;
            }
            ptx::tc_commit2_elect(emptyA(sc));
            if (++sc == NCH) { sc = 0; phA ^= 1; }
          }
        } else {
          int ub = 0, u = 0;
          for (int ch = 0; ch < gm.loopC; ++ch) {
            for (int m0 = 0; m0 < LM; m0 += Mp) {
              const int m1 = min(LM, m0 + Mp);
              ptx::mbar_wait(fullA(sc), phA);
              ptx::tc_fence_after();
              uint32_t a_lo = a_lo0 + (uint32_t)sc * a_slot_step;
              for (int q = m0; q < m1; ++q, ++u) {
                if (ub == 0 && wait_b) {
                  ptx::mbar_wait(fullB(sb), phB);
                  ptx::tc_fence_after();
                }
                ptx::mma2_bf16_ss_lohi_elect(tacc, a_lo, a_hi, b_lo0 + (uint32_t)(sb * S + ub) * b_unit_step, b_hi, idesc, acc);
                acc = 1;
                a_lo += a_unit;
                if (ub + 1 == S || u + 1 == n_units) {
                  if (free_b) ptx::tc_commit2_elect(emptyB(sb));
                  ub = 0;
                  if (++sb == NB) { sb = 0; phB ^= 1; }
                } else {
                  ++ub;
                }
              }
              ptx::tc_commit2_elect(emptyA(sc));
              if (++sc == NCH) { sc = 0; phA ^= 1; }
            }
          }
        }
        ptx::tc_commit2_elect(tmemFull(buf));
      }
    } else if (lane == 0) {
      // ================= peer CTA: tell the leader when our half of each B stage has landed ===========
      int s = 0;
      uint32_t ph = 0;
      if (resident) {
        for (int it = 0; it < n_my; ++it) {
          if (!reload(it)) continue;
          for (int st = 0; st < n_stages; ++st) {
            ptx::mbar_wait(fullB(st), ph);
            ptx::mbar_arrive_cluster(ptx::mapa(fullB(st), 0));
          }
          ph ^= 1;
        }
      } else {
        for (int st = 0; st < n_stages * n_my; ++st) {
          ptx::mbar_wait(fullB(s), ph);
          ptx::mbar_arrive_cluster(ptx::mapa(fullB(s), 0));
          if (++s == NB) { s = 0; ph ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp < 2 + P_NPW) {
    // ================= A producers: basis prologue, tile after tile =================
    const int ptid = threadIdx.x - 64;
    const int r = ptid & (BM - 1), wg = ptid >> 7;  // row of the 128-row tile, k-half (8 features) of the chunk
    const int n_slices = d.n_groups / d.x_share;
    const int in = d.in_features;
    const int row_off = (r >> 3) * 128 + (r & 7) * 16;
    int slot_i = 0;   // running ring-slot counter (parts produced so far, all tiles)
    // X rows of this thread in tile `it`
    struct TileX { const float* xrow; bool row_ok, vec; int g, t, slice; };
    auto tile_x = [&](int it) {
      int g, mp, n_tile;
      tile_of(it, g, mp, n_tile);
      TileX x;
      x.g = g;
      x.slice = g / d.x_share;
      const int xoff = x.slice * in;
      x.t = (mp * 2 + (int)rank) * BM + r;
      x.row_ok = x.t < d.T;
      x.vec = ((d.ldx & 3) == 0) && ((xoff & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.X) & 15) == 0);
      x.xrow = p.X + (size_t)(x.row_ok ? x.t : 0) * d.ldx + xoff;
      return x;
    };
    if constexpr (F == KV_LINEAR) {
      for (int it = 0; it < n_my; ++it) {
        const TileX x = tile_x(it);
        // Linear as the input gradient of a ReLU layer: dY is gated by the layer's own forward output
        const bool mvec = x.vec && ((reinterpret_cast<uintptr_t>(p.in_mask) & 15) == 0);
        const int t0t = x.t - r;                        // first row of this CTA's 128-row tile
        const size_t toff = (size_t)t0t * d.ldx + (size_t)x.slice * in;
        linear_produce_tile(gm, p.X + toff, p.in_mask != nullptr ? p.in_mask + toff : nullptr, min(BM, d.T - t0t), d.ldx,
                            x.vec, mvec, in, sA, bar0, NCH, ptid, lane, slot_i);
      }
    } else {
      // per-group prologue constants: reloaded only when the group changes
      int g_cached = -1, layout = LAYOUT_BASIS_MAJOR;
      float gt0 = 0.f, ginv_h = 0.f;
      float cur[8];
      TileX x = tile_x(0);
      if (n_my > 0) load_x8(x.xrow, wg * 8, in, x.row_ok, x.vec, cur);
      for (int it = 0; it < n_my; ++it) {
        if (x.g != g_cached) {
          g_cached = x.g;
          layout = group_layout(d, gm, p.fparams, x.g);
          if (layout == LAYOUT_EFFKAN8) {
            const float* probe = p.fparams + (size_t)x.g * d.fparam_stride +
                                 (size_t)in * (d.grid_size + 2 * d.spline_order + 1);
            gt0 = probe[1];
            ginv_h = probe[2];
          }
        }
        float mu = 0.f, rs = 1.f;
        if (F == KV_FASTKAN && !(d.flags & KV_FLAG_NO_LAYERNORM) && x.row_ok) {
          mu = p.mean[(size_t)x.t * n_slices + x.slice];
          rs = p.rstd[(size_t)x.t * n_slices + x.slice];
        }
        TileX xn = x;
        // two copies of the chunk loop (closed-form default grid / everything else): the register-hungry arbitrary-knot
        // prologue must not push spills into the default path
        auto run_tile = [&](auto fast_tag) {
          constexpr bool FAST = decltype(fast_tag)::value;
          // (the context is rebuilt per tile rather than kept across tiles: it would stay live through the fast copy)
          BasisCtx c;
          if constexpr (!FAST) c = make_ctx_tc(d, p.fparams, x.g);
          for (int ch = 0; ch < gm.n_chunks; ++ch) {
            float nxt[8];
            if (ch + 1 < gm.n_chunks) {
              load_x8(x.xrow, (ch + 1) * KU + wg * 8, in, x.row_ok, x.vec, nxt);
            } else if (it + 1 < n_my) {           // the next tile's first chunk rides under this tile's last one
              xn = tile_x(it + 1);
              load_x8(xn.xrow, wg * 8, in, xn.row_ok, xn.vec, nxt);
            } else {
#pragma unroll
              for (int q = 0; q < 8; ++q) nxt[q] = 0.0f;
            }
            const int i0 = ch * KU + wg * 8;
            EffkanOctet eo;    // effkan on a non-default grid: once per chunk
            if constexpr (!FAST && F == KV_EFFKAN) effkan_octet_basis(c, i0, cur, eo);
            for (int pt = 0; pt < gm.npart; ++pt, ++slot_i) {   // one ring slot per part: bases [m0, m0 + Mt) of the chunk
              const int sc = slot_i % NCH;
              const int m0 = pt * Mp, Mt = min(Mp, M - m0);
              ptx::mbar_wait(emptyA(sc), ((slot_i / NCH) & 1) ^ 1);
              uint8_t* base = sA + sc * Mp * A_UNIT_BYTES + row_off;
              if constexpr (FAST) {
                effkan8_chunk(base, wg, cur, gt0, ginv_h);   // M = 9: one part
              } else {
                if constexpr (F == KV_EFFKAN) effkan_scatter_part(base + wg * 2048, A_UNIT_BYTES, m0, Mt, c, i0, cur, eo);
                else basis_octet<F>(base + wg * 2048, A_UNIT_BYTES, m0, Mt, cur, c, i0, mu, rs);
              }
              ptx::fence_proxy_async_smem();
              __syncwarp();
              if (lane == 0) ptx::mbar_arrive_cluster(ptx::mapa(fullA(sc), 0));  // the leader's barrier
            }
#pragma unroll
            for (int q = 0; q < 8; ++q) cur[q] = nxt[q];
          }
        };
        if (F == KV_EFFKAN && layout == LAYOUT_EFFKAN8) {
          if constexpr (F == KV_EFFKAN) run_tile(std::true_type{});
        } else {
          run_tile(std::false_type{});
        }
        x = xn;
      }
    }
  } else {
    // ================= epilogue warps: TMEM -> registers -> smem transpose -> coalesced global stores =========
    const int q4 = warp & 3;                 // TMEM lane quadrant this warp may read
    const int ew = warp - (2 + P_NPW);
    constexpr int LDT = 36;                  // staging row stride in floats (conflict-free float4 access)
    float* stg = sStg + ew * (32 * LDT);
    const int out = d.out_features;
    const int ncols_total = d.n_groups * out;
    const int c4 = (lane & 7) * 4, rsub = lane >> 3;
    for (int it = 0; it < n_my; ++it) {
      int g, mp, n_tile;
      tile_of(it, g, mp, n_tile);
      const int buf = it & 1;
      const int t0 = (mp * 2 + (int)rank) * BM;
      ptx::mbar_wait(tmemFull(buf), (uint32_t)((it >> 1) & 1));
      ptx::tc_fence_after();
      const uint32_t trow = tmem_base + ((uint32_t)(q4 * 32) << 16) + buf * 256;
      if (p.attn_img != nullptr) {
        // ---- attention-image epilogue: thread = token row (TMEM lane), 16-byte chunks of 8 consecutive d
        const int te = t0 + q4 * 32 + lane;
        const bool ok = te < d.T;
        const int Ntok = p.attn_N, dh = out / 3;
        const int NPAD = (Ntok + 15) / 16 * 16, DHP = (dh + 15) / 16 * 16;
        const int bs = ok ? te / Ntok : 0, n = ok ? te - bs * Ntok : 0;
        const size_t part_elems = (size_t)NPAD * DHP;
        __nv_bfloat16* hb = p.attn_img + ((size_t)bs * d.n_groups + g) * 3 * part_elems;
        auto chunk_at = [&](int part, int nn, int dd) {
          return reinterpret_cast<uint4*>(hb + part * part_elems + ((size_t)(nn >> 3) * (DHP >> 3) + (dd >> 3)) * 64 + (nn & 7) * 8);
        };
        const uint4 zero4 = make_uint4(0u, 0u, 0u, 0u);
        for (int cb = ew >> 2; cb * 32 < BN; cb += P_NEW / 4) {
          uint32_t rr[32];
          ptx::tmem_ld32(trow + cb * 32, rr);
          ptx::tmem_ld_wait();
          if (!ok) continue;
#pragma unroll
          for (int k8 = 0; k8 < 4; ++k8) {
            const int o = n_tile * BN + cb * 32 + 8 * k8;   // first output of the chunk (out % 8 == 0, d_head % 8 == 0)
            if (cb * 32 + 8 * k8 >= BN || o >= out) continue;
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(rr[8 * k8 + e]) + (p.bias ? p.bias[g * out + o + e] : 0.0f);
            const int part = o / dh, dd = o - part * dh;
            *chunk_at(part, n, dd) = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]),
                                                pack_bf16(v[6], v[7]));
            const bool last_d = (dd + 8 == dh);
            if (last_d)
              for (int dp = dh; dp < DHP; dp += 8) *chunk_at(part, n, dp) = zero4;          // d padding of this row
            if (n == Ntok - 1) {                                                            // row padding of the sample
              for (int nn = Ntok; nn < NPAD; ++nn) {
                *chunk_at(part, nn, dd) = zero4;
                if (last_d)
                  for (int dp = dh; dp < DHP; dp += 8) *chunk_at(part, nn, dp) = zero4;
              }
            }
          }
        }
      } else {
        // the 8 rows this lane stores in the row-wise phase
        long long yoff[8];
        int posrow[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int te = t0 + q4 * 32 + k * 4 + rsub;
          if (te < d.T) {
            yoff[k] = (long long)map_row(te, d.rows_in, d.rows_out, d.row_shift) * d.ldy;
            posrow[k] = d.rows_in ? (d.row_shift + te % d.rows_in) : te;
          } else {
            yoff[k] = -1;
            posrow[k] = 0;
          }
        }
        const int colbase = g * out + n_tile * BN;
        const bool vec_y = ((d.ldy & 3) == 0) && ((colbase & 3) == 0) && ((ncols_total & 3) == 0) &&
                           ((reinterpret_cast<uintptr_t>(p.Y) & 15) == 0) &&
                           (!p.residual || (reinterpret_cast<uintptr_t>(p.residual) & 15) == 0) &&
                           (!p.pos || (reinterpret_cast<uintptr_t>(p.pos) & 15) == 0) &&
                           (!p.bias || (reinterpret_cast<uintptr_t>(p.bias) & 15) == 0);
        for (int cb = ew >> 2; cb * 32 < BN; cb += P_NEW / 4) {
          uint32_t rr[32];
          ptx::tmem_ld32(trow + cb * 32, rr);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int k = 0; k < 8; ++k)
            *reinterpret_cast<uint4*>(stg + lane * LDT + 4 * k) = make_uint4(rr[4 * k], rr[4 * k + 1], rr[4 * k + 2], rr[4 * k + 3]);
          __syncwarp();
          const int cl = cb * 32 + c4;           // column inside the tile
          const int o = n_tile * BN + cl;        // output feature of the group
          const int col = colbase + cl;          // column of Y
          float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
          const bool full4 = vec_y && (cl + 3 < BN) && (o + 3 < out);
          if (p.bias && full4) bv = *reinterpret_cast<const float4*>(p.bias + col);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const float4 a = *reinterpret_cast<const float4*>(stg + (k * 4 + rsub) * LDT + c4);
            if (yoff[k] < 0) continue;
            float v[4] = {a.x, a.y, a.z, a.w};
            if (full4) {
              v[0] += bv.x; v[1] += bv.y; v[2] += bv.z; v[3] += bv.w;
              if (p.pos) {
                const float4 pv = *reinterpret_cast<const float4*>(p.pos + (size_t)posrow[k] * ncols_total + col);
                v[0] += pv.x; v[1] += pv.y; v[2] += pv.z; v[3] += pv.w;
              }
              if (d.flags & KV_FLAG_RELU) {
#pragma unroll
                for (int e = 0; e < 4; ++e) v[e] = fmaxf(v[e], 0.0f);
              }
              if (p.residual) {
                const float4 rv = *reinterpret_cast<const float4*>(p.residual + yoff[k] + col);
                v[0] += rv.x; v[1] += rv.y; v[2] += rv.z; v[3] += rv.w;
              }
              *reinterpret_cast<float4*>(p.Y + yoff[k] + col) = make_float4(v[0], v[1], v[2], v[3]);
            } else {
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                if (cl + e < BN && o + e < out) {
                  float w = v[e];
                  if (p.bias) w += p.bias[col + e];
                  if (p.pos) w += p.pos[(size_t)posrow[k] * ncols_total + col + e];
                  if (d.flags & KV_FLAG_RELU) w = fmaxf(w, 0.0f);
                  if (p.residual) w += p.residual[yoff[k] + col + e];
                  p.Y[yoff[k] + col + e] = w;
                }
              }
            }
          }
          __syncwarp();
        }
      }
      // this warp has drained its lanes of the accumulator: the leader may overwrite it two tiles from now
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(ptx::mapa(tmemEmpty(buf), 0));
    }
  }
  ptx::tc_fence_before();
  ptx::cluster_sync_all();  // the peer's smem / TMEM stay alive until every MMA and epilogue is done
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc2(tmem_base, TMEM_COLS);
  }
}


// Forward kernel choice.  Layers whose output tile fits 256 TMEM columns (q/k/v maps, heads, d <= 256 MLPs) run the
// persistent kernel: same weight image, same tile, epilogue and set-up overlapped.  Wider layers keep the 256 x 512
// tile of kan_tc_fwd_kernel: measured on 768 -> 3072 / 3072 -> 768 (T = 66192) the persistent kernel on 256-wide
// tiles reaches 1127 / 1106 TFLOP/s against 1381 / 1246 -- halving BN doubles the basis work per FLOP and raises the
// shared-memory traffic of the MMA operand reads from 64 to 96 B/cycle/SM -- while the 64 -> 192 q/k/v layer drops
// from 0.911 to 0.574 ms.  Timing experiments: KV_FWD_BN=256 tiles every layer for the persistent kernel,
// KV_FWD_PERSISTENT=0 turns it off.
inline int fwd_max_bn() {
  static const int v = [] {
    const char* e = getenv("KV_FWD_BN");
    return (e != nullptr && atoi(e) == 256) ? 256 : 512;
  }();
  return v;
}
inline bool fwd_persistent_enabled() {
  static const bool v = [] {
    const char* e = getenv("KV_FWD_PERSISTENT");
    return !(e != nullptr && atoi(e) == 0);
  }();
  return v;
}
// nn.Linear (K = in: a 256 x 512 tile's main loop is shorter than its epilogue) is always tiled for the persistent kernel
inline Geom fwd_geom(const kv_kan_desc_t& d) {
  return make_geom(d, (d.family == KV_LINEAR && fwd_persistent_enabled()) ? 256 : fwd_max_bn());
}

template <int F>
static int launch_fwd_persistent(const FwdParams& p, cudaStream_t st) {
  KV_SMEM_OPT_IN((kan_tc_fwd_persistent_kernel<F>), P_SMEM, "tc_fwd persistent smem attribute");
  const long long tiles = (long long)ceil_div(p.d.T, 2 * BM) * p.gm.n_tiles * p.d.n_groups;
  if (tiles > 0x7fffffffLL) return KV_ERR_BAD_SHAPE;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  static int sm_count[64] = {0};
  if (dev < 64) {
    if (sm_count[dev] == 0) cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
    sms = sm_count[dev];
  }
  const int pairs = sms / 2;
  dim3 grid(2, (unsigned)(tiles < pairs ? tiles : pairs), 1);  // x = CTA pair; each pair walks tiles y, y + gridDim.y, ...
  kan_tc_fwd_persistent_kernel<F><<<grid, p_threads(F), P_SMEM, st>>>(p);
  KV_CHECK_LAUNCH("kan_tc_fwd_persistent_kernel");
  return KV_OK;
}

template <int F>
static int launch_fwd(const FwdParams& p, cudaStream_t st) {
  // 16 producer warps (4 features per thread) were measured SLOWER than 8 on the MLP shapes (768->3072: 966 vs
  // 1012 TFLOP/s): the producers are not the pacing stage, and their stores compete with the MMA operand reads
  constexpr int NPW = 8;
  if (p.gm.BN <= 256 && fwd_persistent_enabled() && !(p.d.reserved & 0xfff)) return launch_fwd_persistent<F>(p, st);
  KV_SMEM_OPT_IN((kan_tc_fwd_kernel<F, NPW>), SMEM_BYTES, "tc_fwd smem attribute");
  const int n_pairs = ceil_div(p.d.T, 2 * BM);
  if ((long long)n_pairs * p.gm.n_tiles > 0x3fffffffLL || p.d.n_groups > 65535) return KV_ERR_BAD_SHAPE;
  // x = 2 * (row pair, output tile): consecutive CTAs form the pair (grid.y / z stop at 65535, x does not); output
  // tiles of one row pair adjacent
  dim3 grid(2 * n_pairs * p.gm.n_tiles, 1, p.d.n_groups);
  kan_tc_fwd_kernel<F, NPW><<<grid, 32 * (2 + NPW), SMEM_BYTES, st>>>(p);
  KV_CHECK_LAUNCH("kan_tc_fwd_kernel");
  return KV_OK;
}


// =========================================================================================
// Input gradient on the tensor cores (default effkan on a uniform grid, LAYOUT_EFFKAN8, x_share == 1):
//   S[T, K'] = dY[T, out] . Wp[out, K']   (tcgen05, contraction over out)
//   dX[t, i] = sum_m Phi'_m(x[t,i]) * S[t, (i,m)]   (epilogue, basis derivative recomputed)
// Both MMA operands arrive by bulk copies: A = dY from a bf16 core-matrix-tiled copy made once per
// backward pass (dy_image_kernel, shared with the weight-gradient kernel), B = the transposed weight
// image of tc_pack.  One CTA pair = 256 rows x (n_sub chunks of 16 features = n_sub x 144 columns).
// =========================================================================================
constexpr int BI_KS = 4;                       // K=16 steps per stage (64 outputs)
constexpr int BI_A_STAGE = BM * BI_KS * KU * 2;  // 16 KB
constexpr int BI_B_STAGE = 32 * 1024;          // >= n_sub * SW * 64 bytes
constexpr int BI_STAGES = 4;
constexpr int BI_SMEM = BI_STAGES * (BI_A_STAGE + BI_B_STAGE) + BAR_BYTES;
constexpr int BI_SMEM_E8 = BI_SMEM + 14 * 256 * 4;  // + the epilogue's window scratch (default-effkan kernel)

struct BwdGeom {
  int SW;        // columns (= B rows) per sub-MMA: one chunk = 16 * M
  int n_sub;     // chunks per tile
  int n_tiles;   // column tiles
  int n_stages;  // contraction stages of 64 outputs
  int T8, O8;    // dY image extents in core matrices (T padded to 256, out padded to 64)
  int n_chunks;
  // The tile's n_sub * SW columns are contiguous in TMEM and are issued as ONE MMA (<= 256 columns) or TWO
  // (256 + W1): few wide instructions instead of n_sub narrow ones (a 144-column instruction runs at ~60 % of the
  // rate of a 256-column one, scripts/mma_pace.cu)
  int W1;        // columns of the second MMA, 0 = one MMA
  // More than 32 basis functions (16 * M > 512 TMEM columns): the generic input-gradient kernel is launched nsplit
  // times, each launch on Ms consecutive bases with its own transposed weight image, the later ones adding to dX
  int Ms, nsplit;
  __host__ __device__ int tile_cols() const { return n_sub * SW; }
  __host__ __device__ int W0() const { return n_sub * SW > 256 ? 256 : n_sub * SW; }
};

__host__ __device__ inline BwdGeom make_bwd_geom(const kv_kan_desc_t& d) {
  BwdGeom b;
  b.n_chunks = (d.in_features + KU - 1) / KU;
  b.nsplit = (d.M + MAX_M_TC_DX - 1) / MAX_M_TC_DX;
  b.Ms = (d.M + b.nsplit - 1) / b.nsplit;
  b.SW = KU * b.Ms;
  b.n_sub = 512 / b.SW;
  if (b.n_sub < 1) b.n_sub = 1;
  if (b.n_sub > b.n_chunks) b.n_sub = b.n_chunks;
  if (b.n_sub * b.SW * 64 > BI_B_STAGE) b.n_sub = BI_B_STAGE / (b.SW * 64);
  if (b.n_sub < 1) b.n_sub = 1;
  b.n_tiles = (b.n_chunks + b.n_sub - 1) / b.n_sub;
  b.n_stages = (d.out_features + 63) / 64;
  {
    const Geom gm = make_geom(d);
    const int cover = gm.n_tiles * gm.BN;  // the weight-gradient kernel reads whole output tiles
    b.O8 = (cover > b.n_stages * 64 ? (cover + 7) / 8 : b.n_stages * 8);
  }
  b.T8 = (d.T + 255) / 256 * 32;
  b.W1 = b.n_sub * b.SW > 256 ? b.n_sub * b.SW - 256 : 0;
  return b;
}

// dY fp32 [rows][ldy] (x ReLU mask) -> bf16 image [g][T8][O8][8 rows][8 outputs], zero padded
__global__ void dy_image_kernel(const kv_kan_desc_t d, const float* __restrict__ dY,
                                const float* __restrict__ act_mask, __nv_bfloat16* __restrict__ img, BwdGeom b) {
  const size_t total = (size_t)d.n_groups * b.T8 * b.O8 * 8;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (size_t)gridDim.x * blockDim.x) {
    size_t r = idx;
    const int tr = r % 8; r /= 8;
    const int o8 = r % b.O8; r /= b.O8;
    const int t8 = r % b.T8; r /= b.T8;
    const int g = (int)r;
    const int t = t8 * 8 + tr;
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = 0.0f;
    if (t < d.T) {
      const size_t row = (size_t)map_row(t, d.rows_in, d.rows_out, d.row_shift) * d.ldy + (size_t)g * d.out_features;
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int o = o8 * 8 + e;
        if (o < d.out_features) {
          float w = dY[row + o];
          if ((d.flags & KV_FLAG_RELU) && !(act_mask[row + o] > 0.0f)) w = 0.0f;
          v[e] = w;
        }
      }
    }
    *reinterpret_cast<uint4*>(img + idx * 8) =
        make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
  }
}

// transposed weight image for the input gradient (default effkan on a uniform grid only):
//   [g][tile][stage][CTA rank][sub j][o8 (8)][n8 (SW/16)][8 columns][8 outputs]
// One thread = one core matrix: 8 columns (the 8 B-spline coefficients of one feature, or 8 features of the
// SiLU branch) x 8 outputs, read as 8 x 32 contiguous bytes and written as 128 contiguous bytes.
__global__ void tc_pack_t_kernel(const kv_kan_desc_t d, const kv_kan_desc_t d0, const uint64_t* __restrict__ tbl,
                                 __nv_bfloat16* __restrict__ wtt, BwdGeom b, int layout) {
  const int xs = d.n_groups == d0.n_groups ? 1 : d0.x_share;
  const size_t total64 = (size_t)b.nsplit * d.n_groups * b.n_tiles * b.n_stages * b.tile_cols();  // core matrices
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total64;
       idx += (size_t)gridDim.x * blockDim.x) {
    size_t r = idx;
    // per (stage, CTA rank): [MMA 0: o8 (8)][n8 (W0 / 16)] then [MMA 1: o8 (8)][n8 (W1 / 16)] core matrices; CTA `rank`
    // stages rows [rank * W / 2, (rank + 1) * W / 2) of each MMA's W columns
    const int W0 = b.W0(), W1 = b.W1, per = (W0 + W1) / 2;
    int qi = (int)(r % per); r /= per;
    const int rank = r % 2; r /= 2;
    const int st = r % b.n_stages; r /= b.n_stages;
    const int tile = r % b.n_tiles; r /= b.n_tiles;
    const int g = (int)(r % d.n_groups);
    const int sp = (int)(r / d.n_groups);   // basis split (M > 32): bases [sp * Ms, (sp + 1) * Ms)
    int n8, o8, C;   // C = first tile column of this core matrix
    if (qi < W0 / 2) { n8 = qi % (W0 / 16); o8 = qi / (W0 / 16); C = rank * (W0 / 2) + n8 * 8; }
    else { qi -= W0 / 2; n8 = qi % (W1 / 16); o8 = qi / (W1 / 16); C = W0 + rank * (W1 / 2) + n8 * 8; }
    const int j = C / b.SW, cc = C - j * b.SW;
    const int c = tile * b.n_sub + j;
    const int q = sp * b.Ms + cc / 16, h = (cc / 8) & 1;
    const bool q_ok = b.nsplit == 1 || q < d.M;
    uint32_t w[8][4];  // [column r8][4 output pairs]
    float v[8][8];     // [output oe][column r8]
#pragma unroll
    for (int oe = 0; oe < 8; ++oe) {
      const int o = st * 64 + o8 * 8 + oe;
      if (o < d.out_features && c < b.n_chunks && q_ok) {
        load_w8(d0, tbl, g * xs + o / d0.out_features, o % d0.out_features, layout, c, q, h, v[oe]);
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) v[oe][e] = 0.0f;
      }
    }
#pragma unroll
    for (int r8 = 0; r8 < 8; ++r8)
#pragma unroll
      for (int k = 0; k < 4; ++k) w[r8][k] = pack_bf16_pair(v[2 * k][r8], v[2 * k + 1][r8]);
    uint4* dst = reinterpret_cast<uint4*>(wtt + idx * 64);
#pragma unroll
    for (int r8 = 0; r8 < 8; ++r8) dst[r8] = make_uint4(w[r8][0], w[r8][1], w[r8][2], w[r8][3]);
  }
}

struct BwdInParams {
  CUtensorMap tm_dy;     // dY image, box {32 words, 8 o8 (one 64-output stage), 16 t8 (this CTA's 128 rows)}
  kv_kan_desc_t d;
  const float* X;
  const __nv_bfloat16* dyimg;
  const __nv_bfloat16* wtt;
  const float* fparams;
  float* dX;
  BwdGeom b;
  // generic kernel only
  const float* mean;
  const float* rstd;
  float* dXn;
  float* dfparams;
  int tpc;  // effkan8 kernel: column tiles per CTA pair
  int m_lo, m_cnt, accum;   // generic kernel: this launch's bases [m_lo, m_lo + m_cnt); accum: add to dX / dXn
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NTHREADS, 1) effkan8_bwd_input_kernel(const __grid_constant__ BwdInParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;
  uint8_t* sB = smem + BI_STAGES * BI_A_STAGE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + BI_STAGES * (BI_A_STAGE + BI_B_STAGE));
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto full = [&](int s) { return bar0 + 8u * s; };
  auto empty = [&](int s) { return bar0 + 8u * (BI_STAGES + s); };
  const uint32_t tmem_full = bar0 + 8u * (2 * BI_STAGES);
  const uint32_t tmem_empty = bar0 + 8u * (2 * BI_STAGES + 1);  // (leader) the 16 epilogue warps of the pair have drained TMEM
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * BI_STAGES + 2);
  float* win_base = reinterpret_cast<float*>(smem + BI_STAGES * (BI_A_STAGE + BI_B_STAGE) + BAR_BYTES);  // [14][256]

  const kv_kan_desc_t& d = p.d;
  const BwdGeom b = p.b;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  // one CTA pair = one 256-row block x `tpc` consecutive column tiles (set-up, TMEM allocation and pipeline fill
  // are paid once; the loader runs ahead into the next tile while the epilogue drains the accumulator)
  const int n_tgroups = (b.n_tiles + p.tpc - 1) / p.tpc;
  const int tile0 = ((blockIdx.x >> 1) % n_tgroups) * p.tpc;
  const int tile1 = min(tile0 + p.tpc, b.n_tiles);
  const int m_tile = ((blockIdx.x >> 1) / n_tgroups) * 2 + (int)rank;
  const int g = blockIdx.z;
  const int t0 = m_tile * BM;
  const int SW = b.SW, n_sub = b.n_sub;
  const uint32_t b_stage_bytes = (uint32_t)n_sub * SW * 64;  // this CTA's half: n_sub * SW/2 rows * 64 * 2 B

  if (warp == 1 && lane == 0) {
    for (int s = 0; s < BI_STAGES; ++s) {
      ptx::mbar_init(full(s), rank == 0 ? 2 : 1);
      ptx::mbar_init(empty(s), 1);
    }
    ptx::mbar_init(tmem_full, 1);
    ptx::mbar_init(tmem_empty, 2 * N_PRODUCER_WARPS);
    ptx::fence_mbar_init();
  }
  if (warp == 0) {
    ptx::tmem_alloc2(ptx::smem_u32(tmem_slot), TMEM_COLS);
    ptx::tmem_relinquish2();
  }
  ptx::tc_fence_before();
  ptx::cluster_sync_all();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // ---- loader: per stage 16 x 1 KB of the dY image (this CTA's 128 rows) + one copy of the weight image
      int s = 0;
      uint32_t ph = 1;
      for (int tile = tile0; tile < tile1; ++tile) {
      const uint8_t* bsrc = reinterpret_cast<const uint8_t*>(p.wtt) +
                            (((size_t)g * b.n_tiles + tile) * b.n_stages * 2 + rank) * (size_t)b_stage_bytes;
      for (int st = 0; st < b.n_stages; ++st) {
        ptx::mbar_wait(empty(s), ph);
        ptx::mbar_arrive_expect_tx(full(s), BI_A_STAGE + b_stage_bytes);
        const uint32_t adst = ptx::smem_u32(sA) + s * BI_A_STAGE;
        ptx::tma_load_3d(adst, &p.tm_dy, 0, st * 8, g * b.T8 + m_tile * 16, full(s));
        ptx::bulk_g2s(ptx::smem_u32(sB) + s * BI_B_STAGE, bsrc + (size_t)st * 2 * b_stage_bytes, b_stage_bytes,
                      full(s));
        if (++s == BI_STAGES) { s = 0; ph ^= 1; }
      }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (rank == 0) {
      {   // the whole converged warp walks the issue loop, election inside the asm blocks (see kan_tc_fwd_kernel)
      // ---- MMA issuer: S[256 rows, n_sub x SW] += dY[256, 16] . Wt^T, the tile's columns as one or two wide MMAs
      const int W0 = b.W0();
      const uint32_t idesc = ptx::make_idesc_bf16(2 * BM, W0);
      const uint32_t idesc1 = b.W1 ? ptx::make_idesc_bf16(2 * BM, b.W1) : 0u;
      const uint64_t da0 = ptx::make_smem_desc(ptx::smem_u32(sA), 128, 1024);
      const uint64_t db0 = ptx::make_smem_desc(ptx::smem_u32(sB), (uint32_t)W0 * 8, 128);
      const uint64_t db1_0 = ptx::make_smem_desc(ptx::smem_u32(sB) + W0 * 64, (uint32_t)b.W1 * 8, 128);
      const uint32_t a_hi = (uint32_t)(da0 >> 32), b_hi = (uint32_t)(db0 >> 32), b1_hi = (uint32_t)(db1_0 >> 32);
      const uint32_t a_lo0 = (uint32_t)da0, b_lo0 = (uint32_t)db0, b1_lo0 = (uint32_t)db1_0;
      const uint32_t b_k_step = (uint32_t)(2 * W0 * 8) >> 4;   // next K=16 step: two 8-wide core columns
      const uint32_t b_k_step1 = (uint32_t)(2 * b.W1 * 8) >> 4;
      const bool two = b.W1 != 0;
      int s = 0;
      uint32_t ph = 0;
      for (int tile = tile0; tile < tile1; ++tile) {
      if (tile > tile0) {  // the accumulator of the previous tile must have been read out by both CTAs
        ptx::mbar_wait_cluster(tmem_empty, (uint32_t)((tile - tile0 - 1) & 1));
        ptx::tc_fence_after();
      }
      uint32_t acc = 0;
      for (int st = 0; st < b.n_stages; ++st) {
        ptx::mbar_wait(full(s), ph);
        ptx::tc_fence_after();
        const uint32_t a_lo = a_lo0 + (uint32_t)((s * BI_A_STAGE) >> 4);
        const uint32_t b_lo = b_lo0 + (uint32_t)((s * BI_B_STAGE) >> 4);
        const uint32_t b1_lo = b1_lo0 + (uint32_t)((s * BI_B_STAGE) >> 4);
#pragma unroll
        for (int ks = 0; ks < BI_KS; ++ks) {
          ptx::mma2_bf16_ss_lohi_elect(tmem_base, a_lo + ks * 16, a_hi, b_lo + ks * b_k_step, b_hi, idesc, acc);
          if (two) ptx::mma2_bf16_ss_lohi_elect(tmem_base + 256, a_lo + ks * 16, a_hi, b1_lo + ks * b_k_step1, b1_hi, idesc1, acc);
          acc = 1;
        }
        ptx::tc_commit2_elect(empty(s));
        if (st + 1 == b.n_stages) ptx::tc_commit2_elect(tmem_full);
        if (++s == BI_STAGES) { s = 0; ph ^= 1; }
      }
      }
      }
    } else if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      for (int st = 0; st < b.n_stages * (tile1 - tile0); ++st) {
        ptx::mbar_wait(full(s), ph);
        ptx::mbar_arrive_cluster(ptx::mapa(full(s), 0));
        if (++s == BI_STAGES) { s = 0; ph ^= 1; }
      }
    }
    __syncwarp();
  } else {
    // ---- epilogue: dX[t, i] = sum_e S[.., e] * B'_e(x) + S_base * silu'(x)
    const int pw = warp - 2;
    const int q4 = warp & 3, half = pw >> 2;
    const int te = t0 + q4 * 32 + lane;
    const bool row_ok = te < d.T;
    const int in = d.in_features;
    const float* probe = p.fparams + (size_t)g * d.fparam_stride + (size_t)in * (d.grid_size + 2 * d.spline_order + 1);
    const float gt0 = probe[1], ginv_h = probe[2];
    const int xoff = g * in;  // x_share == 1: group g reads its own slice
    const float* xrow = p.X + (size_t)(row_ok ? te : 0) * d.ldx + xoff;
    float* dxrow = p.dX + (size_t)(row_ok ? te : 0) * d.ldx + xoff;
    const bool vec = ((d.ldx & 3) == 0) && ((xoff & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.X) & 15) == 0) &&
                     ((reinterpret_cast<uintptr_t>(p.dX) & 15) == 0);
    float* win = win_base + (pw * 32 + lane);  // this thread's zero-padded window [14][256]
    win[0] = win[256] = win[512] = 0.0f;
    win[11 * 256] = win[12 * 256] = win[13 * 256] = 0.0f;
    for (int tile = tile0; tile < tile1; ++tile) {
    // prefetch the inputs of every chunk of this tile while the MMAs run
    float xs[3][8];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const int c = tile * n_sub + j;
      if (j < n_sub && c < b.n_chunks) load_x8(xrow, c * KU + half * 8, in, row_ok, vec, xs[j]);
    }
    ptx::mbar_wait(tmem_full, (uint32_t)((tile - tile0) & 1));
    ptx::tc_fence_after();
    const uint32_t trow = tmem_base + ((uint32_t)(q4 * 32) << 16);
    constexpr bool kSmemWindow = true;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const int c = tile * n_sub + j;
      if (j >= n_sub || c >= b.n_chunks || (d.reserved & 512)) continue;   // warp-uniform
      uint32_t sa[32], sb2[32], sbase[8];
      ptx::tmem_ld32(trow + j * SW + half * 64, sa);        // units 4*half, 4*half+1
      ptx::tmem_ld32(trow + j * SW + half * 64 + 32, sb2);  // units 4*half+2, 4*half+3
      ptx::tmem_ld8(trow + j * SW + 128 + half * 8, sbase);
      ptx::tmem_ld_wait();
      float dx[8];
#pragma unroll
      for (int f = 0; f < 8; ++f) {
        const float xv = xs[j][f];
        const float u = (xv - gt0) * ginv_h;
        const float jf = floorf(u);
        const float s = u - jf;
        const int j3 = (int)jf - 3;
        const bool ok = (u >= 0.0f) && (u < 11.0f);
        const float om = 1.0f - s, s2 = s * s;
        // d/dx of the four cubic pieces (times 1/h)
        const float w0 = ok ? -0.5f * om * om * ginv_h : 0.0f;
        const float w1 = ok ? (1.5f * s2 - 2.0f * s) * ginv_h : 0.0f;
        const float w2 = ok ? (-1.5f * s2 + s + 0.5f) * ginv_h : 0.0f;
        const float w3 = ok ? 0.5f * s2 * ginv_h : 0.0f;
        float acc = 0.0f;
        if (kSmemWindow) {
          // the 8 spline sums go into this thread's zero-padded window [3, 11) and the four that meet the
          // non-zero derivative pieces are read back at j3 + 3 .. j3 + 6: 12 conflict-free shared-memory
          // accesses instead of a 32-way select network
#pragma unroll
          for (int e = 0; e < 8; ++e)
            win[(3 + e) * 256] = __uint_as_float((f < 4) ? sa[f * 8 + e] : sb2[(f - 4) * 8 + e]);
          const int jc = min(max(j3, -3), 7) + 3;
          const float* wp = win + jc * 256;
          acc = fmaf(w0, wp[0], fmaf(w1, wp[256], fmaf(w2, wp[512], w3 * wp[768])));
        } else {
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            // coefficient of spline e: piece e - j3 of the derivative (branch-free selects)
            const float coef = sel_eq(j3, e, w0, sel_eq(j3, e - 1, w1, sel_eq(j3, e - 2, w2, sel_eq(j3, e - 3, w3, 0.0f))));
            const uint32_t sv = (f < 4) ? sa[f * 8 + e] : sb2[(f - 4) * 8 + e];
            acc = fmaf(coef, __uint_as_float(sv), acc);
          }
        }
        const float sg = __fdividef(1.0f, 1.0f + __expf(-xv));
        acc = fmaf(sg * (1.0f + xv * (1.0f - sg)), __uint_as_float(sbase[f]), acc);
        dx[f] = acc;
      }
      if (row_ok) {
        const int i0 = c * KU + half * 8;
        if (vec && i0 + 8 <= in) {
          *reinterpret_cast<float4*>(dxrow + i0) = make_float4(dx[0], dx[1], dx[2], dx[3]);
          *reinterpret_cast<float4*>(dxrow + i0 + 4) = make_float4(dx[4], dx[5], dx[6], dx[7]);
        } else {
#pragma unroll
          for (int f = 0; f < 8; ++f)
            if (i0 + f < in) dxrow[i0 + f] = dx[f];
        }
      }
    }
    ptx::tc_fence_before();
    __syncwarp();
    if (lane == 0) ptx::mbar_arrive_cluster(ptx::mapa(tmem_empty, 0));  // this warp has drained its TMEM lanes
    }
  }
  ptx::cluster_sync_all();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc2(tmem_base, TMEM_COLS);
  }
}


// S values staged in shared memory, element m at s[m * stride]
// (only the bases mlo <= m < mhi are staged: the scratch holds at most 24 of them, larger M take two passes)
struct StridedDotSink {
  static constexpr bool kDense = false;
  const float* s;
  int stride, mlo, mhi;
  float acc;
  __device__ __forceinline__ void put(int m, float v) {
    if (m >= mlo && m < mhi) acc = fmaf(v, s[(m - mlo) * stride], acc);
  }
};
struct StridedFastkanDotSink {  // RBF part (m < G) and SiLU part (m == G) kept apart (LayerNorm backward needs the first)
  static constexpr bool kDense = false;
  const float* s;
  int stride, G, mlo, mhi;
  float acc_rbf, acc_base;
  __device__ __forceinline__ void put(int m, float v) {
    if (m < mlo || m >= mhi) return;
    if (m < G) acc_rbf = fmaf(v, s[(m - mlo) * stride], acc_rbf); else acc_base = fmaf(v, s[(m - mlo) * stride], acc_base);
  }
};

template <int F>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NTHREADS, 1) kan_tc_bwd_input_generic_kernel(const __grid_constant__ BwdInParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;
  uint8_t* sB = smem + BI_STAGES * BI_A_STAGE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + BI_STAGES * (BI_A_STAGE + BI_B_STAGE));
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto full = [&](int s) { return bar0 + 8u * s; };
  auto empty = [&](int s) { return bar0 + 8u * (BI_STAGES + s); };
  const uint32_t tmem_full = bar0 + 8u * (2 * BI_STAGES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * BI_STAGES + 1);

  const kv_kan_desc_t& d = p.d;
  const BwdGeom b = p.b;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const int tile = (blockIdx.x >> 1) % b.n_tiles;
  const int m_tile = ((blockIdx.x >> 1) / b.n_tiles) * 2 + (int)rank;
  const int g = blockIdx.z;
  const int t0 = m_tile * BM;
  const int SW = b.SW, n_sub = b.n_sub;
  const uint32_t b_stage_bytes = (uint32_t)n_sub * SW * 64;  // this CTA's half: n_sub * SW/2 rows * 64 * 2 B

  if (warp == 1 && lane == 0) {
    for (int s = 0; s < BI_STAGES; ++s) {
      ptx::mbar_init(full(s), rank == 0 ? 2 : 1);
      ptx::mbar_init(empty(s), 1);
    }
    ptx::mbar_init(tmem_full, 1);
    ptx::fence_mbar_init();
  }
  float* dfreq_s = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 512);  // SineKAN: d freq of this CTA
  if (warp == 2) dfreq_s[lane] = dfreq_s[lane + 32] = 0.0f;
  if (warp == 0) {
    ptx::tmem_alloc2(ptx::smem_u32(tmem_slot), TMEM_COLS);
    ptx::tmem_relinquish2();
  }
  ptx::tc_fence_before();
  ptx::cluster_sync_all();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // ---- loader: per stage 16 x 1 KB of the dY image (this CTA's 128 rows) + one copy of the weight image
      const uint8_t* bsrc = reinterpret_cast<const uint8_t*>(p.wtt) +
                            (((size_t)g * b.n_tiles + tile) * b.n_stages * 2 + rank) * (size_t)b_stage_bytes;
      int s = 0;
      uint32_t ph = 1;
      for (int st = 0; st < b.n_stages; ++st) {
        ptx::mbar_wait(empty(s), ph);
        ptx::mbar_arrive_expect_tx(full(s), BI_A_STAGE + b_stage_bytes);
        const uint32_t adst = ptx::smem_u32(sA) + s * BI_A_STAGE;
        ptx::tma_load_3d(adst, &p.tm_dy, 0, st * 8, g * b.T8 + m_tile * 16, full(s));
        ptx::bulk_g2s(ptx::smem_u32(sB) + s * BI_B_STAGE, bsrc + (size_t)st * 2 * b_stage_bytes, b_stage_bytes,
                      full(s));
        if (++s == BI_STAGES) { s = 0; ph ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (rank == 0) {
      {   // the whole converged warp walks the issue loop, election inside the asm blocks (see kan_tc_fwd_kernel)
      // ---- MMA issuer: S[256 rows, tile columns] += dY[256, 16] . Wt^T, the tile's columns as one or two wide MMAs
      const int W0 = b.W0();
      const uint32_t idesc = ptx::make_idesc_bf16(2 * BM, W0);
      const uint32_t idesc1 = b.W1 ? ptx::make_idesc_bf16(2 * BM, b.W1) : 0u;
      const uint64_t da0 = ptx::make_smem_desc(ptx::smem_u32(sA), 128, 1024);
      const uint64_t db0 = ptx::make_smem_desc(ptx::smem_u32(sB), (uint32_t)W0 * 8, 128);
      const uint64_t db1_0 = ptx::make_smem_desc(ptx::smem_u32(sB) + W0 * 64, (uint32_t)b.W1 * 8, 128);
      const uint32_t a_hi = (uint32_t)(da0 >> 32), b_hi = (uint32_t)(db0 >> 32), b1_hi = (uint32_t)(db1_0 >> 32);
      const uint32_t a_lo0 = (uint32_t)da0, b_lo0 = (uint32_t)db0, b1_lo0 = (uint32_t)db1_0;
      const uint32_t b_k_step = (uint32_t)(2 * W0 * 8) >> 4;   // next K=16 step: two 8-wide core columns
      const uint32_t b_k_step1 = (uint32_t)(2 * b.W1 * 8) >> 4;
      const bool two = b.W1 != 0;
      int s = 0;
      uint32_t ph = 0, acc = 0;
      for (int st = 0; st < b.n_stages; ++st) {
        ptx::mbar_wait(full(s), ph);
        ptx::tc_fence_after();
        const uint32_t a_lo = a_lo0 + (uint32_t)((s * BI_A_STAGE) >> 4);
        const uint32_t b_lo = b_lo0 + (uint32_t)((s * BI_B_STAGE) >> 4);
        const uint32_t b1_lo = b1_lo0 + (uint32_t)((s * BI_B_STAGE) >> 4);
#pragma unroll
        for (int ks = 0; ks < BI_KS; ++ks) {
          ptx::mma2_bf16_ss_lohi_elect(tmem_base, a_lo + ks * 16, a_hi, b_lo + ks * b_k_step, b_hi, idesc, acc);
          if (two) ptx::mma2_bf16_ss_lohi_elect(tmem_base + 256, a_lo + ks * 16, a_hi, b1_lo + ks * b_k_step1, b1_hi, idesc1, acc);
          acc = 1;
        }
        ptx::tc_commit2_elect(empty(s));
        if (st + 1 == b.n_stages) ptx::tc_commit2_elect(tmem_full);
        if (++s == BI_STAGES) { s = 0; ph ^= 1; }
      }
      }
    } else if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      for (int st = 0; st < b.n_stages; ++st) {
        ptx::mbar_wait(full(s), ph);
        ptx::mbar_arrive_cluster(ptx::mapa(full(s), 0));
        if (++s == BI_STAGES) { s = 0; ph ^= 1; }
      }
    }
    __syncwarp();
  } else {
    // ---- epilogue (any family): dX[t, i] = sum_m Phi'_m(x[t,i]) * S[t, (i, m)], basis-major columns m * 16 + f.
    // The S values of one chunk go through shared memory (the operand rings are idle by now), laid out
    // [m][f][thread] so that the family's derivative routine can index them by m.
    const int pw = warp - 2;
    const int q4 = warp & 3, half = pw >> 2;
    const int etid = pw * 32 + lane;
    const int te = t0 + q4 * 32 + lane;
    const bool row_ok = te < d.T;
    const int in = d.in_features, M = d.M;
    const BasisCtx c = make_ctx_tc(d, p.fparams, g);
    const int n_slices = d.n_groups / d.x_share, slice = g / d.x_share;
    const int xoff = slice * in;
    const bool shared_slice = d.x_share > 1;  // several groups add into the same dX columns (zeroed by the host)
    const float* xrow = p.X + (size_t)(row_ok ? te : 0) * d.ldx + xoff;
    float* dxrow = p.dX + (size_t)(row_ok ? te : 0) * d.ldx + xoff;
    const bool vec = ((d.ldx & 3) == 0) && ((xoff & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.X) & 15) == 0) &&
                     ((reinterpret_cast<uintptr_t>(p.dX) & 15) == 0);
    float mu = 0.f, rs = 1.f;
    const bool ln = (F == KV_FASTKAN) && !(d.flags & KV_FLAG_NO_LAYERNORM);
    if (ln && row_ok) {
      mu = p.mean[(size_t)te * n_slices + slice];
      rs = p.rstd[(size_t)te * n_slices + slice];
    }
    float* scratch = reinterpret_cast<float*>(smem);
    constexpr int MS = 8 * 256;  // scratch stride between consecutive m
    ptx::mbar_wait(tmem_full, 0);
    ptx::tc_fence_after();
    const uint32_t trow = tmem_base + ((uint32_t)(q4 * 32) << 16);
    for (int j = 0; j < n_sub; ++j) {
      const int cch = tile * n_sub + j;
      if (cch >= b.n_chunks) break;  // warp-uniform
      float x8[8];
      load_x8(xrow, cch * KU + half * 8, in, row_ok, vec, x8);
      float dx[8], dxn_acc[8];
#pragma unroll
      for (int f = 0; f < 8; ++f) dx[f] = dxn_acc[f] = 0.0f;
      // this launch's bases [m_lo, m_lo + Mloc) sit at TMEM columns (m - m_lo) * 16 + f of the chunk
      const int m_lo = p.m_lo, Mloc = p.m_cnt;
      const int n_pass = Mloc > 24 ? 2 : 1, Mh = (Mloc + n_pass - 1) / n_pass;
      for (int pass = 0; pass < n_pass; ++pass) {
      const int mlo = m_lo + pass * Mh, mhi = min(m_lo + Mloc, mlo + Mh);
      for (int m = mlo; m < mhi; ++m) {
        uint32_t r8[8];
        ptx::tmem_ld8(trow + j * SW + (m - m_lo) * 16 + half * 8, r8);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int f = 0; f < 8; ++f) scratch[(m - mlo) * MS + f * 256 + etid] = __uint_as_float(r8[f]);
      }
      if (F == KV_SINE) {
        // basis outermost: d freq_q (models/sinekan.py:87) is summed over the thread's 8 features before the warp
        // reduction -- one shuffle tree and one shared-memory atomic per basis instead of one per (feature, basis)
        const float* freq = c.fp;
        const int i0s = cch * KU + half * 8;
        for (int q = mlo; q < min(mhi, c.G); ++q) {
          const float fr = freq[q];
          float contrib = 0.0f;
#pragma unroll
          for (int f = 0; f < 8; ++f) {
            const int ic = min(i0s + f, in - 1);
            const float cs = cosf(fmaf(x8[f], fr, c.fp[c.G + (size_t)ic * c.G + q]));
            const float sv = scratch[(q - mlo) * MS + f * 256 + etid];
            dx[f] = fmaf(fr * cs, sv, dx[f]);
            if (row_ok && i0s + f < in) contrib = fmaf(x8[f] * cs, sv, contrib);
          }
          contrib = warp_sum(contrib);
          if (lane == 0) atomicAdd(&dfreq_s[q], contrib);
        }
      } else
#pragma unroll
      for (int f = 0; f < 8; ++f) {
        const int i = cch * KU + half * 8 + f;
        const int ic = min(i, in - 1);
        const bool live = row_ok && i < in;
        const float xv = x8[f];
        const float* srow = scratch + f * 256 + etid;
        float acc = 0.0f;
        (void)live;
        if (F == KV_FASTKAN) {
          const float xn = fastkan_xn(c, ic, xv, mu, rs);
          StridedFastkanDotSink sk{srow, MS, c.G, mlo, mhi, 0.0f, 0.0f};
          basis_deriv<F>(c, ic, xv, xn, sk);
          acc = sk.acc_base;
          if (!ln) acc += sk.acc_rbf;
          else dxn_acc[f] += sk.acc_rbf;
        } else {
          StridedDotSink sk{srow, MS, mlo, mhi, 0.0f};
          basis_deriv<F>(c, ic, xv, xv, sk);
          acc = sk.acc;
        }
        dx[f] += acc;
      }
      }
      if (ln) {
#pragma unroll
        for (int f = 0; f < 8; ++f) {
          const int i = cch * KU + half * 8 + f;
          if (row_ok && i < in) {
            float* dst = &p.dXn[(size_t)te * ((size_t)d.n_groups * in) + (size_t)g * in + i];
            *dst = p.accum ? *dst + dxn_acc[f] : dxn_acc[f];
          }
        }
      }
      if (row_ok && shared_slice) {
        const int i0 = cch * KU + half * 8;
#pragma unroll
        for (int f = 0; f < 8; ++f)
          if (i0 + f < in) atomicAdd(dxrow + i0 + f, dx[f]);
      } else if (row_ok) {
        const int i0 = cch * KU + half * 8;
        if (p.accum) {     // a later basis split: this thread owns the element, the launches are stream-ordered
#pragma unroll
          for (int f = 0; f < 8; ++f)
            if (i0 + f < in) dxrow[i0 + f] += dx[f];
        } else if (vec && i0 + 8 <= in) {
          *reinterpret_cast<float4*>(dxrow + i0) = make_float4(dx[0], dx[1], dx[2], dx[3]);
          *reinterpret_cast<float4*>(dxrow + i0 + 4) = make_float4(dx[4], dx[5], dx[6], dx[7]);
        } else {
#pragma unroll
          for (int f = 0; f < 8; ++f)
            if (i0 + f < in) dxrow[i0 + f] = dx[f];
        }
      }
    }
    if (F == KV_SINE && p.dfparams != nullptr) {
      asm volatile("bar.sync 1, 256;" ::: "memory");  // the eight epilogue warps
      if (etid < c.G) atomicAdd(&p.dfparams[(size_t)g * d.fparam_stride + etid], dfreq_s[etid]);
    }
    ptx::tc_fence_before();
  }
  ptx::cluster_sync_all();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc2(tmem_base, TMEM_COLS);
  }
}



// =========================================================================================
// Weight gradient on the tensor cores:  dWp^T[K', out] = Phi^T[K', T] . dY[T, out]  (contraction over T)
// The expensive, recomputed operand (the basis) is the M side: one CTA owns 128 basis columns K'
// (default effkan: the 8 B-splines of one 16-feature chunk, or the SiLU column of 128 features; Linear:
// 128 features), the pair 256, and reuses every value against up to 512 outputs.  Both operands are
// "MN-major" (the contraction index t is the strided one): A = the same 16-byte basis vectors the forward
// producers build, B = the bf16 dY image.  T is split over CTAs; partial sums are added with fp32 RED.
// =========================================================================================
constexpr int BW_A_STAGE = 16 * 1024;   // 64 t x 128 K' x 2 B
constexpr int BW_B_STAGE = 32 * 1024;   // 64 t x BN/2 x 2 B
constexpr int BW_NA = 3, BW_NB = 4;
constexpr int BW_NX = 4;                // X ring (spline tiles): TMA-staged fp32 tiles of 64 rows x 16 features
constexpr int BW_X_STAGE = 64 * 16 * 4; // 4 KB, SWIZZLE_64B
constexpr int BW_SMEM = BW_NA * BW_A_STAGE + BW_NB * BW_B_STAGE + BW_NX * BW_X_STAGE + BAR_BYTES;

struct BwdWParams {
  CUtensorMap tm_dy1, tm_dy2;   // dY image boxes {32 words, n1/16 | n2/16 o8 (this CTA's half of a sub-MMA), 8 t8 (64 rows)}
  CUtensorMap tm_x;             // X [T][ldx] fp32, box {16 features, 64 rows}, SWIZZLE_64B (spline tiles; use_tma_x)
  int use_tma_x;
  kv_kan_desc_t d;
  const float* X;
  const __nv_bfloat16* dyimg;
  const float* fparams;
  float* dWp;
  BwdGeom b;
  Geom gm;       // BN / n_tiles of the output dimension
  int KT;        // number of 128-wide K' tiles
  int n_kpairs;  // ceil(KT / 2)
  int tsplits, rows_per_split;
  // generic tiles (every family but the default effkan / Linear): a tile = `no` feature octets x `Mt` bases,
  // block b (8 columns) = octet b / Mt, basis m0 + b % Mt;  M > 16 splits the bases over `tpo` tiles per octet
  int generic, no, tpo;
  const float* mean;
  const float* rstd;
};

// NPW = producer / epilogue warps: 8, or 16 (twice the warps per scheduler: the producers' dependent cubic / shift /
// pack chains, not the tensor pipe, pace this kernel -- ~2400 cycles per 64-row stage against 768-1024 of MMA work)
template <int F, int NPW>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(32 * (2 + NPW), 1) kan_tc_bwd_weight_kernel(const __grid_constant__ BwdWParams p) {
  constexpr int NFB = NPW / 2;          // feature blocks: a stage's 64 rows x NFB thread columns
  constexpr int SPF = 16 / NFB;         // spline tile: features per thread (4 | 2)
  constexpr int SLO = 16 / NFB;         // SiLU / Linear tile: 8-feature octets per thread (4 | 2)
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;
  uint8_t* sB = smem + BW_NA * BW_A_STAGE;
  uint8_t* sX = smem + BW_NA * BW_A_STAGE + BW_NB * BW_B_STAGE;   // 1024-byte aligned (swizzle atom = 512 B)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + BW_NA * BW_A_STAGE + BW_NB * BW_B_STAGE + BW_NX * BW_X_STAGE);
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto fullA = [&](int s) { return bar0 + 8u * s; };
  auto emptyA = [&](int s) { return bar0 + 8u * (BW_NA + s); };
  auto fullB = [&](int s) { return bar0 + 8u * (2 * BW_NA + s); };
  auto emptyB = [&](int s) { return bar0 + 8u * (2 * BW_NA + BW_NB + s); };
  const uint32_t tmem_full = bar0 + 8u * (2 * BW_NA + 2 * BW_NB);
  auto fullX = [&](int s) { return bar0 + 8u * (2 * BW_NA + 2 * BW_NB + 1 + s); };
  auto emptyX = [&](int s) { return bar0 + 8u * (2 * BW_NA + 2 * BW_NB + 1 + BW_NX + s); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * BW_NA + 2 * BW_NB + 1 + 2 * BW_NX);

  const kv_kan_desc_t& d = p.d;
  const BwdGeom b = p.b;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const int n_tile = (blockIdx.x >> 1) % p.gm.n_tiles;
  const int kt = ((blockIdx.x >> 1) / p.gm.n_tiles) * 2 + (int)rank;   // this CTA's K' tile
  const int g = blockIdx.z / p.tsplits, ts = blockIdx.z % p.tsplits;
  const int BN = p.gm.BN;
  const int n1 = BN > 256 ? 256 : BN, n2 = BN - n1;
  const int row0 = ts * p.rows_per_split;
  const int row1 = min(d.T, row0 + p.rows_per_split);
  const int n_stages = (row1 - row0 + 63) / 64;
  const uint32_t b_stage_bytes = (uint32_t)BN * 64;
  const int in = d.in_features;
  const bool tile_live = kt < p.KT;
  // tile kind: effkan spline tile (chunk kt) | effkan SiLU tile | linear feature tile
  const bool generic = p.generic != 0;
  const bool spline_tile = !generic && (F == KV_EFFKAN) && kt < b.n_chunks;
  const int feat0 = spline_tile ? kt * KU : ((F == KV_EFFKAN) ? (kt - b.n_chunks) * 128 : kt * 128);
  // generic tile: octets [oc0, oc0 + no), bases [m0, m0 + Mt)
  const int g_oc0 = generic ? (p.tpo > 1 ? kt / p.tpo : kt * p.no) : 0;
  const int g_m0 = (generic && p.tpo > 1) ? (kt % p.tpo) * 16 : 0;
  const int g_Mt = generic ? (p.tpo > 1 ? min(16, d.M - g_m0) : d.M) : 1;

  if (warp == 1 && lane == 0) {
    for (int s = 0; s < BW_NA; ++s) {
      ptx::mbar_init(fullA(s), 2 * NPW);
      ptx::mbar_init(emptyA(s), 1);
    }
    for (int s = 0; s < BW_NB; ++s) {
      ptx::mbar_init(fullB(s), rank == 0 ? 2 : 1);
      ptx::mbar_init(emptyB(s), 1);
    }
    ptx::mbar_init(tmem_full, 1);
    for (int s = 0; s < BW_NX; ++s) {
      ptx::mbar_init(fullX(s), 1);
      ptx::mbar_init(emptyX(s), NPW);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 0) {
    ptx::tmem_alloc2(ptx::smem_u32(tmem_slot), TMEM_COLS);
    ptx::tmem_relinquish2();
  }
  ptx::tc_fence_before();
  ptx::cluster_sync_all();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // ---- loader: this CTA's half of the dY rows of each sub-MMA, 8 row blocks per stage
      const int o8_1 = (n_tile * BN + (int)rank * (n1 / 2)) / 8;
      const int o8_2 = (n_tile * BN + n1 + (int)rank * (n2 / 2)) / 8;
      int s = 0, sx = 0;
      uint32_t ph = 1, phx = 1;
      const bool tma_x = p.use_tma_x && tile_live && spline_tile;
      const int xcol = (g / d.x_share) * in + feat0;
      for (int st = 0; st < n_stages; ++st) {
        if (tma_x) {   // this CTA's X tile of the stage: 64 rows x the chunk's 16 features, one tensor copy
          ptx::mbar_wait(emptyX(sx), phx);
          ptx::mbar_arrive_expect_tx(fullX(sx), BW_X_STAGE);
          ptx::tma_load_2d(ptx::smem_u32(sX) + sx * BW_X_STAGE, &p.tm_x, xcol, row0 + st * 64, fullX(sx));
          if (++sx == BW_NX) { sx = 0; phx ^= 1; }
        }
        ptx::mbar_wait(emptyB(s), ph);
        if (d.reserved & 16) {   // timing experiments: no dY traffic
          ptx::mbar_arrive(fullB(s));
        } else {
          ptx::mbar_arrive_expect_tx(fullB(s), b_stage_bytes);
          const uint32_t dst = ptx::smem_u32(sB) + s * BW_B_STAGE;
          const int t8g = g * b.T8 + (row0 + st * 64) / 8;
          ptx::tma_load_3d(dst, &p.tm_dy1, 0, o8_1, t8g, fullB(s));
          if (n2) ptx::tma_load_3d(dst + n1 * 64, &p.tm_dy2, 0, o8_2, t8g, fullB(s));
        }
        if (++s == BW_NB) { s = 0; ph ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (rank == 0) {
      {   // the whole converged warp walks the issue loop, election inside the asm blocks (see kan_tc_fwd_kernel)
      // ---- MMA issuer: D[256 K', BN] += Phi^T[256, 16 t] . dY^T[BN, 16 t]^T, both operands MN-major
      const uint32_t idesc1 = ptx::make_idesc_bf16(2 * BM, n1, 1, 1);
      const uint32_t idesc2 = n2 ? ptx::make_idesc_bf16(2 * BM, n2, 1, 1) : 0u;
      // A step: [t8 (2)][K' block (16)][t%8][16 B]: K-block stride (LBO) 2048, MN-block stride (SBO) 128
      const uint64_t da0 = ptx::make_smem_desc(ptx::smem_u32(sA), 2048, 128);
      // B sub-tile: [t8 (8)][o8 (nj/16)][t%8][16 B]: LBO = nj*8, SBO = 128
      const uint64_t db1_0 = ptx::make_smem_desc(ptx::smem_u32(sB), (uint32_t)n1 * 8, 128);
      const uint64_t db2_0 = ptx::make_smem_desc(ptx::smem_u32(sB) + n1 * 64, (uint32_t)n2 * 8, 128);
      const uint32_t a_hi = (uint32_t)(da0 >> 32), b1_hi = (uint32_t)(db1_0 >> 32), b2_hi = (uint32_t)(db2_0 >> 32);
      const uint32_t a_lo0 = (uint32_t)da0, b1_lo0 = (uint32_t)db1_0, b2_lo0 = (uint32_t)db2_0;
      const uint32_t k1 = (uint32_t)(2 * n1 * 8) >> 4, k2 = (uint32_t)(2 * n2 * 8) >> 4;   // next K=16 step of B
      const bool two = n2 != 0;
      int sa = 0, sb = 0;
      uint32_t pha = 0, phb = 0, acc = 0;
      for (int st = 0; st < n_stages; ++st) {
        ptx::mbar_wait(fullA(sa), pha);
        ptx::mbar_wait(fullB(sb), phb);
        ptx::tc_fence_after();
        const uint32_t a_lo = a_lo0 + (uint32_t)((sa * BW_A_STAGE) >> 4);
        const uint32_t b1 = b1_lo0 + (uint32_t)((sb * BW_B_STAGE) >> 4), b2 = b2_lo0 + (uint32_t)((sb * BW_B_STAGE) >> 4);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          if (d.reserved & 4) break;   // timing experiments: no MMAs
          ptx::mma2_bf16_ss_lohi_elect(tmem_base, a_lo + ks * 256, a_hi, b1 + ks * k1, b1_hi, idesc1, acc);
          if (two) ptx::mma2_bf16_ss_lohi_elect(tmem_base + 256, a_lo + ks * 256, a_hi, b2 + ks * k2, b2_hi, idesc2, acc);
          acc = 1;
        }
        ptx::tc_commit2_elect(emptyA(sa));
        ptx::tc_commit2_elect(emptyB(sb));
        if (st + 1 == n_stages) ptx::tc_commit2_elect(tmem_full);
        if (++sa == BW_NA) { sa = 0; pha ^= 1; }
        if (++sb == BW_NB) { sb = 0; phb ^= 1; }
      }
      }
    } else if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      for (int st = 0; st < n_stages; ++st) {
        ptx::mbar_wait(fullB(s), ph);
        ptx::mbar_arrive_cluster(ptx::mapa(fullB(s), 0));
        if (++s == BW_NB) { s = 0; ph ^= 1; }
      }
    }
    __syncwarp();
  } else {
    // ---- A producers: basis values of 64 rows x this tile's features per stage
    const int ptid = threadIdx.x - 64;
    const int tl = ptid & 63, fb = ptid >> 6;          // row inside the stage, feature block (0..NFB-1)
    const int xoff = (g / d.x_share) * in;
    const bool vec = ((d.ldx & 3) == 0) && ((xoff & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.X) & 15) == 0);
    float gt0 = 0.f, ginv_h = 0.f;
    if (F == KV_EFFKAN) {
      const float* probe = p.fparams + (size_t)g * d.fparam_stride + (size_t)in * (d.grid_size + 2 * d.spline_order + 1);
      gt0 = probe[1];
      ginv_h = probe[2];
    }
    // byte offset of this row inside a K=16 step image: [t8 (2)][block (16)][t%8][16 B]
    const int row_off = (tl >> 4) * 4096 + ((tl >> 3) & 1) * 2048 + (tl & 7) * 16;
    int s = 0;
    uint32_t ph = 1;
    if (generic) {
      // ---- generic producers: thread = (row tl, part fb); part fb evaluates octets fb, fb + 4, ... of the tile
      const BasisCtx c = make_ctx_tc(d, p.fparams, g);
      const int n_slices = d.n_groups / d.x_share, slice = g / d.x_share;
      const bool ln = (F == KV_FASTKAN) && !(d.flags & KV_FLAG_NO_LAYERNORM);
      int s = 0;
      uint32_t ph = 1;
      // A tile with fewer octets than thread parts (M >= 5: 3, 2 or 1 octets for 4 parts) splits the BASES of an octet
      // over npo parts instead of leaving parts idle: part fb owns octet fb % no and the bases [ms, me) of the tile's
      // range (the families' routines emit only the rows of the range they are given).  Such a thread has exactly one
      // octet per stage, whose X values are loaded one stage ahead.
      const int npo = p.no < NFB ? NFB / p.no : 1;
      const bool one_octet = p.no <= NFB;
      const int my_oq = npo > 1 ? fb % p.no : fb;
      const int sub = npo > 1 ? fb / p.no : 0;
      const int per = (g_Mt + npo - 1) / npo;
      const int ms = g_m0 + sub * per, me = min(g_m0 + g_Mt, ms + per);
      const bool part_live = tile_live && my_oq < p.no && sub < npo && ms < me;
      auto rows_of = [&](int st, bool& row_ok) {
        const int t = row0 + st * 64 + tl;
        row_ok = t < row1;
        return p.X + (size_t)(row_ok ? t : 0) * d.ldx + xoff;
      };
      float xcur[8], xnxt[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) xcur[q] = xnxt[q] = 0.0f;
      if (one_octet && part_live && n_stages > 0) {
        bool ok;
        const float* xr = rows_of(0, ok);
        load_x8(xr, (g_oc0 + my_oq) * 8, in, ok && (g_oc0 + my_oq) * 8 < in, vec, xcur);
      }
      for (int st = 0; st < n_stages; ++st) {
        bool row_ok;
        const float* xrow = rows_of(st, row_ok);
        const int t = row0 + st * 64 + tl;
        float mu = 0.f, rs = 1.f;
        if (ln && row_ok) {
          mu = p.mean[(size_t)t * n_slices + slice];
          rs = p.rstd[(size_t)t * n_slices + slice];
        }
        if (one_octet && part_live && st + 1 < n_stages) {
          bool ok;
          const float* xr = rows_of(st + 1, ok);
          load_x8(xr, (g_oc0 + my_oq) * 8, in, ok && (g_oc0 + my_oq) * 8 < in, vec, xnxt);
        }
        ptx::mbar_wait(emptyA(s), ph);
        uint8_t* base = sA + s * BW_A_STAGE + row_off;
        if (one_octet) {
          if (part_live && !(d.reserved & 8))   // (bit 8, timing experiments: no basis evaluation)
            basis_octet<F>(base + my_oq * g_Mt * 128 + (ms - g_m0) * 128, 128, ms, me - ms, xcur, c, (g_oc0 + my_oq) * 8,
                           mu, rs);
#pragma unroll
          for (int q = 0; q < 8; ++q) xcur[q] = xnxt[q];
        } else if (tile_live) {
          for (int oq = fb; oq < p.no; oq += NFB) {
            const int i0 = (g_oc0 + oq) * 8;
            float x8[8];
            load_x8(xrow, i0, in, row_ok && i0 < in, vec, x8);
            basis_octet<F>(base + oq * g_Mt * 128, 128, g_m0, g_Mt, x8, c, i0, mu, rs);
          }
        }
        ptx::fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive_cluster(ptx::mapa(fullA(s), 0));
        if (++s == BW_NA) { s = 0; ph ^= 1; }
      }
    } else {
    // The tile reads a 64-byte (spline) or 512-byte (SiLU / Linear) stripe of every row: a latency-bound gather.
    // Two levels of prefetch: the stripe of stage st + PF is pulled into L2, and the registers of stage st + 1 are
    // loaded BEFORE stage st is evaluated, so the L2 round trip hides under a whole stage instead of stalling
    // every stage (the producers, not the tensor pipe, paced the out = 768 layers).
    constexpr int XN = 8 * SLO;
    auto load_stage = [&](int st, float (&xv)[XN]) {
      if (d.reserved & 256) return;   // timing experiments: no X traffic
      const int t = row0 + st * 64 + tl;
      const bool row_ok = t < row1;
      const float* xrow = p.X + (size_t)(row_ok ? t : 0) * d.ldx + xoff;
      {
        constexpr int PF = 8;
        const int tp = t + PF * 64;
        if (tile_live && tp < row1) {
          const float* pf = p.X + (size_t)tp * d.ldx + xoff + (spline_tile ? feat0 + SPF * fb : feat0 + 8 * SLO * fb);
          if ((spline_tile ? feat0 + SPF * fb : feat0 + 8 * SLO * fb) < in) asm volatile("prefetch.global.L2 [%0];" ::"l"(pf));
        }
      }
      if (!tile_live) return;
      if (spline_tile) {
        // SPF features per thread: features feat0 + SPF*fb .. + SPF-1
        const int i0 = feat0 + SPF * fb;
        if (SPF == 4 && row_ok && vec && i0 + 4 <= in) {
          const float4 a = __ldg(reinterpret_cast<const float4*>(xrow + i0));
          xv[0] = a.x; xv[1] = a.y; xv[2] = a.z; xv[3] = a.w;
        } else if (SPF == 2 && row_ok && vec && i0 + 2 <= in) {
          const float2 a = __ldg(reinterpret_cast<const float2*>(xrow + i0));
          xv[0] = a.x; xv[1] = a.y;
        } else {
#pragma unroll
          for (int q = 0; q < SPF; ++q) xv[q] = (row_ok && i0 + q < in) ? __ldg(xrow + i0 + q) : 0.0f;
        }
      } else {
#pragma unroll
        for (int k = 0; k < SLO; ++k) {
          float t8v[8];
          load_x8(xrow, feat0 + 8 * SLO * fb + 8 * k, in, row_ok, vec, t8v);
#pragma unroll
          for (int q = 0; q < 8; ++q) xv[8 * k + q] = t8v[q];
        }
      }
    };
    // One stage of the A operand from the stage's inputs xv
    auto produce_stage = [&](const float (&xv)[XN]) {
      ptx::mbar_wait(emptyA(s), ph);
      uint8_t* base = sA + s * BW_A_STAGE + row_off;
      if (tile_live && !(d.reserved & 8)) {
        if (spline_tile) {
#pragma unroll
          for (int q = 0; q < SPF; ++q) {
            const float u = (xv[q] - gt0) * ginv_h;
            const float jf = floorf(u);
            const float sfr = u - jf;
            const int j = (int)jf;
            const bool ok = (u >= 0.0f) && (u < 11.0f);
            const float s2 = sfr * sfr, s3 = s2 * sfr, om = 1.0f - sfr;
            const float w0 = om * om * om * (1.0f / 6.0f);
            const float w1 = fmaf(0.5f, s3, fmaf(-1.0f, s2, 2.0f / 3.0f));
            const float w2 = fmaf(-0.5f, s3, fmaf(0.5f, s2, fmaf(0.5f, sfr, 1.0f / 6.0f)));
            const float w3 = s3 * (1.0f / 6.0f);
            uint64_t W = (uint64_t)pack_bf16(w0, w1) | ((uint64_t)pack_bf16(w2, w3) << 32);
            if (!ok) W = 0;
            const int sh = (j - 3) * 16;
            const uint64_t lo = shl64(W, sh) | shr64(W, -sh);
            const uint64_t hi = shl64(W, sh - 64) | shr64(W, 64 - sh);
            // feature f = SPF*fb + q of the chunk owns K' block f (8 splines)
            sts128(base + (SPF * fb + q) * 128, (uint32_t)lo, (uint32_t)(lo >> 32), (uint32_t)hi, (uint32_t)(hi >> 32));
          }
        } else {
#pragma unroll
          for (int k = 0; k < SLO; ++k) {
            float v[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) v[q] = (F == KV_EFFKAN) ? fast_silu(xv[8 * k + q]) : xv[8 * k + q];
            sts128(base + (SLO * fb + k) * 128, pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]),
                   pack_bf16(v[6], v[7]));
          }
        }
      }
      ptx::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(ptx::mapa(fullA(s), 0));
      if (++s == BW_NA) { s = 0; ph ^= 1; }
    };
    float xv[XN], xnext[XN];
#pragma unroll
    for (int q = 0; q < XN; ++q) xv[q] = xnext[q] = 0.0f;
    // Spline tiles with a TMA-able X: the loader thread stages the stage's 64 x 16 fp32 tile in shared memory
    // (cp.async.bulk.tensor, BW_NX stages deep) and every thread picks its SPF values from there.  The per-thread
    // LDG route issues one 32-byte sector request per thread and stage (plus an L2 prefetch each): measured, that
    // request stream -- not the evaluation, not the MMAs -- took 0.8 of this kernel's 2.6 ms (768 -> 3072, T = 66 192).
    const bool tma_x = p.use_tma_x && tile_live && spline_tile;
    int sx = 0;
    uint32_t phx = 0;
    // byte offset of this thread's values inside a stage: row tl, 16-byte chunk XOR-swizzled (SWIZZLE_64B)
    const int xoff_b = tl * 64 + ((((fb * SPF * 4) >> 4) ^ ((tl >> 1) & 3)) << 4) + ((fb * SPF * 4) & 15);
    if (n_stages > 0 && !tma_x) load_stage(0, xv);
    for (int st = 0; st < n_stages; ++st) {
      if (tma_x) {
        ptx::mbar_wait(fullX(sx), phx);
        const uint8_t* xs = sX + sx * BW_X_STAGE + xoff_b;
        if (SPF == 4) {
          const float4 a4 = *reinterpret_cast<const float4*>(xs);
          xv[0] = a4.x; xv[1] = a4.y; xv[2] = a4.z; xv[3] = a4.w;
        } else {
          const float2 a2 = *reinterpret_cast<const float2*>(xs);
          xv[0] = a2.x; xv[1] = a2.y;
        }
#pragma unroll
        for (int q = 0; q < SPF; ++q)
          if (feat0 + SPF * fb + q >= in) xv[q] = 0.0f;   // beyond this group's features (a neighbour's columns)
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(emptyX(sx));
        if (++sx == BW_NX) { sx = 0; phx ^= 1; }
      } else if (st + 1 < n_stages) {
        load_stage(st + 1, xnext);
      }
      produce_stage(xv);
      if (!tma_x) {
#pragma unroll
        for (int q = 0; q < XN; ++q) xv[q] = xnext[q];
      }
    }
    }
    // ---- epilogue: TMEM lane = basis column K', column = output; fp32 RED into dWp[g][o][i*M+m]
    ptx::mbar_wait(tmem_full, 0);
    ptx::tc_fence_after();
    if (tile_live && !(d.reserved & 64)) {
      const int pw = warp - 2;
      const int q4 = warp & 3, half = pw >> 2;
      const int l = q4 * 32 + lane;  // K' index inside the tile
      int i, m;
      if (generic) {
        const int blk = l >> 3;  // block = (octet, basis)
        i = (g_oc0 + blk / g_Mt) * 8 + (l & 7);
        m = g_m0 + blk % g_Mt;
        if (blk >= p.no * g_Mt) i = in;  // unused block of the tile
      } else if (spline_tile) {
        i = feat0 + (l >> 3);  // K' block f of the tile = feature f of the chunk, 8 splines each
        m = l & 7;
      } else {
        i = feat0 + l;
        m = (F == KV_EFFKAN) ? 8 : 0;
      }
      const int M = d.M, K = in * M, out = d.out_features;
      float* dst = p.dWp + (size_t)g * out * K + (size_t)i * M + m;
      for (int cb = half; cb * 32 < BN; cb += NPW / 4) {
        uint32_t rr[32];
        ptx::tmem_ld32(tmem_base + ((uint32_t)(q4 * 32) << 16) + cb * 32, rr);
        ptx::tmem_ld_wait();
        if (i < in) {
#pragma unroll
          for (int jn = 0; jn < 32; ++jn) {
            const int o = n_tile * BN + cb * 32 + jn;
            if (cb * 32 + jn < BN && o < out) atomicAdd(dst + (size_t)o * K, __uint_as_float(rr[jn]));
          }
        }
      }
    }
    ptx::tc_fence_before();
  }
  ptx::cluster_sync_all();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc2(tmem_base, TMEM_COLS);
  }
}

}  // namespace tc

// ---- TMA tensor map of the bf16 dY image ---------------------------------------------------------------------------
// The image [g][t8][o8][8 rows][8 outputs] is a 3-D tensor of 32-bit words: dim0 = one 128-byte core matrix (32 words),
// dim1 = o8, dim2 = g * T8 + t8.  A box {32, n_o8, n_t8} lands in shared memory as [t8][o8][128 B] -- the operand
// layout both gradient kernels read -- with ONE cp.async.bulk.tensor instead of n_t8 (x2) bulk copies: the loader
// thread issued 16-17 copies per stage before and was the pacing stage of both kernels (scripts/bulk_pace.cu:
// ~60 cycles per extra copy on top of ~500 per stage).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static int encode_dy_image_map(CUtensorMap* map, const void* image, int n_groups, int T8, int O8, int box_o8, int box_t8) {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    KV_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres), "cuTensorMapEncodeTiled lookup");
    if (qres != cudaDriverEntryPointSuccess || sym == nullptr) return KV_ERR_CUDA;
    fn = reinterpret_cast<EncodeTiledFn>(sym);
  }
  const cuuint64_t dims[3] = {32, (cuuint64_t)O8, (cuuint64_t)n_groups * T8};
  const cuuint64_t strides[2] = {128, (cuuint64_t)O8 * 128};   // bytes, dims 1 and 2
  const cuuint32_t box[3] = {32, (cuuint32_t)box_o8, (cuuint32_t)box_t8};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, const_cast<void*>(image), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    snprintf(g_cuda_err, sizeof(g_cuda_err), "cuTensorMapEncodeTiled failed (%d)", (int)r);
    return KV_ERR_CUDA;
  }
  return KV_OK;
}
// X [T][ldx] fp32 as a 2-D tensor, box {16 features, 64 rows} with the 64-byte swizzle (the consumer threads read one
// row each: unswizzled, 32 lanes would hit two banks)
static int encode_x_map(CUtensorMap* map, const float* X, int ldx, int T) {
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult qres;
  KV_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres), "cuTensorMapEncodeTiled lookup");
  if (qres != cudaDriverEntryPointSuccess || sym == nullptr) return KV_ERR_CUDA;
  const EncodeTiledFn fn = reinterpret_cast<EncodeTiledFn>(sym);
  const cuuint64_t dims[2] = {(cuuint64_t)ldx, (cuuint64_t)T};
  const cuuint64_t strides[1] = {(cuuint64_t)ldx * 4};
  const cuuint32_t box[2] = {16, 64};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(X), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    snprintf(g_cuda_err, sizeof(g_cuda_err), "cuTensorMapEncodeTiled (X) failed (%d)", (int)r);
    return KV_ERR_CUDA;
  }
  return KV_OK;
}

int tc_supported(const kv_kan_desc_t& d) {
  if (d.M > tc::MAX_M_TC) return KV_ERR_UNSUPPORTED;
  return KV_OK;
}

size_t tc_wtiles_bytes(const kv_kan_desc_t& d) {
  const tc::Geom gm = tc::fwd_geom(d);
  return (size_t)d.n_groups * gm.n_tiles * gm.tile_elems() * 2;
}

// the tensor-core gradient kernels exist for the default effkan on a uniform grid and for nn.Linear
// (without the class-token row map), one group per slice
static bool tc_bwd_linear(const kv_kan_desc_t& d) {
  return d.family == KV_LINEAR && d.x_share == 1 && d.rows_in == 0 && (d.ldx % 4) == 0 && (d.ldy % 4) == 0 &&
         d.in_features % 4 == 0 && d.out_features % 4 == 0;
}
static bool tc_bwd_effkan8(const kv_kan_desc_t& d) { return tc::make_geom(d).effkan8 && d.x_share == 1 && d.M == 9; }
// any basis family through the generic kernels (basis-major columns / octet x basis tiles), one group per slice
// (groups that share an input slice -- q/k/v maps with per-module prologue parameters -- add their dX with fp32 RED)
static bool tc_bwd_generic(const kv_kan_desc_t& d) {
  return d.family != KV_LINEAR && !tc_bwd_effkan8(d) && d.M <= tc::MAX_M_TC;
}
// input-gradient kernel: one chunk (16 features x M bases) is one MMA of N = 16 M <= 256 columns
static bool tc_bwd_ok(const kv_kan_desc_t& d) {
  return tc_bwd_effkan8(d) || tc_bwd_linear(d) || tc_bwd_generic(d);   // M > 32: basis splits (BwdGeom::nsplit)
}
static bool tc_bwd_w_ok(const kv_kan_desc_t& d) { return tc_bwd_effkan8(d) || tc_bwd_linear(d) || tc_bwd_generic(d); }
// Linear: dX = dY . W runs as the forward kernel of the transposed layer
static kv_kan_desc_t linear_transposed(const kv_kan_desc_t& d) {
  kv_kan_desc_t t = d;
  t.in_features = d.out_features;
  t.out_features = d.in_features;
  t.ldx = d.ldy;
  t.ldy = d.ldx;
  t.flags = d.flags & ~(uint32_t)KV_FLAG_RELU;
  t.fparam_stride = 0;
  return t;
}
size_t tc_wtiles_t_bytes(const kv_kan_desc_t& d) {
  if (!tc_bwd_ok(d)) return 0;
  if (d.family == KV_LINEAR) return tc_wtiles_bytes(linear_transposed(d));
  const tc::BwdGeom b = tc::make_bwd_geom(d);
  return (size_t)b.nsplit * d.n_groups * b.n_tiles * b.n_stages * 2 * b.n_sub * b.SW * 64;
}
void tc_dy_image_geom(const kv_kan_desc_t& d, int* t8, int* o8) {
  const tc::BwdGeom b = tc::make_bwd_geom(d);
  *t8 = b.T8;
  *o8 = b.O8;
}
size_t tc_dy_image_bytes(const kv_kan_desc_t& d) {
  if (!tc_bwd_w_ok(d)) return 0;
  const tc::BwdGeom b = tc::make_bwd_geom(d);
  return (size_t)d.n_groups * b.T8 * b.O8 * 128;
}
bool tc_bwd_weight_available(const kv_kan_desc_t& d) { return tc_bwd_w_ok(d); }
bool tc_bwd_input_available(const kv_kan_desc_t& d) { return tc_bwd_ok(d); }

static int grid_for(size_t total, int threads) {
  const size_t blocks = (total + threads - 1) / threads;
  return (int)(blocks > 148 * 32 ? 148 * 32 : (blocks ? blocks : 1));
}

int tc_pack(const kv_kan_desc_t& d, const kv_kan_desc_t& d0, const uint64_t* ptr_table, const float* fparams,
            void* wtiles, void* wtiles_t, cudaStream_t st) {
  const tc::Geom gm = tc::fwd_geom(d);
  if (wtiles != nullptr) {
    const size_t total = (size_t)d.n_groups * gm.n_tiles * gm.tile_elems() / 8;
    tc::tc_pack_kernel<<<grid_for(total, 256), 256, 0, st>>>(d, d0, ptr_table, fparams,
                                                             reinterpret_cast<__nv_bfloat16*>(wtiles), gm);
    KV_CHECK_LAUNCH("tc_pack_kernel");
  }
  if (wtiles_t != nullptr && d.family == KV_LINEAR) {
    if (!tc_bwd_ok(d)) return KV_ERR_UNSUPPORTED;
    const kv_kan_desc_t dt = linear_transposed(d);
    const tc::Geom gmt = tc::fwd_geom(dt);
    const size_t total = (size_t)dt.n_groups * gmt.n_tiles * gmt.tile_elems() / 8;
    tc::tc_pack_linear_t_kernel<<<grid_for(total, 256), 256, 0, st>>>(dt, d, d0, ptr_table,
                                                                      reinterpret_cast<__nv_bfloat16*>(wtiles_t), gmt);
    KV_CHECK_LAUNCH("tc_pack_linear_t_kernel");
  } else if (wtiles_t != nullptr) {
    if (!tc_bwd_ok(d)) return KV_ERR_UNSUPPORTED;
    const tc::BwdGeom b = tc::make_bwd_geom(d);
    const size_t total = tc_wtiles_t_bytes(d) / 128;
    tc::tc_pack_t_kernel<<<grid_for(total, 128), 128, 0, st>>>(d, d0, ptr_table, reinterpret_cast<__nv_bfloat16*>(wtiles_t), b,
                                                               tc_bwd_effkan8(d) ? tc::LAYOUT_EFFKAN8 : tc::LAYOUT_BASIS_MAJOR);
    KV_CHECK_LAUNCH("tc_pack_t_kernel");
  }
  return KV_OK;
}

int tc_dy_image(const kv_kan_desc_t& d, const float* dY, const float* act_mask, void* image, cudaStream_t st) {
  if (!tc_bwd_w_ok(d)) return KV_ERR_UNSUPPORTED;
  const tc::BwdGeom b = tc::make_bwd_geom(d);
  const size_t total = (size_t)d.n_groups * b.T8 * b.O8 * 8;
  tc::dy_image_kernel<<<grid_for(total, 256), 256, 0, st>>>(d, dY, act_mask, reinterpret_cast<__nv_bfloat16*>(image), b);
  KV_CHECK_LAUNCH("dy_image_kernel");
  return KV_OK;
}

int tc_fwd(const kv_kan_desc_t& d, const float* X, const void* wtiles, const float* fparams,
           const float* bias, const float* mean, const float* rstd, const float* residual,
           const float* pos, float* Y, cudaStream_t st, void* attn_image, int attn_tokens) {
  if (tc_supported(d) != KV_OK) return KV_ERR_UNSUPPORTED;
  if (attn_image != nullptr) {
    // image epilogue: merged q/k/v view (one group per head, 3 * d_head outputs), d_head a multiple of 8
    if (attn_tokens <= 0 || d.T % attn_tokens != 0 || d.out_features % 24 != 0 || d.x_share != 1 || d.rows_in != 0 ||
        (d.flags & KV_FLAG_RELU) || residual != nullptr || pos != nullptr)
      return KV_ERR_UNSUPPORTED;
  }
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
  p.gm = tc::fwd_geom(d);
  p.in_mask = nullptr;
  p.attn_img = reinterpret_cast<__nv_bfloat16*>(attn_image);
  p.attn_N = attn_tokens;
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

template <int F>
static int launch_bwd_input_generic(const tc::BwdInParams& p, dim3 grid, cudaStream_t st) {
  KV_SMEM_OPT_IN(tc::kan_tc_bwd_input_generic_kernel<F>, tc::BI_SMEM, "tc_bwd_input smem attribute");
  tc::kan_tc_bwd_input_generic_kernel<F><<<grid, tc::NTHREADS, tc::BI_SMEM, st>>>(p);
  KV_CHECK_LAUNCH("kan_tc_bwd_input_generic_kernel");
  return KV_OK;
}

int tc_bwd_input(const kv_kan_desc_t& d, const float* X, const float* dY, const float* act_mask, const void* dy_image,
                 const void* wtiles_t, const float* fparams, const float* mean, const float* rstd, float* dX,
                 float* dXn, float* dfparams, cudaStream_t st) {
  if (!tc_bwd_ok(d)) return KV_ERR_UNSUPPORTED;
  if (d.family == KV_LINEAR) {
    tc::FwdParams q;
    q.d = linear_transposed(d);
    q.X = dY;
    q.wt = reinterpret_cast<const __nv_bfloat16*>(wtiles_t);
    q.fparams = nullptr;
    q.bias = q.mean = q.rstd = q.residual = q.pos = nullptr;
    q.Y = dX;
    q.gm = tc::fwd_geom(q.d);
    q.in_mask = (d.flags & KV_FLAG_RELU) ? act_mask : nullptr;
    q.attn_img = nullptr;
    q.attn_N = 0;
    return tc::launch_fwd<KV_LINEAR>(q, st);
  }
  KV_SMEM_OPT_IN(tc::effkan8_bwd_input_kernel, tc::BI_SMEM_E8, "tc_bwd_input smem attribute");
  tc::BwdInParams p;
  p.d = d;
  p.X = X;
  p.dyimg = reinterpret_cast<const __nv_bfloat16*>(dy_image);
  p.wtt = reinterpret_cast<const __nv_bfloat16*>(wtiles_t);
  p.fparams = fparams;
  p.dX = dX;
  p.mean = mean;
  p.rstd = rstd;
  p.dXn = dXn;
  p.dfparams = dfparams;
  p.b = tc::make_bwd_geom(d);
  p.m_lo = 0;
  p.m_cnt = d.M;
  p.accum = 0;
  {
    const int rc = encode_dy_image_map(&p.tm_dy, dy_image, d.n_groups, p.b.T8, p.b.O8, 8, 16);
    if (rc != KV_OK) return rc;
  }
  const int n_pairs = ceil_div(d.T, 2 * tc::BM);
  if ((long long)n_pairs * p.b.n_tiles > 0x3fffffffLL || d.n_groups > 65535) return KV_ERR_BAD_SHAPE;
  dim3 grid(2 * n_pairs * p.b.n_tiles, 1, d.n_groups);   // x = 2 * (row pair, column tile)
  if (tc_bwd_generic(d)) {
    if (d.x_share > 1)
      KV_CUDA(cudaMemset2DAsync(dX, (size_t)d.ldx * sizeof(float), 0,
                                (size_t)(d.n_groups / d.x_share) * d.in_features * sizeof(float), (size_t)d.T, st),
              "tc_bwd_input zero dX");
    // one launch per basis split (M <= 32: one), each with its own transposed weight image
    const size_t split_elems = tc_wtiles_t_bytes(d) / 2 / p.b.nsplit;
    const __nv_bfloat16* wtt0 = p.wtt;
    for (int sp = 0; sp < p.b.nsplit; ++sp) {
      p.wtt = wtt0 + (size_t)sp * split_elems;
      p.m_lo = sp * p.b.Ms;
      p.m_cnt = d.M - p.m_lo < p.b.Ms ? d.M - p.m_lo : p.b.Ms;
      p.accum = sp > 0 ? 1 : 0;
      int rc = KV_ERR_BAD_FAMILY;
      switch (d.family) {
        case KV_EFFKAN: rc = launch_bwd_input_generic<KV_EFFKAN>(p, grid, st); break;
        case KV_CHEBY: rc = launch_bwd_input_generic<KV_CHEBY>(p, grid, st); break;
        case KV_FASTKAN: rc = launch_bwd_input_generic<KV_FASTKAN>(p, grid, st); break;
        case KV_SINE: rc = launch_bwd_input_generic<KV_SINE>(p, grid, st); break;
        case KV_FOURIER: rc = launch_bwd_input_generic<KV_FOURIER>(p, grid, st); break;
      }
      if (rc != KV_OK) return rc;
    }
    return KV_OK;
  }
  // column tiles per CTA pair: as many (up to 8) as still leave ~8 waves of CTA pairs
  p.tpc = 1;
  {
    // measured at T = 50 432: 4 tiles per pair for 48-stage main loops (+4 %), 16 for 12-stage ones (+20 %)
    int want = 192 / (p.b.n_stages > 0 ? p.b.n_stages : 1);
    if (want < 4) want = 4;
    if (const char* e = getenv("KV_TPC")) want = atoi(e);    // timing experiments only
    for (int t = want > 16 ? 16 : want; t > 1; --t) {
      const long long ctas = (long long)n_pairs * ((p.b.n_tiles + t - 1) / t) * d.n_groups;
      if (ctas >= 74 * 8) { p.tpc = t; break; }
    }
  }
  grid = dim3(2 * n_pairs * ((p.b.n_tiles + p.tpc - 1) / p.tpc), 1, d.n_groups);
  tc::effkan8_bwd_input_kernel<<<grid, tc::NTHREADS, tc::BI_SMEM_E8, st>>>(p);
  KV_CHECK_LAUNCH("effkan8_bwd_input_kernel");
  return KV_OK;
}

template <int F>
static int launch_bwd_weight(const tc::BwdWParams& p, dim3 grid, cudaStream_t st) {
  // effkan default layout / Linear tiles: 16 producer warps; generic octet x basis tiles: 8
  if (!p.generic && (F == KV_EFFKAN || F == KV_LINEAR)) {
    KV_SMEM_OPT_IN((tc::kan_tc_bwd_weight_kernel<F, 16>), tc::BW_SMEM, "tc_bwd_weight smem attribute");
    tc::kan_tc_bwd_weight_kernel<F, 16><<<grid, 32 * 18, tc::BW_SMEM, st>>>(p);
  } else {
    KV_SMEM_OPT_IN((tc::kan_tc_bwd_weight_kernel<F, 8>), tc::BW_SMEM, "tc_bwd_weight smem attribute");
    tc::kan_tc_bwd_weight_kernel<F, 8><<<grid, 32 * 10, tc::BW_SMEM, st>>>(p);
  }
  KV_CHECK_LAUNCH("kan_tc_bwd_weight_kernel");
  return KV_OK;
}

int tc_bwd_weight(const kv_kan_desc_t& d, const float* X, const void* dy_image, const float* fparams,
                  const float* mean, const float* rstd, float* dWp, cudaStream_t st) {
  if (!tc_bwd_w_ok(d)) return KV_ERR_UNSUPPORTED;
  tc::BwdWParams p;
  p.d = d;
  p.X = X;
  p.dyimg = reinterpret_cast<const __nv_bfloat16*>(dy_image);
  p.fparams = fparams;
  p.dWp = dWp;
  p.b = tc::make_bwd_geom(d);
  p.gm = tc::make_geom(d);
  {
    const int n1 = p.gm.BN > 256 ? 256 : p.gm.BN, n2 = p.gm.BN - n1;
    int rc = encode_dy_image_map(&p.tm_dy1, dy_image, d.n_groups, p.b.T8, p.b.O8, n1 / 16, 8);
    if (rc == KV_OK) rc = encode_dy_image_map(&p.tm_dy2, dy_image, d.n_groups, p.b.T8, p.b.O8, n2 ? n2 / 16 : 1, 8);
    if (rc != KV_OK) return rc;
  }
  p.use_tma_x = 0;
  if (!tc_bwd_generic(d) && d.family == KV_EFFKAN && (d.ldx % 4) == 0 && (reinterpret_cast<uintptr_t>(X) & 15) == 0) {
    const int rc = encode_x_map(&p.tm_x, X, d.ldx, d.T);
    if (rc != KV_OK) return rc;
    p.use_tma_x = 1;
  }
  // K' tiles of 128 basis columns: effkan = one per 16-feature chunk (8 splines each) + one per 128 features
  // (SiLU column); Linear = one per 128 features
  p.mean = mean;
  p.rstd = rstd;
  p.generic = tc_bwd_generic(d) ? 1 : 0;
  p.no = 1;
  p.tpo = 1;
  if (p.generic) {
    const int n_oct = (d.in_features + 7) / 8;
    if (d.M <= 16) {
      p.no = 16 / d.M;
      p.KT = (n_oct + p.no - 1) / p.no;
    } else {
      p.tpo = (d.M + 15) / 16;
      p.KT = n_oct * p.tpo;
    }
  } else {
    p.KT = d.family == KV_LINEAR ? (d.in_features + 127) / 128 : p.b.n_chunks + (p.b.n_chunks + 7) / 8;
  }
  p.n_kpairs = (p.KT + 1) / 2;
  // split T so that the grid fills the 74 CTA-pair slots evenly; every split keeps >= 512 rows
  const int P = p.n_kpairs * p.gm.n_tiles * d.n_groups;
  int best = 1;
  double best_eff = 0.0;
  for (int s = 1; s <= 32; ++s) {
    if (s > 1 && d.T / s < 512) break;
    const double waves = (double)P * s / 74.0;
    const double eff = waves / (double)(long long)(waves + 0.999999);
    if (eff > best_eff + 0.03) { best_eff = eff; best = s; }
  }
  p.rows_per_split = (ceil_div(d.T, best) + 63) / 64 * 64;
  p.tsplits = ceil_div(d.T, p.rows_per_split);  // rounding to whole 64-row stages may empty the last splits
  best = p.tsplits;
  if ((long long)p.n_kpairs * p.gm.n_tiles > 0x3fffffffLL || (long long)d.n_groups * best > 65535) return KV_ERR_BAD_SHAPE;
  dim3 grid(2 * p.n_kpairs * p.gm.n_tiles, 1, d.n_groups * best);   // x = 2 * (K' pair, output tile)
  switch (d.family) {
    case KV_LINEAR: return launch_bwd_weight<KV_LINEAR>(p, grid, st);
    case KV_EFFKAN: return launch_bwd_weight<KV_EFFKAN>(p, grid, st);
    case KV_CHEBY: return launch_bwd_weight<KV_CHEBY>(p, grid, st);
    case KV_FASTKAN: return launch_bwd_weight<KV_FASTKAN>(p, grid, st);
    case KV_SINE: return launch_bwd_weight<KV_SINE>(p, grid, st);
    case KV_FOURIER: return launch_bwd_weight<KV_FOURIER>(p, grid, st);
  }
  return KV_ERR_BAD_FAMILY;
}

}  // namespace kv
