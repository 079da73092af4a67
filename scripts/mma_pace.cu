// Microbenchmark (timing experiment, not product code): how fast does one CTA pair retire tcgen05.mma K=16 steps
// when nothing else is in the way?  No producers, no loads: operands are whatever shared memory holds.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/mma_pace scripts/mma_pace.cu
// Variants: N per instruction, instructions per "unit", commit frequency, layout type (no swizzle / 128B swizzle),
// cta_group 1 or 2, with / without other warps hammering shared memory with 16-byte stores.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../kan_vit_b200/csrc/tc_ptx.cuh"

using namespace kv;

struct Cfg {
  int n1, n2;        // N of the first / second MMA of a unit (n2 = 0: one MMA)
  int units;         // K=16 units per kernel
  int commit_every;  // commit to a dummy barrier every this many units (0 = only at the end)
  int layout;        // descriptor layout type bits [61,64)
  int two_cta;       // cta_group::2 (M = 256) or ::1 (M = 128)
  int writers;       // 8 extra warps writing 16-byte vectors into the A region
  int lbo, sbo;      // descriptor strides (bytes)
  int a_rot;         // rotate the A start address over this many 4 KB slots (1 = fixed)
  int ckind;         // 0 = multicast commit (both CTAs), 1 = commit to the leader's barrier only, 2 = two multicast commits
  int wwork;         // writers: dependent FMAs between two 16-byte stores (0 = back-to-back stores)
  int lean;          // 1 = single-thread issue loop without per-iteration election / division
  int data;          // operands: 0 = zeros, 1 = pseudo-random bf16 values in (-2, 2)
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(320, 1) pace_kernel(Cfg c, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bars[4];
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  if (threadIdx.x == 0) {
    ptx::mbar_init(ptx::smem_u32(&bars[0]), 1);
    ptx::mbar_init(ptx::smem_u32(&bars[1]), 1);
    ptx::mbar_init(ptx::smem_u32(&bars[2]), 1);
    ptx::fence_mbar_init();
  }
  for (int i = threadIdx.x * 16; i < 200 * 1024; i += 320 * 16) {
    uint4 v = make_uint4(0, 0, 0, 0);
    if (c.data) {   // bf16 pairs: sign and mantissa from a hash, exponent 126 / 127 (|x| in [0.5, 2))
      uint32_t h = (uint32_t)i * 2654435761u + blockIdx.y * 40503u + rank * 977u;
      uint32_t w[4];
      for (int k = 0; k < 4; ++k) {
        h = h * 1664525u + 1013904223u;
        const uint32_t lo = ((h >> 3) & 0x80ffu) | (0x3f00u + ((h >> 20) & 0x80u));
        h = h * 1664525u + 1013904223u;
        const uint32_t hi = ((h >> 3) & 0x80ffu) | (0x3f00u + ((h >> 20) & 0x80u));
        w[k] = lo | (hi << 16);
      }
      v = make_uint4(w[0], w[1], w[2], w[3]);
    }
    *reinterpret_cast<uint4*>(smem + i) = v;
  }
  if (warp == 0) {
    if (c.two_cta) { ptx::tmem_alloc2(ptx::smem_u32(&tmem_slot), 512); ptx::tmem_relinquish2(); }
    else { ptx::tmem_alloc(ptx::smem_u32(&tmem_slot), 512); ptx::tmem_relinquish(); }
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  ptx::cluster_sync_all();
  ptx::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  __shared__ volatile int stop;
  if (threadIdx.x == 0) stop = 0;
  __syncthreads();
  if (warp == 1 && (rank == 0 || !c.two_cta)) {
    const int M = c.two_cta ? 256 : 128;
    const uint32_t id1 = ptx::make_idesc_bf16(M, c.n1), id2 = c.n2 ? ptx::make_idesc_bf16(M, c.n2) : 0u;
    const uint64_t lt = (uint64_t)c.layout << 61;
    const uint64_t da0 = ptx::make_smem_desc(ptx::smem_u32(smem), c.lbo, c.sbo) | lt;
    const uint64_t db0 = ptx::make_smem_desc(ptx::smem_u32(smem) + 112 * 1024, c.lbo, c.sbo) | lt;
    long long t0 = clock64();
    if (c.lean) {
      // one thread, no per-iteration election / warp sync / division: the tensor pipe (and the commits) set the pace
      if (lane == 0) {
        int rot = 0, cc = 0, bq = 0;
        for (int u = 0; u < c.units; ++u) {
          const uint64_t da = da0 + (uint64_t)(rot * 256);
          const uint64_t db = db0 + (uint64_t)(bq * 1024);
          if (c.two_cta) {
            if (c.n1) ptx::mma2_bf16_ss(tmem, da, db, id1, u > 0);
            if (c.n2) ptx::mma2_bf16_ss(tmem + 256, da, db + (8192 >> 4), id2, u > 0);
          } else {
            if (c.n1) ptx::mma_bf16_ss(tmem, da, db, id1, u > 0);
            if (c.n2) ptx::mma_bf16_ss(tmem + 256, da, db + (8192 >> 4), id2, u > 0);
          }
          if (++cc == c.commit_every) {
            cc = 0;
            if (c.two_cta) {
              if (c.ckind == 1) {
                asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                                 ptx::smem_u32(&bars[1])) : "memory");
              } else {
                ptx::tc_commit2(ptx::smem_u32(&bars[1]));
                if (c.ckind == 2) ptx::tc_commit2(ptx::smem_u32(&bars[2]));
              }
            } else {
              ptx::tc_commit(ptx::smem_u32(&bars[1]));
            }
          }
          if (++rot == c.a_rot) rot = 0;
          bq = (bq + 1) & 3;
        }
      }
      __syncwarp();
    } else
    for (int u = 0; u < c.units; ++u) {
      if (ptx::elect_one()) {
        const uint64_t da = da0 + (uint64_t)(((u % c.a_rot) * 4096) >> 4);
        const uint64_t db = db0 + (uint64_t)(((u & 3) * 16384) >> 4);
        if (c.two_cta) {
          if (c.n1) ptx::mma2_bf16_ss(tmem, da, db, id1, u > 0);
          if (c.n2) ptx::mma2_bf16_ss(tmem + 256, da, db + (8192 >> 4), id2, u > 0);
          if (c.commit_every && (u % c.commit_every) == c.commit_every - 1) {
            if (c.ckind == 1) {
              asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                               ptx::smem_u32(&bars[1])) : "memory");
            } else {
              ptx::tc_commit2(ptx::smem_u32(&bars[1]));
              if (c.ckind == 2) ptx::tc_commit2(ptx::smem_u32(&bars[2]));
            }
          }
        } else {
          if (c.n1) ptx::mma_bf16_ss(tmem, da, db, id1, u > 0);
          if (c.n2) ptx::mma_bf16_ss(tmem + 256, da, db + (8192 >> 4), id2, u > 0);
          if (c.commit_every && (u % c.commit_every) == c.commit_every - 1) ptx::tc_commit(ptx::smem_u32(&bars[1]));
        }
      }
      __syncwarp();
    }
    long long t1 = clock64();
    if (ptx::elect_one()) {
      if (c.two_cta) ptx::tc_commit2(ptx::smem_u32(&bars[0])); else ptx::tc_commit(ptx::smem_u32(&bars[0]));
    }
    __syncwarp();
    ptx::mbar_wait(ptx::smem_u32(&bars[0]), 0);
    long long t2 = clock64();
    if (lane == 0 && blockIdx.y == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    stop = 1;
  } else if (warp >= 2 && c.writers) {
    // producer-like traffic: every thread stores 16-byte vectors, conflict-free, into the A region
    uint8_t* base = smem + ((threadIdx.x - 64) & 127) * 16 + ((threadIdx.x - 64) >> 7) * 2048;
    int k = 0;
    float f = (float)threadIdx.x;
    while (!stop) {
#pragma unroll 1
      for (int q = 0; q < 9; ++q) {
        for (int w = 0; w < c.wwork; ++w) f = fmaf(f, 1.0001f, 0.5f);
        *reinterpret_cast<uint4*>(base + ((k + q) % 24) * 4096) = make_uint4(k, q, __float_as_uint(f), q);
      }
      ptx::fence_proxy_async_smem();
      ++k;
    }
  }
  if (c.two_cta && rank == 1 && warp == 1) {  // the multicast commit lands here too
    ptx::mbar_wait(ptx::smem_u32(&bars[0]), 0);
    stop = 1;
  }
  ptx::tc_fence_before();
  ptx::cluster_sync_all();
  if (warp == 0) {
    ptx::tc_fence_after();
    if (c.two_cta) ptx::tmem_dealloc2(tmem, 512); else ptx::tmem_dealloc(tmem, 512);
  }
}

int main() {
  unsigned long long* d;
  cudaMalloc(&d, 16);
  cudaFuncSetAttribute(pace_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  struct Named { const char* name; Cfg c; };
  const int U = 40000;   // ~5 ms per launch: long enough for the power management to react
  Named runs[] = {
      {"lean 2cta N=256+256 zeros", {256, 256, U, 0, 0, 1, 0, 2048, 128, 1, 0, 0, 1, 0}},
      {"lean 2cta N=256+256 random data", {256, 256, U, 0, 0, 1, 0, 2048, 128, 1, 0, 0, 1, 1}},
      {"lean 2cta N=256+256 random data A rot 27 commit/3", {256, 256, U, 3, 0, 1, 0, 2048, 128, 27, 0, 0, 1, 1}},
      {"lean 2cta N=256+128 random data", {256, 128, U, 0, 0, 1, 0, 2048, 128, 1, 0, 0, 1, 1}},
      {"lean 2cta N=256+256 random data + writers wwork=16", {256, 256, U, 0, 0, 1, 1, 2048, 128, 27, 0, 16, 1, 1}},
      {"lean 1cta N=256+256 random data", {256, 256, U, 0, 0, 0, 0, 2048, 128, 1, 0, 0, 1, 1}},
  };
  for (auto& r : runs) {
    for (int grid : {1, 16, 74}) {
      cudaMemset(d, 0, 16);
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0);
      cudaEventCreate(&e1);
      pace_kernel<<<dim3(2, grid, 1), 320, 200 * 1024>>>(r.c, d);   // warm-up
      cudaEventRecord(e0);
      pace_kernel<<<dim3(2, grid, 1), 320, 200 * 1024>>>(r.c, d);
      cudaEventRecord(e1);
      cudaError_t e = cudaDeviceSynchronize();
      float ms = 0.f;
      cudaEventElapsedTime(&ms, e0, e1);
      unsigned long long h[2] = {0, 0};
      cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      const double per = (double)h[1] / r.c.units;
      const int M = r.c.two_cta ? 256 : 128;
      const double flop = 2.0 * M * (r.c.n1 + r.c.n2) * 16;
      // TFLOP/s if every SM ran at this pace at 1.965 GHz: pairs = 74 (or 148 single CTAs)
      const double tf = flop / per * 1.965e9 * (r.c.two_cta ? 74 : 148) / 1e12;
      const double wall_tf = flop * r.c.units * (r.c.two_cta ? grid : 2 * grid) / (ms * 1e-3) / 1e12;
      printf("%-55s grid %2d: issue %7.1f cyc/unit, retire %7.1f cyc/unit -> %6.0f TFLOP/s chip-equivalent; kernel %.3f ms = %6.0f TFLOP/s wall %s\n",
             r.name, grid, (double)h[0] / r.c.units, per, tf, ms, wall_tf, e == cudaSuccess ? "" : cudaGetErrorString(e));
    }
  }
  return 0;
}
