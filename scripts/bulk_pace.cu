// Microbenchmark (timing experiment, not product code): per-SM throughput of cp.async.bulk global -> shared when every
// CTA streams a weight image the way the KAN GEMM loaders do: a ring of NB stages of `stage` bytes, one thread issuing.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o build/bulk_pace scripts/bulk_pace.cu
// Variants: stage size, ring depth, all CTAs reading the SAME image (L2-resident, hot lines) or each its own
// region, one big copy per stage or the stage split into several copies.
#include <cstdio>
#include <cuda_runtime.h>
#include "../kan_vit_b200/csrc/tc_ptx.cuh"

using namespace kv;

struct Cfg {
  int stage;     // bytes per stage
  int nb;        // ring depth
  int n_stages;  // stages streamed per CTA
  int shared;    // 1 = every CTA reads the same image, 0 = CTA b reads its own region
  int split;     // copies per stage
  int span;      // bytes of the image that is cycled through (per CTA when !shared)
};

__global__ void __launch_bounds__(64, 1) bulk_kernel(Cfg c, const uint8_t* src, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t full[16];
  if (threadIdx.x == 0) {
    for (int i = 0; i < 16; ++i) ptx::mbar_init(ptx::smem_u32(&full[i]), 1);
    ptx::fence_mbar_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint8_t* base = src + (c.shared ? 0 : (size_t)blockIdx.x * c.span);
    const int per = c.stage / c.split;
    long long t0 = clock64();
    size_t off = 0;
    for (int st = 0; st < c.n_stages + c.nb; ++st) {
      const int s = st % c.nb;
      if (st >= c.nb) ptx::mbar_wait(ptx::smem_u32(&full[s]), ((st / c.nb) - 1) & 1);   // consume the stage
      if (st < c.n_stages) {
        ptx::mbar_arrive_expect_tx(ptx::smem_u32(&full[s]), c.stage);
        for (int k = 0; k < c.split; ++k)
          ptx::bulk_g2s(ptx::smem_u32(smem) + s * c.stage + k * per, base + off + (size_t)k * per, per,
                        ptx::smem_u32(&full[s]));
        off += c.stage;
        if (off + c.stage > (size_t)c.span) off = 0;
      }
    }
    long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
  }
}

int main() {
  uint8_t* src;
  const size_t total = (size_t)148 * 64 * 1024 * 1024;   // 148 x 64 MB > L2 for the private-region runs
  cudaMalloc(&src, total);
  cudaMemset(src, 1, total);
  unsigned long long* d;
  cudaMalloc(&d, 16);
  cudaFuncSetAttribute(bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  struct Named { const char* name; Cfg c; };
  const int MB = 1 << 20;
  Named runs[] = {
      {"shared 42MB image, 24KB x4 stages (fwd BN=512 today)", {24576, 4, 2000, 1, 1, 42 * MB}},
      {"shared 42MB image, 24KB x4, 3 copies per stage", {24576, 4, 2000, 1, 3, 42 * MB}},
      {"shared 42MB image, 8KB x12 stages", {8192, 12, 6000, 1, 1, 42 * MB}},
      {"shared 42MB image, 8KB x16 stages", {8192, 16, 6000, 1, 1, 42 * MB}},
      {"shared 42MB image, 48KB x2 stages", {49152, 2, 1000, 1, 1, 42 * MB}},
      {"shared 42MB image, 48KB x3 stages", {49152, 3, 1000, 1, 1, 42 * MB}},
      {"shared 42MB image, 72KB x2 stages", {73728, 2, 700, 1, 1, 42 * MB}},
      {"shared 42MB image, 16KB x8 stages", {16384, 8, 3000, 1, 1, 42 * MB}},
      {"shared 3.5MB image (one tile), 24KB x4", {24576, 4, 2000, 1, 1, 7 * MB / 2}},
      {"private 0.5MB regions (L2 resident), 24KB x4", {24576, 4, 2000, 0, 1, MB / 2}},
      {"private 64MB regions (DRAM), 24KB x4", {24576, 4, 2000, 0, 1, 64 * MB}},
      {"private 64MB regions (DRAM), 8KB x16", {8192, 16, 6000, 0, 1, 64 * MB}},
  };
  for (auto& r : runs) {
    for (int grid : {1, 148}) {
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0);
      cudaEventCreate(&e1);
      bulk_kernel<<<grid, 64, r.c.stage * r.c.nb>>>(r.c, src, d);
      cudaEventRecord(e0);
      bulk_kernel<<<grid, 64, r.c.stage * r.c.nb>>>(r.c, src, d);
      cudaEventRecord(e1);
      cudaError_t e = cudaDeviceSynchronize();
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      unsigned long long h = 0;
      cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
      const double bytes = (double)r.c.stage * r.c.n_stages;
      printf("%-58s grid %3d: %6.1f B/cyc/SM (CTA 0), %7.0f cyc/stage; wall %.3f ms = %6.2f TB/s chip %s\n", r.name, grid,
             bytes / (double)h, (double)h / r.c.n_stages, ms, bytes * grid / (ms * 1e-3) / 1e12,
             e == cudaSuccess ? "" : cudaGetErrorString(e));
    }
  }
  return 0;
}
