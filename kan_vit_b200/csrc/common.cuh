// Shared helpers for libkanvit_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "../../include/kanvit_b200.h"

namespace kv {

// last CUDA launch error text (diagnostic only; written on failure paths)
extern thread_local char g_cuda_err[256];
int record_cuda_error(cudaError_t e, const char* where);

// kernels launched by this library since load (diagnostic; read by kv_launch_count)
extern unsigned long long g_launches;

#define KV_CHECK_LAUNCH(where)                                   \
  do {                                                           \
    __atomic_fetch_add(&kv::g_launches, 1ull, __ATOMIC_RELAXED); \
    cudaError_t _e = cudaGetLastError();                         \
    if (_e != cudaSuccess) return kv::record_cuda_error(_e, where); \
  } while (0)

#define KV_CUDA(call, where)                                     \
  do {                                                           \
    cudaError_t _e = (call);                                     \
    if (_e != cudaSuccess) return kv::record_cuda_error(_e, where); \
  } while (0)

// Opt a kernel in to more than 48 KB of dynamic shared memory, once per DEVICE (the attribute is per device;
// a process may touch several GPUs).
#define KV_SMEM_OPT_IN(kernel, bytes, where)                                                          \
  do {                                                                                                \
    static unsigned long long _done = 0ull; /* bit per device ordinal; same value on every writer */ \
    int _dev = 0;                                                                                     \
    KV_CUDA(cudaGetDevice(&_dev), where);                                                             \
    const unsigned long long _bit = 1ull << (_dev & 63);                                              \
    if (!(__atomic_load_n(&_done, __ATOMIC_RELAXED) & _bit)) {                                        \
      KV_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (bytes)), where); \
      __atomic_fetch_or(&_done, _bit, __ATOMIC_RELAXED);                                              \
    }                                                                                                 \
  } while (0)

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Row map of kv_kan_desc_t (see header): logical row t -> physical row of Y / dY.
__device__ __forceinline__ int map_row(int t, int rows_in, int rows_out, int row_shift) {
  if (rows_in == 0) return t;
  int b = t / rows_in;
  return b * rows_out + row_shift + (t - b * rows_in);
}

}  // namespace kv
