// tcgen05 / TMEM bf16 basis-prologue GEMMs (placeholder until the kernels land).
#include "internal.cuh"
namespace kv {
int tc_supported(const kv_kan_desc_t&) { return KV_ERR_UNSUPPORTED; }
int tc_pack(const kv_kan_desc_t&, const uint64_t*, void*, void*, cudaStream_t) { return KV_ERR_UNSUPPORTED; }
int tc_fwd(const kv_kan_desc_t&, const float*, const void*, const float*, const float*, const float*,
           const float*, const float*, const float*, float*, cudaStream_t) { return KV_ERR_UNSUPPORTED; }
int tc_bwd_input(const kv_kan_desc_t&, const float*, const float*, const float*, const void*, const float*,
                 const float*, const float*, float*, float*, float*, cudaStream_t) { return KV_ERR_UNSUPPORTED; }
int tc_bwd_weight(const kv_kan_desc_t&, const float*, const float*, const float*, const float*, const float*,
                  const float*, float*, float*, cudaStream_t) { return KV_ERR_UNSUPPORTED; }
}  // namespace kv
