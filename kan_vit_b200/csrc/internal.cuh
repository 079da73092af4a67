// Internal entry points shared between the translation units of libkanvit_b200.
#pragma once
#include "common.cuh"

namespace kv {

// kan_simt.cu -- fp32 SIMT basis-prologue GEMMs
int simt_fwd(const kv_kan_desc_t& d, const float* X, const float* Wp, const float* fparams,
             const float* bias, const float* mean, const float* rstd, const float* residual,
             const float* pos, float* Y, cudaStream_t st);
int simt_bwd_input(const kv_kan_desc_t& d, const float* X, const float* dY, const float* mask,
                   const float* Wp, const float* fparams, const float* mean, const float* rstd,
                   float* dX, float* dXn, float* dfparams, cudaStream_t st);
int simt_bwd_weight(const kv_kan_desc_t& d, const float* X, const float* dY, const float* mask,
                    const float* fparams, const float* mean, const float* rstd, float* dWp,
                    float* dbias, cudaStream_t st);

int simt_col_sums(const kv_kan_desc_t& d, const float* dY, const float* mask, float* dbias, cudaStream_t st);

// kan_tc.cu -- tcgen05 / TMEM bf16 basis-prologue GEMMs
int tc_supported(const kv_kan_desc_t& d);
size_t tc_wtiles_bytes(const kv_kan_desc_t& d);    // forward operand image
size_t tc_wtiles_t_bytes(const kv_kan_desc_t& d);  // input-gradient operand image (0 = kernel not available)
bool tc_bwd_weight_available(const kv_kan_desc_t& d);
int tc_pack(const kv_kan_desc_t& d, const float* wp_f32, const float* fparams, void* wtiles, void* wtiles_t,
            cudaStream_t st);
int tc_fwd(const kv_kan_desc_t& d, const float* X, const void* wtiles, const float* fparams,
           const float* bias, const float* mean, const float* rstd, const float* residual,
           const float* pos, float* Y, cudaStream_t st);
size_t tc_dy_image_bytes(const kv_kan_desc_t& d);  // bf16 copy of dY shared by the gradient kernels (0 = n/a)
int tc_dy_image(const kv_kan_desc_t& d, const float* dY, const float* act_mask, void* image, cudaStream_t st);
int tc_bwd_input(const kv_kan_desc_t& d, const float* X, const void* dy_image, const void* wtiles_t,
                 const float* fparams, float* dX, cudaStream_t st);
int tc_bwd_weight(const kv_kan_desc_t& d, const float* X, const void* dy_image, const float* fparams, float* dWp,
                  cudaStream_t st);

// validation shared by every kan entry point
int check_desc(const kv_kan_desc_t* d);

}  // namespace kv
