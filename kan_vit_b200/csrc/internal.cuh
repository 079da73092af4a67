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
// operand images straight from the module tensors (ptr_table of kv_kan_pack): d = the tensor-core view of the
// layer (api.cu tc_view), d0 = the caller's descriptor that indexes the pointer table
int tc_pack(const kv_kan_desc_t& d, const kv_kan_desc_t& d0, const uint64_t* ptr_table, const float* fparams,
            void* wtiles, void* wtiles_t, cudaStream_t st);
// attn_image != nullptr: the epilogue writes the bf16 attention image of attn_tc5.cu instead of Y (q/k/v layer of MSA)
int tc_fwd(const kv_kan_desc_t& d, const float* X, const void* wtiles, const float* fparams,
           const float* bias, const float* mean, const float* rstd, const float* residual,
           const float* pos, float* Y, cudaStream_t st, void* attn_image = nullptr, int attn_tokens = 0);
void tc_dy_image_geom(const kv_kan_desc_t& d, int* t8, int* o8);  // extents of the dY image in 8-row / 8-column groups
size_t tc_dy_image_bytes(const kv_kan_desc_t& d);  // bf16 copy of dY shared by the gradient kernels (0 = n/a)
int tc_dy_image(const kv_kan_desc_t& d, const float* dY, const float* act_mask, void* image, cudaStream_t st);
int tc_bwd_input(const kv_kan_desc_t& d, const float* X, const float* dY, const float* act_mask, const void* dy_image,
                 const void* wtiles_t, const float* fparams, const float* mean, const float* rstd, float* dX,
                 float* dXn, float* dfparams, cudaStream_t st);
int tc_bwd_weight(const kv_kan_desc_t& d, const float* X, const void* dy_image, const float* fparams,
                  const float* mean, const float* rstd, float* dWp, cudaStream_t st);

// attn_tiled.cu -- exact fp32 attention for sequences that do not fit one CTA's shared memory (fallback of attn.cu and
// attn_general.cu).  Head h of sample b = rows [b*N, (b+1)*N), columns [h*hs, h*hs + dh) of a matrix with row stride ld.
struct AttnTiled {
  const float *q, *k, *v;
  int ldq, ldk, ldv, hs;
  const unsigned char* mask;  // nullable; element (b, h, i, j) at b*msb + h*msh + i*msq + j (0 stride = broadcast)
  long long msb, msh, msq;
  int causal;
  int B, H, Nq, Nk, dh;
};
int attn_tiled_fwd(const AttnTiled& a, const float* residual, float* out, int ldo, int hso, float* lse,
                   cudaStream_t st);
// o == nullptr: D_i is recomputed (MSA entry point); dq / dk / dv use the layouts of q / k / v
int attn_tiled_bwd(const AttnTiled& a, const float* o, const float* dout, int ldo, int hso, const float* lse,
                   float* dq, float* dk, float* dv, cudaStream_t st);

// parameter pointer table of kv_kan_pack
__device__ __forceinline__ const float* slot(const uint64_t* tbl, int g, int s) {
  return reinterpret_cast<const float*>(tbl[(size_t)g * KV_PTRS_PER_GROUP + s]);
}
__device__ __forceinline__ float* slot_out(const uint64_t* tbl, int g, int s) {
  return reinterpret_cast<float*>(tbl[(size_t)g * KV_PTRS_PER_GROUP + s]);
}

// packed weight element (g, o, i, m) read from the module's own tensors
__device__ __forceinline__ float load_packed_weight(const kv_kan_desc_t& d, const uint64_t* tbl, int g,
                                                    int o, int i, int m) {
  const int in = d.in_features, out = d.out_features, G = d.grid_size;
  switch (d.family) {
    case KV_LINEAR:
      return slot(tbl, g, 0)[(size_t)o * in + i];
    case KV_EFFKAN: {
      const int nb = G + d.spline_order;
      if (m == nb) return slot(tbl, g, 0)[(size_t)o * in + i];
      float w = slot(tbl, g, 1)[((size_t)o * in + i) * nb + m];
      const float* sc = slot(tbl, g, 2);
      if (sc) w *= sc[(size_t)o * in + i];  // scaled_spline_weight, models/effkan.py:166-172
      return w;
    }
    case KV_CHEBY:
      return slot(tbl, g, 0)[((size_t)i * out + o) * (G + 1) + m];
    case KV_FASTKAN:
      if (m < G) return slot(tbl, g, 3)[(size_t)o * in * G + (size_t)i * G + m];
      return slot(tbl, g, 4)[(size_t)o * in + i];
    case KV_SINE:
      return slot(tbl, g, 0)[((size_t)o * in + i) * G + m];
    case KV_FOURIER: {
      const int part = m / G, k = m % G;
      return slot(tbl, g, 0)[(((size_t)part * out + o) * in + i) * G + k];
    }
  }
  return 0.0f;
}

// validation shared by every kan entry point
int check_desc(const kv_kan_desc_t* d);

}  // namespace kv
