/* kanvit_b200.h -- C ABI of the B200-native KAN-ViT hot path (libkanvit_b200.so).
 *
 * The reference (akshathmangudi/KAN-ViT) has no FFI / operator layer: its hot path is
 * the torch.nn.Module API of model.py / attention.py / models/*.py, every op a stock
 * ATen call.  The entry points below are what a binding for that path replaces; each
 * one cites the reference code whose arithmetic it takes over (paths are relative to
 * the reference tree).  INTEGRATION.md shows the ctypes stub a maintainer adds.
 *
 * Conventions
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer unless the
 *     name ends in _host; all activations are row-major float32; the tensor-core
 *     path additionally takes bf16 copies of the packed weights (made by kv_kan_pack)
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*), does
 *     no allocation and no synchronisation; the caller owns all memory
 *   - return value: 0 on success, a negative kv_status otherwise (never aborts)
 *   - re-entrant: no mutable global state besides a launch counter and a per-thread error string
 */
#ifndef KANVIT_B200_H
#define KANVIT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KV_ABI_VERSION 1

typedef enum {
  KV_OK = 0,
  KV_ERR_BAD_ARG = -1,      /* null pointer / negative size */
  KV_ERR_BAD_SHAPE = -2,    /* shape the kernels cannot serve (e.g. M > KV_MAX_M) */
  KV_ERR_BAD_FAMILY = -3,
  KV_ERR_ALIGNMENT = -4,    /* pointer / leading dimension not aligned for the path */
  KV_ERR_UNSUPPORTED = -5,  /* combination not available in this precision mode */
  KV_ERR_CUDA = -6          /* launch failed; see kv_last_cuda_error() */
} kv_status;

/* Basis family of a KAN layer:  Y[T,out] = Phi(X)[T, in*M] . Wp[out, in*M]^T (+ bias). */
typedef enum {
  KV_LINEAR = 0,   /* nn.Linear (model.py:25-29,68,108; attention.py:137-143): M=1, Phi=x            */
  KV_EFFKAN = 1,   /* models/effkan.py:99-132,166-187: M = G+k (B-splines) + 1 (SiLU base)          */
  KV_CHEBY = 2,    /* models/cheby.py:36-48:           M = degree+1, T_n(tanh x)                    */
  KV_FASTKAN = 3,  /* models/fastkan.py:15-30,66-76:   M = num_grids (RBF of LayerNorm(x)) + 1 base */
  KV_SINE = 4,     /* models/sinekan.py:81-91:         M = G, sin(x*freq_g + phase_{i,g})           */
  KV_FOURIER = 5   /* models/nfkan.py:36-52:           M = 2G, cos(kx) then sin(kx), k=1..G         */
} kv_family;

#define KV_MAX_M 64          /* basis functions per input scalar                         */
#define KV_MAX_SPLINE_ORDER 7
#define KV_PTRS_PER_GROUP 8  /* slots of the parameter pointer table (see kv_kan_pack)   */

/* flags */
#define KV_FLAG_RELU          1u  /* forward: Y = max(Y,0) (model.py:27); backward: dY *= (act_mask>0) */
#define KV_FLAG_NO_LAYERNORM  2u  /* fastkan time_benchmark=True (models/fastkan.py:69-70)            */
#define KV_FLAG_NO_BASE       4u  /* fastkan use_base_update=False (models/fastkan.py:60-63)          */
#define KV_FLAG_BF16_TC       8u  /* tcgen05 bf16 tensor-core path (fp32 accumulate) instead of fp32 SIMT */
#define KV_FLAG_UNIFORM_GRID 16u  /* effkan: the caller has checked that every feature shares one uniform knot
                                     vector (true until KANLinear.update_grid, models/effkan.py:189-242, is used);
                                     enables the closed-form B-spline prologue and the tensor-core gradient kernels */

#define KV_FLAG_ACCUM_GRADS  32u  /* kv_kan_unpack_grads adds into the gradient tensors instead of overwriting them */

/* One KAN (or Linear) layer, or a batch of `n_groups` independent layers of equal
 * shape applied to column slices of the same activation matrix (the per-head Q/K/V
 * maps of MSA, attention.py:136-170,196-197).
 *
 *   group g reads  X[t, (g / x_share) * in_features + i]      (ldx = row stride)
 *   group g writes Y[row(t), g * out_features + o]            (ldy = row stride)
 *   row(t) = t                                  if rows_in == 0
 *          = (t / rows_in) * rows_out + row_shift + t % rows_in   otherwise
 * (the row map writes patch tokens behind the class token, model.py:148-150; `pos`,
 *  if given, is added with row index row_shift + t % rows_in, model.py:152).
 */
typedef struct {
  int32_t family;        /* kv_family */
  int32_t T;             /* rows (all leading dims flattened) */
  int32_t in_features;   /* per group */
  int32_t out_features;  /* per group */
  int32_t n_groups;      /* 1 = dense layer */
  int32_t x_share;       /* consecutive groups sharing one input slice (3 for q,k,v); >=1 */
  int32_t M;             /* basis functions per input incl. base branch */
  int32_t grid_size;     /* G | degree | num_grids */
  int32_t spline_order;  /* effkan only */
  int32_t ldx, ldy;      /* row strides in elements */
  int32_t rows_in, rows_out, row_shift;
  uint32_t flags;
  float rbf_denom;       /* fastkan: rbf.denominator (models/fastkan.py:26-27) */
  int32_t fparam_stride; /* floats of prologue parameters per group (kv_kan_fparam_count) */
  int32_t reserved;
} kv_kan_desc_t;

int kv_version(void);
const char* kv_status_string(int status);
const char* kv_last_cuda_error(void);
/* Kernels launched by this library since it was loaded (diagnostic counter; bench.py reports it). */
unsigned long long kv_launch_count(void);

/* Number of fp32 prologue parameters per group for a family:
 *   effkan  in*(G+2k+1)+4  knot vectors `grid` (models/effkan.py:45-54) + {uniform?, t0, 1/h, 0}
 *   fastkan 2*in+G       layernorm weight, bias, rbf.grid (models/fastkan.py:24-27,57)
 *   sine    G+in*G       freq, phase (models/sinekan.py:59-75)
 *   others  0                                                                     */
int kv_kan_fparam_count(const kv_kan_desc_t* d);

/* Gather the reference-layout parameters of n_groups modules into the packed form
 * the kernels read.  ptr_table: device array [n_groups][KV_PTRS_PER_GROUP] of device
 * pointers (uint64), slot meaning per family:
 *   linear  0 weight(out,in)            1 bias(out)|0
 *   effkan  0 base_weight(out,in)       1 spline_weight(out,in,G+k) 2 spline_scaler(out,in)|0 3 grid(in,G+2k+1)
 *   cheby   0 cheby_coeffs(in,out,deg+1)
 *   fastkan 0 layernorm.weight 1 layernorm.bias 2 rbf.grid 3 spline_linear.weight(out,in*G)
 *           4 base_linear.weight|0 5 base_linear.bias|0
 *   sine    0 amplitudes(out,in,G) 1 freq(G) 2 phase(in,G) 3 bias(out)|0
 *   fourier 0 fouriercoeffs(2,out,in,G) 1 bias(out)|0
 * Outputs (any may be NULL to skip): wp_f32 [g][out][in*M] (K index i*M+m, effkan's
 * spline_scaler folded in -- models/effkan.py:166-172,180-183); wtiles_bf16 / wtiles_t_bf16:
 * the bf16 tcgen05 operand images of the same weights for the forward / input-gradient
 * tensor-core kernels (opaque pre-tiled layouts, sizes from kv_kan_tc_bytes; cut straight from the module tensors);
 * fparams [g][fparam_stride], bias [g][out] (zeros where the module has none). */
int kv_kan_pack(const kv_kan_desc_t* d, const uint64_t* ptr_table, float* wp_f32,
                void* wtiles_bf16, void* wtiles_t_bf16, float* fparams, float* bias, void* stream);

/* kv_kan_pack + kv_kan_fwd (exact fp32 kernels) of one layer in ONE call that allocates nothing.
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The scratch that
 * kv_kan_pack otherwise writes into caller-allocated buffers (packed weights [g][out][in*M], prologue parameters
 * [g][fparam_stride], bias [g][out], each rounded up to 64 floats) lives in a file-scope device array of
 * KV_NOALLOC_SCRATCH_FLOATS floats (4 MiB) that every GPU loading the library holds until the process exits.
 * A layer that needs more returns KV_ERR_BAD_SHAPE (nothing launched; kv_last_cuda_error() names both sizes);
 * KV_FLAG_BF16_TC returns KV_ERR_UNSUPPORTED.  ptr_table as for kv_kan_pack; X, mean, rstd, residual, pos, Y as
 * for kv_kan_fwd; d->fparam_stride >= kv_kan_fparam_count(d).  Every call shares the one array and it keeps the
 * last call's contents: each call rewrites all it reads, but calls that OVERLAP on the GPU (two streams, two
 * host threads) race -- serialise them. */
#define KV_NOALLOC_SCRATCH_FLOATS (1 << 20)
int kv_kan_fwd_noalloc(const kv_kan_desc_t* d, const uint64_t* ptr_table, const float* X, const float* mean,
                       const float* rstd, const float* residual, const float* pos, float* Y, void* stream);

/* Bytes of the tensor-core operand images of kv_kan_pack (KV_FLAG_BF16_TC path).  Returns
 * KV_ERR_UNSUPPORTED when the layer cannot run on the tensor-core kernels (M > KV_MAX_M);
 * *wtiles_t_bytes == 0 (input gradient: more than 32 basis functions) / *dy_image_bytes == 0 mean that gradient runs on
 * the fp32 kernels for this layer. */
int kv_kan_tc_bytes(const kv_kan_desc_t* d, size_t* wtiles_bytes, size_t* wtiles_t_bytes,
                    size_t* dy_image_bytes);

/* Tensor-core gradient path: bf16 copy of dY (times the ReLU mask when KV_FLAG_RELU is set) in the
 * operand tiling both gradient kernels read; made once per backward pass of a layer.
 * dy_image: *dy_image_bytes of kv_kan_tc_bytes (0 = this layer's gradients run on the fp32 kernels). */
int kv_kan_dy_image(const kv_kan_desc_t* d, const float* dY, const float* act_mask, void* dy_image,
                    void* stream);

/* Extents of that image: [n_groups][t8_groups][o8_groups][8 rows][8 outputs] bf16, zero padded (the tensor-core view of
 * the layer: the merged q/k/v maps of one head count as one group with 3 * d_head outputs).  For callers that write the
 * image themselves (kv_attn_image_bwd). */
int kv_kan_dy_image_geom(const kv_kan_desc_t* d, int* t8_groups, int* o8_groups);

/* ---- MSA on the tensor cores (attention.py:181-202): q/k/v GEMM -> bf16 attention image -> tcgen05 attention ----
 * Attention image: img[b][h][part (q, k, v)][n8][d8][n % 8][d % 8] bf16, n < roundup(N, 16), d < roundup(d_head, 16),
 * zero padded: every 8 x 8 block is one core matrix of the tcgen05 shared-memory descriptors, so a part is one
 * bulk copy and serves as K-major or MN-major operand alike. */
size_t kv_attn_image_bytes(int B, int N, int H, int d_head);

/* Forward of the grouped per-head q/k/v maps (the layer of kv_kan_fwd with x_share = 3, bf16 tensor-core path) whose
 * epilogue writes the attention image instead of an fp32 Y.  T = B * tokens_per_sample rows. */
int kv_kan_fwd_attn_image(const kv_kan_desc_t* d, const float* X, const void* wtiles_bf16, const float* fparams,
                          const float* bias, const float* mean, const float* rstd, void* attn_image,
                          int tokens_per_sample, void* stream);

/* softmax(q k^T / sqrt(d_head)) v per (sample, head) from the image (attention.py:199-200): S and O accumulate in
 * TMEM (tcgen05.mma), out [B*N][H*d_head] fp32 (+ residual, model.py:34), lse [B][H][N] (utils.py:222).
 * N <= 256, d_head <= 64, d_head % 8 == 0; B = 0 probes the shape. */
int kv_attn_image_fwd(const void* attn_image, const float* residual, float* out, float* lse, int B, int N, int H,
                      int d_head, void* stream);

/* Backward (recompute scheme of utils.py:229-295).  The gradients w.r.t. q, k, v leave as the bf16 dY image of the
 * q/k/v layer (geometry from kv_kan_dy_image_geom), ready for kv_kan_bwd_input / kv_kan_bwd_weight with dY = NULL. */
int kv_attn_image_bwd(const void* attn_image, const float* dout, const float* lse, void* dy_image, int B, int N, int H,
                      int d_head, int t8_groups, int o8_groups, void* stream);

/* Per-(row, slice) mean and 1/sqrt(var+eps) over `width` consecutive columns:
 * the LayerNorm statistics of FastKANLayer.layernorm (models/fastkan.py:57,68).
 * mean/rstd: [T][n_slices]. */
int kv_row_stats(const float* X, int T, int ldx, int n_slices, int width, float eps,
                 float* mean, float* rstd, void* stream);

/* Forward of the layer(s): replaces KANLinear.forward (models/effkan.py:174-187),
 * ChebyKANLayer.forward (models/cheby.py:36-48), FastKANLayer.forward
 * (models/fastkan.py:66-76), SineKANLayer.forward (models/sinekan.py:81-91),
 * NaiveFourierKANLayer.forward (models/nfkan.py:36-52) and F.linear.  The basis
 * expansion is evaluated on the fly in the GEMM prologue and never stored.
 * residual (nullable, same addressing as Y) and pos (nullable, [rows_out][n_groups*out])
 * are added in the epilogue; mean/rstd only for fastkan. */
int kv_kan_fwd(const kv_kan_desc_t* d, const float* X, const float* wp_f32, const void* wtiles_bf16,
               const float* fparams, const float* bias, const float* mean, const float* rstd,
               const float* residual, const float* pos, float* Y, void* stream);

/* Input gradient: dX[t, slice+i] = sum_g sum_m Phi'_m(x) * (dY_g . Wp_g)[t, i*M+m]
 * (what autograd derives from the modules above; basis recomputed, not stored).
 * act_mask (nullable): the layer's own forward output when KV_FLAG_RELU is set.
 * dXn (fastkan only): [T][n_groups*in] gradient w.r.t. the LayerNorm output, consumed
 * by kv_fastkan_ln_bwd.  dfparams [g][fparam_stride] (sine: d freq accumulated with
 * atomics, must be zeroed by the caller; others untouched).  dX is OVERWRITTEN. */
int kv_kan_bwd_input(const kv_kan_desc_t* d, const float* X, const float* dY, const float* act_mask,
                     const float* wp_f32, const void* wtiles_t_bf16, const void* dy_image,
                     const float* fparams, const float* mean, const float* rstd, float* dX, float* dXn,
                     float* dfparams, void* stream);

/* Weight gradient in packed layout: dWp[g][o][i*M+m] = sum_t dY[t,o] * Phi_m(x[t,i]),
 * dbias[g][o] = sum_t dY[t,o] (nullable).  dWp and dbias must be ZEROED by the caller
 * (split over row ranges, accumulated with fp32 atomics). */
int kv_kan_bwd_weight(const kv_kan_desc_t* d, const float* X, const float* dY, const float* act_mask,
                      const void* dy_image, const float* fparams, const float* mean, const float* rstd,
                      float* dWp, float* dbias, void* stream);

/* FastKAN: LayerNorm backward of the RBF branch (models/fastkan.py:57,68):
 * dX[t, slice+i] += LN'(dXn_g)[t,i] for every group on that slice; d layernorm.weight
 * / bias accumulated into dfparams[g][0:in] / [in:2in] (zeroed by the caller). */
int kv_fastkan_ln_bwd(const kv_kan_desc_t* d, const float* X, const float* dXn, const float* fparams,
                      const float* mean, const float* rstd, float* dX, float* dfparams, void* stream);

/* Scatter packed gradients back to the reference parameter layouts.  grad_table:
 * device array [n_groups][KV_PTRS_PER_GROUP] of OUTPUT pointers with the same slot
 * meaning as kv_kan_pack (0 = skip); ptr_table = the forward parameters (effkan
 * needs spline_weight and spline_scaler for the product rule, models/effkan.py:166-172).
 * With KV_FLAG_ACCUM_GRADS the values are added to what the gradient tensors hold (the caller's
 * p.grad buffers), which is how autograd accumulates. */
int kv_kan_unpack_grads(const kv_kan_desc_t* d, const float* dWp, const float* dfparams,
                        const float* dbias, const uint64_t* ptr_table, const uint64_t* grad_table,
                        void* stream);

/* softmax(q k^T / sqrt(d_h)) v per (sample, head): attention.py:199-200, with the
 * online-softmax / log-sum-exp bookkeeping of utils.py:137-227.
 * qkv: [B*N][H][3][d_h] (the grouped layer's output), out: [B*N][H*d_h] (heads
 * concatenated in order, attention.py:201), residual nullable (model.py:34),
 * lse: [B][H][N] saved for the backward.  Any N: a head that does not fit one CTA's shared memory
 * (N > ~280 forward, ~215 backward at d_h 64) runs on the tiled kernels of attn_tiled.cu; d_h <= 128. */
int kv_attn_fwd(const float* qkv, const float* residual, float* out, float* lse,
                int B, int N, int H, int d_head, void* stream);

/* Recompute backward (utils.py:229-295): p = exp(s - lse), ds = p * scale * (dp - D);
 * D_i = sum_j p_ij dp_ij is recomputed in the row pass (equal to sum(do * o),
 * utils.py:277), so the forward output is not needed.  dqkv: same layout as qkv,
 * overwritten.  dout: [B*N][H*d_h]. */
int kv_attn_bwd(const float* qkv, const float* dout, const float* lse, float* dqkv,
                int B, int N, int H, int d_head, void* stream);

/* Tensor-core variants of the two calls above (bf16 operands, fp32 softmax and accumulation; same
 * arguments, layouts and saved lse, tolerance 2e-2): flash-style online softmax over key chunks in
 * the forward, two-pass recompute in the backward.  Return KV_ERR_UNSUPPORTED (nothing launched) when
 * d_head > 64, d_head % 4 != 0 or a head does not fit in shared memory; the caller then uses the
 * fp32 entry points. */
int kv_attn_fwd_bf16(const float* qkv, const float* residual, float* out, float* lse,
                     int B, int N, int H, int d_head, void* stream);
int kv_attn_bwd_bf16(const float* qkv, const float* dout, const float* lse, float* dqkv,
                     int B, int N, int H, int d_head, void* stream);

/* Attention of the FlashAttention module (attention.py:61-109 -> FlashAttentionFunction, utils.py:137-295),
 * the block of type="flash-attn" (model.py:93-95,156-159): q/k/v are separate fp32 matrices in
 * 'b n (h d)' layout with row strides ldq/ldk/ldv (head h = columns [h*d_head, (h+1)*d_head); k and v may
 * be the two halves of the to_kv output), Nk may differ from Nq (context), out has row stride ldo.
 * mask: nullable bytes (torch.bool), element (b,h,i,j) at b*mask_stride_b + h*mask_stride_h +
 * i*mask_stride_q + j, stride 0 = broadcast (a 'b n' key mask has strides (Nk,0,0)); false = hidden.
 * causal != 0: key j is visible to query i iff j <= i - max(Nk-Nq,0) (the reference's rule, utils.py:142,
 * 175,192-194).  Hidden scores are -FLT_MAX (not -inf), mask-false weights are zeroed after the
 * exponential and the row sum is clamped at 1e-10, so fully hidden rows behave as in the reference.
 * lse [B][H][Nq] and out are saved for the backward (D_i = sum(dO_i * O_i), utils.py:277).
 * The q/k bucket sizes of the reference only change the summation order and have no counterpart.
 * Any Nq / Nk: heads that fit one CTA's shared memory (roughly Nk <= 400 forward, Nq + Nk <= 430 backward at
 * d_head 64) run resident, longer ones on the tiled kernels of attn_tiled.cu (same arithmetic, keys / queries
 * streamed in tiles of 64).  KV_ERR_BAD_SHAPE only for d_head > 128. */
typedef struct kv_attn_general {
  int32_t B, H, Nq, Nk, d_head;
  int32_t ldq, ldk, ldv, ldo;
  int32_t causal;
  int64_t mask_stride_b, mask_stride_h, mask_stride_q;
} kv_attn_general_t;
int kv_attn_general_fwd(const kv_attn_general_t* a, const float* q, const float* k, const float* v,
                        const unsigned char* mask, float* out, float* lse, void* stream);
/* dq/dk/dv use the layouts (row strides) of q/k/v and are overwritten. */
int kv_attn_general_bwd(const kv_attn_general_t* a, const float* q, const float* k, const float* v,
                        const unsigned char* mask, const float* out, const float* dout, const float* lse,
                        float* dq, float* dk, float* dv, void* stream);

/* nn.LayerNorm(d), eps 1e-5, affine (model.py:22,24,107).  mean/rstd [T] saved. */
int kv_layernorm_fwd(const float* X, const float* gamma, const float* beta, float* Y,
                     float* mean, float* rstd, int T, int d, int ldx, float eps, void* stream);
/* dX = LN'(dY) (+ dresidual if non-null: the skip path of model.py:34,36);
 * dgamma/dbeta accumulated with atomics (zeroed by the caller). */
int kv_layernorm_bwd(const float* X, const float* dY, const float* gamma, const float* mean,
                     const float* rstd, const float* dresidual, float* dX, float* dgamma,
                     float* dbeta, int T, int d, int ldx, float eps, void* stream);

/* VisionTransformer.patchify (model.py:111-126), bit-exact index map:
 * patches[b, i*np+j, c*ph*pw + y*pw + x] = images[b, c, i*ph+y, j*pw+x]. */
int kv_patchify(const float* images, float* patches, int B, int C, int H, int W, int n_patches,
                void* stream);
/* its adjoint (gradient w.r.t. the images) */
int kv_patchify_bwd(const float* dpatches, float* dimages, int B, int C, int H, int W,
                    int n_patches, void* stream);

/* Class-token rows: tokens[b, 0, :] = v_class + pos[0] (model.py:148-152). */
int kv_cls_rows_fwd(const float* v_class, const float* pos, float* tokens, int B, int N, int d,
                    void* stream);
/* d v_class = sum_b dtokens[b,0,:] */
int kv_cls_rows_bwd(const float* dtokens, float* dv_class, int B, int N, int d, void* stream);

/* Strided row gather / scatter-add used for out[:,0] (model.py:165): Y[b,:] = X[b*stride_rows,:]. */
int kv_gather_rows(const float* X, float* Y, int B, int row_stride, int d, void* stream);
int kv_scatter_rows(const float* dY, float* dX, int B, int row_stride, int d, void* stream);

/* ---- the two ops either side of the path in every iteration ---- */

/* nn.CrossEntropyLoss() of train.py:22,36 (class-index targets, mean reduction, ignore_index -100 by
 * default): loss[0] = mean_i (logsumexp(z_i) - z_i[y_i]) over the rows with y_i != ignore_index, and, when
 * dlogits is non-null, d loss / d logits = (softmax(z_i) - onehot(y_i)) / n_valid (zero rows for ignored
 * labels) in the same pass.  logits [B][n_classes] fp32, labels [B] int64.  workspace: device floats,
 * kv_cross_entropy_workspace_floats() of them, zeroed ONCE by the caller (the kernel leaves it zeroed);
 * the block partials are added in block order, so the loss is bit-reproducible.  A label outside
 * [0, n_classes) makes the loss NaN. */
int kv_cross_entropy_workspace_floats(void);
int kv_cross_entropy(const float* logits, const long long* labels, int B, int n_classes,
                     long long ignore_index, float* loss, float* dlogits, float* workspace, float* probs,
                     long long* pred, int* confusion, void* stream);
/* The three nullable outputs serve the per-step metric collection of train.py:42-44,74-76 without a host
 * round trip: probs [B][n_classes] = softmax(logits) (y_pred_proba), pred [B] = argmax (lowest index on
 * ties, like torch.argmax), confusion [n_classes][n_classes] int32 += 1 at (label, pred) -- accuracy,
 * balanced accuracy and weighted F1 (utils.py:38-40) are functions of that matrix, read once per epoch.
 *
 * The training transform of train.py:99-104 (RandomHorizontalFlip -> RandomCrop(H, padding=pad) -> ToTensor ->
 * Normalize(mean, std)) from the dataset's uint8 [B][H][W][C] pixels to the fp32 [B][C][H][W] batch:
 * flip [B] bytes (nullable = none), top/left [B] crop offsets in [0, 2*pad] (nullable = centred, i.e. the
 * test transform of train.py:106-109), mean/stdv [C] device floats.  The random draws are the caller's. */
int kv_image_augment(const unsigned char* images_hwc, const unsigned char* flip, const int* top, const int* left,
                     int B, int H, int W, int C, int pad, const float* mean, const float* stdv, float* out,
                     void* stream);

/* torch.optim.Adam(params, lr) of train.py:23,39-40 (betas/eps arguments, no weight decay, no amsgrad)
 * over FLAT buffers holding every parameter of the model: one launch for the whole model instead of
 * one per parameter tensor.
 *   g = grad * grad_scale  (grad_scale = 1/world turns the all-reduced sum into the data-parallel mean)
 *   m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2;  p -= lr/(1-b1^t) * m / (sqrt(v)/sqrt(1-b2^t) + eps)
 * n: elements, a multiple of 4, all four buffers 16-byte aligned.  zero_grad != 0 clears grad after it
 * is read (the next iteration's zero_grad()).  state: 3 device doubles {t, lr/(1-b1^t), sqrt(1-b2^t)},
 * zero-initialised by the caller; t is advanced on the device so a captured CUDA graph keeps counting. */
int kv_adam_step(float* param, float* grad, float* exp_avg, float* exp_avg_sq, long long n,
                 float lr, float beta1, float beta2, float eps, float grad_scale, int zero_grad,
                 double* state, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* KANVIT_B200_H */
