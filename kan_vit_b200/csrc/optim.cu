// The two ops that bracket the hot path in every training iteration (train.py:22-23,35-40 of the
// reference): nn.CrossEntropyLoss on the logits and the Adam update.  Both are HBM-bound:
//   * kv_cross_entropy: loss and d loss / d logits in one pass over the logits (one warp per row);
//   * kv_adam_step: ONE launch over the flat parameter / gradient / moment buffers of the whole model
//     (16 B read + 16 B written per parameter), with the 1/world gradient scale of the data-parallel
//     mean and the zeroing of the gradient buffer for the next iteration folded in.
#include "internal.cuh"

namespace kv {

// ------------------------------------------------------------------------------
// Adam (torch.optim.Adam defaults of train.py:23: no weight decay, no amsgrad)
//   m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2
//   p -= (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
// The step counter lives on the device (state[0]) so that a captured CUDA graph advances it on replay.
// ------------------------------------------------------------------------------
__global__ void adam_advance_kernel(double* __restrict__ state, double lr, double beta1, double beta2) {
  const double t = state[0] + 1.0;
  state[0] = t;
  state[1] = lr / (1.0 - pow(beta1, t));   // step size
  state[2] = sqrt(1.0 - pow(beta2, t));    // sqrt of the second bias correction
}

constexpr int ADAM_THREADS = 256;

__global__ void __launch_bounds__(ADAM_THREADS)
adam_kernel(float4* __restrict__ P, float4* __restrict__ G, float4* __restrict__ M, float4* __restrict__ V,
            long long n4, const double* __restrict__ state, float beta1, float beta2, float eps,
            float grad_scale, int zero_grad) {
  const float step_size = (float)state[1];
  const float inv_bc2s = (float)(1.0 / state[2]);
  const float c1 = 1.0f - beta1, c2 = 1.0f - beta2;
  const long long stride = (long long)gridDim.x * ADAM_THREADS;
  for (long long i = (long long)blockIdx.x * ADAM_THREADS + threadIdx.x; i < n4; i += stride) {
    float4 p = P[i], m = M[i], v = V[i];
    const float4 g4 = __ldcs(G + i);
    float* pp = reinterpret_cast<float*>(&p);
    float* mp = reinterpret_cast<float*>(&m);
    float* vp = reinterpret_cast<float*>(&v);
    const float* gp = reinterpret_cast<const float*>(&g4);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float g = gp[k] * grad_scale;
      mp[k] = fmaf(beta1, mp[k], c1 * g);
      vp[k] = fmaf(beta2, vp[k], c2 * g * g);
      const float denom = sqrtf(vp[k]) * inv_bc2s + eps;
      pp[k] -= step_size * (mp[k] / denom);
    }
    P[i] = p;
    M[i] = m;
    V[i] = v;
    if (zero_grad) G[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// ------------------------------------------------------------------------------
// nn.CrossEntropyLoss(): mean over the rows whose label is not ignore_index of
//   lse(z_i) - z_i[y_i];   d loss / d z_i = (softmax(z_i) - onehot(y_i)) / n_valid
// One warp per row.  Every block counts the valid labels itself (B labels, L2-resident), writes the
// sum of its rows' losses to workspace[1 + block]; the block that finishes last adds the partials in
// block order (deterministic) and writes the mean.  workspace[0] is the arrival counter (reset here).
// ------------------------------------------------------------------------------
constexpr int CE_THREADS = 256;
constexpr int CE_WARPS = CE_THREADS / 32;
constexpr int CE_MAX_BLOCKS = 148;

__global__ void __launch_bounds__(CE_THREADS)
cross_entropy_kernel(const float* __restrict__ Z, const long long* __restrict__ labels, int B, int Cn,
                     long long ignore_index, float* __restrict__ loss, float* __restrict__ dZ,
                     float* __restrict__ workspace, float* __restrict__ probs, long long* __restrict__ pred,
                     int* __restrict__ confusion) {
  __shared__ float red[CE_WARPS];
  __shared__ int cnt[CE_WARPS];
  __shared__ int is_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int c = 0;
  for (int i = threadIdx.x; i < B; i += CE_THREADS) c += labels[i] != ignore_index;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if (lane == 0) cnt[warp] = c;
  __syncthreads();
  int n_valid = 0;
#pragma unroll
  for (int w = 0; w < CE_WARPS; ++w) n_valid += cnt[w];
  const float inv_n = n_valid > 0 ? 1.0f / (float)n_valid : 0.0f;

  float acc = 0.0f;  // lane 0: sum of this warp's row losses (rows visited in increasing order)
  for (int r = blockIdx.x * CE_WARPS + warp; r < B; r += gridDim.x * CE_WARPS) {
    const float* z = Z + (size_t)r * Cn;
    const long long y = labels[r];
    float mx = -INFINITY;
    int arg = 0x7fffffff;
    for (int j = lane; j < Cn; j += 32) {
      if (z[j] > mx) { mx = z[j]; arg = j; }   // strict: the lowest index of a lane's maximum
    }
    const float lane_mx = mx;
    mx = warp_max(mx);
    if (pred || confusion) {
      // torch.argmax (train.py:43): lowest index attaining the row maximum
      int cand = lane_mx == mx ? arg : 0x7fffffff;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) cand = min(cand, __shfl_xor_sync(0xffffffffu, cand, o));
      if (lane == 0) {
        if (pred) pred[r] = cand;
        if (confusion && y >= 0 && y < Cn && cand < Cn) atomicAdd(confusion + (size_t)y * Cn + cand, 1);
      }
    }
    float s = 0.0f;
    for (int j = lane; j < Cn; j += 32) s += __expf(z[j] - mx);
    s = warp_sum(s);
    const bool valid = y != ignore_index;
    const float inv_s = 1.0f / s;
    if (probs) {  // softmax(y_hat, dim=1), train.py:44
      float* pr = probs + (size_t)r * Cn;
      for (int j = lane; j < Cn; j += 32) pr[j] = __expf(z[j] - mx) * inv_s;
    }
    if (dZ) {
      float* dz = dZ + (size_t)r * Cn;
      for (int j = lane; j < Cn; j += 32) {
        const float p = __expf(z[j] - mx) * inv_s;
        dz[j] = valid ? (p - (j == y ? 1.0f : 0.0f)) * inv_n : 0.0f;
      }
    }
    // a label outside [0, n_classes) poisons the loss (torch raises a device assert there)
    if (valid && lane == 0) acc += (y >= 0 && y < Cn) ? (logf(s) + mx) - z[y] : __int_as_float(0x7fc00000);
  }
  if (lane == 0) red[warp] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float part = 0.0f;
#pragma unroll
    for (int w = 0; w < CE_WARPS; ++w) part += red[w];
    workspace[1 + blockIdx.x] = part;
    __threadfence();
    const unsigned arrived = atomicAdd(reinterpret_cast<unsigned*>(workspace), 1u);
    is_last = arrived == gridDim.x - 1;
  }
  __syncthreads();
  if (is_last && threadIdx.x == 0) {
    __threadfence();
    float total = 0.0f;
    for (unsigned b = 0; b < gridDim.x; ++b) total += __ldcg(workspace + 1 + b);
    // torch: an all-ignored batch gives nan (0 / 0)
    *loss = n_valid > 0 ? total * inv_n : __int_as_float(0x7fc00000);
    *reinterpret_cast<unsigned*>(workspace) = 0u;
  }
}

// ------------------------------------------------------------------------------
// The training transform of train.py:99-104 on the GPU: RandomHorizontalFlip -> RandomCrop(size, padding)
// -> ToTensor -> Normalize, from the dataset's uint8 HWC pixels to the fp32 CHW batch the model reads.
// The random draws (flip, crop offsets in [0, 2*pad]) are made by the caller.  Padding is zero BEFORE
// ToTensor, so a padded pixel becomes (0 - mean) / std.
//   out[b,c,y,x] = (P(y + top_b, x + left_b) / 255 - mean_c) / std_c,
//   P(r, q) = 0 outside [pad, pad+H) x [pad, pad+W), else src[b, r-pad, flip ? W-1-(q-pad) : q-pad, c]
// ------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
augment_kernel(const unsigned char* __restrict__ src, const unsigned char* __restrict__ flip,
               const int* __restrict__ top, const int* __restrict__ left, int B, int H, int W, int Cc, int pad,
               const float* __restrict__ mean, const float* __restrict__ stdv, float* __restrict__ out) {
  const long long total = (long long)B * Cc * H * W;
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < total;
       p += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(p % W);
    const int y = (int)((p / W) % H);
    const int c = (int)((p / ((long long)W * H)) % Cc);
    const int b = (int)(p / ((long long)W * H * Cc));
    const int r = y + (top ? top[b] : pad) - pad;
    int q = x + (left ? left[b] : pad) - pad;
    float v = 0.0f;
    if (r >= 0 && r < H && q >= 0 && q < W) {
      if (flip && flip[b]) q = W - 1 - q;
      v = (float)src[(((size_t)b * H + r) * W + q) * Cc + c] / 255.0f;
    }
    out[p] = (v - mean[c]) / stdv[c];
  }
}

}  // namespace kv

using namespace kv;

extern "C" {

int kv_adam_step(float* param, float* grad, float* exp_avg, float* exp_avg_sq, long long n,
                 float lr, float beta1, float beta2, float eps, float grad_scale, int zero_grad,
                 double* state, void* stream) {
  if (n < 0 || (n & 3)) return KV_ERR_BAD_ARG;
  if (n == 0) return KV_OK;
  if (!param || !grad || !exp_avg || !exp_avg_sq || !state) return KV_ERR_BAD_ARG;
  if (((uintptr_t)param | (uintptr_t)grad | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq) & 15) return KV_ERR_ALIGNMENT;
  cudaStream_t st = (cudaStream_t)stream;
  adam_advance_kernel<<<1, 1, 0, st>>>(state, (double)lr, (double)beta1, (double)beta2);
  KV_CHECK_LAUNCH("adam_advance_kernel");
  const long long n4 = n / 4;
  // 148 SMs x 8 resident blocks of 256 threads; grid-stride over the rest
  const long long want = (n4 + ADAM_THREADS - 1) / ADAM_THREADS;
  const int grid = (int)(want < 148 * 8 ? want : 148 * 8);
  adam_kernel<<<grid, ADAM_THREADS, 0, st>>>(reinterpret_cast<float4*>(param), reinterpret_cast<float4*>(grad),
                                             reinterpret_cast<float4*>(exp_avg), reinterpret_cast<float4*>(exp_avg_sq),
                                             n4, state, beta1, beta2, eps, grad_scale, zero_grad);
  KV_CHECK_LAUNCH("adam_kernel");
  return KV_OK;
}

int kv_cross_entropy_workspace_floats(void) { return 1 + CE_MAX_BLOCKS; }

int kv_cross_entropy(const float* logits, const long long* labels, int B, int n_classes,
                     long long ignore_index, float* loss, float* dlogits, float* workspace, float* probs,
                     long long* pred, int* confusion, void* stream) {
  if (B <= 0 || n_classes <= 0) return KV_ERR_BAD_ARG;
  if (!logits || !labels || !loss || !workspace) return KV_ERR_BAD_ARG;
  int grid = ceil_div(B, CE_WARPS);
  if (grid > CE_MAX_BLOCKS) grid = CE_MAX_BLOCKS;
  cross_entropy_kernel<<<grid, CE_THREADS, 0, (cudaStream_t)stream>>>(logits, labels, B, n_classes, ignore_index,
                                                                     loss, dlogits, workspace, probs, pred,
                                                                     confusion);
  KV_CHECK_LAUNCH("cross_entropy_kernel");
  return KV_OK;
}

int kv_image_augment(const unsigned char* images_hwc, const unsigned char* flip, const int* top, const int* left,
                     int B, int H, int W, int C, int pad, const float* mean, const float* stdv, float* out,
                     void* stream) {
  if (B < 0 || H <= 0 || W <= 0 || C <= 0 || pad < 0) return KV_ERR_BAD_ARG;
  if (B == 0) return KV_OK;
  if (!images_hwc || !mean || !stdv || !out) return KV_ERR_BAD_ARG;
  const long long total = (long long)B * C * H * W;
  const long long want = (total + 255) / 256;
  const int grid = (int)(want < 148 * 8 ? want : 148 * 8);
  augment_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(images_hwc, flip, top, left, B, H, W, C, pad, mean, stdv, out);
  KV_CHECK_LAUNCH("augment_kernel");
  return KV_OK;
}

}  // extern "C"
