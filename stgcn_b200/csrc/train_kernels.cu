// The two steps of the reference's training_step that sit right after the block stack:
//   * F.cross_entropy(output, y), mean reduction (src/lightning_model.py:202)  -> one kernel forward (loss AND d loss / d logits),
//     one kernel backward (scale by the incoming gradient);
//   * optim.Adam(self.parameters(), lr) (src/lightning_model.py:196-197)       -> ONE launch per 64 parameter tensors that walks
//     2048-element chunks of every tensor (torch's multi-tensor kernel uses 65536-element chunks: 47 CTAs for this model's 3.06 M
//     parameters on a 148-SM part).
#include "kernels.cuh"

namespace stgcn {

// ---- cross entropy ------------------------------------------------------------------------------------------------------------
// one warp per row; loss = mean over rows with label != ignore_index of (logsumexp(row) - row[label]);
// dlogits[r][j] = (softmax(row)[j] - [j == label]) / n_valid   (0 for ignored rows).  Single CTA: the row sums are reduced in a
// fixed order, so the loss is deterministic.
constexpr int kCeThreads = 512;
__global__ void __launch_bounds__(kCeThreads) cross_entropy_fwd_kernel(const float* __restrict__ logits, const long long* __restrict__ labels,
                                                                       int n, int classes, long long ignore_index,
                                                                       float* __restrict__ loss, float* __restrict__ dlogits) {
  pdl_enter();
  __shared__ float s_loss[kCeThreads / 32];
  __shared__ int s_cnt[kCeThreads / 32];
  __shared__ float s_inv;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32, nw = kCeThreads / 32;
  // pass 0: number of rows that count
  int cnt = 0;
  for (int r = threadIdx.x; r < n; r += kCeThreads) cnt += labels[r] != ignore_index;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  if (lane == 0) s_cnt[warp] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int c = 0;
    for (int w = 0; w < nw; ++w) c += s_cnt[w];
    s_inv = c > 0 ? 1.f / (float)c : 0.f;        // all rows ignored: torch returns nan; the gradient here is 0 and the loss nan
    s_cnt[0] = c;
  }
  __syncthreads();
  const float inv = s_inv;
  float acc = 0.f;
  for (int r = warp; r < n; r += nw) {
    const float* row = logits + (size_t)r * classes;
    const long long lab = labels[r];
    float mx = -INFINITY;
    for (int j = lane; j < classes; j += 32) mx = fmaxf(mx, row[j]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float se = 0.f;
    for (int j = lane; j < classes; j += 32) se += expf(row[j] - mx);
    se = warp_sum(se);
    // a label outside [0, classes) that is not ignore_index (torch raises a device assert): no out-of-bounds read here --
    // the row contributes a NaN to the loss and a zero gradient
    const bool bad = lab != ignore_index && (lab < 0 || lab >= classes);
    const bool valid = lab != ignore_index && !bad;
    const float rinv = valid ? inv / se : 0.f;
    for (int j = lane; j < classes; j += 32) {
      float g = expf(row[j] - mx) * rinv;
      if (valid && j == (int)lab) g -= inv;
      dlogits[(size_t)r * classes + j] = g;
    }
    if (valid && lane == 0) acc += (logf(se) + mx) - row[lab];
    if (bad && lane == 0) acc = __int_as_float(0x7fc00000);
  }
  if (lane == 0) s_loss[warp] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < nw; ++w) t += s_loss[w];
    loss[0] = s_cnt[0] > 0 ? t * inv : __int_as_float(0x7fc00000);
  }
}

// dst = src * g[0]
__global__ void __launch_bounds__(256) scale_by_scalar_kernel(const float* __restrict__ src, const float* __restrict__ g, size_t n,
                                                              float* __restrict__ dst) {
  pdl_enter();
  const float s = g[0];
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i] * s;
}

// ---- Adam -----------------------------------------------------------------------------------------------------------------------
constexpr int kAdamMaxTensors = 64;
constexpr int kAdamChunk = 2048;          // elements per CTA: 256 threads x 2 x float4
struct AdamTable {
  float* p[kAdamMaxTensors];
  const float* g[kAdamMaxTensors];
  float* m[kAdamMaxTensors];
  float* v[kAdamMaxTensors];
  unsigned n[kAdamMaxTensors];            // elements
  unsigned first_chunk[kAdamMaxTensors + 1];
  int count;
};
struct AdamHyper { float beta1, beta2, eps, weight_decay; };
// device-resident step state (CUDA-graph replays must advance the step without the host): {int64 step, float lr/bc1, float 1/sqrt(bc2)}
struct AdamDevState { long long step; float lr_over_bc1, inv_sqrt_bc2; };
__global__ void adam_tick_kernel(AdamDevState* s, long long set_step, double lr, double beta1, double beta2) {
  pdl_enter();
  if (threadIdx.x == 0) {
    const long long t = set_step > 0 ? set_step : s->step + 1;
    s->step = t;
    const double bc1 = 1.0 - pow(beta1, (double)t), bc2 = 1.0 - pow(beta2, (double)t);
    s->lr_over_bc1 = (float)(lr / bc1);
    s->inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
  }
}

struct AdamCoef { float lr_over_bc1, inv_sqrt_bc2, beta1, beta2, eps, weight_decay; };
__device__ __forceinline__ void adam_update(float& p, float g, float& m, float& v, const AdamCoef& h) {
  g = fmaf(h.weight_decay, p, g);                 // L2 (torch.optim.Adam, not AdamW)
  m = fmaf(1.f - h.beta1, g - m, m);              // exp_avg.lerp_(grad, 1 - beta1)
  v = fmaf(h.beta2, v, (1.f - h.beta2) * g * g);  // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
  const float denom = sqrtf(v) * h.inv_sqrt_bc2 + h.eps;
  p -= h.lr_over_bc1 * (m / denom);
}

__global__ void __launch_bounds__(256) adam_multi_kernel(const __grid_constant__ AdamTable tb, const __grid_constant__ AdamHyper h_in, const float2* __restrict__ dev_coef) {
  pdl_enter();
  // the two step-dependent coefficients come from the device-resident step state (written by adam_tick_kernel just before)
  AdamCoef h;
  { const float2 c = __ldcg(dev_coef); h.lr_over_bc1 = c.x; h.inv_sqrt_bc2 = c.y; }
  h.beta1 = h_in.beta1; h.beta2 = h_in.beta2; h.eps = h_in.eps; h.weight_decay = h_in.weight_decay;
  // which tensor does this chunk belong to (binary search over the chunk prefix, warp-uniform)
  int lo = 0, hi = tb.count;
  const unsigned chunk = blockIdx.x;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (tb.first_chunk[mid] <= chunk) lo = mid; else hi = mid;
  }
  const int t = lo;
  float* __restrict__ p = tb.p[t]; const float* __restrict__ g = tb.g[t];
  float* __restrict__ m = tb.m[t]; float* __restrict__ v = tb.v[t];
  const unsigned n = tb.n[t];
  const unsigned base = (chunk - tb.first_chunk[t]) * kAdamChunk;
  const unsigned end = min(n, base + kAdamChunk);
  const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                     reinterpret_cast<uintptr_t>(v)) & 15) == 0;
  if (vec) {
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const unsigned i = base + (it * 256 + threadIdx.x) * 4;
      if (i + 4 <= end) {
        float4 P = *reinterpret_cast<float4*>(p + i), M = *reinterpret_cast<float4*>(m + i), V = *reinterpret_cast<float4*>(v + i);
        const float4 G = *reinterpret_cast<const float4*>(g + i);
        adam_update(P.x, G.x, M.x, V.x, h); adam_update(P.y, G.y, M.y, V.y, h);
        adam_update(P.z, G.z, M.z, V.z, h); adam_update(P.w, G.w, M.w, V.w, h);
        *reinterpret_cast<float4*>(p + i) = P; *reinterpret_cast<float4*>(m + i) = M; *reinterpret_cast<float4*>(v + i) = V;
      } else {
        for (unsigned j = i; j < end; ++j) { float P = p[j], M = m[j], V = v[j]; adam_update(P, g[j], M, V, h); p[j] = P; m[j] = M; v[j] = V; }
      }
    }
  } else {
    for (unsigned j = base + threadIdx.x; j < end; j += 256) { float P = p[j], M = m[j], V = v[j]; adam_update(P, g[j], M, V, h); p[j] = P; m[j] = M; v[j] = V; }
  }
}

}  // namespace stgcn

using namespace stgcn;
extern "C" {

int32_t stgcn_cross_entropy_forward(int32_t n, int32_t classes, const float* logits, const int64_t* labels, int64_t ignore_index,
                                    float* loss, float* dlogits, void* stream) {
  if (n <= 0 || classes <= 0) { set_error("cross_entropy: bad shape n=%d classes=%d", n, classes); return STGCN_ERR_INVALID; }
  if (!logits || !labels || !loss || !dlogits) { set_error("cross_entropy: null pointer"); return STGCN_ERR_INVALID; }
  cudaStream_t st = (cudaStream_t)stream;
  KernelTimer kt_(KC_HEAD, 0.0, st);
  launch_k(cross_entropy_fwd_kernel, 1, kCeThreads, 0, st, logits, (const long long*)labels, n, classes, (long long)ignore_index, loss,
           dlogits);
  STGCN_CHECK_LAUNCH("cross_entropy_fwd");
  return STGCN_OK;
}

int32_t stgcn_cross_entropy_backward(int32_t n, int32_t classes, const float* dlogits_unit, const float* grad_loss, float* dlogits,
                                     void* stream) {
  if (n <= 0 || classes <= 0 || !dlogits_unit || !grad_loss || !dlogits) { set_error("cross_entropy_backward: bad argument"); return STGCN_ERR_INVALID; }
  cudaStream_t st = (cudaStream_t)stream;
  KernelTimer kt_(KC_HEAD, 0.0, st);
  const size_t total = (size_t)n * classes;
  launch_k(scale_by_scalar_kernel, (int)std::min<size_t>((total + 255) / 256, 148 * 4), 256, 0, st, dlogits_unit, grad_loss, total, dlogits);
  STGCN_CHECK_LAUNCH("cross_entropy_bwd");
  return STGCN_OK;
}

int32_t stgcn_adam_step(int32_t count, float* const* params, const float* const* grads, float* const* exp_avg,
                        float* const* exp_avg_sq, const int64_t* sizes, double lr, double beta1, double beta2, double eps,
                        double weight_decay, int64_t step, void* step_state, void* stream) {
  if (count < 0 || step < 0 || !step_state) { set_error("adam_step: count=%d step=%lld step_state=%p", count, (long long)step, step_state); return STGCN_ERR_INVALID; }
  if (count > 0 && (!params || !grads || !exp_avg || !exp_avg_sq || !sizes)) { set_error("adam_step: null table"); return STGCN_ERR_INVALID; }
  cudaStream_t st = (cudaStream_t)stream;
  KernelTimer kt_(KC_PACK, 0.0, st);
  // the step number lives on the device: advance (or set) it there; the bias corrections are formed in double by that kernel
  // and the update kernels read them from the state
  AdamDevState* ds = (AdamDevState*)step_state;
  launch_k(adam_tick_kernel, 1, 32, 0, st, ds, (long long)step, lr, beta1, beta2);
  STGCN_CHECK_LAUNCH("adam_tick");
  AdamHyper h;
  h.beta1 = (float)beta1; h.beta2 = (float)beta2; h.eps = (float)eps; h.weight_decay = (float)weight_decay;
  const float2* dev_coef = reinterpret_cast<const float2*>(&ds->lr_over_bc1);
  int i = 0;
  while (i < count) {
    AdamTable tb; tb.count = 0;
    unsigned chunks = 0;
    while (i < count && tb.count < kAdamMaxTensors) {
      const int64_t n = sizes[i];
      if (n < 0 || n > 0xfff00000ll) { set_error("adam_step: tensor %d has %lld elements", i, (long long)n); return STGCN_ERR_UNSUPPORTED; }
      if (n > 0) {
        if (!params[i] || !grads[i] || !exp_avg[i] || !exp_avg_sq[i]) { set_error("adam_step: null pointer in tensor %d", i); return STGCN_ERR_INVALID; }
        const int k = tb.count++;
        tb.p[k] = params[i]; tb.g[k] = grads[i]; tb.m[k] = exp_avg[i]; tb.v[k] = exp_avg_sq[i]; tb.n[k] = (unsigned)n;
        tb.first_chunk[k] = chunks;
        chunks += (unsigned)((n + kAdamChunk - 1) / kAdamChunk);
      }
      ++i;
    }
    if (tb.count == 0) break;
    tb.first_chunk[tb.count] = chunks;
    launch_k(adam_multi_kernel, chunks, 256, 0, st, tb, h, dev_coef);
    STGCN_CHECK_LAUNCH("adam_step");
  }
  return STGCN_OK;
}

}  // extern "C"
