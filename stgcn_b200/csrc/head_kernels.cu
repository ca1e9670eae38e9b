// Classification head of L_STGCN.forward (lightning_model.py:105-107):
// F.avg_pool2d over the whole (T,V) plane, view, nn.Linear(C, classes).
// Tiny (2*N*C*classes flops): one pooling pass over the final activation,
// one warp per logit.
#include "kernels.cuh"

namespace stgcn {

// pooled[n][c] = mean_{t,v} feat[n][t][v][c]; grid (n, row-splits), atomics into zeroed pooled
template <class T>
__global__ void __launch_bounds__(256) pool_kernel(const T* __restrict__ feat, int rows, int C, float inv_rows, float* __restrict__ pooled) {
  pdl_enter();
  const int n = blockIdx.x;
  const T* f = feat + (size_t)n * rows * C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s = 0.f;
    for (int r = blockIdx.y; r < rows; r += gridDim.y) s += ldf(f + (size_t)r * C + c);
    atomicAdd(pooled + (size_t)n * C + c, s * inv_rows);
  }
}

// vector form: a thread owns one 16-byte channel group of one sample and strides over rows, 4 loads in flight
template <class T>
__global__ void __launch_bounds__(256) pool_vec_kernel(const T* __restrict__ feat, int rows, int C, float inv_rows, float* __restrict__ pooled) {
  pdl_enter();
  constexpr int VEC = VecN<T>::N;
  const int n = blockIdx.x, cv = C / VEC, lanes = 256 / cv;        // host guarantees cv <= 256
  const int myc = (threadIdx.x % cv) * VEC, myr = threadIdx.x / cv;
  __shared__ float part[256 * 8];                  // [lanes][C] partial sums of this CTA (lanes * C = 256 * VEC <= 2048)
  const bool active = myr < lanes;
  const T* f = feat + (size_t)n * rows * C + myc;
  float s[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) s[i] = 0.f;
  if (active) {
    const int step = gridDim.y * lanes;
    int r = blockIdx.y * lanes + myr;
    for (; r + 3 * step < rows; r += 4 * step) {
      float a[4][VEC];
#pragma unroll
      for (int u = 0; u < 4; ++u) ldv(f + (size_t)(r + u * step) * C, a[u]);
#pragma unroll
      for (int i = 0; i < VEC; ++i) s[i] += (a[0][i] + a[1][i]) + (a[2][i] + a[3][i]);
    }
    for (; r < rows; r += step) {
      float a[VEC]; ldv(f + (size_t)r * C, a);
#pragma unroll
      for (int i = 0; i < VEC; ++i) s[i] += a[i];
    }
#pragma unroll
    for (int i = 0; i < VEC; ++i) part[myr * C + myc + i] = s[i];
  }
  __syncthreads();
  // one atomic per channel per CTA (not per thread): the row lanes of the CTA are folded in shared memory first
  for (int c = threadIdx.x; c < C; c += 256) {
    float t = 0.f;
    for (int l = 0; l < lanes; ++l) t += part[l * C + c];
    atomicAdd(pooled + (size_t)n * C + c, t * inv_rows);
  }
}
template <class T>
__global__ void __launch_bounds__(256) pool_bwd_vec_kernel(const float* __restrict__ dpooled, int rows, int C, float inv_rows, T* __restrict__ dfeat) {
  pdl_enter();
  constexpr int VEC = VecN<T>::N;
  const int n = blockIdx.x, cv = C / VEC, lanes = 256 / cv;
  const int myc = (threadIdx.x % cv) * VEC, myr = threadIdx.x / cv;
  if (myr >= lanes) return;
  float v[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) v[i] = dpooled[(size_t)n * C + myc + i] * inv_rows;
  T* d = dfeat + (size_t)n * rows * C + myc;
  for (int r = blockIdx.y * lanes + myr; r < rows; r += gridDim.y * lanes) stv(d + (size_t)r * C, v);
}

__global__ void __launch_bounds__(128) fc_kernel(const float* __restrict__ pooled, const float* __restrict__ w, const float* __restrict__ b,
                                                 int N, int C, int classes, float* __restrict__ logits) {
  pdl_enter();
  int warp = (blockIdx.x * blockDim.x + threadIdx.x) / 32, lane = threadIdx.x % 32;
  if (warp >= N * classes) return;
  int n = warp / classes, j = warp % classes;
  float s = 0.f;
  for (int c = lane; c < C; c += 32) s = fmaf(pooled[(size_t)n * C + c], w[(size_t)j * C + c], s);
  s = warp_sum(s);
  if (lane == 0) logits[warp] = s + b[j];
}

int32_t head_forward(int dtype, int n, int t, int v, int c, int classes, const void* feat, const float* fcw, const float* fcb,
                     float* pooled, float* logits, cudaStream_t st) {
  KernelTimer kt_(KC_HEAD, 0.0, st);
  int rows = t * v;
  cudaMemsetAsync(pooled, 0, sizeof(float) * (size_t)n * c, st);
  dim3 grid((unsigned)n, (unsigned)std::min(rows, 16));
  const int vec = dtype == STGCN_F32 ? 4 : 8;
  if (c % vec == 0 && c / vec <= 256) {
    dim3 gv((unsigned)n, (unsigned)std::max(1, std::min(rows / 16, cdiv(148 * 8, n))));
    if (dtype == STGCN_F32) launch_k(pool_vec_kernel<float>, gv, 256, 0, st, (const float*)feat, rows, c, 1.f / rows, pooled);
    else launch_k(pool_vec_kernel<bf16>, gv, 256, 0, st, (const bf16*)feat, rows, c, 1.f / rows, pooled);
  } else if (dtype == STGCN_F32) launch_k(pool_kernel<float>, grid, 256, 0, st, (const float*)feat, rows, c, 1.f / rows, pooled);
  else launch_k(pool_kernel<bf16>, grid, 256, 0, st, (const bf16*)feat, rows, c, 1.f / rows, pooled);
  STGCN_CHECK_LAUNCH("pool");
  launch_k(fc_kernel, cdiv((long long)n * classes * 32, 128), 128, 0, st, pooled, fcw, fcb, n, c, classes, logits);
  STGCN_CHECK_LAUNCH("fc");
  return STGCN_OK;
}

// dpooled[n][c] = sum_j dlogits[n][j] w[j][c]
__global__ void __launch_bounds__(256) fc_bwd_data_kernel(const float* __restrict__ dlogits, const float* __restrict__ w, int N, int C,
                                                          int classes, float* __restrict__ dpooled) {
  pdl_enter();
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N * C) return;
  int n = i / C, c = i % C;
  float s = 0.f;
  for (int j = 0; j < classes; ++j) s = fmaf(dlogits[(size_t)n * classes + j], w[(size_t)j * C + c], s);
  dpooled[i] = s;
}
// dw[j][c] = sum_n dlogits[n][j] pooled[n][c] ; db[j] = sum_n dlogits[n][j]
__global__ void __launch_bounds__(256) fc_bwd_weight_kernel(const float* __restrict__ dlogits, const float* __restrict__ pooled, int N, int C,
                                                            int classes, float* __restrict__ dw, float* __restrict__ db) {
  pdl_enter();
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < classes * C) {
    int j = i / C, c = i % C;
    float s = 0.f;
    for (int n = 0; n < N; ++n) s = fmaf(dlogits[(size_t)n * classes + j], pooled[(size_t)n * C + c], s);
    dw[i] = s;
  }
  if (i < classes) {
    float s = 0.f;
    for (int n = 0; n < N; ++n) s += dlogits[(size_t)n * classes + i];
    db[i] = s;
  }
}
template <class T>
__global__ void __launch_bounds__(256) pool_bwd_kernel(const float* __restrict__ dpooled, int rows, int C, float inv_rows, size_t total,
                                                       T* __restrict__ dfeat) {
  pdl_enter();
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int c = (int)(i % C);
    size_t n = i / ((size_t)rows * C);
    stf(dfeat + i, dpooled[n * C + c] * inv_rows);
  }
}

// pooled doubles as scratch for dpooled after dw is formed (stream order)
int32_t head_backward(int dtype, int n, int t, int v, int c, int classes, const float* pooled, const float* fcw,
                      const float* dlogits, void* dfeat, float* dfcw, float* dfcb, cudaStream_t st) {
  KernelTimer kt_(KC_HEAD, 0.0, st);
  int rows = t * v;
  launch_k(fc_bwd_weight_kernel, cdiv((long long)classes * c, 256), 256, 0, st, dlogits, pooled, n, c, classes, dfcw, dfcb);
  STGCN_CHECK_LAUNCH("fc_bwd_weight");
  float* dpooled = const_cast<float*>(pooled);
  launch_k(fc_bwd_data_kernel, cdiv((long long)n * c, 256), 256, 0, st, dlogits, fcw, n, c, classes, dpooled);
  STGCN_CHECK_LAUNCH("fc_bwd_data");
  size_t total = (size_t)n * rows * c;
  int blocks = (int)std::min<size_t>((total + 255) / 256, 148 * 16);
  const int vec = dtype == STGCN_F32 ? 4 : 8;
  if (c % vec == 0 && c / vec <= 256) {
    dim3 gv((unsigned)n, (unsigned)std::max(1, std::min(rows / 16, cdiv(148 * 8, n))));
    if (dtype == STGCN_F32) launch_k(pool_bwd_vec_kernel<float>, gv, 256, 0, st, dpooled, rows, c, 1.f / rows, (float*)dfeat);
    else launch_k(pool_bwd_vec_kernel<bf16>, gv, 256, 0, st, dpooled, rows, c, 1.f / rows, (bf16*)dfeat);
  } else if (dtype == STGCN_F32) launch_k(pool_bwd_kernel<float>, blocks, 256, 0, st, dpooled, rows, c, 1.f / rows, total, (float*)dfeat);
  else launch_k(pool_bwd_kernel<bf16>, blocks, 256, 0, st, dpooled, rows, c, 1.f / rows, total, (bf16*)dfeat);
  STGCN_CHECK_LAUNCH("pool_bwd");
  return STGCN_OK;
}

}  // namespace stgcn
