// BatchNorm statistics / apply / backward, ReLU and dropout passes.
// These are the HBM-bound passes of the block chain (reference:
// nn.BatchNorm2d at st_graphconv.py:34,46 applied at :73,:77,:98,:100;
// ReLU :74,:78,:105; Dropout :108).  One read (+ one write) of a
// channels-last [rows, C] tensor each, 16-byte vector accesses, fp32 math,
// grid sized to the 148 SMs; column sums go through per-CTA partials and a
// double-precision finalize so the result is deterministic.
#include <stdlib.h>

#include "kernels.cuh"

namespace stgcn {

constexpr int kStatThreads = 256;
constexpr int kMaxStatBlocks = 148 * 4;

// ---------------------------------------------------------------------------
// generic per-channel two-moment reduction.
// Functor F(row, c, out g, out h) yields the two values whose column sums are
// wanted (forward stats: x, x*x ; backward: g, g*zhat).
// ---------------------------------------------------------------------------
template <int VEC, class F>
__global__ void __launch_bounds__(kStatThreads) colreduce_kernel(long long rows, int C, F f, float* __restrict__ partials) {
  pdl_enter();
  extern __shared__ float sm[];               // [2][rows_per_step][C] folded below
  const int cv = C / VEC;                     // vectors per row
  const int rps = kStatThreads / cv;          // rows per step (>=1, host guarantees cv <= 256)
  const int tid = threadIdx.x;
  const bool active = tid < rps * cv;
  const int myc = (tid % cv) * VEC;
  const int myr = tid / cv;
  float s1[VEC], s2[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) { s1[i] = 0.f; s2[i] = 0.f; }
  if (active) {
    for (long long r = (long long)blockIdx.x * rps + myr; r < rows; r += (long long)gridDim.x * rps) {
      float g[VEC], h[VEC];
      f(r, myc, g, h);
#pragma unroll
      for (int i = 0; i < VEC; ++i) { s1[i] += g[i]; s2[i] += h[i]; }
    }
  }
  // reduce over the rps row-lanes through shared memory
  float* a1 = sm;                 // [rps][C]
  float* a2 = sm + rps * C;
  if (active) {
#pragma unroll
    for (int i = 0; i < VEC; ++i) { a1[myr * C + myc + i] = s1[i]; a2[myr * C + myc + i] = s2[i]; }
  }
  __syncthreads();
  for (int c = tid; c < C; c += kStatThreads) {
    float t1 = 0.f, t2 = 0.f;
    for (int r = 0; r < rps; ++r) { t1 += a1[r * C + c]; t2 += a2[r * C + c]; }
    partials[((size_t)blockIdx.x * 2 + 0) * C + c] = t1;
    partials[((size_t)blockIdx.x * 2 + 1) * C + c] = t2;
  }
}

template <class T, int VEC>
struct FwdStatF {
  const T* x; int C;
  __device__ void operator()(long long r, int c, float (&g)[VEC], float (&h)[VEC]) const {
    if constexpr (VEC == 4) { ld4(x + r * C + c, g); } else { g[0] = ldf(x + r * C + c); }
#pragma unroll
    for (int i = 0; i < VEC; ++i) h[i] = g[i] * g[i];
  }
};

static int stat_blocks(long long rows, int C, int vec) {
  int cv = C / vec, rps = kStatThreads / cv;
  long long b = (rows + rps - 1) / rps;
  // at least ~8 row-steps per block so the partial buffer stays small
  b = (b + 7) / 8;
  if (b < 1) b = 1;
  if (b > kMaxStatBlocks) b = kMaxStatBlocks;
  return (int)b;
}

// ---------------------------------------------------------------------------
// fast path of the two-moment reduction: 16-byte loads, a thread keeps ONE column group for the whole
// kernel (so per-column constants live in registers), four independent rows in flight per thread.
//   MODE 0: g = x, h = x*x                                (forward statistics)
//   MODE 1: g = dy*[a*z+b > 0 if relu], h = g*(z-mean)*invstd   (BatchNorm backward sums; no dropout)
// ---------------------------------------------------------------------------
template <class T, int MODE>
__global__ void __launch_bounds__(kStatThreads, 2) colreduce_fast_kernel(const T* __restrict__ x, const T* __restrict__ z,
                                                                      const float* __restrict__ stats, int relu, long long rows,
                                                                      int C, float* __restrict__ partials, int rev, const __grid_constant__ ZeroJobs zj) {
  pdl_enter();
  for (int j = 0; j < zj.count; ++j) {           // side job: clear the accumulators later kernels of the chain add into
    float* zp = zj.p[j];
    for (unsigned i = blockIdx.x * kStatThreads + threadIdx.x; i < zj.n[j]; i += gridDim.x * kStatThreads) zp[i] = 0.f;
  }
  constexpr int VEC = VecN<T>::N;
  extern __shared__ float sm[];
  const int cv = C / VEC, rps = kStatThreads / cv, tid = threadIdx.x;
  const bool active = tid < rps * cv;
  const int myc = (tid % cv) * VEC, myr = tid / cv;
  float s1[VEC], s2[VEC], mean[VEC], invstd[VEC], ca[VEC], cb[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) { s1[i] = 0.f; s2[i] = 0.f; mean[i] = 0.f; invstd[i] = 0.f; ca[i] = 0.f; cb[i] = 0.f; }
  if (MODE == 1 && active) {
#pragma unroll
    for (int i = 0; i < VEC; ++i) { mean[i] = stats[myc + i]; invstd[i] = stats[C + myc + i]; ca[i] = stats[2 * C + myc + i]; cb[i] = stats[3 * C + myc + i]; }
  }
  auto accum = [&](const float (&g0)[VEC], const float (&z0)[VEC]) {
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
      if (MODE == 0) { s1[i] += g0[i]; s2[i] = fmaf(g0[i], g0[i], s2[i]); }
      else {
        float g = g0[i];
        if (relu && !(fmaf(ca[i], z0[i], cb[i]) > 0.f)) g = 0.f;
        s1[i] += g; s2[i] = fmaf(g, (z0[i] - mean[i]) * invstd[i], s2[i]);
      }
    }
  };
  if (active) {
    const long long step = (long long)gridDim.x * rps;
    auto R = [&](long long q) { return rev ? rows - 1 - q : q; };       // ping-pong traversal (common.cuh)
    long long r = (long long)blockIdx.x * rps + myr;
    for (; r + 3 * step < rows; r += 4 * step) {
      float g[4][VEC], zz[4][VEC];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        ldv(x + R(r + u * step) * C + myc, g[u]);
        if (MODE == 1) ldv(z + R(r + u * step) * C + myc, zz[u]);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) accum(g[u], MODE == 1 ? zz[u] : g[u]);
    }
    for (; r < rows; r += step) {
      float g[VEC], zz[VEC];
      ldv(x + R(r) * C + myc, g);
      if (MODE == 1) ldv(z + R(r) * C + myc, zz);
      accum(g, MODE == 1 ? zz : g);
    }
  }
  float* a1 = sm; float* a2 = sm + rps * C;
  if (active) {
#pragma unroll
    for (int i = 0; i < VEC; ++i) { a1[myr * C + myc + i] = s1[i]; a2[myr * C + myc + i] = s2[i]; }
  }
  __syncthreads();
  for (int c = tid; c < C; c += kStatThreads) {
    float t1 = 0.f, t2 = 0.f;
    for (int rr = 0; rr < rps; ++rr) { t1 += a1[rr * C + c]; t2 += a2[rr * C + c]; }
    partials[((size_t)blockIdx.x * 2 + 0) * C + c] = t1;
    partials[((size_t)blockIdx.x * 2 + 1) * C + c] = t2;
  }
}
template <class T> static bool colreduce_fast_ok(int C) { return C % VecN<T>::N == 0 && C / VecN<T>::N <= kStatThreads; }
template <class T, int MODE>
static void launch_colreduce_fast(const T* x, const T* z, const float* stats, int relu, long long rows, int C, float* partials,
                                  int* nblk, cudaStream_t st, ZeroJobs* zero = nullptr) {
  const int cv = C / VecN<T>::N, rps = kStatThreads / cv;
  long long b = (rows + rps - 1) / rps; b = (b + 7) / 8;
  const int red_cap = 148 * 2;      // MODE 1 holds 128 registers: two CTAs per SM, so 296 CTAs are one wave (measured: -17 us per step vs 592)
  if (b < 1) b = 1; if (b > kMaxStatBlocks) b = kMaxStatBlocks;
  if (MODE == 1 && b > red_cap) b = red_cap;
  *nblk = (int)b;
  launch_k(colreduce_fast_kernel<T, MODE>, (int)b, kStatThreads, 2 * rps * C * sizeof(float), st, x, z, stats, relu, rows, C, partials, next_direction(),
           zero ? *zero : ZeroJobs());
  if (zero) zero->done = true;
}

size_t stat_partials_bytes(int C) { return (size_t)kMaxStatBlocks * 2 * C * sizeof(float); }

template <class T>
int32_t launch_colstats(const T* x, long long rows, int C, float* partials, int* nblk, cudaStream_t st) {
  KernelTimer kt_(KC_BN_STATS, (double)rows * C, st);
  if (C > 1024) { set_error("colstats: C=%d > 1024 unsupported", C); return STGCN_ERR_UNSUPPORTED; }
  if (colreduce_fast_ok<T>(C)) {
    launch_colreduce_fast<T, 0>(x, nullptr, nullptr, 0, rows, C, partials, nblk, st);
  } else if (C % 4 == 0) {
    int nb = stat_blocks(rows, C, 4); *nblk = nb;
    int rps = kStatThreads / (C / 4);
    launch_k(colreduce_kernel<4, FwdStatF<T, 4>>, nb, kStatThreads, 2 * rps * C * sizeof(float), st, rows, C, FwdStatF<T, 4>{x, C}, partials);
  } else {
    if (C > 256) { set_error("colstats: C=%d not a multiple of 4 and > 256", C); return STGCN_ERR_UNSUPPORTED; }
    int nb = stat_blocks(rows, C, 1); *nblk = nb;
    int rps = kStatThreads / C;
    launch_k(colreduce_kernel<1, FwdStatF<T, 1>>, nb, kStatThreads, 2 * rps * C * sizeof(float), st, rows, C, FwdStatF<T, 1>{x, C}, partials);
  }
  STGCN_CHECK_LAUNCH("colstats");
  return STGCN_OK;
}
template int32_t launch_colstats<float>(const float*, long long, int, float*, int*, cudaStream_t);
template int32_t launch_colstats<bf16>(const bf16*, long long, int, float*, int*, cudaStream_t);

// ---------------------------------------------------------------------------
// finalize forward statistics -> stats[4][C] = {mean, invstd, a, b}
//   a = gamma*invstd, b = beta - a*mean ; running stats updated like torch
// ---------------------------------------------------------------------------
// 32 channels per CTA; 32 partial-slices per channel reduced in double through shared memory, so the ~600
// per-CTA partials are read with 32-way memory parallelism instead of one dependent chain per channel.
constexpr int kFinThreads = 1024;
__device__ __forceinline__ bool reduce_partials_32x32(const float* __restrict__ partials, int nblk, int C, double& s1, double& s2) {
  __shared__ double red[2][32][33];
  const int cl = threadIdx.x & 31, bs = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cl;
  double a1 = 0, a2 = 0;
  if (c < C) {
    float f1 = 0.f, f2 = 0.f;        // <= 19 fp32 terms per slice (nblk <= 592); the cross-slice sum below is in double
#pragma unroll 8
    for (int b = bs; b < nblk; b += 32) {
      f1 += partials[((size_t)b * 2 + 0) * C + c];
      f2 += partials[((size_t)b * 2 + 1) * C + c];
    }
    a1 = f1; a2 = f2;
  }
  red[0][bs][cl] = a1; red[1][bs][cl] = a2;
  __syncthreads();
  if (bs != 0 || c >= C) return false;
  s1 = 0; s2 = 0;
#pragma unroll 8
  for (int i = 0; i < 32; ++i) { s1 += red[0][i][cl]; s2 += red[1][i][cl]; }
  return true;
}

__global__ void __launch_bounds__(kFinThreads)
bn_finalize_kernel(const float* __restrict__ partials, int nblk, int C, double count,
                   const float* __restrict__ gamma, const float* __restrict__ beta,
                   float* running_mean, float* running_var, long long* nbt,
                   float eps, float momentum, float* __restrict__ stats) {
  pdl_enter();
  if (blockIdx.x == 0 && threadIdx.x == 0 && nbt) *nbt += 1;
  double s1, s2;
  if (!reduce_partials_32x32(partials, nblk, C, s1, s2)) return;
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  double mean = s1 / count;
  double var = s2 / count - mean * mean;
  if (var < 0) var = 0;
  float invstd = (float)(1.0 / sqrt(var + (double)eps));
  float a = gamma[c] * invstd;
  stats[0 * C + c] = (float)mean;
  stats[1 * C + c] = invstd;
  stats[2 * C + c] = a;
  stats[3 * C + c] = beta[c] - a * (float)mean;
  if (running_mean) {
    double unbiased = count > 1 ? var * count / (count - 1) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)mean;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
  }
}

__global__ void bn_eval_coeffs_kernel(int C, const float* __restrict__ gamma, const float* __restrict__ beta,
                                      const float* __restrict__ rm, const float* __restrict__ rv, float eps,
                                      float* __restrict__ stats, const float* __restrict__ pre_bias) {
  pdl_enter();
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float invstd = rsqrtf(rv[c] + eps);
  float a = gamma[c] * invstd;
  stats[0 * C + c] = rm[c];
  stats[1 * C + c] = invstd;
  stats[2 * C + c] = a;
  // pre_bias: a per-channel bias added BEFORE the normalisation (the producing conv's) folded into the shift
  stats[3 * C + c] = beta[c] - a * rm[c] + (pre_bias ? a * pre_bias[c] : 0.f);
}

int32_t launch_bn_finalize(const float* partials, int nblk, int C, double count, const float* gamma,
                           const float* beta, float* rm, float* rv, long long* nbt, float eps, float momentum,
                           float* stats, cudaStream_t st) {
  launch_k(bn_finalize_kernel, cdiv(C, 32), kFinThreads, 0, st, partials, nblk, C, count, gamma, beta, rm, rv, nbt, eps, momentum, stats);
  STGCN_CHECK_LAUNCH("bn_finalize");
  return STGCN_OK;
}
int32_t launch_bn_eval_coeffs(int C, const float* gamma, const float* beta, const float* rm, const float* rv,
                              float eps, float* stats, cudaStream_t st, const float* pre_bias) {
  launch_k(bn_eval_coeffs_kernel, cdiv(C, 128), 128, 0, st, C, gamma, beta, rm, rv, eps, stats, pre_bias);
  STGCN_CHECK_LAUNCH("bn_eval_coeffs");
  return STGCN_OK;
}

// ---------------------------------------------------------------------------
// y = drop(act(a*x + b))   (a,b may be null -> identity affine)
// ---------------------------------------------------------------------------
template <class T, int VEC>
__global__ void __launch_bounds__(256) affine_act_kernel(const T* __restrict__ x, long long total, int C,
                                                         const float* __restrict__ a, const float* __restrict__ b,
                                                         int relu, float p, float inv_keep, uint64_t seed,
                                                         uint64_t offset, T* __restrict__ y, const uint64_t* __restrict__ rng) {
  pdl_enter();
  if (rng) { seed = rng[0]; offset = rng[1]; }      // device-resident Philox state (replayable CUDA graphs)
  long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * VEC;
  const long long step = (long long)gridDim.x * blockDim.x * VEC;
  for (; i < total; i += step) {
    int c = (int)(i % C);
    float v[VEC];
    if constexpr (VEC == 4) ld4(x + i, v); else v[0] = ldf(x + i);
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      float t = a ? fmaf(a[c + j], v[j], b[c + j]) : v[j];
      if (relu) t = fmaxf(t, 0.f);
      if (p > 0.f) t *= dropout_scale(seed, offset, (uint64_t)(i + j), p, inv_keep);
      v[j] = t;
    }
    if constexpr (VEC == 4) st4(y + i, v); else stf(y + i, v[0]);
  }
}

// fast path: 16-byte accesses; the grid stride is a multiple of C, so a thread keeps one column group and its
// affine coefficients in registers; two independent vectors in flight per thread.  No dropout on this path.
template <class T>
__global__ void __launch_bounds__(256) affine_act_fast_kernel(const T* __restrict__ x, long long total, int C,
                                                              const float* __restrict__ a, const float* __restrict__ b,
                                                              int relu, T* __restrict__ y, int rev) {
  pdl_enter();
  constexpr int VEC = VecN<T>::N;
  long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * VEC;
  const long long step = (long long)gridDim.x * blockDim.x * VEC;
  auto P = [&](long long e) { return rev ? total - VEC - e : e; };       // ping-pong traversal (common.cuh)
  // column group of this thread: total % C == 0, so the mirrored position has column C - VEC - c0 (computed without P(i),
  // which is negative for threads beyond the data -- their slot must stay their own)
  const int c0 = (int)(i % C);
  const int c = rev ? C - VEC - c0 : c0;
  float ca[VEC], cb[VEC];
#pragma unroll
  for (int j = 0; j < VEC; ++j) { ca[j] = a ? a[c + j] : 1.f; cb[j] = a ? b[c + j] : 0.f; }
  auto apply = [&](float (&v)[VEC]) {
#pragma unroll
    for (int j = 0; j < VEC; ++j) { float t = fmaf(ca[j], v[j], cb[j]); v[j] = relu ? fmaxf(t, 0.f) : t; }
  };
  for (; i + step < total; i += 2 * step) {
    float v0[VEC], v1[VEC];
    ldv(x + P(i), v0); ldv(x + P(i + step), v1);
    apply(v0); apply(v1);
    stv(y + P(i), v0); stv(y + P(i + step), v1);
  }
  if (i < total) { float v0[VEC]; ldv(x + P(i), v0); apply(v0); stv(y + P(i), v0); }
}
// Same pass, additionally producing the column sums / sums of squares of the STORED output as per-CTA partials
// [blockIdx][2][C] (the layout bn_finalize reads): the BatchNorm-1 batch statistics of the next block, without its own pass.
template <class T>
__global__ void __launch_bounds__(256) affine_act_stats_kernel(const T* __restrict__ x, long long total, int C,
                                                               const float* __restrict__ a, const float* __restrict__ b,
                                                               int relu, T* __restrict__ y, float* __restrict__ partials, int rev) {
  pdl_enter();
  constexpr int VEC = VecN<T>::N;
  extern __shared__ float sm[];                       // [2][rows per sweep][C]
  long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * VEC;
  const long long step = (long long)gridDim.x * blockDim.x * VEC;
  auto P = [&](long long e) { return rev ? total - VEC - e : e; };       // ping-pong traversal (common.cuh)
  // column group of this thread: total % C == 0, so the mirrored position has column C - VEC - c0 (computed without P(i),
  // which is negative for threads beyond the data -- their slot must stay their own)
  const int c0 = (int)(i % C);
  const int c = rev ? C - VEC - c0 : c0;
  float ca[VEC], cb[VEC], s1[VEC], s2[VEC];
#pragma unroll
  for (int j = 0; j < VEC; ++j) { ca[j] = a ? a[c + j] : 1.f; cb[j] = a ? b[c + j] : 0.f; s1[j] = 0.f; s2[j] = 0.f; }
  auto apply = [&](float (&v)[VEC]) {
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      float t = fmaf(ca[j], v[j], cb[j]);
      t = relu ? fmaxf(t, 0.f) : t;
      t = roundf_as(t, y);                            // statistics of the value as stored
      v[j] = t; s1[j] += t; s2[j] = fmaf(t, t, s2[j]);
    }
  };
  for (; i + step < total; i += 2 * step) {
    float v0[VEC], v1[VEC];
    ldv(x + P(i), v0); ldv(x + P(i + step), v1);
    apply(v0); apply(v1);
    stv(y + P(i), v0); stv(y + P(i + step), v1);
  }
  if (i < total) { float v0[VEC]; ldv(x + P(i), v0); apply(v0); stv(y + P(i), v0); }
  const int rps = (256 * VEC) / C, myr = (threadIdx.x * VEC) / C;
  float* a1 = sm; float* a2 = sm + rps * C;
#pragma unroll
  for (int j = 0; j < VEC; ++j) { a1[myr * C + c + j] = s1[j]; a2[myr * C + c + j] = s2[j]; }
  __syncthreads();
  for (int cc = threadIdx.x; cc < C; cc += 256) {
    float t1 = 0.f, t2 = 0.f;
    for (int rr = 0; rr < rps; ++rr) { t1 += a1[rr * C + cc]; t2 += a2[rr * C + cc]; }
    partials[((size_t)blockIdx.x * 2 + 0) * C + cc] = t1;
    partials[((size_t)blockIdx.x * 2 + 1) * C + cc] = t2;
  }
}
template <class T> static bool elementwise_fast_ok(long long total, int C) {
  return (256 * VecN<T>::N) % C == 0 && C % VecN<T>::N == 0 && total % VecN<T>::N == 0;
}
static int elementwise_blocks(long long total, int vec) {
  return (int)std::max<long long>(1, std::min<long long>((total / vec + 255) / 256, 148 * 8));
}

template <class T>
int32_t launch_affine_act(const T* x, long long rows, int C, const float* a, const float* b, int relu, float p,
                          uint64_t seed, uint64_t offset, T* y, cudaStream_t st, float* out_partials, int* out_nblk,
                          const uint64_t* rng) {
  KernelTimer kt_(KC_ELEMWISE, (double)rows * C, st);
  long long total = rows * C;
  float inv_keep = p > 0.f ? 1.f / (1.f - p) : 1.f;
  if (out_nblk) *out_nblk = 0;
  if (out_partials && out_nblk && p == 0.f && elementwise_fast_ok<T>(total, C)) {
    const int nb = std::min(elementwise_blocks(total, VecN<T>::N), kMaxStatBlocks);
    launch_k(affine_act_stats_kernel<T>, nb, 256, 2 * 256 * VecN<T>::N * sizeof(float), st, x, total, C, a, b, relu, y, out_partials, next_direction());
    *out_nblk = nb;
  } else if (p == 0.f && elementwise_fast_ok<T>(total, C)) {
    launch_k(affine_act_fast_kernel<T>, elementwise_blocks(total, VecN<T>::N), 256, 0, st, x, total, C, a, b, relu, y, next_direction());
  } else if (C % 4 == 0) {
    int blocks = (int)std::min<long long>((total / 4 + 255) / 256, 148 * 16);
    launch_k(affine_act_kernel<T, 4>, blocks, 256, 0, st, x, total, C, a, b, relu, p, inv_keep, seed, offset, y, rng);
  } else {
    int blocks = (int)std::min<long long>((total + 255) / 256, 148 * 16);
    launch_k(affine_act_kernel<T, 1>, blocks, 256, 0, st, x, total, C, a, b, relu, p, inv_keep, seed, offset, y, rng);
  }
  STGCN_CHECK_LAUNCH("affine_act");
  return STGCN_OK;
}
template int32_t launch_affine_act<float>(const float*, long long, int, const float*, const float*, int, float, uint64_t, uint64_t, float*, cudaStream_t, float*, int*, const uint64_t*);
template int32_t launch_affine_act<bf16>(const bf16*, long long, int, const float*, const float*, int, float, uint64_t, uint64_t, bf16*, cudaStream_t, float*, int*, const uint64_t*);

// ---------------------------------------------------------------------------
// backward of  y = drop(relu?(a*z + b))  w.r.t. the BN:
//   g    = dy * drop_scale * [a*z+b > 0]
//   s1   = sum g ; s2 = sum g * zhat,  zhat = (z-mean)*invstd
//   dz   = a * (g - s1/m - zhat*s2/m)      (training)   |   a * g   (eval)
// ---------------------------------------------------------------------------
template <class T, int VEC>
struct BwdRedF {
  const T* dy; const T* z; const float* stats; int C; int relu; float p; float inv_keep; uint64_t seed, offset; const uint64_t* rng;
  __device__ void operator()(long long r, int c, float (&g)[VEC], float (&h)[VEC]) const {
    float zz[VEC];
    const uint64_t seed = rng ? rng[0] : this->seed, offset = rng ? rng[1] : this->offset;
    long long i = r * C + c;
    if constexpr (VEC == 4) { ld4(dy + i, g); ld4(z + i, zz); } else { g[0] = ldf(dy + i); zz[0] = ldf(z + i); }
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      float mean = stats[c + j], invstd = stats[C + c + j], a = stats[2 * C + c + j], b = stats[3 * C + c + j];
      if (relu && !(fmaf(a, zz[j], b) > 0.f)) g[j] = 0.f;
      if (p > 0.f) g[j] *= dropout_scale(seed, offset, (uint64_t)(i + j), p, inv_keep);
      h[j] = g[j] * (zz[j] - mean) * invstd;
    }
  }
};

template <class T>
int32_t launch_bn_bwd_reduce(const T* dy, const T* z, const float* stats, long long rows, int C, int relu, float p,
                             uint64_t seed, uint64_t offset, float* partials, int* nblk, cudaStream_t st, ZeroJobs* zero,
                             const uint64_t* rng) {
  KernelTimer kt_(KC_BN_STATS, (double)rows * C, st);
  float inv_keep = p > 0.f ? 1.f / (1.f - p) : 1.f;
  if (p == 0.f && colreduce_fast_ok<T>(C)) {
    launch_colreduce_fast<T, 1>(dy, z, stats, relu, rows, C, partials, nblk, st, zero);
  } else if (C % 4 == 0) {
    int nb = stat_blocks(rows, C, 4); *nblk = nb;
    int rps = kStatThreads / (C / 4);
    launch_k(colreduce_kernel<4, BwdRedF<T, 4>>, nb, kStatThreads, 2 * rps * C * sizeof(float), st,
        rows, C, BwdRedF<T, 4>{dy, z, stats, C, relu, p, inv_keep, seed, offset, rng}, partials);
  } else {
    int nb = stat_blocks(rows, C, 1); *nblk = nb;
    int rps = kStatThreads / C;
    launch_k(colreduce_kernel<1, BwdRedF<T, 1>>, nb, kStatThreads, 2 * rps * C * sizeof(float), st,
        rows, C, BwdRedF<T, 1>{dy, z, stats, C, relu, p, inv_keep, seed, offset, rng}, partials);
  }
  STGCN_CHECK_LAUNCH("bn_bwd_reduce");
  return STGCN_OK;
}
template int32_t launch_bn_bwd_reduce<float>(const float*, const float*, const float*, long long, int, int, float, uint64_t, uint64_t, float*, int*, cudaStream_t, ZeroJobs*, const uint64_t*);
template int32_t launch_bn_bwd_reduce<bf16>(const bf16*, const bf16*, const float*, long long, int, int, float, uint64_t, uint64_t, float*, int*, cudaStream_t, ZeroJobs*, const uint64_t*);

// sums[2][C] = {s1, s2} reduced in double; also emits dgamma = s2, dbeta = s1
__global__ void __launch_bounds__(kFinThreads)
bn_bwd_finalize_kernel(const float* __restrict__ partials, int nblk, int C, float* __restrict__ sums,
                       float* __restrict__ dgamma, float* __restrict__ dbeta) {
  pdl_enter();
  double s1, s2;
  if (!reduce_partials_32x32(partials, nblk, C, s1, s2)) return;
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  sums[c] = (float)s1; sums[C + c] = (float)s2;
  if (dgamma) dgamma[c] = (float)s2;
  if (dbeta) dbeta[c] = (float)s1;
}
int32_t launch_bn_bwd_finalize(const float* partials, int nblk, int C, float* sums, float* dgamma, float* dbeta, cudaStream_t st) {
  launch_k(bn_bwd_finalize_kernel, cdiv(C, 32), kFinThreads, 0, st, partials, nblk, C, sums, dgamma, dbeta);
  STGCN_CHECK_LAUNCH("bn_bwd_finalize");
  return STGCN_OK;
}

template <class T, int VEC>
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const T* __restrict__ dy, const T* __restrict__ z,
                                                           const float* __restrict__ stats, const float* __restrict__ sums,
                                                           long long total, int C, int relu, int training, float inv_m,
                                                           float p, float inv_keep, uint64_t seed, uint64_t offset,
                                                           T* __restrict__ dz, const uint64_t* __restrict__ rng, const ResAddend<T> ra) {
  pdl_enter();
  if (rng) { seed = rng[0]; offset = rng[1]; }
  long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * VEC;
  const long long step = (long long)gridDim.x * blockDim.x * VEC;
  for (; i < total; i += step) {
    int c = (int)(i % C);
    float g[VEC], zz[VEC];
    if constexpr (VEC == 4) { ld4(dy + i, g); ld4(z + i, zz); } else { g[0] = ldf(dy + i); zz[0] = ldf(z + i); }
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      float mean = stats[c + j], invstd = stats[C + c + j], a = stats[2 * C + c + j], b = stats[3 * C + c + j];
      if (relu && !(fmaf(a, zz[j], b) > 0.f)) g[j] = 0.f;
      if (p > 0.f) g[j] *= dropout_scale(seed, offset, (uint64_t)(i + j), p, inv_keep);
      float r = g[j];
      if (training) r = r - sums[c + j] * inv_m - (zz[j] - mean) * invstd * sums[C + c + j] * inv_m;
      g[j] = a * r;
    }
    if (ra.ptr) {
      const long long src = ra.source(i, C);
      if (src >= 0) {
        float ad[VEC];
        if constexpr (VEC == 4) ld4(ra.ptr + src, ad); else ad[0] = ldf(ra.ptr + src);
#pragma unroll
        for (int j = 0; j < VEC; ++j) g[j] += ad[j];
      }
    }
    if constexpr (VEC == 4) st4(dz + i, g); else stf(dz + i, g[0]);
  }
}

template <class T>
__global__ void __launch_bounds__(256, 3) bn_bwd_apply_fast_kernel(const T* __restrict__ dy, const T* __restrict__ z,
                                                                const float* __restrict__ stats, const float* __restrict__ sums,
                                                                long long total, int C, int relu, int training, float inv_m,
                                                                T* __restrict__ dz, int rev, const ResAddend<T> ra) {
  pdl_enter();
  constexpr int VEC = VecN<T>::N;
  long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * VEC;
  const long long step = (long long)gridDim.x * blockDim.x * VEC;
  auto P = [&](long long e) { return rev ? total - VEC - e : e; };       // ping-pong traversal (common.cuh)
  // column group of this thread: total % C == 0, so the mirrored position has column C - VEC - c0 (computed without P(i),
  // which is negative for threads beyond the data -- their slot must stay their own)
  const int c0 = (int)(i % C);
  const int c = rev ? C - VEC - c0 : c0;
  float mean[VEC], ca[VEC], cb[VEC], k1[VEC], k2[VEC];
#pragma unroll
  for (int j = 0; j < VEC; ++j) {
    mean[j] = stats[c + j]; ca[j] = stats[2 * C + c + j]; cb[j] = stats[3 * C + c + j];
    k1[j] = training ? sums[c + j] * inv_m : 0.f;
    k2[j] = training ? stats[C + c + j] * sums[C + c + j] * inv_m : 0.f;
  }
  auto apply = [&](float (&g)[VEC], const float (&zz)[VEC]) {
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      float r = g[j];
      if (relu && !(fmaf(ca[j], zz[j], cb[j]) > 0.f)) r = 0.f;
      r = r - k1[j] - (zz[j] - mean[j]) * k2[j];
      g[j] = ca[j] * r;
    }
  };
  auto addend = [&](float (&g)[VEC], long long e) {          // skip connection's gradient rides in this pass
    if (!ra.ptr) return;
    const long long src = ra.source(e, C);
    if (src < 0) return;
    float ad[VEC];
    ldv(ra.ptr + src, ad);
#pragma unroll
    for (int j = 0; j < VEC; ++j) g[j] += ad[j];
  };
  for (; i + step < total; i += 2 * step) {
    float g0[VEC], z0[VEC], g1[VEC], z1[VEC];
    ldv(dy + P(i), g0); ldv(z + P(i), z0); ldv(dy + P(i + step), g1); ldv(z + P(i + step), z1);
    apply(g0, z0); apply(g1, z1);
    addend(g0, P(i)); addend(g1, P(i + step));
    stv(dz + P(i), g0); stv(dz + P(i + step), g1);
  }
  if (i < total) { float g0[VEC], z0[VEC]; ldv(dy + P(i), g0); ldv(z + P(i), z0); apply(g0, z0); addend(g0, P(i)); stv(dz + P(i), g0); }
}

template <class T>
int32_t launch_bn_bwd_apply(const T* dy, const T* z, const float* stats, const float* sums, long long rows, int C,
                            int relu, int training, float p, uint64_t seed, uint64_t offset, T* dz, cudaStream_t st,
                            const uint64_t* rng, const ResAddend<T>* addend, double count_scale) {
  const ResAddend<T> ra = addend ? *addend : ResAddend<T>();
  KernelTimer kt_(KC_ELEMWISE, (double)rows * C, st);
  long long total = rows * C;
  float inv_keep = p > 0.f ? 1.f / (1.f - p) : 1.f;
  float inv_m = (float)(1.0 / ((double)rows * count_scale));      // count_scale = ranks when `sums` are global (SyncBN)
  if (p == 0.f && elementwise_fast_ok<T>(total, C)) {
    // one wave: the kernel keeps 3 CTAs per SM (launch bounds), and every extra wave repeats the per-column constant loads
    // (measured on B200: 444 CTAs 0.666 ms per step for the ten launches, 888: 0.681, 1184: 0.702)
    const int apply_cap = 148 * 3;
    launch_k(bn_bwd_apply_fast_kernel<T>, std::min(elementwise_blocks(total, VecN<T>::N), apply_cap), 256, 0, st, dy, z, stats, sums, total, C, relu, training,
                                                                                         inv_m, dz, next_direction(), ra);
  } else if (C % 4 == 0) {
    int blocks = (int)std::min<long long>((total / 4 + 255) / 256, 148 * 16);
    launch_k(bn_bwd_apply_kernel<T, 4>, blocks, 256, 0, st, dy, z, stats, sums, total, C, relu, training, inv_m, p, inv_keep, seed, offset, dz, rng, ra);
  } else {
    int blocks = (int)std::min<long long>((total + 255) / 256, 148 * 16);
    launch_k(bn_bwd_apply_kernel<T, 1>, blocks, 256, 0, st, dy, z, stats, sums, total, C, relu, training, inv_m, p, inv_keep, seed, offset, dz, rng, ra);
  }
  STGCN_CHECK_LAUNCH("bn_bwd_apply");
  return STGCN_OK;
}
template int32_t launch_bn_bwd_apply<float>(const float*, const float*, const float*, const float*, long long, int, int, int, float, uint64_t, uint64_t, float*, cudaStream_t, const uint64_t*, const ResAddend<float>*, double);
template int32_t launch_bn_bwd_apply<bf16>(const bf16*, const bf16*, const float*, const float*, long long, int, int, int, float, uint64_t, uint64_t, bf16*, cudaStream_t, const uint64_t*, const ResAddend<bf16>*, double);

// ---------------------------------------------------------------------------
// dz = dy * [out > 0] * (1/(1-p) if dropout)    -- backward of the trailing
// ReLU (+dropout) of the residual path (st_graphconv.py:105-108): out > 0
// implies the element was kept.
// ---------------------------------------------------------------------------
template <class T, int VEC>
__global__ void __launch_bounds__(256) relu_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ out, long long total,
                                                       float inv_keep, T* __restrict__ dz) {
  pdl_enter();
  long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * VEC;
  const long long step = (long long)gridDim.x * blockDim.x * VEC;
  for (; i < total; i += step) {
    float g[VEC], o[VEC];
    if constexpr (VEC == 4) { ld4(dy + i, g); ld4(out + i, o); } else { g[0] = ldf(dy + i); o[0] = ldf(out + i); }
#pragma unroll
    for (int j = 0; j < VEC; ++j) g[j] = o[j] > 0.f ? g[j] * inv_keep : 0.f;
    if constexpr (VEC == 4) st4(dz + i, g); else stf(dz + i, g[0]);
  }
}
template <class T>
int32_t launch_relu_bwd(const T* dy, const T* out, long long total, float p, T* dz, cudaStream_t st) {
  KernelTimer kt_(KC_ELEMWISE, (double)total, st);
  float inv_keep = p > 0.f ? 1.f / (1.f - p) : 1.f;
  if (total % 4 == 0) {
    int blocks = (int)std::min<long long>((total / 4 + 255) / 256, 148 * 16);
    launch_k(relu_bwd_kernel<T, 4>, blocks, 256, 0, st, dy, out, total, inv_keep, dz);
  } else {
    int blocks = (int)std::min<long long>((total + 255) / 256, 148 * 16);
    launch_k(relu_bwd_kernel<T, 1>, blocks, 256, 0, st, dy, out, total, inv_keep, dz);
  }
  STGCN_CHECK_LAUNCH("relu_bwd");
  return STGCN_OK;
}
template int32_t launch_relu_bwd<float>(const float*, const float*, long long, float, float*, cudaStream_t);
template int32_t launch_relu_bwd<bf16>(const bf16*, const bf16*, long long, float, bf16*, cudaStream_t);

// dst = a + b (residual identity gradient merge), and cast
template <class TS, class TD>
__global__ void __launch_bounds__(256) cast_kernel(const TS* __restrict__ s, TD* __restrict__ d, size_t n) {
  pdl_enter();
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i < n; i += (size_t)gridDim.x * blockDim.x) stf(d + i, ldf(s + i));
}
template <class TS, class TD>
int32_t launch_cast(const TS* s, TD* d, size_t n, cudaStream_t st) {
  int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 16);
  if (blocks < 1) blocks = 1;
  launch_k(cast_kernel<TS, TD>, blocks, 256, 0, st, s, d, n);
  STGCN_CHECK_LAUNCH("cast");
  return STGCN_OK;
}
template int32_t launch_cast<float, bf16>(const float*, bf16*, size_t, cudaStream_t);
template int32_t launch_cast<bf16, float>(const bf16*, float*, size_t, cudaStream_t);
template int32_t launch_cast<float, float>(const float*, float*, size_t, cudaStream_t);
template int32_t launch_cast<bf16, bf16>(const bf16*, bf16*, size_t, cudaStream_t);

template <class T>
__global__ void __launch_bounds__(256) add_inplace_kernel(T* __restrict__ d, const T* __restrict__ s, size_t n) {
  pdl_enter();
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i < n; i += (size_t)gridDim.x * blockDim.x) stf(d + i, ldf(d + i) + ldf(s + i));
}
template <class T>
int32_t launch_add_inplace(T* d, const T* s, size_t n, cudaStream_t st) {
  KernelTimer kt_(KC_ELEMWISE, (double)n, st);
  int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 16);
  if (blocks < 1) blocks = 1;
  launch_k(add_inplace_kernel<T>, blocks, 256, 0, st, d, s, n);
  STGCN_CHECK_LAUNCH("add_inplace");
  return STGCN_OK;
}
template int32_t launch_add_inplace<float>(float*, const float*, size_t, cudaStream_t);
template int32_t launch_add_inplace<bf16>(bf16*, const bf16*, size_t, cudaStream_t);

// ---------------------------------------------------------------------------
// SyncBN support (global-batch BatchNorm statistics under data parallel; no reference counterpart): the sums a
// BatchNorm needs are brought into ONE [2][C] row that the caller's callback all-reduces over the ranks.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kFinThreads) reduce_partials_kernel(const float* __restrict__ partials, int nblk, int C, float* __restrict__ out) {
  pdl_enter();
  double s1, s2;
  if (!reduce_partials_32x32(partials, nblk, C, s1, s2)) return;
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  out[c] = (float)s1; out[C + c] = (float)s2;
}
int32_t launch_reduce_partials(const float* partials, int nblk, int C, float* out, cudaStream_t st) {
  launch_k(reduce_partials_kernel, cdiv(C, 32), kFinThreads, 0, st, partials, nblk, C, out);
  STGCN_CHECK_LAUNCH("reduce_partials");
  return STGCN_OK;
}
// out = [a | b]  (two C-vectors side by side)
__global__ void pack2_kernel(const float* __restrict__ a, const float* __restrict__ b, int C, float* __restrict__ out) {
  pdl_enter();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < C) { out[c] = a[c]; out[C + c] = b[c]; }
}
int32_t launch_pack2(const float* a, const float* b, int C, float* out, cudaStream_t st) {
  launch_k(pack2_kernel, cdiv(C, 128), 128, 0, st, a, b, C, out);
  STGCN_CHECK_LAUNCH("pack2");
  return STGCN_OK;
}
// folded BatchNorm-1 backward (agg_kernels.cu, bn1_fold_grads): the dx-correction coefficients from GLOBAL sums
//   g = [sum dbeta | sum dgamma] over all ranks, inv_m = 1 / global row count:  coef = { mean, a*db/m, a*invstd*dg/m }
__global__ void bn1_coef_global_kernel(const float* __restrict__ stats, const float* __restrict__ g, int C, float inv_m, float* __restrict__ coef) {
  pdl_enter();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float a = stats[2 * C + c], invstd = stats[C + c];
  coef[c] = stats[c];
  coef[C + c] = a * g[c] * inv_m;
  coef[2 * C + c] = a * invstd * g[C + c] * inv_m;
}
int32_t launch_bn1_coef_global(const float* stats, const float* g, int C, double rows_global, float* coef, cudaStream_t st) {
  launch_k(bn1_coef_global_kernel, cdiv(C, 128), 128, 0, st, stats, g, C, (float)(1.0 / rows_global), coef);
  STGCN_CHECK_LAUNCH("bn1_coef_global");
  return STGCN_OK;
}

// ---------------------------------------------------------------------------
// fp32 -> split bf16: x = hi + mid + lo with hi = bf16(x), mid = bf16(x - hi), lo = bf16(x - hi - mid) (24 significand bits
// together; the subtractions are exact in fp32); rows of [hi | mid | lo]
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) split_bf16_kernel(const float* __restrict__ src, long long rows, int C, bf16* __restrict__ dst) {
  pdl_enter();
  const int c4 = C / 4;
  const long long total = rows * c4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / c4; const int c = (int)(i - r * c4) * 4;
    float v[4], hi[4], mid[4], lo[4];
    ld4(src + r * C + c, v);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      hi[j] = __bfloat162float(__float2bfloat16_rn(v[j]));
      const float r1 = v[j] - hi[j];
      mid[j] = __bfloat162float(__float2bfloat16_rn(r1));
      lo[j] = r1 - mid[j];                 // rounded to bf16 by the store
    }
    bf16* d = dst + r * 3 * C + c;
    st4(d, hi); st4(d + C, mid); st4(d + 2 * C, lo);
  }
}
int32_t launch_split_bf16(const float* src, long long rows, int C, bf16* dst, cudaStream_t st) {
  KernelTimer kt_(KC_ELEMWISE, (double)rows * C, st);
  if (C % 4 != 0) { set_error("split_bf16: C %% 4 != 0"); return STGCN_ERR_UNSUPPORTED; }
  int blocks = (int)std::max<long long>(1, std::min<long long>((rows * (C / 4) + 255) / 256, 148 * 16));
  launch_k(split_bf16_kernel, blocks, 256, 0, st, src, rows, C, dst);
  STGCN_CHECK_LAUNCH("split_bf16");
  return STGCN_OK;
}

// ---------------------------------------------------------------------------
// dropout random state on the device: {seed, offset}.  Forward copies the state it will use next to `slot` (kept with the
// tensors saved for backward, so backward replays the same mask) and advances the caller's state by the number of Philox
// calls the block consumes; a CUDA graph of the step therefore draws a fresh mask at every replay.
// ---------------------------------------------------------------------------
__global__ void rng_fetch_kernel(uint64_t* state, uint64_t seed, uint64_t offset, uint64_t advance, uint64_t* slot) {
  pdl_enter();
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    if (state) { slot[0] = state[0]; slot[1] = state[1]; state[1] += advance; }
    else { slot[0] = seed; slot[1] = offset; }
  }
}
int32_t launch_rng_fetch(uint64_t* state, uint64_t seed, uint64_t offset, uint64_t advance, uint64_t* slot, cudaStream_t st) {
  launch_k(rng_fetch_kernel, 1, 32, 0, st, state, seed, offset, advance, slot);
  STGCN_CHECK_LAUNCH("rng_fetch");
  return STGCN_OK;
}
__global__ void __launch_bounds__(256) dropout_mask_kernel(uint64_t seed, uint64_t offset, float p, float inv_keep, size_t n,
                                                           float* __restrict__ out) {
  pdl_enter();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    out[i] = dropout_scale(seed, offset, (uint64_t)i, p, inv_keep);
}
int32_t launch_dropout_mask(uint64_t seed, uint64_t offset, float p, size_t n, float* out, cudaStream_t st) {
  const float inv_keep = p > 0.f ? 1.f / (1.f - p) : 1.f;
  int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 16);
  if (blocks < 1) blocks = 1;
  launch_k(dropout_mask_kernel, blocks, 256, 0, st, seed, offset, p, inv_keep, n, out);
  STGCN_CHECK_LAUNCH("dropout_mask");
  return STGCN_OK;
}

}  // namespace stgcn
