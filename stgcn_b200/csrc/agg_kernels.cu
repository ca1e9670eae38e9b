// K-partition adjacency aggregation on CUDA cores (generic path: any V<=32,
// K<=4, any C).  Reference: the einsum 'kvw,nkctw->nctv' at
// st_graphconv.py:150, applied on the C_in side of the 1x1 convolution
// (aggregation and channel mixing commute):
//     XA[f, v, k*C + c] = sum_w A[k,v,w] * h[f, w, c],   h = act(a*x + b)
// so that S = XA . W2 is a single GEMM with reduction K*C_in.  The adjacency
// (K,V,V) lives in shared memory, each frame tile is staged in shared memory
// once, outputs are written channel-contiguous.
#include "kernels.cuh"

namespace stgcn {

constexpr int kAggThreads = 256;
constexpr int kVP = 32;   // padded joint count in shared memory

// ---------------------------------------------------------------------------
// forward aggregation
// ---------------------------------------------------------------------------
template <class T>
__global__ void __launch_bounds__(kAggThreads) agg_fwd_kernel(const T* __restrict__ x, long long frames, int V, int C, int K,
                                                              const float* __restrict__ A, const float* __restrict__ aff_a,
                                                              const float* __restrict__ aff_b, int relu, int FB,
                                                              T* __restrict__ xa) {
  pdl_enter();
  extern __shared__ float sm[];
  float* As = sm;                         // [K][V][kVP]
  float* hs = sm + K * V * kVP;           // [FB][V][C]
  for (int i = threadIdx.x; i < K * V * kVP; i += blockDim.x) {
    int w = i % kVP, kv = i / kVP;
    As[i] = w < V ? A[kv * V + w] : 0.f;
  }
  const int KC = K * C;
  for (long long f0 = (long long)blockIdx.x * FB; f0 < frames; f0 += (long long)gridDim.x * FB) {
    int nf = (int)min((long long)FB, frames - f0);
    __syncthreads();
    const T* xp = x + f0 * V * C;
    for (int i = threadIdx.x; i < nf * V * C; i += blockDim.x) {
      float t = ldf(xp + i);
      if (aff_a) { int c = i % C; t = fmaf(aff_a[c], t, aff_b[c]); }
      if (relu) t = fmaxf(t, 0.f);
      hs[i] = t;
    }
    __syncthreads();
    // work item = (frame_local, k, c): 25 staged inputs in registers, V outputs
    for (int it = threadIdx.x; it < nf * KC; it += blockDim.x) {
      int c = it % C, k = (it / C) % K, fl = it / KC;
      float h[kVP];
#pragma unroll
      for (int w = 0; w < kVP; ++w) h[w] = w < V ? hs[(fl * V + w) * C + c] : 0.f;
      T* op = xa + ((f0 + fl) * V) * KC + k * C + c;
      const float* Ak = As + k * V * kVP;
      for (int v = 0; v < V; ++v) {
        const float4* ar = reinterpret_cast<const float4*>(Ak + v * kVP);
        float acc = 0.f;
#pragma unroll
        for (int q = 0; q < kVP / 4; ++q) {
          float4 a4 = ar[q];
          acc = fmaf(a4.x, h[4 * q], acc); acc = fmaf(a4.y, h[4 * q + 1], acc);
          acc = fmaf(a4.z, h[4 * q + 2], acc); acc = fmaf(a4.w, h[4 * q + 3], acc);
        }
        stf(op + (size_t)v * KC, acc);
      }
    }
  }
}

static int agg_fb(int V, int C, int K) {
  // frames staged per loop iteration: enough for ~8 work items (frame, k, c) per thread between two block-wide barriers (one frame
  // per iteration, as before round 2, left the kernel latency-bound: 2.3 ms for a 0.15 ms FMA / memory bound at C = 64)
  int fb = (8 * kAggThreads) / (K * C);
  if (fb < 1) fb = 1;
  // keep the staged frames under ~64 KB
  while (fb > 1 && (size_t)fb * V * C * 4 > 64 * 1024) --fb;
  return fb;
}

template <class T>
int32_t launch_agg_fwd(const T* x, long long frames, int V, int C, int K, const float* A, const float* a, const float* b,
                       int relu, T* xa, cudaStream_t st) {
  KernelTimer kt_(KC_AGG_FWD, 2.0 * frames * V * V * (double)K * C, st);
  if (V > kVP || K > 4) { set_error("agg: V=%d K=%d unsupported (V<=32, K<=4)", V, K); return STGCN_ERR_UNSUPPORTED; }
  int FB = agg_fb(V, C, K);
  size_t smem = ((size_t)K * V * kVP + (size_t)FB * V * C) * sizeof(float);
  if (smem > 200 * 1024) { set_error("agg: C=%d too large for shared staging", C); return STGCN_ERR_UNSUPPORTED; }
  static bool attr_f = false, attr_b = false;
  bool& attr = std::is_same<T, float>::value ? attr_f : attr_b;
  if (!attr) { cudaFuncSetAttribute(agg_fwd_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); attr = true; }
  int blocks = (int)std::min<long long>((frames + FB - 1) / FB, 148 * 4);
  launch_k(agg_fwd_kernel<T>, blocks, kAggThreads, smem, st, x, frames, V, C, K, A, a, b, relu, FB, xa);
  STGCN_CHECK_LAUNCH("agg_fwd");
  return STGCN_OK;
}
template int32_t launch_agg_fwd<float>(const float*, long long, int, int, int, const float*, const float*, const float*, int, float*, cudaStream_t);
template int32_t launch_agg_fwd<bf16>(const bf16*, long long, int, int, int, const float*, const float*, const float*, int, bf16*, cudaStream_t);

// ---------------------------------------------------------------------------
// transposed aggregation:  dh[f,w,c] = sum_k sum_v A[k,v,w] * dxa[f,v,k*C+c]
// optionally masked by the ReLU of h = a*x+b (residual path, st_graphconv.py:74)
// ---------------------------------------------------------------------------
template <class T>
__global__ void __launch_bounds__(kAggThreads) agg_bwd_kernel(const T* __restrict__ dxa, long long frames, int V, int C, int K,
                                                              const float* __restrict__ A, const T* __restrict__ x,
                                                              const float* __restrict__ aff_a, const float* __restrict__ aff_b,
                                                              int relu, T* __restrict__ dh) {
  pdl_enter();
  extern __shared__ float sm[];
  float* At = sm;                         // [K][W=V][kVP over v]  (transposed)
  for (int i = threadIdx.x; i < K * V * kVP; i += blockDim.x) {
    int v = i % kVP, kw = i / kVP, w = kw % V, k = kw / V;
    At[i] = v < V ? A[(k * V + v) * V + w] : 0.f;
  }
  __syncthreads();
  const int KC = K * C;
  const long long items = frames * C;
  for (long long it = (long long)blockIdx.x * blockDim.x + threadIdx.x; it < items; it += (long long)gridDim.x * blockDim.x) {
    int c = (int)(it % C);
    long long f = it / C;
    float acc[kVP];
#pragma unroll
    for (int w = 0; w < kVP; ++w) acc[w] = 0.f;
    for (int k = 0; k < K; ++k) {
      float g[kVP];
#pragma unroll
      for (int v = 0; v < kVP; ++v) g[v] = v < V ? ldf(dxa + (f * V + v) * KC + k * C + c) : 0.f;
      const float* Ak = At + k * V * kVP;
#pragma unroll
      for (int w = 0; w < kVP; ++w) {
        if (w < V) {
          const float4* ar = reinterpret_cast<const float4*>(Ak + w * kVP);
          float s = acc[w];
#pragma unroll
          for (int q = 0; q < kVP / 4; ++q) {
            float4 a4 = ar[q];
            s = fmaf(a4.x, g[4 * q], s); s = fmaf(a4.y, g[4 * q + 1], s);
            s = fmaf(a4.z, g[4 * q + 2], s); s = fmaf(a4.w, g[4 * q + 3], s);
          }
          acc[w] = s;
        }
      }
    }
#pragma unroll
    for (int w = 0; w < kVP; ++w) {
      if (w < V) {
        float r = acc[w];
        if (relu) {
          float xv = ldf(x + (f * V + w) * C + c);
          if (!(fmaf(aff_a[c], xv, aff_b[c]) > 0.f)) r = 0.f;
        }
        stf(dh + (f * V + w) * C + c, r);
      }
    }
  }
}

template <class T>
int32_t launch_agg_bwd(const T* dxa, long long frames, int V, int C, int K, const float* A, const T* x, const float* a,
                       const float* b, int relu, T* dh, cudaStream_t st) {
  KernelTimer kt_(KC_AGG_BWD, 2.0 * frames * V * V * (double)K * C, st);
  if (V > kVP || K > 4) { set_error("agg_bwd: V=%d K=%d unsupported", V, K); return STGCN_ERR_UNSUPPORTED; }
  size_t smem = (size_t)K * V * kVP * sizeof(float);
  int blocks = (int)std::min<long long>((frames * C + kAggThreads - 1) / kAggThreads, 148 * 8);
  launch_k(agg_bwd_kernel<T>, blocks, kAggThreads, smem, st, dxa, frames, V, C, K, A, x, a, b, relu, dh);
  STGCN_CHECK_LAUNCH("agg_bwd");
  return STGCN_OK;
}
template int32_t launch_agg_bwd<float>(const float*, long long, int, int, int, const float*, const float*, const float*, const float*, int, float*, cudaStream_t);
template int32_t launch_agg_bwd<bf16>(const bf16*, long long, int, int, int, const float*, const bf16*, const float*, const float*, int, bf16*, cudaStream_t);

// ---------------------------------------------------------------------------
// adjacency gradient (the "dA-mask" of the north star; in this reference the
// trained leaf is spatialConv.A, st_graphconv.py:129):
//     dA[k,v,w] += sum_f sum_c dxa[f,v,k*C+c] * h[f,w,c]
// One CTA walks a strip of frames, keeps the K*V*V accumulators in registers
// (8 per thread) and flushes them with one atomicAdd each at the end.
// ---------------------------------------------------------------------------
constexpr int kDaPer = 8;   // accumulators per thread: 256*8 = 2048 >= 4*... see host check

template <class T>
__global__ void __launch_bounds__(kAggThreads) agg_dA_kernel(const T* __restrict__ dxa, const T* __restrict__ x, long long frames,
                                                             int V, int C, int K, const float* __restrict__ aff_a,
                                                             const float* __restrict__ aff_b, int relu, int CC, int KC,
                                                             float* __restrict__ dA) {
  pdl_enter();
  extern __shared__ float sm[];
  const int KV = K * V;
  float* hs = sm;                          // [V][CC+1]
  float* gs = sm + V * (CC + 1);           // [K*V][CC+1]
  const int nout = K * V * V;
  float acc[kDaPer];
  int okv[kDaPer], ow[kDaPer];
#pragma unroll
  for (int j = 0; j < kDaPer; ++j) {
    acc[j] = 0.f;
    int o = threadIdx.x + j * kAggThreads;
    okv[j] = o < nout ? o / V : -1;
    ow[j] = o < nout ? o % V : 0;
  }
  for (long long f = blockIdx.x; f < frames; f += gridDim.x) {
    for (int c0 = 0; c0 < C; c0 += CC) {
      int cc = min(CC, C - c0);
      __syncthreads();
      for (int i = threadIdx.x; i < V * cc; i += blockDim.x) {
        int c = i % cc, w = i / cc;
        float t = ldf(x + (f * V + w) * C + c0 + c);
        if (aff_a) t = fmaf(aff_a[c0 + c], t, aff_b[c0 + c]);
        if (relu) t = fmaxf(t, 0.f);
        hs[w * (CC + 1) + c] = t;
      }
      for (int i = threadIdx.x; i < KV * cc; i += blockDim.x) {
        int c = i % cc, kv = i / cc, k = kv / V, v = kv % V;
        gs[kv * (CC + 1) + c] = ldf(dxa + (f * V + v) * KC + k * C + c0 + c);
      }
      __syncthreads();
#pragma unroll
      for (int j = 0; j < kDaPer; ++j) {
        if (okv[j] >= 0) {
          const float* gp = gs + okv[j] * (CC + 1);
          const float* hp = hs + ow[j] * (CC + 1);
          float s = acc[j];
          for (int c = 0; c < cc; ++c) s = fmaf(gp[c], hp[c], s);
          acc[j] = s;
        }
      }
    }
  }
#pragma unroll
  for (int j = 0; j < kDaPer; ++j)
    if (okv[j] >= 0) atomicAdd(dA + okv[j] * V + ow[j], acc[j]);
}

// Small-C_in form (first block, C_in <= 4, no affine): per frame only K*V*C + V*C values matter, so a batch of 8 frames is staged
// per barrier pair as zero-padded float4 rows and every thread runs 4-wide dot products for its <= 8 outputs.
constexpr int kDaFrames = 8;
template <class T>
__global__ void __launch_bounds__(kAggThreads) agg_dA_small_kernel(const T* __restrict__ dxa, const T* __restrict__ x, long long frames,
                                                                   int V, int C, int K, int KC, float* __restrict__ dA) {
  pdl_enter();
  extern __shared__ float4 sm4[];
  const int KV = K * V;
  float4* hs = sm4;                          // [frames][V]
  float4* gs = sm4 + kDaFrames * V;          // [frames][K*V]
  const int nout = K * V * V;
  float acc[kDaPer];
  int okv[kDaPer], ow[kDaPer];
#pragma unroll
  for (int j = 0; j < kDaPer; ++j) {
    acc[j] = 0.f;
    int o = threadIdx.x + j * kAggThreads;
    okv[j] = o < nout ? o / V : -1;
    ow[j] = o < nout ? o % V : 0;
  }
  for (long long f0 = (long long)blockIdx.x * kDaFrames; f0 < frames; f0 += (long long)gridDim.x * kDaFrames) {
    const int nf = (int)min((long long)kDaFrames, frames - f0);
    __syncthreads();
    for (int i = threadIdx.x; i < nf * V; i += blockDim.x) {
      const T* xp = x + (f0 * V + i) * C;
      float t[4] = {0.f, 0.f, 0.f, 0.f};
      for (int c = 0; c < C; ++c) t[c] = ldf(xp + c);
      hs[i] = make_float4(t[0], t[1], t[2], t[3]);
    }
    for (int i = threadIdx.x; i < nf * KV; i += blockDim.x) {
      const int fb = i / KV, kv = i - fb * KV, k = kv / V, v = kv - k * V;
      const T* gp = dxa + ((f0 + fb) * V + v) * KC + k * C;
      float t[4] = {0.f, 0.f, 0.f, 0.f};
      for (int c = 0; c < C; ++c) t[c] = ldf(gp + c);
      gs[i] = make_float4(t[0], t[1], t[2], t[3]);
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < kDaPer; ++j) {
      if (okv[j] >= 0) {
        float s = acc[j];
        for (int fb = 0; fb < nf; ++fb) {
          const float4 g = gs[fb * KV + okv[j]], h = hs[fb * V + ow[j]];
          s = fmaf(g.x, h.x, fmaf(g.y, h.y, fmaf(g.z, h.z, fmaf(g.w, h.w, s))));
        }
        acc[j] = s;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < kDaPer; ++j)
    if (okv[j] >= 0) atomicAdd(dA + okv[j] * V + ow[j], acc[j]);
}

template <class T>
int32_t launch_agg_dA(const T* dxa, const T* x, long long frames, int V, int C, int K, const float* a, const float* b,
                      int relu, float* dA, cudaStream_t st, int ld) {
  KernelTimer kt_(KC_AGG_DA, 2.0 * frames * V * V * (double)K * C, st);
  if (K * V * V > kDaPer * kAggThreads) { set_error("agg_dA: K*V*V=%d too large", K * V * V); return STGCN_ERR_UNSUPPORTED; }
  if (C <= 4 && !a && !relu) {
    const long long batches = (frames + kDaFrames - 1) / kDaFrames;
    const int nblk = (int)std::min<long long>(batches, 148 * 4);
    launch_k(agg_dA_small_kernel<T>, nblk, kAggThreads, (size_t)kDaFrames * (V + K * V) * sizeof(float4), st, dxa, x, frames, V, C, K,
                                                                                                       ld > 0 ? ld : K * C, dA);
    STGCN_CHECK_LAUNCH("agg_dA_small");
    return STGCN_OK;
  }
  int CC = std::min(C, 64);
  size_t smem = (size_t)(V + K * V) * (CC + 1) * sizeof(float);
  int blocks = (int)std::min<long long>(frames, 148 * 4);
  launch_k(agg_dA_kernel<T>, blocks, kAggThreads, smem, st, dxa, x, frames, V, C, K, a, b, relu, CC, ld > 0 ? ld : K * C, dA);
  STGCN_CHECK_LAUNCH("agg_dA");
  return STGCN_OK;
}
template int32_t launch_agg_dA<float>(const float*, const float*, long long, int, int, int, const float*, const float*, int, float*, cudaStream_t, int);
template int32_t launch_agg_dA<bf16>(const bf16*, const bf16*, long long, int, int, int, const float*, const float*, int, float*, cudaStream_t, int);

// ---------------------------------------------------------------------------
// per-joint column sums  t1[v][c] = sum_f ds[f,v,c]   (feeds dW_bias and the
// bias part of dA).  grid.y tiles the V*C axis, grid.x strips of frames.
// ---------------------------------------------------------------------------
template <class T>
__global__ void __launch_bounds__(256) vcolsum_kernel(const T* __restrict__ ds, long long frames, int VC, float* __restrict__ t1) {
  pdl_enter();
  int e = blockIdx.y * 256 + threadIdx.x;
  if (e >= VC) return;
  float s = 0.f;
  for (long long f = blockIdx.x; f < frames; f += gridDim.x) s += ldf(ds + f * VC + e);
  atomicAdd(t1 + e, s);
}
// vector form: one thread owns one 16-byte group of consecutive (v,c) elements and strides over frames with 4 independent loads
// in flight
template <class T>
__global__ void __launch_bounds__(256) vcolsum_vec_kernel(const T* __restrict__ ds, long long frames, int VC, float* __restrict__ t1, int rev) {
  pdl_enter();
  auto F = [&](long long q) { return rev ? frames - 1 - q : q; };         // ping-pong traversal (common.cuh)
  constexpr int VEC = VecN<T>::N;
  const int e = (blockIdx.y * 256 + threadIdx.x) * VEC;
  if (e >= VC) return;
  float s[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) s[i] = 0.f;
  const long long step = gridDim.x;
  long long f = blockIdx.x;
  for (; f + 7 * step < frames; f += 8 * step) {
    float a[8][VEC];
#pragma unroll
    for (int u = 0; u < 8; ++u) ldv(ds + F(f + u * step) * VC + e, a[u]);
#pragma unroll
    for (int i = 0; i < VEC; ++i) s[i] += ((a[0][i] + a[1][i]) + (a[2][i] + a[3][i])) + ((a[4][i] + a[5][i]) + (a[6][i] + a[7][i]));
  }
  for (; f < frames; f += step) {
    float a[VEC]; ldv(ds + F(f) * VC + e, a);
#pragma unroll
    for (int i = 0; i < VEC; ++i) s[i] += a[i];
  }
  // 16-byte vector reductions: a quarter of the L2 atomic operations of scalar atomicAdd
#pragma unroll
  for (int i = 0; i < VEC; i += 4)
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(t1 + e + i), "f"(s[i]), "f"(s[i + 1]), "f"(s[i + 2]), "f"(s[i + 3])
                 : "memory");
}
template <class T>
int32_t launch_vcolsum(const T* ds, long long frames, int V, int C, float* t1, cudaStream_t st) {
  KernelTimer kt_(KC_BN_STATS, 0.0, st);
  int VC = V * C;
  constexpr int VEC = VecN<T>::N;
  if (VC % VEC == 0) {
    const int by = cdiv(VC / VEC, 256);
    // few, long-running CTAs: the final atomics contend bx-way per address
    const int bx = (int)std::max<long long>(1, std::min<long long>(frames / 16, (148 * 2 + by - 1) / by));
    launch_k(vcolsum_vec_kernel<T>, dim3((unsigned)bx, (unsigned)by), 256, 0, st, ds, frames, VC, t1, next_direction());
  } else {
    dim3 grid((unsigned)std::min<long long>(frames, 64), (unsigned)cdiv(VC, 256));
    launch_k(vcolsum_kernel<T>, grid, 256, 0, st, ds, frames, VC, t1);
  }
  STGCN_CHECK_LAUNCH("vcolsum");
  return STGCN_OK;
}
template int32_t launch_vcolsum<float>(const float*, long long, int, int, float*, cudaStream_t);
template int32_t launch_vcolsum<bf16>(const bf16*, long long, int, int, float*, cudaStream_t);

// row sums of A into shared memory: rs[k*V+v] = sum_w A[k,v,w]   (K*V <= 128 entries, one warp-sliced pass)
__device__ __forceinline__ void adjacency_rowsums(const float* __restrict__ A, int K, int V, float* rs) {
  for (int i = threadIdx.x; i < K * V; i += blockDim.x) {
    float s = 0.f;
    for (int w = 0; w < V; ++w) s += A[i * V + w];
    rs[i] = s;
  }
  __syncthreads();
}

// bias2d[v][co] = sum_k rowsum_k[v] * bias[k*Co+co]   (bias of W passes through A, st_graphconv.py:148-150)
__global__ void __launch_bounds__(256) bias2d_kernel(const float* __restrict__ A, const float* __restrict__ bias, int K, int V, int Co,
                                                     float* __restrict__ out) {
  pdl_enter();
  __shared__ float rs[128];
  adjacency_rowsums(A, K, V, rs);
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= V * Co) return;
  int co = i % Co, v = i / Co;
  float s = 0.f;
  for (int k = 0; k < K; ++k) s = fmaf(rs[k * V + v], bias[k * Co + co], s);
  out[i] = s;
}
int32_t launch_bias2d(const float* A, const float* bias, int K, int V, int Co, float* out, cudaStream_t st) {
  launch_k(bias2d_kernel, cdiv(V * Co, 256), 256, 0, st, A, bias, K, V, Co, out);
  STGCN_CHECK_LAUNCH("bias2d");
  return STGCN_OK;
}

// dbias[k*Co+co] = sum_v t1[v][co]*rowsum_k[v] ;  dA[k,v,w] += sum_co t1[v][co]*bias[k*Co+co]
// blocks [0, nb_bias): dbias, 256 outputs each;  blocks [nb_bias, nb_bias + K*V): one (k,v) row of dA each
__global__ void __launch_bounds__(256) spatial_bias_grads_kernel(const float* __restrict__ t1, const float* __restrict__ A,
                                                                 const float* __restrict__ bias, const float* __restrict__ bias_dA,
                                                                 int K, int V, int Co, int nb_bias,
                                                                 float* __restrict__ dbias, float* __restrict__ dA) {
  pdl_enter();
  __shared__ float rs[128];
  __shared__ float wsum[8];
  if ((int)blockIdx.x < nb_bias) {
    adjacency_rowsums(A, K, V, rs);
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < K * Co && dbias) {
      int co = i % Co, k = i / Co;
      float s = 0.f;
      for (int v = 0; v < V; ++v) s = fmaf(t1[v * Co + co], rs[k * V + v], s);
      dbias[i] = s;
    }
    return;
  }
  bias = bias_dA ? bias_dA : bias;
  if (!dA) return;
  const int kv = blockIdx.x - nb_bias, v = kv % V, k = kv / V;
  float s = 0.f;
  for (int co = threadIdx.x; co < Co; co += blockDim.x) s = fmaf(t1[v * Co + co], bias[k * Co + co], s);
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < V) {
    float tot = 0.f;
    for (int i = 0; i < 8; ++i) tot += wsum[i];
    dA[(k * V + v) * V + threadIdx.x] += tot;
  }
}
int32_t launch_spatial_bias_grads(const float* t1, const float* A, const float* bias, int K, int V, int Co, float* dbias,
                                  float* dA, cudaStream_t st, const float* bias_dA) {
  const int nb_bias = cdiv(K * Co, 256);
  launch_k(spatial_bias_grads_kernel, nb_bias + K * V, 256, 0, st, t1, A, bias, bias_dA, K, V, Co, nb_bias, dbias, dA);
  STGCN_CHECK_LAUNCH("spatial_bias_grads");
  return STGCN_OK;
}

// ---------------------------------------------------------------------------
// BatchNorm-1 folded into the 1x1 weights (tensor-core path, default block):
//   S = sum_k (A_k (a.x + b)) W_k^T + ...  =  sum_k (A_k x) (W_k diag(a))^T + bias2d'
// so the aggregation runs on the RAW input and BN-1 costs two tiny kernels:
//   w2p [co][k*Ci+ci] = bf16(W[(k*Co+co)*Ci+ci] * a[ci])        GEMM-1 operand
//   w2pT[k*Ci+ci][co] = same value                                 operand of dXA' = dS . W2'
//   wb  [k*Co+co]     = sum_ci W[(k*Co+co)*Ci+ci] * b[ci]          BN shift through the weights
//   bias2d'[v][co]    = sum_k rowsum(A)_k[v]*Wb[k,co] + rowsum(bf16 A)_k[v]*wb[k,co]
// (reference: batch_n -> spatialConv at st_graphconv.py:98-99,148-150)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(128) bn1_fold_weights_kernel(const float* __restrict__ W, const float* __restrict__ a,
                                                               const float* __restrict__ b, int K, int Co, int Ci, int ld,
                                                               bf16* __restrict__ w2p, bf16* __restrict__ w2pT, float* __restrict__ wb) {
  pdl_enter();
  __shared__ float red[4];
  const int r = blockIdx.x, k = r / Co, co = r % Co;
  float part = 0.f;
  for (int ci = threadIdx.x; ci < Ci; ci += blockDim.x) {
    const float w = W[(size_t)r * Ci + ci];
    const bf16 q = __float2bfloat16_rn(w * a[ci]);
    w2p[(size_t)co * ld + k * Ci + ci] = q;
    w2pT[(size_t)(k * Ci + ci) * Co + co] = q;
    part = fmaf(w, b[ci], part);
  }
  part = warp_sum(part);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = part;
  __syncthreads();
  if (threadIdx.x == 0) wb[r] = red[0] + red[1] + red[2] + red[3];
}
__global__ void __launch_bounds__(256) bias2d_fold_kernel(const float* __restrict__ A, const float* __restrict__ rowsum_q,
                                                          const float* __restrict__ Wb, const float* __restrict__ wb, int K, int V, int Co,
                                                          float* __restrict__ out, float* __restrict__ wbb,
                                                          float* __restrict__ rsum2, float* __restrict__ bvec2) {
  pdl_enter();
  __shared__ float rs[128];
  adjacency_rowsums(A, K, V, rs);
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < K * Co) wbb[i] = Wb[i] + wb[i];
  if (rsum2) {          // rank-2K factors of bias2d' for the fused spatial kernel: [2K][V] row sums, [2K][Co] bias vectors
    if (i < K * V) { rsum2[i] = rs[i]; rsum2[K * V + i] = rowsum_q[i]; }
    if (i < K * Co) { bvec2[i] = Wb[i]; bvec2[K * Co + i] = wb[i]; }
  }
  if (i >= V * Co) return;
  const int co = i % Co, v = i / Co;
  float s = 0.f;
  for (int k = 0; k < K; ++k) s += rs[k * V + v] * Wb[k * Co + co] + rowsum_q[k * V + v] * wb[k * Co + co];
  out[i] = s;
}
int32_t launch_bn1_fold(const float* W, const float* Wb, const float* A, const float* rowsum_q, const float* a, const float* b, int K,
                        int V, int Co, int Ci, bf16* w2p, bf16* w2pT, float* wb, float* wbb, float* bias2d, cudaStream_t st, int ld,
                        float* rsum2, float* bvec2) {
  KernelTimer kt_(KC_PACK, 0.0, st);
  if (ld <= 0) ld = K * Ci;
  if (ld > K * Ci) {      // zero-padded reduction dimension (small C_in): rows / columns K*Ci..ld-1 stay zero
    cudaMemsetAsync(w2p, 0, sizeof(bf16) * (size_t)Co * ld, st);
    cudaMemsetAsync(w2pT, 0, sizeof(bf16) * (size_t)Co * ld, st);
  }
  launch_k(bn1_fold_weights_kernel, K * Co, 128, 0, st, W, a, b, K, Co, Ci, ld, w2p, w2pT, wb);
  STGCN_CHECK_LAUNCH("bn1_fold_weights");
  launch_k(bias2d_fold_kernel, cdiv(std::max(V, K) * Co, 256), 256, 0, st, A, rowsum_q, Wb, wb, K, V, Co, bias2d, wbb, rsum2, bvec2);
  STGCN_CHECK_LAUNCH("bias2d_fold");
  return STGCN_OK;
}

// ---------------------------------------------------------------------------
// small C_in (block 0: 3 coordinates): XA[r][k*C+c] = sum_w A[k,v,w] * x[f,w,c] written as zero-padded rows of
// `ld` channels (ld = 64) so that GEMM-1 / its gradients run on the tcgen05 engine.  The frames of a CTA and A are staged in
// shared memory; one thread per (row, partition) forms the C sums from 16-byte coordinate loads, and the rows leave as
// coalesced 16-byte groups.
// ---------------------------------------------------------------------------
template <class T>
__global__ void __launch_bounds__(256) agg_small_fwd_kernel(const T* __restrict__ x, long long frames, int V, int C, int K, int ld,
                                                            const float* __restrict__ A, int FB, T* __restrict__ xa) {
  pdl_enter();
  extern __shared__ float4 sm4s[];
  // host guarantees C <= 4 and K*C <= 16
  float4* xs = sm4s;                                     // [FB][V]     one joint's C coordinates, zero padded to 4
  float* os = reinterpret_cast<float*>(sm4s + FB * V);   // [FB][V][16] the K*C aggregated values of a row, zero padded to 16
  float* As = os + FB * V * 16;                          // [K][V][V]
  for (int i = threadIdx.x; i < K * V * V; i += blockDim.x) As[i] = A[i];
  constexpr int VEC = VecN<T>::N;
  const int gpr = ld / VEC;                              // 16-byte groups per output row
  for (long long f0 = (long long)blockIdx.x * FB; f0 < frames; f0 += (long long)gridDim.x * FB) {
    const int nf = (int)min((long long)FB, frames - f0);
    __syncthreads();
    for (int i = threadIdx.x; i < nf * V; i += blockDim.x) {
      const T* xp = x + (f0 * V + i) * C;
      float t[4] = {0.f, 0.f, 0.f, 0.f};
      for (int c = 0; c < C; ++c) t[c] = ldf(xp + c);
      xs[i] = make_float4(t[0], t[1], t[2], t[3]);
    }
    for (int i = threadIdx.x; i < nf * V * 16; i += blockDim.x) os[i] = 0.f;
    __syncthreads();
    // one thread per (row, partition): V steps of one adjacency load and one 16-byte coordinate load
    for (int it = threadIdx.x; it < nf * V * K; it += blockDim.x) {
      const int r = it / K, k = it - r * K;
      const int fl = r / V, v = r - fl * V;
      const float* ar = As + (k * V + v) * V;
      const float4* xr = xs + fl * V;
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
      for (int w = 0; w < V; ++w) {
        const float aw = ar[w]; const float4 q = xr[w];
        a0 = fmaf(aw, q.x, a0); a1 = fmaf(aw, q.y, a1); a2 = fmaf(aw, q.z, a2); a3 = fmaf(aw, q.w, a3);
      }
      float* o = os + r * 16 + k * C;
      o[0] = a0; if (C > 1) o[1] = a1; if (C > 2) o[2] = a2; if (C > 3) o[3] = a3;
    }
    __syncthreads();
    // coalesced store: consecutive threads write consecutive 16-byte groups, i.e. whole zero-padded rows
    for (int idx = threadIdx.x; idx < nf * V * gpr; idx += blockDim.x) {
      const int r = idx / gpr, j0 = (idx - r * gpr) * VEC;
      float o[VEC];
#pragma unroll
      for (int u = 0; u < VEC; ++u) o[u] = (j0 + u < 16) ? os[r * 16 + j0 + u] : 0.f;
      stv(xa + (f0 * V + r) * (long long)ld + j0, o);
    }
  }
}
// any C with K*C <= ld: one thread per (frame, joint) row
template <class T>
__global__ void __launch_bounds__(256) agg_small_fwd_generic_kernel(const T* __restrict__ x, long long frames, int V, int C, int K, int ld,
                                                            const float* __restrict__ A, int FB, T* __restrict__ xa) {
  pdl_enter();
  extern __shared__ float sm[];
  float* As = sm;                   // [K][V][V]
  float* xs = sm + K * V * V;       // [FB][V][C]
  for (int i = threadIdx.x; i < K * V * V; i += blockDim.x) As[i] = A[i];
  constexpr int VEC = VecN<T>::N;
  const int KC = K * C;
  for (long long f0 = (long long)blockIdx.x * FB; f0 < frames; f0 += (long long)gridDim.x * FB) {
    const int nf = (int)min((long long)FB, frames - f0);
    __syncthreads();
    for (int i = threadIdx.x; i < nf * V * C; i += blockDim.x) xs[i] = ldf(x + f0 * V * C + i);
    __syncthreads();
    for (int r = threadIdx.x; r < nf * V; r += blockDim.x) {
      const int fl = r / V, v = r % V;
      T* op = xa + ((f0 + fl) * V + v) * (long long)ld;
      for (int j0 = 0; j0 < ld; j0 += VEC) {
        float o[VEC];
#pragma unroll
        for (int u = 0; u < VEC; ++u) {
          const int j = j0 + u;
          float acc = 0.f;
          if (j < KC) {
            const int k = j / C, c = j % C;
            const float* ar = As + (k * V + v) * V;
            const float* xr = xs + fl * V * C + c;
            for (int w = 0; w < V; ++w) acc = fmaf(ar[w], xr[w * C], acc);
          }
          o[u] = acc;
        }
        stv(op + j0, o);
      }
    }
  }
}
template <class T>
int32_t launch_agg_small_fwd(const T* x, long long frames, int V, int C, int K, int ld, const float* A, T* xa, cudaStream_t st) {
  KernelTimer kt_(KC_AGG_FWD, 2.0 * frames * V * V * (double)K * C, st);
  if (K * C > ld || ld % VecN<T>::N != 0) { set_error("agg_small: K*C=%d does not fit ld=%d", K * C, ld); return STGCN_ERR_UNSUPPORTED; }
  const int FB = std::max(1, 256 / V);
  const int blocks = (int)std::min<long long>((frames + FB - 1) / FB, 148 * 8);
  if (C <= 4 && K * C <= 16) {
    const size_t smem = (size_t)FB * V * sizeof(float4) + ((size_t)FB * V * 16 + (size_t)K * V * V) * sizeof(float);
    launch_k(agg_small_fwd_kernel<T>, blocks, 256, smem, st, x, frames, V, C, K, ld, A, FB, xa);
  } else {
    const size_t smem = ((size_t)K * V * V + (size_t)FB * V * C) * sizeof(float);
    launch_k(agg_small_fwd_generic_kernel<T>, blocks, 256, smem, st, x, frames, V, C, K, ld, A, FB, xa);
  }
  STGCN_CHECK_LAUNCH("agg_small_fwd");
  return STGCN_OK;
}
template int32_t launch_agg_small_fwd<float>(const float*, long long, int, int, int, int, const float*, float*, cudaStream_t);
template int32_t launch_agg_small_fwd<bf16>(const bf16*, long long, int, int, int, int, const float*, bf16*, cudaStream_t);

__global__ void adjacency_rowsum_kernel(const float* __restrict__ A, int KV, int V, float* __restrict__ rs) {
  pdl_enter();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= KV) return;
  float s = 0.f;
  for (int w = 0; w < V; ++w) s += A[i * V + w];
  rs[i] = s;
}
int32_t launch_adjacency_rowsum(const float* A, int K, int V, float* rs, cudaStream_t st) {
  launch_k(adjacency_rowsum_kernel, cdiv(K * V, 128), 128, 0, st, A, K * V, V, rs);
  STGCN_CHECK_LAUNCH("adjacency_rowsum");
  return STGCN_OK;
}

// gradients of the folded BatchNorm-1 (no pass over activations):
//   q[k,co]  = sum_v rowsum_q_k[v] * t1[v,co]
//   dbeta_c  = sum_{k,co} q[k,co] * W[(k,co),c]
//   dgamma_c = invstd_c * ( sum_{k,co} W[(k,co),c] * dWraw[(k,co),c]  -  mean_c * dbeta_c )
//   dW       = dWraw (.) a  +  q (x) b        (in place; the second term is the BN shift travelling through bias2d')
//   coef     = { mean, a*dbeta/m, a*invstd*dgamma/m }  for the fused  dx = dh' - c1 - (x-mean)*c2   epilogue
// 8 channels per CTA (C_in / 8 CTAs keep the launch latency-bound part short), 128 row-slices reduced through shared memory.
__global__ void __launch_bounds__(1024) bn1_fold_grads_kernel(const float* __restrict__ W, float* __restrict__ dW,
                                                              const float* __restrict__ stats, const float* __restrict__ t1,
                                                              const float* __restrict__ rowsum_q, int K, int V, int Co, int Ci,
                                                              int training, float inv_m, float* __restrict__ dgamma,
                                                              float* __restrict__ dbeta, float* __restrict__ coef) {
  pdl_enter();
  __shared__ float q[1024];
  __shared__ float red[2][128][9];
  const int KCo = K * Co;
  for (int i = threadIdx.x; i < KCo; i += blockDim.x) {
    const int k = i / Co, co = i % Co;
    float s = 0.f;
    for (int v = 0; v < V; ++v) s = fmaf(rowsum_q[k * V + v], t1[v * Co + co], s);
    q[i] = s;
  }
  __syncthreads();
  const int cl = threadIdx.x & 7, sl = threadIdx.x >> 3, c = blockIdx.x * 8 + cl;
  float s1 = 0.f, s2 = 0.f;
  float a = 0.f, bsh = 0.f;
  if (c < Ci) {
    a = stats[2 * Ci + c]; bsh = stats[3 * Ci + c];
    for (int r = sl; r < KCo; r += 128) {
      const size_t o = (size_t)r * Ci + c;
      const float w = W[o], g = dW[o];
      s1 = fmaf(q[r], w, s1);
      s2 = fmaf(w, g, s2);
      dW[o] = fmaf(g, a, q[r] * bsh);            // raw-weight gradient: scale + the shift term through bias2d'
    }
  }
  red[0][sl][cl] = s1; red[1][sl][cl] = s2;
  __syncthreads();
  if (c < Ci && sl < 2) {                         // slice 0 sums the dbeta partials, slice 1 the dgamma partials, then they meet
    float acc = 0.f;
    for (int i = 0; i < 128; ++i) acc += red[sl][i][cl];
    red[sl][0][cl] = acc;
  }
  __syncthreads();
  if (c < Ci && sl == 0) {
    const float db = red[0][0][cl], sw = red[1][0][cl];
    const float mean = stats[c], invstd = stats[Ci + c];
    const float dg = invstd * (sw - mean * db);
    dgamma[c] = dg; dbeta[c] = db;
    coef[c] = mean;
    coef[Ci + c] = training ? a * db * inv_m : 0.f;
    coef[2 * Ci + c] = training ? a * invstd * dg * inv_m : 0.f;
  }
}
int32_t launch_bn1_fold_grads(const float* W, float* dW, const float* stats, const float* t1, const float* rowsum_q, int K, int V,
                              int Co, int Ci, int training, double rows, float* dgamma, float* dbeta, float* coef, cudaStream_t st) {
  KernelTimer kt_(KC_PACK, 0.0, st);
  if (K * Co > 1024) { set_error("bn1_fold_grads: K*C_out=%d > 1024", K * Co); return STGCN_ERR_UNSUPPORTED; }
  launch_k(bn1_fold_grads_kernel, cdiv(Ci, 8), 1024, 0, st, W, dW, stats, t1, rowsum_q, K, V, Co, Ci, training, (float)(1.0 / rows), dgamma,
                                                       dbeta, coef);
  STGCN_CHECK_LAUNCH("bn1_fold_grads");
  return STGCN_OK;
}

// ---------------------------------------------------------------------------
// adjacency normalisation on the device (reference: normalize, src/data/adjacency.py:161-183):
//   Lambda = (diag(rowsum(A_k)) + alpha) ** expo     -- alpha added to EVERY entry (dense_alpha, the reference) or to the
//                                                       diagonal only (the textbook D^-1/2 A D^-1/2)
//   out_k  = Lambda . A_k . Lambda
// One CTA per partition, V <= 32: the three V x V matrices live in shared memory; fp32, sums in index order.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) adjacency_normalize_kernel(const float* __restrict__ raw, int V, float alpha, float expo,
                                                                    int dense_alpha, float* __restrict__ out) {
  pdl_enter();
  __shared__ float As[32 * 32], Ls[32 * 32], Ts[32 * 32];
  __shared__ float rs[32];
  const int k = blockIdx.x, i = threadIdx.x / 32, j = threadIdx.x % 32;
  const bool in = i < V && j < V;
  As[i * 32 + j] = in ? raw[((size_t)k * V + i) * V + j] : 0.f;
  __syncthreads();
  if (j == 0 && i < V) { float s = 0.f; for (int w = 0; w < V; ++w) s += As[i * 32 + w]; rs[i] = s; }
  __syncthreads();
  float lam = 0.f;
  if (in) {
    if (dense_alpha) lam = powf((i == j ? rs[i] : 0.f) + alpha, expo);
    else lam = i == j ? powf(rs[i] + alpha, expo) : 0.f;
  }
  Ls[i * 32 + j] = lam;
  __syncthreads();
  float t = 0.f;
  if (in) for (int w = 0; w < V; ++w) t = fmaf(Ls[i * 32 + w], As[w * 32 + j], t);
  Ts[i * 32 + j] = t;
  __syncthreads();
  if (in) {
    float o = 0.f;
    for (int w = 0; w < V; ++w) o = fmaf(Ts[i * 32 + w], Ls[w * 32 + j], o);
    out[((size_t)k * V + i) * V + j] = o;
  }
}
int32_t launch_adjacency_normalize(const float* raw, int K, int V, float alpha, float expo, int dense_alpha, float* out, cudaStream_t st) {
  if (K < 1 || V < 1 || V > 32) { set_error("adjacency_normalize: K=%d V=%d unsupported (V <= 32)", K, V); return STGCN_ERR_UNSUPPORTED; }
  launch_k(adjacency_normalize_kernel, K, 1024, 0, st, raw, V, alpha, expo, dense_alpha, out);
  STGCN_CHECK_LAUNCH("adjacency_normalize");
  return STGCN_OK;
}

}  // namespace stgcn
