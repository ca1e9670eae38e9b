// Shared device/host helpers for the stgcn_b200 kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>

#include "../../include/stgcn_b200.h"

namespace stgcn {

typedef __nv_bfloat16 bf16;

// ---- error plumbing -------------------------------------------------------
void set_error(const char* fmt, ...);
extern std::atomic<uint64_t> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

#define STGCN_CHECK_LAUNCH(what)                                                         \
  do {                                                                                   \
    ::stgcn::count_launch();                                                             \
    cudaError_t e__ = cudaGetLastError();                                                \
    if (e__ != cudaSuccess) {                                                            \
      ::stgcn::set_error("%s: launch failed: %s", what, cudaGetErrorString(e__));        \
      return STGCN_ERR_CUDA;                                                             \
    }                                                                                    \
  } while (0)

#define STGCN_TRY(expr)              \
  do {                               \
    int32_t rc__ = (expr);           \
    if (rc__ != STGCN_OK) return rc__; \
  } while (0)

// ---- programmatic dependent launch ------------------------------------------------------------------------------
// Every kernel of the library is launched with cudaLaunchAttributeProgrammaticStreamSerialization (launch_k) and starts with
// pdl_trigger() -- "my dependents may be scheduled as soon as all my CTAs are running" -- followed, before its first global
// memory access, by pdl_wait(), which returns once the preceding grid in the stream has completed and flushed.  A kernel's
// launch latency, CTA ramp and on-chip prologue (barrier init, TMEM allocation, descriptor prefetch) then overlap the tail
// of its predecessor.  Since every kernel waits before touching memory, completion order stays transitive along the stream.
// STGCN_NO_PDL=1 launches without the attribute (both instructions are then no-ops).
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_enter() { pdl_trigger(); pdl_wait(); }
bool pdl_enabled();
// Ping-pong traversal: consecutive large kernels of a chain walk their tiles in OPPOSITE directions, so each one starts with the
// rows its predecessor touched last -- the part of a 61 MB activation tensor that is still resident in the 126 MB L2.  Every
// participating launcher takes the next direction from this per-thread toggle (STGCN_NO_PINGPONG=1: always ascending).
int next_direction();
template <class... KArgs, class... Args>
inline void launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);       // errors surface through STGCN_CHECK_LAUNCH (cudaGetLastError)
}

// ---- element access ---------------------------------------------------------
__device__ __forceinline__ float ldf(const float* p) { return *p; }
__device__ __forceinline__ float ldf(const bf16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ void stf(float* p, float v) { *p = v; }
__device__ __forceinline__ void stf(bf16* p, float v) { *p = __float2bfloat16_rn(v); }
// value after a round trip through the storage type (statistics are taken on what is stored)
__device__ __forceinline__ float roundf_as(float v, const float*) { return v; }
__device__ __forceinline__ float roundf_as(float v, const bf16*) { return __bfloat162float(__float2bfloat16_rn(v)); }

// 4-wide vector access (C is always a multiple of 4 on the vector paths)
__device__ __forceinline__ void ld4(const float* p, float (&v)[4]) {
  float4 t = *reinterpret_cast<const float4*>(p);
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
__device__ __forceinline__ void ld4(const bf16* p, float (&v)[4]) {
  uint2 t = *reinterpret_cast<const uint2*>(p);
  __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&t.x);
  __nv_bfloat162 b = *reinterpret_cast<__nv_bfloat162*>(&t.y);
  v[0] = __low2float(a); v[1] = __high2float(a); v[2] = __low2float(b); v[3] = __high2float(b);
}
__device__ __forceinline__ void st4(float* p, const float (&v)[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ void st4(bf16* p, const float (&v)[4]) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]);
  __nv_bfloat162 b = __floats2bfloat162_rn(v[2], v[3]);
  uint2 t;
  t.x = *reinterpret_cast<uint32_t*>(&a);
  t.y = *reinterpret_cast<uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = t;
}

// 16-byte vector access: VecN<T>::N elements per access (4 floats / 8 bf16)
template <class T> struct VecN;
template <> struct VecN<float> { static constexpr int N = 4; };
template <> struct VecN<bf16> { static constexpr int N = 8; };
__device__ __forceinline__ void ldv(const float* p, float (&v)[4]) { ld4(p, v); }
__device__ __forceinline__ void stv(float* p, const float (&v)[4]) { st4(p, v); }
__device__ __forceinline__ void ldv(const bf16* p, float (&v)[8]) {
  const uint4 t = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
  for (int i = 0; i < 4; ++i) { v[2 * i] = __low2float(h[i]); v[2 * i + 1] = __high2float(h[i]); }
}
__device__ __forceinline__ void stv(bf16* p, const float (&v)[8]) {
  uint4 t;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&t);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
  *reinterpret_cast<uint4*>(p) = t;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- Philox4x32-10 (dropout mask replayable in backward) -------------------
__device__ __forceinline__ uint4 philox4x32(uint64_t seed, uint64_t ctr_lo, uint64_t ctr_hi) {
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  uint32_t c0 = (uint32_t)ctr_lo, c1 = (uint32_t)(ctr_lo >> 32), c2 = (uint32_t)ctr_hi, c3 = (uint32_t)(ctr_hi >> 32);
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}
// keep-scale for element index i (1/(1-p) if kept else 0); one Philox call serves 4 consecutive elements
__device__ __forceinline__ float dropout_scale(uint64_t seed, uint64_t offset, uint64_t i, float p, float inv_keep) {
  uint4 r = philox4x32(seed, (i >> 2) + offset, 0x5747434eull);
  uint32_t w = (i & 3) == 0 ? r.x : (i & 3) == 1 ? r.y : (i & 3) == 2 ? r.z : r.w;
  float u = (float)(w >> 8) * (1.0f / 16777216.0f);
  return u >= p ? inv_keep : 0.0f;
}

// ---- geometry ----------------------------------------------------------------
struct ConvGeom {      // implicit GEMM over frames: see gemm_simt.cu / umma_gemm.cu
  int nb;              // batch
  int t_in, t_out;     // frames of the input / output tensor
  int v;               // joints
  int c;               // reduction channels per tap
  int n_out;           // output channels
  int taps;
  int stride, pad;     // forward: t_in_frame = j*stride - pad + tap
  int dgrad;           // 1: transposed conv: q = j + pad - tap, valid iff q%stride==0, frame = q/stride
  int c_valid = 0;     // reduction / output channels that carry data when the operand is zero-padded to c / n_out (0 = all):
  int n_valid = 0;     // only used for the ALGORITHMIC work the timing hooks report
};
// algorithmic multiply-adds x 2 of one implicit GEMM: forward 2 rows_out taps c n; the data gradient of a
// strided conv does the forward's work (its rows are the forward's OUTPUT rows = this geometry's input rows), not taps x its rows
__host__ inline double conv_algorithmic_flops(const ConvGeom& g) {
  const double rows = (double)g.nb * (g.dgrad ? g.t_in : g.t_out) * g.v;
  return 2.0 * rows * g.taps * (double)(g.c_valid > 0 ? g.c_valid : g.c) * (double)(g.n_valid > 0 ? g.n_valid : g.n_out);
}

__host__ __device__ __forceinline__ int conv_in_frame(const ConvGeom& g, int j, int tap) {
  if (!g.dgrad) {
    int f = j * g.stride - g.pad + tap;
    return (f >= 0 && f < g.t_in) ? f : -1;
  }
  int q = j + g.pad - tap;
  if (q < 0 || (q % g.stride) != 0) return -1;
  q /= g.stride;
  return q < g.t_in ? q : -1;
}

static inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }
static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace stgcn
