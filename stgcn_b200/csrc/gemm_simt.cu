// fp32 CUDA-core implicit-GEMM engine: the exact-arithmetic (1e-4) path and
// the engine for shapes the tcgen05 path does not take (C_in = 2/3 in block 0).
//
//   forward / dgrad :  out[r][n] = sum_tap sum_c in[row(r,tap)][c] * w[tap][n][c]  (+epilogue)
//   wgrad           :  dw[tap][n][c] = sum_r dout[r][n] * in[row(r,tap)][c]
//
// rows are (batch, frame, joint) of a channels-last tensor; row(r,tap) shifts
// the frame by the temporal tap with zero padding (nn.Conv2d(C,C,(9,1),
// stride=(s,1),padding=(4,0)) at st_graphconv.py:41-43; taps=1 gives the 1x1
// convolutions at :28 and :136).  64x64x16 shared-memory tiles, 4x4 register
// micro-tiles, fp32 FMA throughout.
#include <stdlib.h>
#include <string.h>

#include "kernels.cuh"

namespace stgcn {

constexpr int BM = 64, BN = 64, BK = 16, LDS_ = 68;

template <class T>
__device__ __forceinline__ void load_row_vec(const T* base, long long row, int C, int c0, bool row_ok, float (&v)[4]) {
  // loads channels c0..c0+3 of `row` (zero when the row is padding or c >= C)
  if (row_ok && (C % 4 == 0) && c0 + 3 < C) { ld4(base + row * C + c0, v); return; }
#pragma unroll
  for (int i = 0; i < 4; ++i) v[i] = (row_ok && c0 + i < C) ? ldf(base + row * C + c0 + i) : 0.f;
}

template <class T>
__global__ void __launch_bounds__(256) conv_gemm_simt_kernel(ConvGeom g, const T* __restrict__ in, const float* __restrict__ w,
                                                             GemmEpilogue ep, T* __restrict__ out) {
  pdl_enter();
  __shared__ __align__(16) float As[BK][LDS_];
  __shared__ __align__(16) float Bs[BK][LDS_];
  const int tid = threadIdx.x, tx = tid % 16, ty = tid / 16;
  const long long rows_out = (long long)g.nb * g.t_out * g.v;
  const long long m0 = (long long)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  // loader coordinates
  const int lrow = tid / 4, lk = (tid % 4) * 4;
  const long long r_load = m0 + lrow;
  int lb = 0, lj = 0, lv = 0;
  const bool r_ok = r_load < rows_out;
  if (r_ok) { lv = (int)(r_load % g.v); long long fo = r_load / g.v; lj = (int)(fo % g.t_out); lb = (int)(fo / g.t_out); }
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int tap = 0; tap < g.taps; ++tap) {
    int fi = r_ok ? conv_in_frame(g, lj, tap) : -1;
    const bool a_ok = fi >= 0;
    const long long a_row = a_ok ? ((long long)lb * g.t_in + fi) * g.v + lv : 0;
    const float* wt = w + (size_t)tap * g.n_out * g.c;
    for (int c0 = 0; c0 < g.c; c0 += BK) {
      float av[4], bv[4];
      load_row_vec<T>(in, a_row, g.c, c0 + lk, a_ok, av);
      const int bn = n0 + lrow;
      load_row_vec<float>(wt, bn, g.c, c0 + lk, bn < g.n_out, bv);
      __syncthreads();
#pragma unroll
      for (int i = 0; i < 4; ++i) { As[lk + i][lrow] = av[i]; Bs[lk + i][lrow] = bv[i]; }
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < BK; ++kk) {
        float4 a4 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
        float4 b4 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
        float a[4] = {a4.x, a4.y, a4.z, a4.w}, b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
      }
    }
  }
  // epilogue
  const float inv_keep = ep.dropout_p > 0.f ? 1.f / (1.f - ep.dropout_p) : 1.f;
  const uint64_t d_seed = ep.rng ? ep.rng[0] : ep.seed, d_off = ep.rng ? ep.rng[1] : ep.offset;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    long long r = m0 + ty * 4 + i;
    if (r >= rows_out) continue;
    int v = (int)(r % g.v);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n = n0 + tx * 4 + j;
      if (n >= g.n_out) continue;
      float t = acc[i][j];
      if (ep.bias) t += ep.bias[n];
      if (ep.bias2d) t += ep.bias2d[v * g.n_out + n];
      if (ep.scale) t = fmaf(ep.scale[n], t, ep.shift[n]);
      size_t o = (size_t)r * g.n_out + n;
      if (ep.residual) t += ldf(reinterpret_cast<const T*>(ep.residual) + o);
      if (ep.relu) t = fmaxf(t, 0.f);
      if (ep.dropout_p > 0.f) t *= dropout_scale(d_seed, d_off, o, ep.dropout_p, inv_keep);
      if (ep.accumulate) t += ldf(out + o);
      stf(out + o, t);
    }
  }
}

template <class T>
int32_t launch_conv_gemm_simt(const ConvGeom& g, const T* in, const float* w, const GemmEpilogue& ep, T* out, cudaStream_t st) {
  KernelTimer kt_(KC_SIMT_GEMM, conv_algorithmic_flops(g), st);
  long long rows = (long long)g.nb * g.t_out * g.v;
  dim3 grid((unsigned)cdiv(rows, BM), (unsigned)cdiv(g.n_out, BN));
  launch_k(conv_gemm_simt_kernel<T>, grid, 256, 0, st, g, in, w, ep, out);
  STGCN_CHECK_LAUNCH("conv_gemm_simt");
  return STGCN_OK;
}
template int32_t launch_conv_gemm_simt<float>(const ConvGeom&, const float*, const float*, const GemmEpilogue&, float*, cudaStream_t);
template int32_t launch_conv_gemm_simt<bf16>(const ConvGeom&, const bf16*, const float*, const GemmEpilogue&, bf16*, cudaStream_t);

// ---------------------------------------------------------------------------
// weight gradient: reduce over rows.  grid = (row chunks, n-tiles * c-tiles, taps)
// `g` describes the FORWARD conv (in -> dout rows); the result is scattered with
// the 4-stride WIndex so it lands directly in the reference's parameter layout.
// ---------------------------------------------------------------------------
template <class T>
__global__ void __launch_bounds__(256) wgrad_simt_kernel(ConvGeom g, const T* __restrict__ dout, const T* __restrict__ in,
                                                         long long rows_per_chunk, int c_tiles, WIndex wi, float* __restrict__ dw) {
  pdl_enter();
  __shared__ __align__(16) float Ds[BK][LDS_];   // [row k][n]
  __shared__ __align__(16) float Is[BK][LDS_];   // [row k][c]
  const int tid = threadIdx.x, tx = tid % 16, ty = tid / 16;
  const int tap = blockIdx.z;
  const int n0 = (blockIdx.y / c_tiles) * BN, c0 = (blockIdx.y % c_tiles) * BN;
  const long long rows_out = (long long)g.nb * g.t_out * g.v;
  const long long r_begin = (long long)blockIdx.x * rows_per_chunk;
  const long long r_end = min(rows_out, r_begin + rows_per_chunk);
  const int lrow = tid / 16, lq = (tid % 16) * 4;     // 16 rows x 16 quads of 4 channels
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (long long r0 = r_begin; r0 < r_end; r0 += BK) {
    long long r = r0 + lrow;
    bool ok = r < r_end;
    float dv[4], iv[4];
    load_row_vec<T>(dout, r, g.n_out, n0 + lq, ok, dv);
    long long a_row = 0; bool a_ok = false;
    if (ok) {
      int v = (int)(r % g.v); long long fo = r / g.v; int j = (int)(fo % g.t_out); int b = (int)(fo / g.t_out);
      int fi = conv_in_frame(g, j, tap);
      a_ok = fi >= 0;
      a_row = a_ok ? ((long long)b * g.t_in + fi) * g.v + v : 0;
    }
    load_row_vec<T>(in, a_row, g.c, c0 + lq, a_ok, iv);
    __syncthreads();
    *reinterpret_cast<float4*>(&Ds[lrow][lq]) = make_float4(dv[0], dv[1], dv[2], dv[3]);
    *reinterpret_cast<float4*>(&Is[lrow][lq]) = make_float4(iv[0], iv[1], iv[2], iv[3]);
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float4 a4 = *reinterpret_cast<const float4*>(&Ds[kk][ty * 4]);
      float4 b4 = *reinterpret_cast<const float4*>(&Is[kk][tx * 4]);
      float a[4] = {a4.x, a4.y, a4.z, a4.w}, b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int n = n0 + ty * 4 + i;
    if (n >= g.n_out) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int c = c0 + tx * 4 + j;
      if (c >= g.c) continue;
      atomicAdd(dw + wi.at(tap, n, c), acc[i][j]);
    }
  }
}

template <class T>
int32_t launch_wgrad_simt(const ConvGeom& g, const T* dout, const T* in, const WIndex& wi, float* dw, cudaStream_t st) {
  KernelTimer kt_(KC_WGRAD, 2.0 * g.nb * g.t_out * g.v * (double)g.c * g.n_out * g.taps, st);
  long long rows = (long long)g.nb * g.t_out * g.v;
  int n_tiles = cdiv(g.n_out, BN), c_tiles = cdiv(g.c, BN);
  int tiles = n_tiles * c_tiles * g.taps;
  long long chunks = std::max<long long>(1, std::min<long long>((rows + 255) / 256, (148 * 4 + tiles - 1) / tiles));
  long long rpc = ((rows + chunks - 1) / chunks + BK - 1) / BK * BK;
  chunks = (rows + rpc - 1) / rpc;
  dim3 grid((unsigned)chunks, (unsigned)(n_tiles * c_tiles), (unsigned)g.taps);
  launch_k(wgrad_simt_kernel<T>, grid, 256, 0, st, g, dout, in, rpc, c_tiles, wi, dw);
  STGCN_CHECK_LAUNCH("wgrad_simt");
  return STGCN_OK;
}
template int32_t launch_wgrad_simt<float>(const ConvGeom&, const float*, const float*, const WIndex&, float*, cudaStream_t);
template int32_t launch_wgrad_simt<bf16>(const ConvGeom&, const bf16*, const bf16*, const WIndex&, float*, cudaStream_t);

// ---------------------------------------------------------------------------
// weight packing: dst[tap][n][c] = src[wi.at(tap,n,c)] * (colscale ? colscale[c % cs_mod] : 1)
// ---------------------------------------------------------------------------
template <class TD>
__global__ void __launch_bounds__(256) pack_weights_kernel(const float* __restrict__ src, WIndex wi, int taps, int n_out, int c,
                                                           int round_bf16, TD* __restrict__ dst) {
  pdl_enter();
  size_t total = (size_t)taps * n_out * c;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int cc = (int)(i % c); size_t r = i / c; int n = (int)(r % n_out); int tap = (int)(r / n_out);
    float wv = src[wi.at(tap, n, cc)];
    if (round_bf16) wv = __bfloat162float(__float2bfloat16_rn(wv));   // bf16 mode: every GEMM operand is bf16-valued
    stf(dst + i, wv);
  }
}
template <class TD>
int32_t launch_pack_weights(const float* src, const WIndex& wi, int taps, int n_out, int c, TD* dst, cudaStream_t st, int round_bf16) {
  KernelTimer kt_(KC_PACK, 0.0, st);
  size_t total = (size_t)taps * n_out * c;
  int blocks = (int)std::min<size_t>((total + 255) / 256, 148 * 4);
  launch_k(pack_weights_kernel<TD>, blocks, 256, 0, st, src, wi, taps, n_out, c, round_bf16, dst);
  STGCN_CHECK_LAUNCH("pack_weights");
  return STGCN_OK;
}
template int32_t launch_pack_weights<float>(const float*, const WIndex&, int, int, int, float*, cudaStream_t, int);
template int32_t launch_pack_weights<bf16>(const float*, const WIndex&, int, int, int, bf16*, cudaStream_t, int);

// column sums (bias gradients): out[c] = scale * sum_r x[r][c], via the stats partials
__global__ void colsum_finalize_kernel(const float* __restrict__ partials, int nblk, int C, float* __restrict__ out) {
  pdl_enter();
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double s = 0;
  for (int b = 0; b < nblk; ++b) s += (double)partials[((size_t)b * 2) * C + c];
  out[c] = (float)s;
}
template <class T>
int32_t launch_colsum(const T* x, long long rows, int C, float* partials, float* out, cudaStream_t st) {
  KernelTimer kt_(KC_BN_STATS, 0.0, st);
  int nblk = 0;
  STGCN_TRY(launch_colstats<T>(x, rows, C, partials, &nblk, st));
  launch_k(colsum_finalize_kernel, cdiv(C, 128), 128, 0, st, partials, nblk, C, out);
  STGCN_CHECK_LAUNCH("colsum_finalize");
  return STGCN_OK;
}
template int32_t launch_colsum<float>(const float*, long long, int, float*, float*, cudaStream_t);
template int32_t launch_colsum<bf16>(const bf16*, long long, int, float*, float*, cudaStream_t);

}  // namespace stgcn
