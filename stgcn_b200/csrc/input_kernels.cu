// Input pipeline in front of block 0, one kernel:
//   * loop-padding collate  -- loopy_pad_collate_fn / pad_array_with_loops, src/data/util.py:12-47: every sequence of the batch
//     is extended to the longest one by np.pad(mode='wrap') on the time axis, i.e. frame t >= T_i repeats frame t mod T_i;
//   * the 2-D affine augmentation -- augment_data, src/data/augmentation.py:8-69: each frame's (x, y, 0) row vector is multiplied by
//     a 3x3 matrix T and the first two components are kept.  The homogeneous coordinate is 0 there, so only T[:2,:2] acts
//     (the translation entries of T never reach the output); the kernel takes that 2x2 block per sequence;
//   * the cast to the storage type of the block stack, written in the channels-last (N, T, V, C) layout block 0 reads.
// The ragged batch arrives as ONE packed fp32 buffer (sequence i = lengths[i] frames of V*C values at element offsets[i]).
#include "kernels.cuh"

namespace stgcn {

template <class T>
__global__ void __launch_bounds__(256) collate_pad_augment_kernel(const float* __restrict__ packed, const long long* __restrict__ offsets,
                                                                  const int* __restrict__ lengths, const float* __restrict__ xform,
                                                                  int n, int t_out, int v, int c, T* __restrict__ out) {
  pdl_enter();
  const long long joints = (long long)n * t_out * v;
  for (long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x; j < joints; j += (long long)gridDim.x * blockDim.x) {
    const int vv = (int)(j % v);
    const long long r = j / v;
    const int t = (int)(r % t_out), i = (int)(r / t_out);
    const int len = lengths[i];
    T* dst = out + j * c;
    if (len <= 0) {                                  // an empty sequence has nothing to wrap: zeros
      for (int k = 0; k < c; ++k) stf(dst + k, 0.f);
      continue;
    }
    const float* src = packed + offsets[i] + ((long long)(t % len) * v + vv) * c;
    if (c >= 2) {
      const float x = src[0], y = src[1];
      float m00 = 1.f, m01 = 0.f, m10 = 0.f, m11 = 1.f;
      if (xform) { const float* m = xform + (size_t)i * 4; m00 = m[0]; m01 = m[1]; m10 = m[2]; m11 = m[3]; }
      // row vector times matrix, like np.dot(frames, T): x' = x T00 + y T10, y' = x T01 + y T11
      stf(dst + 0, xform ? fmaf(y, m10, x * m00) : x);
      stf(dst + 1, xform ? fmaf(y, m11, x * m01) : y);
      for (int k = 2; k < c; ++k) stf(dst + k, src[k]);
    } else {
      for (int k = 0; k < c; ++k) stf(dst + k, src[k]);
    }
  }
}

}  // namespace stgcn

using namespace stgcn;
extern "C" int32_t stgcn_collate_pad_augment(int32_t dtype, int32_t n, int32_t t_out, int32_t v, int32_t c, const float* packed,
                                             const int64_t* offsets, const int32_t* lengths, const float* xform, void* out,
                                             void* stream) {
  if (n < 0 || t_out < 0 || v <= 0 || c <= 0) { set_error("collate_pad_augment: bad shape n=%d t=%d v=%d c=%d", n, t_out, v, c); return STGCN_ERR_INVALID; }
  if ((long long)n * t_out == 0) return STGCN_OK;                   // empty batch: nothing to write
  if (!packed || !offsets || !lengths || !out) { set_error("collate_pad_augment: null pointer"); return STGCN_ERR_INVALID; }
  cudaStream_t st = (cudaStream_t)stream;
  KernelTimer kt_(KC_PACK, 0.0, st);
  const long long joints = (long long)n * t_out * v;
  const int blocks = (int)std::min<long long>((joints + 255) / 256, 148 * 8);
  if (dtype == STGCN_F32)
    launch_k(collate_pad_augment_kernel<float>, blocks, 256, 0, st, packed, (const long long*)offsets, lengths, xform, n, t_out, v, c, (float*)out);
  else if (dtype == STGCN_BF16)
    launch_k(collate_pad_augment_kernel<bf16>, blocks, 256, 0, st, packed, (const long long*)offsets, lengths, xform, n, t_out, v, c, (bf16*)out);
  else { set_error("collate_pad_augment: bad dtype %d", dtype); return STGCN_ERR_INVALID; }
  STGCN_CHECK_LAUNCH("collate_pad_augment");
  return STGCN_OK;
}
