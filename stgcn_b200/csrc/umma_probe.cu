// Single-CTA tcgen05 probe used by tests/test_umma_probe.py to pin down descriptor
// semantics that the fused kernels rely on and that cannot be checked without a GPU:
//   mode 0  K-major A ([rows][64] bf16, 128B swizzle): does a descriptor whose start address is
//           shifted by `row_off` rows (not a multiple of the 8-row swizzle atom) read rows
//           row_off .. row_off+127?   (temporal taps as pure descriptor offsets into one slab)
//   mode 1  MN-major A and B ([k rows][64 + 64 channels]): D[m][n] = sum_k A[k+row_off][m] * B[k][n]
//           (the weight-gradient GEMM reduces over rows, so both operands are MN-major)
// D (128 x 64 fp32) is written to global memory; the test compares with a CPU product.
#include "kernels.cuh"
#include "umma_common.cuh"

namespace stgcn {
using namespace umma;

struct ProbeParams { int mode; int row_off; int base_off; int a_rows; };

__global__ void __launch_bounds__(128, 1)
umma_probe_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                  const __grid_constant__ ProbeParams p, float* __restrict__ d_out) {
  pdl_trigger();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  // A region: up to 512 rows x 128 B (mode 1: two slabs of 256 rows); B region: 128 rows x 128 B
  uint8_t* sa = smem;
  uint8_t* sb = smem + 512 * 128;
  uint64_t* bar = reinterpret_cast<uint64_t*>(sb + 128 * 128);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 2);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  if (threadIdx.x == 0) { mbar_init(bar, 1); mbar_init(bar + 1, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(tmem_slot, 64);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();                       // everything above (barriers, TMEM, descriptor prefetch) overlaps the previous kernel's tail
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) {
    if (p.mode == 0) {
      mbar_expect_tx(bar, (uint32_t)(p.a_rows * 128 + 64 * 128));
      tma_load_2d(&map_a, sa, bar, 0, 0);            // a_rows x 64
      tma_load_2d(&map_b, sb, bar, 0, 0);            // 64 x 64
    } else {
      mbar_expect_tx(bar, (uint32_t)(2 * p.a_rows * 128 + 64 * 128));
      tma_load_2d(&map_a, sa, bar, 0, 0);            // channels 0..63,   a_rows k-rows
      tma_load_2d(&map_a, sa + 256 * 128, bar, 64, 0);   // channels 64..127
      tma_load_2d(&map_b, sb, bar, 0, 0);            // 64 k-rows x 64 n
    }
    mbar_wait(bar, 0);
    tc_fence_after();
    const uint32_t a_addr = smem_u32(sa) + (uint32_t)p.row_off * 128u;
    uint64_t bo = p.base_off >= 0 ? ((uint64_t)(p.base_off & 7) << 49) : 0;
    if (p.mode == 0) {
      const uint32_t idesc = make_idesc_bf16(128, 64, 0, 0);
      const uint64_t ad = make_smem_desc_sw128(a_addr, 16, 1024) | bo;
      const uint64_t bd = make_smem_desc_sw128(smem_u32(sb), 16, 1024);
      for (int k = 0; k < 4; ++k) umma_bf16_ss(tmem_base, ad + (uint64_t)(k * 2), bd + (uint64_t)(k * 2), idesc, k != 0);
    } else {
      const uint32_t idesc = make_idesc_bf16(128, 64, 1, 1);
      // MN-major: LBO = stride between 64-wide M/N slabs, SBO = stride between 8-row k groups
      const uint64_t ad = make_smem_desc_sw128(a_addr, 256 * 128, 1024) | bo;
      const uint64_t bd = make_smem_desc_sw128(smem_u32(sb), 64 * 128, 1024);
      for (int k = 0; k < 4; ++k)   // K = 16 rows = 2048 bytes per step
        umma_bf16_ss(tmem_base, ad + (uint64_t)(k * 128), bd + (uint64_t)(k * 128), idesc, k != 0);
    }
    umma_commit(bar + 1);
  }
  __syncwarp();
  mbar_wait(bar + 1, 0);
  tc_fence_after();
  for (int half = 0; half < 2; ++half) {
    uint32_t r[32];
    tmem_ld_32x32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(half * 32), r);
    tmem_ld_wait();
    const int row = warp * 32 + lane;
    for (int i = 0; i < 32; ++i) d_out[row * 64 + half * 32 + i] = __uint_as_float(r[i]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem_base, 64); }
}

}  // namespace stgcn

using namespace stgcn;

// a: (a_rows_total x a_cols) bf16 row-major; b: (64 x 64) bf16 row-major; d: 128 x 64 fp32
extern "C" int32_t stgcn_umma_probe(int32_t mode, int32_t row_off, int32_t base_off, int32_t a_rows, const void* a,
                                    const void* b, float* d, void* stream) {
  if (!a || !b || !d) { set_error("probe: null pointer"); return STGCN_ERR_INVALID; }
  if (a_rows < 1 || a_rows > 256 || row_off < 0) { set_error("probe: bad rows"); return STGCN_ERR_INVALID; }
  CUtensorMap ma, mb;
  const uint64_t a_cols = mode == 0 ? 64 : 128;
  { uint64_t dims[2] = {a_cols, (uint64_t)a_rows}; uint64_t str[1] = {a_cols * 2}; uint32_t box[2] = {64, (uint32_t)a_rows};
    STGCN_TRY(make_tmap_bf16(&ma, a, 2, dims, str, box)); }
  { uint64_t dims[2] = {64, 64}; uint64_t str[1] = {128}; uint32_t box[2] = {64, 64};
    STGCN_TRY(make_tmap_bf16(&mb, b, 2, dims, str, box)); }
  ProbeParams p{mode, row_off, base_off, a_rows};
  const size_t smem = 512 * 128 + 128 * 128 + 1024 + 256;
  cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  launch_k(umma_probe_kernel, 1, 128, smem, (cudaStream_t)stream, ma, mb, p, d);
  STGCN_CHECK_LAUNCH("umma_probe");
  return STGCN_OK;
}
