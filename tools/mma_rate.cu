// tcgen05.mma issue/execute rate microbenchmark (one CTA per SM, one issuing warp):  tools/_bin/mma_rate
// For each variant a warp issues R back-to-back MMAs (128 x N x 16, bf16, fp32 accumulate), commits, and waits; cycles per MMA are
// (t_end - t_start) / R by %clock of SM 0.  Operand contents are irrelevant (shared memory zeroed).
//   build: nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a --expt-relaxed-constexpr -I stgcn_b200/csrc tools/mma_rate.cu -o tools/_bin/mma_rate -lcuda
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
typedef __nv_bfloat16 bf16;
#include "umma_common.cuh"
using namespace stgcn::umma;

struct Var { int n; int a_tmem; int b_mn; int alt; const char* name; };

__device__ __forceinline__ uint32_t clk() { uint32_t c; asm volatile("mov.u32 %0, %%clock;" : "=r"(c)); return c; }

// pattern: alt = 0 plain sequence of R MMAs into D0.
//          alt = 1: groups of [8 x (variant A1: N=64, B MN-major, into D0)] [4 x (this variant, into D1)]  (the spatial kernel's chunk)
//          alt = 2: same with D0 MMAs reading A from shared memory (SS) instead of TMEM
__global__ void __launch_bounds__(128, 1) rate_kernel(int n, int a_tmem, int b_mn, int alt, int reps, unsigned* out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sa = smem;                 // 128 rows x 128 B x 2 (K = 128 for the MN-major case)
  uint8_t* sb = smem + 64 * 1024;     // up to 256 rows x 128 B
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 160 * 1024);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
  for (int i = threadIdx.x; i < 160 * 1024 / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  const int warp = threadIdx.x / 32;
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(slot, 512);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *slot;
  if (warp == 1) {
    const uint32_t idesc = make_idesc_bf16(128, n, 0, b_mn);
    const uint32_t idesc1 = make_idesc_bf16(128, 64, 0, 1);
    const uint64_t ad = make_smem_desc_sw128(smem_u32(sa), 16, 1024);
    const uint64_t bd = make_smem_desc_sw128(smem_u32(sb), b_mn ? 16384 : 16, 1024);
    const uint64_t bd1 = make_smem_desc_sw128(smem_u32(sb), 16384, 1024);
    const uint32_t d0 = tm, d1 = tm + 256, at = tm + 448;        // A operand columns (packed bf16): 64 columns
    uint32_t t0 = clk();
    int issued = 0;
    if (alt == 0) {
      for (int r = 0; r < reps; ++r) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const uint64_t b = bd + (uint64_t)(b_mn ? kk * 128 : kk * 2);
          if (a_tmem) umma_bf16_ts_elect(d0, at + kk * 8, b, idesc, 1u);
          else umma_bf16_ss_elect(d0, ad + (uint64_t)(kk * 2), b, idesc, 1u);
        }
        issued += 4;
      }
    } else {
      for (int r = 0; r < reps; ++r) {
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          if (alt == 1) umma_bf16_ts_elect(d0, at + kk * 8, bd1 + (uint64_t)(kk * 128), idesc1, 1u);
          else umma_bf16_ss_elect(d0, ad + (uint64_t)((kk / 4) * 1024 + (kk % 4) * 2), bd1 + (uint64_t)(kk * 128), idesc1, 1u);
        }
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const uint64_t b = bd + (uint64_t)(b_mn ? kk * 128 : kk * 2);
          if (a_tmem) umma_bf16_ts_elect(d1, at + 32 + kk * 8, b, idesc, 1u);
          else umma_bf16_ss_elect(d1, ad + (uint64_t)(kk * 2), b, idesc, 1u);
        }
        issued += 12;
      }
    }
    uint32_t t1 = clk();
    umma_commit_elect(bar);
    mbar_wait(bar, 0);
    uint32_t t2 = clk();
    if (blockIdx.x == 0 && threadIdx.x == 32) { out[0] = t1 - t0; out[1] = t2 - t0; out[2] = issued; }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tm, 512); }
}


// Tight issue sequence: every operand warp-uniform (tensor-memory base through a shuffle, descriptors from the shared-memory window),
// 16 identical MMAs unrolled per loop trip.  MODE 0: elect per MMA (the library's umma_*_elect helpers); MODE 1: one elect, then a
// single-lane branch issues everything.
template <int N, int TS, int MODE>
__global__ void __launch_bounds__(128, 1) tight_kernel(int reps, unsigned* out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 160 * 1024);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
  for (int i = threadIdx.x; i < 160 * 1024 / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  const int warp = uniform_warp_idx();
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(slot, 512);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) {
    const uint32_t tm = __shfl_sync(0xffffffffu, *slot, 0);
    const uint32_t idesc = make_idesc_bf16(128, N, 0, 0);
    const uint64_t ad = make_smem_desc_sw128(smem_u32(smem), 16, 1024);
    const uint64_t bd = make_smem_desc_sw128(smem_u32(smem + 64 * 1024), 16, 1024);
    const uint32_t at = tm + 448;
    const uint32_t t0 = clk();
    if (MODE == 0) {
      for (int r = 0; r < reps; ++r) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
          if (TS) umma_bf16_ts_elect(tm, at + (u & 3) * 8, bd + (uint64_t)((u & 3) * 2), idesc, 1u);
          else umma_bf16_ss_elect(tm, ad + (uint64_t)((u & 3) * 2), bd + (uint64_t)((u & 3) * 2), idesc, 1u);
        }
      }
    } else if (elect_one()) {
      for (int r = 0; r < reps; ++r) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
          if (TS) asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(tm), "r"(at + (u & 3) * 8), "l"(bd + (uint64_t)((u & 3) * 2)), "r"(idesc), "r"(1u) : "memory");
          else umma_bf16_ss(tm, ad + (uint64_t)((u & 3) * 2), bd + (uint64_t)((u & 3) * 2), idesc, 1u);
        }
      }
    }
    __syncwarp();
    const uint32_t t1 = clk();
    umma_commit_elect(bar);
    mbar_wait(bar, 0);
    const uint32_t t2 = clk();
    if (blockIdx.x == 0 && threadIdx.x == 32) { out[0] = t1 - t0; out[1] = t2 - t0; out[2] = reps * 16; }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(__shfl_sync(0xffffffffu, *slot, 0), 512); }
}
template <int N, int TS, int MODE> static void run_tight(const char* name, unsigned* out) {
  cudaFuncSetAttribute(tight_kernel<N, TS, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  unsigned h[3] = {0, 0, 0};
  for (int it = 0; it < 3; ++it) { tight_kernel<N, TS, MODE><<<148, 128, 170 * 1024>>>(32, out); cudaDeviceSynchronize(); }
  cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
  printf("  tight %-40s issue %6.1f  complete %6.1f cycles per MMA (floor %d)\n", name, (double)h[0] / h[2], (double)h[1] / h[2], N / 2);
}

int main() {
  unsigned* out; cudaMalloc(&out, 16);
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const Var vars[] = {
      {64, 0, 0, 0, "SS  N=64  B K-major"},  {64, 0, 1, 0, "SS  N=64  B MN-major"}, {64, 1, 0, 0, "TS  N=64  B K-major"},
      {64, 1, 1, 0, "TS  N=64  B MN-major"}, {128, 0, 0, 0, "SS  N=128 B K-major"},  {128, 1, 0, 0, "TS  N=128 B K-major"},
      {256, 0, 0, 0, "SS  N=256 B K-major"}, {256, 1, 0, 0, "TS  N=256 B K-major"},
      {64, 1, 0, 1, "chunk pattern: 8 x TS(N=64,MN) -> D0, 4 x TS(N=64,K) -> D1"},
      {64, 1, 0, 2, "chunk pattern: 8 x SS(N=64,MN) -> D0, 4 x TS(N=64,K) -> D1"},
      {128, 1, 0, 1, "chunk pattern: 8 x TS(N=64,MN) -> D0, 4 x TS(N=128,K) -> D1"},
      {64, 0, 0, 2, "chunk pattern: 8 x SS(N=64,MN) -> D0, 4 x SS(N=64,K) -> D1"},
  };
  run_tight<64, 0, 0>("SS N=64  elect per MMA", out);   run_tight<64, 0, 1>("SS N=64  single-lane branch", out);
  run_tight<64, 1, 0>("TS N=64  elect per MMA", out);   run_tight<64, 1, 1>("TS N=64  single-lane branch", out);
  run_tight<128, 0, 0>("SS N=128 elect per MMA", out);  run_tight<128, 0, 1>("SS N=128 single-lane branch", out);
  run_tight<128, 1, 0>("TS N=128 elect per MMA", out);  run_tight<256, 0, 0>("SS N=256 elect per MMA", out);
  run_tight<32, 0, 0>("SS N=32  elect per MMA", out);   run_tight<16, 0, 1>("SS N=16  single-lane branch", out);
  for (int grid : {148}) {
    printf("grid = %d CTA(s)\n", grid);
    for (const Var& v : vars) {
      unsigned h[3] = {0, 0, 0};
      for (int it = 0; it < 3; ++it) {
        rate_kernel<<<grid, 128, 170 * 1024>>>(v.n, v.a_tmem, v.b_mn, v.alt, 64, out);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("  %-64s ERROR %s\n", v.name, cudaGetErrorString(e)); return 1; }
      }
      cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
      printf("  %-64s issue %6.1f  complete %6.1f cycles per MMA (floor %d)  [%u MMAs]\n", v.name, (double)h[0] / h[2], (double)h[1] / h[2], v.n / 2, h[2]);
    }
  }
  return 0;
}
