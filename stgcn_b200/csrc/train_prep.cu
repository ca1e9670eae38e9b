// Weight-only operands of the training chain, for ALL blocks of a stack in ONE launch.
//
// A training step used to spend ~40 small launches (two weight packs and one adjacency expansion per block) on operands that
// depend on the parameters only: the temporal weights (nn.Conv2d(C,C,(9,1)), st_graphconv.py:41-43) as bf16 [tap][n][c] for the
// forward GEMM and as [tap][c][n] for the data gradient, and the block-diagonal bf16 expansions of A (st_graphconv.py:129) the
// tensor-core aggregation multiplies by.  Measured on B200 (cfg4): skipping those launches shortens the 6.09 ms step by 0.33 ms,
// far more than their run time, because each one sits in the dependency chain of its block.  Here they become one grid: job j
// (blockIdx.y) is one pack or one adjacency expansion, its CTAs (blockIdx.x) stride over the job's elements.
#include "kernels.cuh"

namespace stgcn {

__global__ void __launch_bounds__(256) train_prep_multi_kernel(const __grid_constant__ PrepTable tab) {
  pdl_enter();
  const PrepJob& j = tab.jobs[blockIdx.y];
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (size_t)gridDim.x * blockDim.x;
  if (j.kind == 0) {                       // dst[tap][n][c] = bf16(src[wi.at(tap, n, c)])
    const size_t total = (size_t)j.taps * j.n_out * j.c;
    for (size_t i = tid; i < total; i += nthr) {
      const int c = (int)(i % j.c); const size_t r = i / j.c;
      const int n = (int)(r % j.n_out), tap = (int)(r / j.n_out);
      j.dst[i] = __float2bfloat16_rn(j.src[j.wi.at(tap, n, c)]);
    }
  } else {                                 // block-diagonal expansions of A and the row sums of the rounded A (umma_agg.cu)
    const int total = j.K * 128 * 128;
    for (size_t i = tid; i < (size_t)total; i += nthr) {
      const int c = (int)(i % 128), r = (int)((i / 128) % 128), k = (int)(i / (128 * 128));
      const int fr = r / j.V, fc = c / j.V;
      float a = 0.f, at = 0.f;
      if (fr == fc && fr < j.FT) {
        a = j.src[(k * j.V + r % j.V) * j.V + c % j.V];
        at = j.src[(k * j.V + c % j.V) * j.V + r % j.V];
      }
      j.big[i] = __float2bfloat16_rn(a);
      j.bigT[i] = __float2bfloat16_rn(at);
    }
    if (tid < (size_t)(j.K * j.V)) {
      float s = 0.f;
      for (int w = 0; w < j.V; ++w) s += __bfloat162float(__float2bfloat16_rn(j.src[tid * j.V + w]));
      j.rowsum[tid] = s;
    }
  }
}

int32_t launch_train_prep_multi(const PrepTable& tab, cudaStream_t st) {
  KernelTimer kt_(KC_PACK, 0.0, st);
  if (tab.count <= 0) return STGCN_OK;
  launch_k(train_prep_multi_kernel, dim3(48, (unsigned)tab.count), 256, 0, st, tab);
  STGCN_CHECK_LAUNCH("train_prep_multi");
  return STGCN_OK;
}

}  // namespace stgcn
