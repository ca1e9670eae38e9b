// K-partition adjacency aggregation on tcgen05 tensor cores.
//
// Reference: torch.einsum('kvw,nkctw->nctv', A, Z) at st_graphconv.py:150, applied on the
// C_in side (aggregation and 1x1 channel mixing commute):
//     XA[f, v, k*C + c] = sum_w A[k,v,w] * h[f, w, c]
// A tile of FT whole frames (FT*V <= 128 rows) is aggregated with ONE dense MMA against a
// block-diagonal 128x128 expansion of A_k that lives in shared memory for the whole kernel:
//     XA_k tile [128 x C] = BigA_k [128 x 128] . h tile [128 x C]
// so the V x V products of all frames of the tile run on the tensor pipe (the 5.1x zero
// padding of the block-diagonal costs far less than CUDA-core FMAs: measured 9.6 ms -> see
// profiles/).  The activation tile is the MMA's B operand in MN-major form, i.e. exactly the
// [rows][64 channels] image TMA delivers; BigA_k is the K-major A operand.
//
//   agg_fwd : XA_k = a (.) (BigA_k . x) + b * rowsum_k      (BatchNorm-1 affine in the epilogue)
//   agg_bwd : dh   = sum_k BigA_k^T . dXA_k                 (optionally masked by ReLU(h) > 0)
//   agg_dA  : G_k += dXA_k tile . h tile^T  accumulated in TMEM over all tiles of the CTA;
//             dA[k] = sum of the FT diagonal V x V blocks of G_k
#include <stdlib.h>
#include <string.h>

#include "kernels.cuh"
#include "umma_common.cuh"

namespace stgcn {
using namespace umma;

constexpr int kAggUThreads = 192;
constexpr int kTile16K = 128 * 128;     // one [128 rows][64 bf16] tile image

// ---------------------------------------------------------------------------
// block-diagonal expansions of A (bf16), their transposes, and row sums of the rounded A
//   big [k][r][c]  : r = f*V+v, c = f*V+w  -> A[k][v][w]      (zero elsewhere)
//   bigT[k][r][c]  : r = f*V+w, c = f*V+v  -> A[k][v][w]
// ---------------------------------------------------------------------------
__global__ void build_big_adjacency_kernel(const float* __restrict__ A, int K, int V, int FT, bf16* __restrict__ big,
                                           bf16* __restrict__ bigT, float* __restrict__ rowsum) {
  pdl_enter();
  const int total = K * 128 * 128;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c = i % 128, r = (i / 128) % 128, k = i / (128 * 128);
    const int fr = r / V, fc = c / V;
    float a = 0.f, at = 0.f;
    if (fr == fc && fr < FT) {
      a = A[(k * V + r % V) * V + c % V];
      at = A[(k * V + c % V) * V + r % V];
    }
    big[i] = __float2bfloat16_rn(a);
    bigT[i] = __float2bfloat16_rn(at);
  }
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < K * V) {
    float s = 0.f;
    for (int w = 0; w < V; ++w) s += __bfloat162float(__float2bfloat16_rn(A[i * V + w]));
    rowsum[i] = s;
  }
}

size_t umma_agg_prep_bytes(int K) { return (size_t)K * 128 * 128 * 2 * 2 + 4096; }

struct AggPrep { bf16* big; bf16* bigT; float* rowsum; };
AggPrep carve_agg_prep(void* mem, int K) {
  AggPrep a;
  a.big = reinterpret_cast<bf16*>(mem);
  a.bigT = a.big + (size_t)K * 128 * 128;
  a.rowsum = reinterpret_cast<float*>(a.bigT + (size_t)K * 128 * 128);
  return a;
}

void umma_agg_prep_pointers(void* prep, int K, bf16** big, bf16** bigT, float** rowsum) {
  const AggPrep a = carve_agg_prep(prep, K);
  *big = a.big; *bigT = a.bigT; *rowsum = a.rowsum;
}

int32_t launch_build_big_adjacency(const float* A, int K, int V, void* prep_mem, cudaStream_t st) {
  KernelTimer kt_(KC_PACK, 0.0, st);
  AggPrep a = carve_agg_prep(prep_mem, K);
  launch_k(build_big_adjacency_kernel, cdiv(K * 128 * 128, 256), 256, 0, st, A, K, V, 128 / V, a.big, a.bigT, a.rowsum);
  STGCN_CHECK_LAUNCH("build_big_adjacency");
  return STGCN_OK;
}

bool umma_agg_supported(int V, int C, int K) { return C % 64 == 0 && C <= 1024 && V >= 4 && V <= 32 && K >= 1 && K <= 3; }

static int num_sms_agg() {
  static int n = 0;
  if (!n) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev); if (n <= 0) n = 148; }
  return n;
}

constexpr int kAggU2Threads = 320;    // agg_umma_kernel: TMA warp, MMA warp, two epilogue groups of four warps
struct AggParams {
  int nb, t, v, ft, c, k, nch;            // nch = channels per work item (64 or 128)
  int c_chunks, m_tiles_per_b, total_items, stages, rev;
  uint32_t tmem_cols, tile_bytes;
  int transpose;                           // 0: fwd (K outputs), 1: bwd (sum over k)
  const float* aff_a; const float* aff_b;  // fwd epilogue affine (may be null)
  const float* rowsum;                     // [K][V]
  const bf16* mask_h;                      // bwd: zero where h <= 0 (may be null), layout [rows][c]
  const bf16* bn_x; const float* bn_coef;  // bwd: fused BatchNorm-1 backward  dx = dh' - c1 - (x - mean)*c2 ; coef = [3][C] {mean, c1, c2}
  float* dA;                               // bwd: also accumulate dA[k][v][w] += sum dXA_k[(f,v),c] * x[(f,w),c]  (null = off)
};

// smem: bigA (K x 32 KB) | stages x (nch/64 x 16 KB) | staging 2 x 16 KB | barriers
// MODE bits (compile time, so unused epilogue options cost no instructions: the epilogue warps are dependent-issue bound):
//   1 transposed (backward)   2 dA products ride along   4 forward affine   8 ReLU mask   16 fused BatchNorm-1 backward
template <int MODE>
__global__ void __launch_bounds__(kAggU2Threads, 1)
agg_umma_kernel(const __grid_constant__ CUtensorMap map_in, const __grid_constant__ CUtensorMap map_big,
                const __grid_constant__ CUtensorMap map_out, const __grid_constant__ CUtensorMap map_x,
                const __grid_constant__ AggParams p) {
  constexpr bool TR = (MODE & 1) != 0, DA = (MODE & 2) != 0, AFF = (MODE & 4) != 0, MASKH = (MODE & 8) != 0, BNB = (MODE & 16) != 0;
  pdl_trigger();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int slabs = p.nch / 64;
  const int stage_bytes = slabs * kTile16K;
  uint8_t* smem_big = smem;                                          // [k][2 halves of 64 cols][128 rows x 128 B]
  uint8_t* smem_x = smem_big + (size_t)p.k * 2 * kTile16K;
  uint8_t* smem_xb = smem_x + (size_t)p.stages * stage_bytes;        // dA mode: 2 x stage_bytes ring of raw-input tiles
  uint8_t* staging = smem_xb + (DA ? 2 * (size_t)stage_bytes : 0);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + 2 * kTile16K);
  uint64_t* empty_bar = full_bar + p.stages;
  uint64_t* tfull_bar = empty_bar + p.stages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* big_bar = tempty_bar + 2;
  uint64_t* xfull_bar = big_bar + 1;       // [2]
  uint64_t* xempty_bar = xfull_bar + 2;    // [2]
  uint64_t* done_bar = xempty_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done_bar + 1);
  float* aff_tab = reinterpret_cast<float*>(full_bar + 32);          // [3][C] per-channel epilogue coefficients
  float* da_tab = aff_tab + 3 * p.c;                                 // dA mode: [K][V][V] CTA-local sums
  if (DA) for (int i = threadIdx.x; i < p.k * p.v * p.v; i += kAggU2Threads) da_tab[i] = 0.f;
  const int warp = uniform_warp_idx(), lane = threadIdx.x % 32;

  {  // padding rows of the activation tiles must be finite (they meet BigA's zero columns)
    uint4* z = reinterpret_cast<uint4*>(smem_x);
    const int n16 = (p.stages + (DA ? 2 : 0)) * stage_bytes / 16;
    for (int i = threadIdx.x; i < n16; i += kAggU2Threads) z[i] = make_uint4(0, 0, 0, 0);
  }
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&map_in); prefetch_tmap(&map_big); prefetch_tmap(&map_out); prefetch_tmap(&map_x);
    for (int s = 0; s < p.stages; ++s) { mbar_init(full_bar + s, 1); mbar_init(empty_bar + s, 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar + a, 1); mbar_init(tempty_bar + a, 256); mbar_init(xfull_bar + a, 1); mbar_init(xempty_bar + a, 129); }
    mbar_init(big_bar, 1); mbar_init(done_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, p.tmem_cols);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();                       // everything above (barriers, TMEM, descriptor prefetch) overlaps the previous kernel's tail
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);      // warp-uniform for the compiler: the MMA operands stay in uniform registers (no R2UR per tcgen05.mma)

  auto decode = [&](int item, int& cch, int& j0, int& b) {
    if (p.rev) item = p.total_items - 1 - item;
    cch = item % p.c_chunks; item /= p.c_chunks;
    j0 = (item % p.m_tiles_per_b) * p.ft; b = item / p.m_tiles_per_b;
  };
  // forward: 1 load, K accumulations (K outputs).  transpose: K loads, 1 output.
  const int loads_per_item = TR ? p.k : 1;
  const int outs_per_item = TR ? 1 : p.k;

  if (warp == 0) {
    {   // whole warp runs the loop, one elected lane arms barriers and issues copies
      mbar_expect_tx_e(big_bar, (uint32_t)(p.k * 2 * kTile16K));
      for (int k = 0; k < p.k; ++k)
        for (int h = 0; h < 2; ++h) tma_load_2d_e(&map_big, smem_big + (size_t)(k * 2 + h) * kTile16K, big_bar, h * 64, k * 128);
      int stage = 0; uint32_t phase = 0; int xs = 0; uint32_t xphase = 0;
      for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
        int cch, j0, b; decode(item, cch, j0, b);
        if (DA) {                                  // raw-input tile of the same rows / channels: B operand of the dA products
          mbar_wait(xempty_bar + xs, xphase ^ 1);
          mbar_expect_tx_e(xfull_bar + xs, (uint32_t)slabs * p.tile_bytes);
          for (int s = 0; s < slabs; ++s)
            tma_load_5d_e(&map_x, smem_xb + (size_t)xs * stage_bytes + s * kTile16K, xfull_bar + xs, cch * p.nch + s * 64, 0, 0, j0, b);
          if (++xs == 2) { xs = 0; xphase ^= 1; }
        }
        for (int l = 0; l < loads_per_item; ++l) {
          mbar_wait(empty_bar + stage, phase ^ 1);
          mbar_expect_tx_e(full_bar + stage, (uint32_t)slabs * p.tile_bytes);
          const int col0 = (TR ? l * p.c : 0) + cch * p.nch;
          for (int s = 0; s < slabs; ++s)
            tma_load_5d_e(&map_in, smem_x + (size_t)stage * stage_bytes + s * kTile16K, full_bar + stage, col0 + s * 64, 0, 0, j0, b);
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    {   // whole warp runs the issue loop, one elected lane issues (umma_common.cuh)
      const uint32_t idesc = make_idesc_bf16(128, p.nch, 0, 1);     // A K-major (BigA), B MN-major (activation tile)
      mbar_wait(big_bar, 0);
      const uint32_t idesc_da = make_idesc_bf16(128, 128, 0, 0);     // dA: both operands K-major (channels contiguous)
      const uint32_t g_tmem = tmem_base + (uint32_t)(2 * p.nch);     // K accumulators of 128 columns, live for the whole kernel
      const uint64_t big_d0 = make_smem_desc_sw128(smem_u32(smem_big), 16, 1024);
      const uint64_t x_d0 = make_smem_desc_sw128(smem_u32(smem_x), (uint32_t)kTile16K, 1024);      // MN-major view (aggregation)
      const uint64_t xk_d0 = make_smem_desc_sw128(smem_u32(smem_x), 16, 1024);                      // K-major view (dA products)
      const uint64_t xb_d0 = make_smem_desc_sw128(smem_u32(smem_xb), 16, 1024);
      int stage = 0; uint32_t phase = 0; int as = 0; uint32_t aphase = 0; int xs = 0; uint32_t xphase = 0; bool g_first = true;
      for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
        if (TR) { mbar_wait(tempty_bar + as, aphase ^ 1); tc_fence_after(); }
        if (DA) { mbar_wait(xfull_bar + xs, xphase); tc_fence_after(); }
        for (int k = 0; k < p.k; ++k) {
          if (!TR) { mbar_wait(tempty_bar + as, aphase ^ 1); tc_fence_after(); }
          if (TR || k == 0) { mbar_wait(full_bar + stage, phase); tc_fence_after(); }
          // shared-memory descriptors are affine in the address (field = addr >> 4): one base each, then constant offsets
          const uint32_t st16 = (uint32_t)(stage * (stage_bytes >> 4));
          const uint64_t ad0 = big_d0 + (uint64_t)(k * (2 * kTile16K >> 4));
          const uint64_t bd0 = x_d0 + (uint64_t)st16;
          const uint32_t d_tmem = tmem_base + (uint32_t)(as * p.nch);
          const uint32_t acc0 = (TR && k != 0) ? 1u : 0u;
#pragma unroll
          for (int kk = 0; kk < 8; ++kk)           // reduction over the 128 tile rows, 16 per instruction
            umma_bf16_ss_elect(d_tmem, ad0 + (uint64_t)((kk / 4) * (kTile16K >> 4) + (kk % 4) * 2), bd0 + (uint64_t)(kk * 128), idesc,
                         kk != 0 ? 1u : acc0);
          if (DA) {
            // G_k[(f,v)][(f',w)] += sum_c dXA_k[(f,v),c] * x[(f',w),c]: the SAME tile images read K-major; the FT diagonal
            // V x V blocks (f == f') are folded into dA at the end of the kernel
            const uint64_t ga0 = xk_d0 + (uint64_t)st16;
            const uint64_t gb0 = xb_d0 + (uint64_t)(xs * (stage_bytes >> 4));
            for (int s2 = 0; s2 < slabs; ++s2) {
#pragma unroll
              for (int kk = 0; kk < 4; ++kk)
                umma_bf16_ss_elect(g_tmem + (uint32_t)(k * 128), ga0 + (uint64_t)(s2 * (kTile16K >> 4) + kk * 2),
                             gb0 + (uint64_t)(s2 * (kTile16K >> 4) + kk * 2), idesc_da, (g_first && s2 == 0 && kk == 0) ? 0u : 1u);
            }
          }
          if (TR || k == p.k - 1) { umma_commit_elect(empty_bar + stage); if (++stage == p.stages) { stage = 0; phase ^= 1; } }
          if (!TR) { umma_commit_elect(tfull_bar + as); if (++as == 2) { as = 0; aphase ^= 1; } }
        }
        if (TR) { umma_commit_elect(tfull_bar + as); if (++as == 2) { as = 0; aphase ^= 1; } }
        if (DA) { umma_commit_elect(xempty_bar + xs); if (++xs == 2) { xs = 0; xphase ^= 1; } g_first = false; }
      }
      if (DA) umma_commit_elect(done_bar);
    }
  } else {
    // two epilogue groups (warps 2..5 and 6..9) take alternate 64-column chunks; each has its own staging buffer and barriers
    const int grp = (warp - 2) >> 2;
    const int q = warp % 4, row = q * 32 + lane, et = threadIdx.x - 64 - grp * 128;
    const int bar_a = 1 + 2 * grp, bar_b = 2 + 2 * grp;
    uint8_t* const stg = staging + grp * kTile16K;
    int gchunk = 0;                                // chunks of the CTA so far (both groups count them)
    const int fl = row / p.v, vv = row % p.v;
    const int c_out_total = TR ? p.c : p.k * p.c;
    int as = 0; uint32_t aphase = 0; int chunk_ctr = 0; int xslot = 0; uint32_t xph = 0;
    constexpr bool affine = AFF;
    constexpr bool bnb = BNB;
    if (affine) {
      for (int i = et; i < p.c; i += 128) { aff_tab[i] = __ldg(p.aff_a + i); aff_tab[p.c + i] = __ldg(p.aff_b + i); }   // both groups write the same values
      asm volatile("bar.sync %0, 128;" ::"r"(bar_a) : "memory");
    } else if (bnb) {
      // dx = dh' - c1 - (x - mean) * c2 = (dh' + d0) - c2 * x  with d0 = mean * c2 - c1: two coefficients per channel
      for (int i = et; i < p.c; i += 128) {
        const float mu = __ldg(p.bn_coef + i), c1 = __ldg(p.bn_coef + p.c + i), c2 = __ldg(p.bn_coef + 2 * p.c + i);
        aff_tab[i] = mu * c2 - c1; aff_tab[p.c + i] = -c2;
      }
      asm volatile("bar.sync %0, 128;" ::"r"(bar_a) : "memory");
    }
    for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
      int cch, j0, b; decode(item, cch, j0, b);
      const bool row_valid = row < p.ft * p.v && (j0 + fl) < p.t;
      const long long grow = ((long long)b * p.t + j0 + fl) * p.v + vv;
      const bool mine_item = ((gchunk & 1) == grp);   // dA mode: an item is one chunk, so one group owns its raw-input tile
      if (DA && mine_item) mbar_wait(xfull_bar + xslot, xph);      // acquire the TMA-written raw-input tile for generic reads
      for (int o = 0; o < outs_per_item; ++o) {
        mbar_wait(tfull_bar + as, aphase);
        tc_fence_after();
        const float rs = (AFF && row_valid) ? p.rowsum[o * p.v + vv] : 0.f;
        for (int ch = 0; ch < slabs; ++ch) {
          if (((gchunk++) & 1) != grp) continue;
          if (chunk_ctr >= 1) {
            if (et == 0) tma_store_wait_read<0>();       // this group's previous store (two chunks ago) has drained
            asm volatile("bar.sync %0, 128;" ::"r"(bar_a) : "memory");
          }
          ++chunk_ctr;
          uint32_t r2[2][32];                           // both 32-column halves in flight before the first use
          tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * p.nch + ch * 64), r2[0]);
          tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * p.nch + ch * 64 + 32), r2[1]);
          tmem_ld_wait();
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            const int col0 = ch * 64 + half * 32;
            const int c0 = cch * p.nch + col0;          // channel of the first value
            float f[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(r2[half][i]);
            if (affine) {
              const float4* ta = reinterpret_cast<const float4*>(aff_tab + c0);
              const float4* tb = reinterpret_cast<const float4*>(aff_tab + p.c + c0);
#pragma unroll
              for (int g = 0; g < 8; ++g) {
                const float4 a4 = ta[g], b4 = tb[g];
                f[g * 4 + 0] = fmaf(a4.x, f[g * 4 + 0], b4.x * rs); f[g * 4 + 1] = fmaf(a4.y, f[g * 4 + 1], b4.y * rs);
                f[g * 4 + 2] = fmaf(a4.z, f[g * 4 + 2], b4.z * rs); f[g * 4 + 3] = fmaf(a4.w, f[g * 4 + 3], b4.w * rs);
              }
            }
            if (MASKH && row_valid) {
              const uint4* hp = reinterpret_cast<const uint4*>(p.mask_h + (size_t)grow * p.c + c0);
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                uint4 u = __ldg(hp + g);
                const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  if (!(__low2float(h2[e]) > 0.f)) f[g * 8 + 2 * e] = 0.f;
                  if (!(__high2float(h2[e]) > 0.f)) f[g * 8 + 2 * e + 1] = 0.f;
                }
              }
            }
            if (bnb && row_valid) {
              // raw input of this row: from the shared-memory tile ring in dA mode (the swizzled image TMA delivered), else global
              const uint4* xp = reinterpret_cast<const uint4*>(p.bn_x + (size_t)grow * p.c + c0);
              const uint8_t* xsm = smem_xb + (size_t)xslot * stage_bytes + (size_t)ch * kTile16K + row * 128;
              const float4* d04 = reinterpret_cast<const float4*>(aff_tab + c0);
              const float4* n24 = reinterpret_cast<const float4*>(aff_tab + p.c + c0);
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                const uint4 u = DA ? *reinterpret_cast<const uint4*>(xsm + (((half * 4 + g) ^ (row & 7)) << 4)) : __ldg(xp + g);
                const __nv_bfloat162* x2 = reinterpret_cast<const __nv_bfloat162*>(&u);
                float xv[8];
#pragma unroll
                for (int e = 0; e < 4; ++e) { xv[2 * e] = __low2float(x2[e]); xv[2 * e + 1] = __high2float(x2[e]); }
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                  const float4 d0 = d04[g * 2 + hh], n2 = n24[g * 2 + hh];
                  float* ff = f + g * 8 + hh * 4; const float* xx = xv + hh * 4;
                  ff[0] = fmaf(n2.x, xx[0], ff[0] + d0.x); ff[1] = fmaf(n2.y, xx[1], ff[1] + d0.y);
                  ff[2] = fmaf(n2.z, xx[2], ff[2] + d0.z); ff[3] = fmaf(n2.w, xx[3], ff[3] + d0.w);
                }
              }
            }
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              uint4 u;
              __nv_bfloat162 h0 = __floats2bfloat162_rn(f[g * 8 + 0], f[g * 8 + 1]);
              __nv_bfloat162 h1 = __floats2bfloat162_rn(f[g * 8 + 2], f[g * 8 + 3]);
              __nv_bfloat162 h2 = __floats2bfloat162_rn(f[g * 8 + 4], f[g * 8 + 5]);
              __nv_bfloat162 h3 = __floats2bfloat162_rn(f[g * 8 + 6], f[g * 8 + 7]);
              u.x = *reinterpret_cast<uint32_t*>(&h0); u.y = *reinterpret_cast<uint32_t*>(&h1);
              u.z = *reinterpret_cast<uint32_t*>(&h2); u.w = *reinterpret_cast<uint32_t*>(&h3);
              const int j = half * 4 + g;
              *reinterpret_cast<uint4*>(stg + row * 128 + ((j ^ (row & 7)) << 4)) = u;
            }
          }
          fence_proxy_async_smem();
          asm volatile("bar.sync %0, 128;" ::"r"(bar_b) : "memory");
          if (et == 0) {
            const int oc = (TR ? 0 : o * p.c) + cch * p.nch + ch * 64;
            tma_store_5d(&map_out, stg, oc, 0, 0, j0, b);
            tma_store_commit();
          }
        }
        tc_fence_before();
        mbar_arrive(tempty_bar + as);
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
      if (DA) { if (mine_item) mbar_arrive(xempty_bar + xslot); if ((xslot ^= 1) == 0) xph ^= 1; }     // done reading the raw-input tile
      (void)c_out_total;
    }
    if (et == 0) tma_store_wait<0>();
    if (DA) asm volatile("bar.sync 7, 256;" ::: "memory");          // every epilogue thread is done with the raw-input ring
    if (DA && grp == 0) {                        // fold the FT diagonal V x V blocks of every G_k into dA
      mbar_wait(done_bar, 0);
      tc_fence_after();
      // every load has landed and every product has completed: the activation ring is free and holds the unfolded blocks,
      // fold_tab[k][f][v][w] = G_k[(f,v)][(f,w)] written with plain stores (each thread owns its row), then summed over f
      float* fold_tab = reinterpret_cast<float*>(smem_x);
      const int vv2 = p.v * p.v;
      for (int k = 0; k < p.k; ++k) {
        float* mine = fold_tab + ((size_t)(k * p.ft + fl) * p.v + vv) * p.v - fl * p.v;   // indexed by the tile column
#pragma unroll
        for (int chunk = 0; chunk < 4; ++chunk) {
          uint32_t r[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(2 * p.nch + k * 128 + chunk * 32), r);
          tmem_ld_wait();
          if (row < p.ft * p.v) {
            const int lo = fl * p.v - chunk * 32, hi = lo + p.v;     // this row's own frame block, in chunk-local columns
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (i >= lo && i < hi) mine[chunk * 32 + i] = __uint_as_float(r[i]);
          }
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      for (int i = et; i < p.k * vv2; i += 128) {                     // one global atomic per entry per CTA
        const int k = i / vv2, rem = i - k * vv2;
        float acc = 0.f;
        for (int f = 0; f < p.ft; ++f) acc += fold_tab[(size_t)(k * p.ft + f) * vv2 + rem];
        atomicAdd(p.dA + i, acc);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, p.tmem_cols); }
}

// ---------------------------------------------------------------------------
// adjacency gradient: G_k[v'][w'] += sum_c dXA_k[v'][c] * h[w'][c] over all tiles of the CTA
// (both operands K-major: channels are contiguous), accumulators stay in TMEM for the
// whole kernel; the FT diagonal V x V blocks are folded into dA with atomics at the end.
// ---------------------------------------------------------------------------
struct DaParams {
  int nb, t, v, ft, c, k, m_tiles_per_b, total_tiles, stages;
  uint32_t tmem_cols, tile_bytes;
  float* dA;
};

__global__ void __launch_bounds__(kAggUThreads, 1)
agg_dA_umma_kernel(const __grid_constant__ CUtensorMap map_dxa, const __grid_constant__ CUtensorMap map_h,
                   const __grid_constant__ DaParams p) {
  pdl_trigger();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int stage_bytes = (p.k + 1) * kTile16K;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)p.stages * stage_bytes);
  uint64_t* empty_bar = full_bar + p.stages;
  uint64_t* done_bar = empty_bar + p.stages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done_bar + 1);
  float* da_tab = reinterpret_cast<float*>(smem + (size_t)p.stages * stage_bytes + 256);   // [K][V][V] CTA-local sums
  const int warp = uniform_warp_idx(), lane = threadIdx.x % 32;
  for (int i = threadIdx.x; i < p.k * p.v * p.v; i += kAggUThreads) da_tab[i] = 0.f;
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&map_dxa); prefetch_tmap(&map_h);
    for (int s = 0; s < p.stages; ++s) { mbar_init(full_bar + s, 1); mbar_init(empty_bar + s, 1); }
    mbar_init(done_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();                       // everything above (barriers, TMEM, descriptor prefetch) overlaps the previous kernel's tail
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);      // warp-uniform for the compiler: the MMA operands stay in uniform registers (no R2UR per tcgen05.mma)
  const int cbs = p.c / 64;
  const bool have_work = (int)blockIdx.x < p.total_tiles;

  if (warp == 0) {
    {   // whole warp runs the loop, one elected lane arms barriers and issues copies
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        const int j0 = (tile % p.m_tiles_per_b) * p.ft, b = tile / p.m_tiles_per_b;
        for (int cb = 0; cb < cbs; ++cb) {
          mbar_wait(empty_bar + stage, phase ^ 1);
          mbar_expect_tx_e(full_bar + stage, (uint32_t)(p.k + 1) * p.tile_bytes);
          uint8_t* sb = smem + (size_t)stage * stage_bytes;
          for (int k = 0; k < p.k; ++k) tma_load_5d_e(&map_dxa, sb + k * kTile16K, full_bar + stage, k * p.c + cb * 64, 0, 0, j0, b);
          tma_load_5d_e(&map_h, sb + p.k * kTile16K, full_bar + stage, cb * 64, 0, 0, j0, b);
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    {   // whole warp runs the issue loop, one elected lane issues (umma_common.cuh)
      const uint32_t idesc = make_idesc_bf16(128, 128, 0, 0);
      int stage = 0; uint32_t phase = 0; bool first = true;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        for (int cb = 0; cb < cbs; ++cb) {
          mbar_wait(full_bar + stage, phase);
          tc_fence_after();
          const uint32_t sb = smem_u32(smem + (size_t)stage * stage_bytes);
          const uint64_t hd = make_smem_desc_sw128(sb + (uint32_t)(p.k * kTile16K), 16, 1024);
          for (int k = 0; k < p.k; ++k) {
            const uint64_t ad = make_smem_desc_sw128(sb + (uint32_t)(k * kTile16K), 16, 1024);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              umma_bf16_ss_elect(tmem_base + (uint32_t)(k * 128), ad + (uint64_t)(kk * 2), hd + (uint64_t)(kk * 2), idesc, (!first || kk) ? 1u : 0u);
          }
          first = false;
          umma_commit_elect(empty_bar + stage);
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
      umma_commit_elect(done_bar);
    }
  } else {
    const int q = warp % 4, row = q * 32 + lane;
    const int fl = row / p.v, vv = row % p.v;
    mbar_wait(done_bar, 0);
    tc_fence_after();
    for (int k = 0; k < p.k; ++k) {
#pragma unroll
      for (int chunk = 0; chunk < 4; ++chunk) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(k * 128 + chunk * 32), r);
        tmem_ld_wait();
        if (have_work && row < p.ft * p.v) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int w = chunk * 32 + i - fl * p.v;          // column inside this row's own frame block
            if (w >= 0 && w < p.v) atomicAdd(da_tab + (k * p.v + vv) * p.v + w, __uint_as_float(r[i]));   // FT-way smem contention only
          }
        }
      }
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (have_work)
      for (int i = threadIdx.x - 64; i < p.k * p.v * p.v; i += 128) atomicAdd(p.dA + i, da_tab[i]);   // one global atomic per entry per CTA
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, p.tmem_cols); }
}

// ---------------------------------------------------------------------------
// host
// ---------------------------------------------------------------------------
static int32_t make_act_map(CUtensorMap* m, const bf16* base, int cols, int v, int t, int nb, int ft) {
  uint64_t dims[5] = {(uint64_t)cols, (uint64_t)v, 1, (uint64_t)t, (uint64_t)nb};
  uint64_t str[4] = {(uint64_t)cols * 2, (uint64_t)v * cols * 2, (uint64_t)v * cols * 2, (uint64_t)t * v * cols * 2};
  uint32_t box[5] = {64, (uint32_t)v, 1, (uint32_t)ft, 1};
  return make_tmap_bf16(m, base, 5, dims, str, box);
}

const bf16* umma_agg_big(const void* prep, int k) { return carve_agg_prep(const_cast<void*>(prep), k).big; }
const float* umma_agg_rowsum_q(const void* prep, int k) { return carve_agg_prep(const_cast<void*>(prep), k).rowsum; }

static int32_t launch_agg_common(int transpose, const bf16* in, int nb, int t, int v, int c, int k, const void* prep_mem,
                                 const float* aff_a, const float* aff_b, const bf16* mask_h, bf16* out, cudaStream_t st,
                                 const bf16* bn_x = nullptr, const float* bn_coef = nullptr, float* dA = nullptr) {
  if (!umma_agg_supported(v, c, k)) { set_error("umma agg: unsupported shape V=%d C=%d K=%d", v, c, k); return STGCN_ERR_UNSUPPORTED; }
  AggPrep prep = carve_agg_prep(const_cast<void*>(prep_mem), k);
  AggParams p; memset(&p, 0, sizeof(p));
  p.rev = next_direction();
  p.nb = nb; p.t = t; p.v = v; p.ft = 128 / v; p.c = c; p.k = k; p.transpose = transpose;
  p.nch = (c % 128 == 0 && !dA) ? 128 : 64;          // dA mode keeps a second tile ring in shared memory: 64-channel items
  p.dA = dA;
  p.c_chunks = c / p.nch; p.m_tiles_per_b = cdiv(t, p.ft); p.total_items = p.c_chunks * p.m_tiles_per_b * nb;
  p.tile_bytes = (uint32_t)(p.ft * v * 128);
  p.aff_a = aff_a; p.aff_b = aff_b; p.rowsum = prep.rowsum; p.mask_h = mask_h; p.bn_x = bn_x; p.bn_coef = bn_coef;
  const size_t stage_bytes = (size_t)(p.nch / 64) * kTile16K;
  const size_t fixed = (size_t)k * 2 * kTile16K + 2 * kTile16K + 1024 + 512 + 3 * (size_t)c * sizeof(float) +
                       (dA ? 2 * (size_t)(p.nch / 64) * kTile16K + (size_t)k * v * v * sizeof(float) : 0);
  int stages = (int)((220 * 1024 - fixed) / stage_bytes);
  if (stages > 4) stages = 4;
  if (stages < 2) { set_error("umma agg: not enough shared memory"); return STGCN_ERR_UNSUPPORTED; }
  p.stages = stages;
  uint32_t cols = 32; while (cols < (uint32_t)(2 * p.nch + (dA ? k * 128 : 0))) cols <<= 1;
  p.tmem_cols = cols;
  CUtensorMap map_in, map_big, map_out, map_x;
  STGCN_TRY(make_act_map(&map_in, in, transpose ? k * c : c, v, t, nb, p.ft));
  STGCN_TRY(make_act_map(&map_out, out, transpose ? c : k * c, v, t, nb, p.ft));
  if (dA) { STGCN_TRY(make_act_map(&map_x, bn_x, c, v, t, nb, p.ft)); } else map_x = map_in;
  {
    uint64_t dims[2] = {128, (uint64_t)k * 128}; uint64_t str[1] = {256}; uint32_t box[2] = {64, 128};
    STGCN_TRY(make_tmap_bf16(&map_big, transpose ? prep.bigT : prep.big, 2, dims, str, box));
  }
  const size_t smem = fixed + (size_t)stages * stage_bytes;
  const int mode = (transpose ? 1 : 0) | (dA ? 2 : 0) | ((!transpose && aff_a) ? 4 : 0) | ((transpose && mask_h) ? 8 : 0) |
                   ((transpose && bn_coef) ? 16 : 0);
  const int grid = std::min(p.total_items, num_sms_agg());
#define STGCN_AGG_CASE(M)                                                                                                  \
  case M: {                                                                                                                \
    static bool attr_set = false;                                                                                          \
    if (!attr_set) {                                                                                                       \
      cudaError_t e = cudaFuncSetAttribute(agg_umma_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);   \
      if (e != cudaSuccess) { set_error("cudaFuncSetAttribute failed: %s", cudaGetErrorString(e)); return STGCN_ERR_CUDA; } \
      attr_set = true;                                                                                                     \
    }                                                                                                                      \
    launch_k(agg_umma_kernel<M>, grid, kAggU2Threads, smem, st, map_in, map_big, map_out, map_x, p);                              \
  } break;
  switch (mode) {
    STGCN_AGG_CASE(0) STGCN_AGG_CASE(4) STGCN_AGG_CASE(1) STGCN_AGG_CASE(9) STGCN_AGG_CASE(17) STGCN_AGG_CASE(19)
    default: set_error("umma agg: epilogue combination %d not instantiated", mode); return STGCN_ERR_UNSUPPORTED;
  }
#undef STGCN_AGG_CASE
  STGCN_CHECK_LAUNCH("agg_umma");
  return STGCN_OK;
}

int32_t launch_agg_fwd_umma(const bf16* x, int nb, int t, int v, int c, int k, const void* prep, const float* aff_a,
                            const float* aff_b, bf16* xa, cudaStream_t st) {
  KernelTimer kt_(KC_AGG_FWD, 2.0 * nb * t * v * v * (double)k * c, st);
  return launch_agg_common(0, x, nb, t, v, c, k, prep, aff_a, aff_b, nullptr, xa, st);
}
int32_t launch_agg_bwd_umma(const bf16* dxa, int nb, int t, int v, int c, int k, const void* prep, const bf16* mask_h, bf16* dh,
                            cudaStream_t st, const bf16* bn_x, const float* bn_coef, float* dA) {
  KernelTimer kt_(KC_AGG_BWD, 2.0 * nb * t * v * v * (double)k * c * (dA ? 2 : 1), st);
  if (dA && !bn_x) { set_error("umma agg bwd: dA accumulation needs the raw input"); return STGCN_ERR_INVALID; }
  return launch_agg_common(1, dxa, nb, t, v, c, k, prep, nullptr, nullptr, mask_h, dh, st, bn_x, bn_coef, dA);
}

int32_t launch_agg_dA_umma(const bf16* dxa, const bf16* h, int nb, int t, int v, int c, int k, float* dA, cudaStream_t st) {
  KernelTimer kt_(KC_AGG_DA, 2.0 * nb * t * v * v * (double)k * c, st);
  if (!umma_agg_supported(v, c, k)) { set_error("umma dA: unsupported shape"); return STGCN_ERR_UNSUPPORTED; }
  DaParams p; memset(&p, 0, sizeof(p));
  p.nb = nb; p.t = t; p.v = v; p.ft = 128 / v; p.c = c; p.k = k;
  p.m_tiles_per_b = cdiv(t, p.ft); p.total_tiles = p.m_tiles_per_b * nb;
  p.tile_bytes = (uint32_t)(p.ft * v * 128);
  const size_t stage_bytes = (size_t)(k + 1) * kTile16K;
  const size_t tail_bytes = 2048 + (size_t)k * v * v * sizeof(float);      // align slack + barriers + CTA-local dA table
  int stages = (int)((220 * 1024 - tail_bytes) / stage_bytes);
  if (stages > 4) stages = 4;
  if (stages < 2) { set_error("umma dA: not enough shared memory"); return STGCN_ERR_UNSUPPORTED; }
  p.stages = stages;
  uint32_t cols = 32; while (cols < (uint32_t)(k * 128)) cols <<= 1;
  p.tmem_cols = cols; p.dA = dA;
  CUtensorMap map_dxa, map_h;
  STGCN_TRY(make_act_map(&map_dxa, dxa, k * c, v, t, nb, p.ft));
  STGCN_TRY(make_act_map(&map_h, h, c, v, t, nb, p.ft));
  const size_t smem = (size_t)stages * stage_bytes + tail_bytes;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(agg_dA_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) { set_error("cudaFuncSetAttribute failed: %s", cudaGetErrorString(e)); return STGCN_ERR_CUDA; }
    attr_set = true;
  }
  launch_k(agg_dA_umma_kernel, std::min(p.total_tiles, num_sms_agg()), kAggUThreads, smem, st, map_dxa, map_h, p);
  STGCN_CHECK_LAUNCH("agg_dA_umma");
  return STGCN_OK;
}

}  // namespace stgcn
