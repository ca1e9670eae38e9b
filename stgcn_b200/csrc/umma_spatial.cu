// Fused spatial graph convolution, forward (tensor-core default block, BatchNorm-1 folded into the weights):
//
//     S[r, co] = sum_k sum_ci ( sum_w A[k,v,w] x[(f,w), ci] ) * W2'[co, k*Ci+ci]  +  bias2d'[v, co]
//
// Reference: SpatialConv.forward, st_graphconv.py:139-152 (1x1 conv :148, einsum :150) behind batch_n (:98).
// One persistent CTA per SM walks tiles of FT whole frames (FT*V <= 128 rows).  Per tile and per
// (64-channel block cc, partition k):
//   MMA-1  T[128 x 64]   = BigA_k[128 x 128] . x_cc[128 x 64]        (block-diagonal A_k resident in smem, x tile via TMA)
//   epi-1  T (TMEM, fp32) -> bf16 -> P (smem, 128B-swizzled K-major tile)      [optionally TMA-stored as XA for backward]
//   MMA-2  S[128 x Co]  += P[128 x 64] . W2'[Co x 64 (k, cc)]^T       (weight chunk via TMA, accumulator stays in TMEM)
// and once per tile  S -> + bias2d' -> bf16 -> TMA store.  The N x K x C x T x V intermediate of the reference (and
// the K*C_in-wide aggregated tensor of the unfused chain) never goes through HBM in inference.
// Warp roles: 0 = TMA producer, 1 = MMA issuer (both products, software-pipelined), 2..5 / 6..9 = two epilogue groups.
#include <stdlib.h>

#include "kernels.cuh"
#include "umma_common.cuh"

namespace stgcn {
using namespace umma;

constexpr int kSpThreads = 320;          // TMA warp, MMA warp, 2 x 4 epilogue warps  (SUB = 1)
// SUB = 2: every 64-column chunk is split between two sets of four warps (32 columns each), 2 x 8 epilogue warps.  The kernel is
// bound by the LATENCY of the TMEM -> bf16 -> shared-memory conversion between its two products, which a
// thread then does for 32 instead of 64 values.
constexpr int sp_threads(int sub) { return 64 + 2 * 128 * sub; }
constexpr int kSpDefaultSub = 1;
constexpr int kSpTile = 128 * 128;          // one [128 rows][64 bf16] tile image

struct SpParams {
  int nb, t, v, ft, ci, co, k;
  int m_tiles_per_b, total_tiles, rev;
  int cblocks;                 // ci / 64
  int wn, wparts;              // weight stage rows (<=128) and stages per (k, cc) chunk (co / wn)
  int xs, wst;                 // ring depths
  uint32_t tmem_cols, tile_bytes;
  int write_xa;
  int ns;                      // output accumulators in TMEM (2 when 128 + 2*Co <= 512 columns)
  const float* b2d;            // [V][Co] joint bias table bias2d' (bias2d_fold_kernel)
  const bf16* big_g;           // block-diagonal adjacency in global memory (tensor-memory-resident variant)
  int wres;                    // all weight chunks stay resident in shared memory (loaded once; wst = their count)
  bf16* xa_g;                  // tensor-memory variant, training: the aggregated rows are ALSO stored (nb,t,v,k*ci) for backward's weight gradient
  long long* trace;            // per-role segment timers of CTA 0 (builds with -DSTGCN_SP_TRACE only; null otherwise)
};

template <int SUB>
__global__ void __launch_bounds__(sp_threads(SUB), 1)
spatial_fused_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_big,
                     const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_xa,
                     const __grid_constant__ CUtensorMap map_out, const __grid_constant__ SpParams p) {
  pdl_trigger();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int w_stage_bytes = p.wn * 128;
  uint8_t* smem_big = smem;                                        // [k][2 halves][128 rows x 128 B]
  uint8_t* smem_x = smem_big + (size_t)p.k * 2 * kSpTile;          // xs x 16 KB
  uint8_t* smem_p = smem_x + (size_t)p.xs * kSpTile;               // 2 x 16 KB  (XA chunk; also the output staging)
  uint8_t* smem_w = smem_p + 2 * kSpTile;                          // wst x wn*128 B
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_w + (size_t)p.wst * w_stage_bytes);
  uint64_t* xfull = bars;            uint64_t* xempty = xfull + p.xs;
  uint64_t* wfull = xempty + p.xs;   uint64_t* wempty = wfull + p.wst;
  uint64_t* tfull = wempty + p.wst;  uint64_t* tempty = tfull + 2;
  uint64_t* pfull = tempty + 2;      uint64_t* pempty = pfull + 2;
  uint64_t* sfull = pempty + 2;      uint64_t* sempty = sfull + 2;
  uint64_t* big_bar = sempty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(big_bar + 1);
  float* b2_tab = reinterpret_cast<float*>(bars + 64);             // [V][Co+4] joint bias table (512 B after the barrier block start)
  const int warp = uniform_warp_idx(), lane = threadIdx.x % 32;

  {  // padding rows of the activation tiles (never written by TMA) must be finite: they meet BigA's zero columns
    uint4* z = reinterpret_cast<uint4*>(smem_x);
    const int n16 = (p.xs + 2) * kSpTile / 16;
    for (int i = threadIdx.x; i < n16; i += sp_threads(SUB)) z[i] = make_uint4(0, 0, 0, 0);
  }
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&map_x); prefetch_tmap(&map_big); prefetch_tmap(&map_w); prefetch_tmap(&map_xa); prefetch_tmap(&map_out);
    for (int s = 0; s < p.xs; ++s) { mbar_init(xfull + s, 1); mbar_init(xempty + s, 1); }
    for (int s = 0; s < p.wst; ++s) { mbar_init(wfull + s, 1); mbar_init(wempty + s, 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(tfull + a, 1); mbar_init(tempty + a, 128 * SUB); mbar_init(pfull + a, 1); mbar_init(pempty + a, 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(sfull + a, 1); mbar_init(sempty + a, 256 * SUB); }
    mbar_init(big_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, p.tmem_cols);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();                       // everything above (barriers, TMEM, descriptor prefetch) overlaps the previous kernel's tail
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);      // warp-uniform for the compiler: the MMA operands stay in uniform registers (no R2UR per tcgen05.mma)
  const uint32_t t_tmem = tmem_base;                 // 2 x 64 columns: aggregation accumulators
  const uint32_t s_tmem = tmem_base + 128;           // Co columns: the output tile
  const int chunks = p.cblocks * p.k;                // (cc, k) pairs per tile, cc outer

  if (warp == 0) {
    // ===================== TMA producer =====================
    {   // whole warp runs the loop, one elected lane arms barriers and issues copies
      mbar_expect_tx_e(big_bar, (uint32_t)(p.k * 2 * kSpTile));
      for (int k = 0; k < p.k; ++k)
        for (int h = 0; h < 2; ++h) tma_load_2d_e(&map_big, smem_big + (size_t)(k * 2 + h) * kSpTile, big_bar, h * 64, k * 128);
      int xs = 0; uint32_t xph = 0; int ws = 0; uint32_t wph = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        const int tl = p.rev ? p.total_tiles - 1 - tile : tile;
      const int j0 = (tl % p.m_tiles_per_b) * p.ft, b = tl / p.m_tiles_per_b;
        for (int cc = 0; cc < p.cblocks; ++cc) {
          mbar_wait(xempty + xs, xph ^ 1);
          mbar_expect_tx_e(xfull + xs, p.tile_bytes);
          tma_load_5d_e(&map_x, smem_x + (size_t)xs * kSpTile, xfull + xs, cc * 64, 0, 0, j0, b);
          if (++xs == p.xs) { xs = 0; xph ^= 1; }
          for (int k = 0; k < p.k; ++k)
            for (int wp = 0; wp < p.wparts; ++wp) {
              mbar_wait(wempty + ws, wph ^ 1);
              mbar_expect_tx_e(wfull + ws, (uint32_t)w_stage_bytes);
              tma_load_2d_e(&map_w, smem_w + (size_t)ws * w_stage_bytes, wfull + ws, k * p.ci + cc * 64, wp * p.wn);
              if (++ws == p.wst) { ws = 0; wph ^= 1; }
            }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    {   // whole warp runs the issue loop, one elected lane issues (umma_common.cuh)
      const uint32_t idesc1 = make_idesc_bf16(128, 64, 0, 1);        // A = BigA (K-major), B = x tile (MN-major)
      const uint32_t idesc2 = make_idesc_bf16(128, p.wn, 0, 0);      // A = P (K-major), B = weights (K-major)
      mbar_wait(big_bar, 0);
      // This single thread is the issue bottleneck of the kernel, so everything below is 32-bit, division-free and works on
      // descriptors precomputed once: a shared-memory descriptor is affine in the address (field = addr >> 4).
      const uint64_t big_d0 = make_smem_desc_sw128(smem_u32(smem_big), 16, 1024);
      const uint64_t x_d0 = make_smem_desc_sw128(smem_u32(smem_x), (uint32_t)kSpTile, 1024);
      const uint64_t p_d0 = make_smem_desc_sw128(smem_u32(smem_p), 16, 1024);
      const uint64_t w_d0 = make_smem_desc_sw128(smem_u32(smem_w), 16, 1024);
      const uint32_t w_step = (uint32_t)(w_stage_bytes >> 4);
      int my_tiles = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) ++my_tiles;
      const int total = my_tiles * chunks;
      int xs = 0; uint32_t xph = 0; int ws = 0; uint32_t wph = 0;
      uint32_t sph = 0; int sb = 0;             // per-accumulator phase bits, current output accumulator
      int k1 = 0;                                // partition index of the next aggregation product
      auto issue1 = [&](int g) {                 // T[g & 1] = BigA_k . x_cc
        const int tb = g & 1;
        mbar_wait(tempty + tb, (uint32_t)(((g >> 1) & 1) ^ 1));
        if (k1 == 0) { mbar_wait(xfull + xs, xph); }
        tc_fence_after();
        const uint64_t ad0 = big_d0 + (uint64_t)(k1 * (2 * kSpTile >> 4));
        const uint64_t bd0 = x_d0 + (uint64_t)(xs * (kSpTile >> 4));
        const uint32_t d = t_tmem + (uint32_t)(tb * 64);
#pragma unroll
        for (int kk = 0; kk < 8; ++kk)
          umma_bf16_ss_elect(d, ad0 + (uint64_t)((kk / 4) * (kSpTile >> 4) + (kk % 4) * 2), bd0 + (uint64_t)(kk * 128), idesc1, kk != 0 ? 1u : 0u);
        umma_commit_elect(tfull + tb);
        if (++k1 == p.k) { k1 = 0; umma_commit_elect(xempty + xs); if (++xs == p.xs) { xs = 0; xph ^= 1; } }
      };
      if (total > 0) issue1(0);
      int c = 0;                                 // chunk index inside the tile of the next mixing product
      for (int g = 0; g < total; ++g) {
        if (g + 1 < total) issue1(g + 1);
        const int pb = g & 1;
        if (c == 0) { mbar_wait(sempty + sb, ((sph >> sb) & 1u) ^ 1u); }   // the tile that last used this accumulator has left TMEM
        mbar_wait(pfull + pb, (uint32_t)((g >> 1) & 1));
        tc_fence_after();
        const uint64_t ad = p_d0 + (uint64_t)(pb * (kSpTile >> 4));
        for (int wp = 0; wp < p.wparts; ++wp) {
          mbar_wait(wfull + ws, wph);
          tc_fence_after();
          const uint64_t bd = w_d0 + (uint64_t)(ws * w_step);
          const uint32_t d = s_tmem + (uint32_t)(sb * p.co + wp * p.wn);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_bf16_ss_elect(d, ad + (uint64_t)(kk * 2), bd + (uint64_t)(kk * 2), idesc2, (c | kk) != 0 ? 1u : 0u);
          umma_commit_elect(wempty + ws);
          if (++ws == p.wst) { ws = 0; wph ^= 1; }
        }
        umma_commit_elect(pempty + pb);
        if (++c == chunks) { c = 0; umma_commit_elect(sfull + sb); sph ^= (1u << sb); if (++sb == p.ns) sb = 0; }
      }
    }
  } else {
    // ===================== epilogue warps: group 0 = warps 2..5 (even chunks, T[0] / P[0]), group 1 = warps 6..9 (odd) ====
    // Two independent 128-thread pipelines, each owning one accumulator / operand buffer pair, so that the latencies of one
    // (TMEM load, proxy fence, barriers) hide behind the other.
    constexpr int GT = 128 * SUB;                    // threads per epilogue group
    const int grp = (warp - 2) / (4 * SUB);
    const int sub = ((warp - 2) / 4) % SUB;           // which 32-column half of a chunk this warp converts (SUB = 2)
    const int q = warp % 4, row = q * 32 + lane, et = threadIdx.x - 64 - grp * GT;
    const int vv = row % p.v;
    const int bar_a = 1 + grp * 2, bar_b = 2 + grp * 2;
    auto gsync = [&](int id) { asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(GT) : "memory"); };
    const int b2_ld = p.co + 4;                     // row pitch (floats): 16-byte rows, conflict-free across joints
    for (int i = et; i < p.v * p.co; i += GT) b2_tab[(i / p.co) * b2_ld + (i % p.co)] = __ldg(p.b2d + i);   // both groups write the same values
    gsync(bar_a);
    const float* const my_b2 = b2_tab + (row < p.ft * p.v ? vv : 0) * b2_ld;
    int g = 0;
    uint32_t sph = 0; int sb = 0; int och = 0;      // output-accumulator phase bits / index, output chunks so far
    uint8_t* const my_p = smem_p + (size_t)grp * kSpTile;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      const int tl = p.rev ? p.total_tiles - 1 - tile : tile;
      const int j0 = (tl % p.m_tiles_per_b) * p.ft, b = tl / p.m_tiles_per_b;
      int cc = 0, k = -1;
      for (int c = 0; c < chunks; ++c, ++g) {
        if (++k == p.k) { k = 0; ++cc; }
        if ((g & 1) != grp) continue;
        const uint32_t use = (uint32_t)((g >> 1) & 1);
        mbar_wait(tfull + grp, use);
        tc_fence_after();
        uint32_t r0[32], r1[32];
        if (SUB == 1) {
          tmem_ld_32x32(t_tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(grp * 64), r0);
          tmem_ld_32x32(t_tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(grp * 64 + 32), r1);
        } else {
          tmem_ld_32x32(t_tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(grp * 64 + sub * 32), r0);
        }
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(tempty + grp);                              // the aggregation accumulator may be overwritten
        // P[grp] is free once the mixing product of chunk g-2 retired and (training) its XA store finished reading
        if (p.write_xa) { if (et == 0) tma_store_wait_read<0>(); gsync(bar_a); }
        mbar_wait(pempty + grp, use ^ 1);
        uint8_t* pb = my_p + row * 128;
#pragma unroll
        for (int jj = 0; jj < 8 / SUB; ++jj) {
          const int j = SUB == 1 ? jj : sub * 4 + jj;          // 16-byte chunk of the row this thread writes
          const uint32_t* r = SUB == 1 ? (jj < 4 ? r0 + jj * 8 : r1 + (jj - 4) * 8) : r0 + jj * 8;
          uint4 u;
          __nv_bfloat162 h0 = __floats2bfloat162_rn(__uint_as_float(r[0]), __uint_as_float(r[1]));
          __nv_bfloat162 h1 = __floats2bfloat162_rn(__uint_as_float(r[2]), __uint_as_float(r[3]));
          __nv_bfloat162 h2 = __floats2bfloat162_rn(__uint_as_float(r[4]), __uint_as_float(r[5]));
          __nv_bfloat162 h3 = __floats2bfloat162_rn(__uint_as_float(r[6]), __uint_as_float(r[7]));
          u.x = *reinterpret_cast<uint32_t*>(&h0); u.y = *reinterpret_cast<uint32_t*>(&h1);
          u.z = *reinterpret_cast<uint32_t*>(&h2); u.w = *reinterpret_cast<uint32_t*>(&h3);
          *reinterpret_cast<uint4*>(pb + ((j ^ (row & 7)) << 4)) = u;
        }
        fence_proxy_async_smem();                               // generic writes -> visible to the tensor core / TMA (async proxy)
        gsync(bar_b);
        if (et == 0) {
          mbar_arrive(pfull + grp);
          if (p.write_xa) { tma_store_5d(&map_xa, my_p, k * p.ci + cc * 64, 0, 0, j0, b); tma_store_commit(); }
        }
      }
      // ---- output tile: S -> + bias2d' -> bf16 -> staging (this group's P buffer, idle now) -> TMA store; the two groups
      //      take alternate 64-column chunks ----
      mbar_wait(sfull + sb, (sph >> sb) & 1u); sph ^= (1u << sb);
      tc_fence_after();
      const uint32_t s_cur = s_tmem + (uint32_t)(sb * p.co);
      for (int ch = 0; ch < p.co / 64; ++ch) {
        if (((och + ch) & 1) != grp) continue;
        if (et == 0) tma_store_wait_read<0>();                  // XA store / previous output chunk done reading the buffer
        gsync(bar_a);
        uint32_t r2[2][32];                                     // both 32-column halves in flight before the first use
        if (SUB == 1) {
          tmem_ld_32x32(s_cur + ((uint32_t)(q * 32) << 16) + (uint32_t)(ch * 64), r2[0]);
          tmem_ld_32x32(s_cur + ((uint32_t)(q * 32) << 16) + (uint32_t)(ch * 64 + 32), r2[1]);
        } else {
          tmem_ld_32x32(s_cur + ((uint32_t)(q * 32) << 16) + (uint32_t)(ch * 64 + sub * 32), r2[0]);       // fixed slot: no dynamic register indexing
        }
        tmem_ld_wait();
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          if (SUB == 2 && half != sub) continue;
          const uint32_t (&rr)[32] = r2[SUB == 1 ? half : 0];
          const float4* b2 = reinterpret_cast<const float4*>(my_b2 + ch * 64 + half * 32);
          float f[32];
#pragma unroll
          for (int g4 = 0; g4 < 8; ++g4) {
            const float4 t4 = b2[g4];
            f[g4 * 4 + 0] = __uint_as_float(rr[g4 * 4 + 0]) + t4.x; f[g4 * 4 + 1] = __uint_as_float(rr[g4 * 4 + 1]) + t4.y;
            f[g4 * 4 + 2] = __uint_as_float(rr[g4 * 4 + 2]) + t4.z; f[g4 * 4 + 3] = __uint_as_float(rr[g4 * 4 + 3]) + t4.w;
          }
#pragma unroll
          for (int g4 = 0; g4 < 4; ++g4) {
            uint4 u;
            __nv_bfloat162 h0 = __floats2bfloat162_rn(f[g4 * 8 + 0], f[g4 * 8 + 1]);
            __nv_bfloat162 h1 = __floats2bfloat162_rn(f[g4 * 8 + 2], f[g4 * 8 + 3]);
            __nv_bfloat162 h2 = __floats2bfloat162_rn(f[g4 * 8 + 4], f[g4 * 8 + 5]);
            __nv_bfloat162 h3 = __floats2bfloat162_rn(f[g4 * 8 + 6], f[g4 * 8 + 7]);
            u.x = *reinterpret_cast<uint32_t*>(&h0); u.y = *reinterpret_cast<uint32_t*>(&h1);
            u.z = *reinterpret_cast<uint32_t*>(&h2); u.w = *reinterpret_cast<uint32_t*>(&h3);
            const int j = half * 4 + g4;
            *reinterpret_cast<uint4*>(my_p + row * 128 + ((j ^ (row & 7)) << 4)) = u;
          }
        }
        fence_proxy_async_smem();
        gsync(bar_b);
        if (et == 0) { tma_store_5d(&map_out, my_p, ch * 64, 0, 0, j0, b); tma_store_commit(); }
      }
      och += p.co / 64;
      tc_fence_before();
      mbar_arrive(sempty + sb); if (++sb == p.ns) sb = 0;                                       // the output accumulator may be overwritten
      // this group's staging store must have finished READING the buffer before its next chunk refills it
      if (et == 0) tma_store_wait_read<0>();
      gsync(bar_a);
    }
    if (et == 0) tma_store_wait<0>();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, p.tmem_cols); }
}

// ---------------------------------------------------------------------------
// Tensor-memory variant (the default for inference): the aggregated chunk never leaves tensor memory.
//
// The kernel above has ONE warp issuing both products and a shared-memory round trip between them.  Per-role cycle counters
// (-DSTGCN_SP_TRACE below) showed that warp busy ~900 cycles per chunk for 384 cycles of tensor work -- every barrier wait it
// passes costs ~50 cycles even when already satisfied -- while the conversion warps idled; and a stand-alone issue-rate
// benchmark (tools/mma_rate.cu, profiles/r2_mma_issue_rate.txt) gave the hardware floors: a 128x64x16 MMA with both operands in
// shared memory takes 48 cycles (6 KB of operand reads at 128 B/clk), with A in tensor memory 32.  Hence:
//   * the bf16 chunk is written back to TENSOR MEMORY (tcgen05.st, 32 columns of packed pairs) and the mixing product takes its A
//     operand from there (tcgen05.mma [d], [a_tmem], b_desc): no shared-memory tile, no proxy fence, no barrier among the
//     threads of a group -- each of the 128 threads arrives on the mbarrier by itself;
//   * TWO issuer warps, one per product, and two TMA producer warps (activation tiles / weight chunks), so no instruction stream
//     waits behind another one's barriers;
//   * G conversion groups (G accumulators T and G operands P in tensor memory) keep G chunks in flight; a separate output group
//     owns the S -> bias -> bf16 -> TMA-store leg with two staging buffers;
//   * BT: the block-diagonal adjacency itself (k x 64 packed columns) is the tensor-memory A operand of the aggregation product
//     when the columns allow (C_out = 64), which frees 96 KB of shared memory; all weight chunks then stay resident.
// Measured (N_eff = 128, one B200, tools/spatial_probe.py): 136 -> 69 us (64->64, T = 300), 145 -> 75 us (128->128, T = 150),
// 187 -> 133 us (256->256, T = 75); cfg3 inference 22.8 k -> 27.4 k sequences/s.  For training (XA wanted by the weight gradient)
// the rows can be stored straight from the conversion registers (STGCN_FUSED_SPATIAL=1), but that measured 6.20 against 6.03 ms
// per step for the two-kernel pair, which stays the default there.
// Warp roles: 0 / 3 = TMA producers (activation tiles / weight chunks), 1 / 2 = MMA issuers (aggregation / mixing products),
// 4..7 = output group, 8.. = G conversion groups of four warps.
// ---------------------------------------------------------------------------
constexpr int sp_ts_threads(int g) { return 128 + 128 + 128 * g; }
// Development aid (compile with -DSTGCN_SP_TRACE, run with STGCN_SPATIAL_TRACE=1, tools/spatial_trace.py): every role of CTA 0 sums the
// cycles it spends in each segment of its loop (waits and issue separately) in registers and the host prints them per chunk.  This is
// how the single-issuer bottleneck above was found; a per-event trace to global memory cost ~140 cycles per event and hid it.
#ifdef STGCN_SP_TRACE
__device__ __forceinline__ uint32_t sp_clk() { uint32_t c; asm volatile("mov.u32 %0, %%clock;" : "=r"(c)); return c; }
#define SPSEG(i) do { if (tr) { const uint32_t now_ = sp_clk(); acc[i] += now_ - last_; last_ = now_; } } while (0)
#define SPSTART() do { if (tr) last_ = sp_clk(); } while (0)
__device__ __forceinline__ long long sp_ns() { long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define SPNS(slot, cond) do { if (tr && (cond)) p.trace[slot] = sp_ns(); } while (0)
#define SPDUMP(cond, first, count, base) do { if (tr && (cond)) { for (int i_ = 0; i_ < (count); ++i_) p.trace[(base) + i_] = acc[(first) + i_]; } } while (0)
#else
#define SPSEG(i) do { } while (0)
#define SPSTART() do { } while (0)
#define SPNS(slot, cond) do { } while (0)
#define SPDUMP(cond, first, count, base) do { } while (0)
#endif

template <int G, bool BT>
__global__ void __launch_bounds__(sp_ts_threads(G), 1)
spatial_fused_ts_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_big,
                        const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_out,
                        const __grid_constant__ SpParams p) {
  pdl_trigger();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int w_stage_bytes = p.wn * 128;
  uint8_t* smem_big = smem;                                        // [k][2 halves][128 rows x 128 B]  (BT: not in shared memory)
  uint8_t* smem_x = smem_big + (BT ? 0 : (size_t)p.k * 2 * kSpTile);    // xs x 16 KB
  uint8_t* smem_o = smem_x + (size_t)p.xs * kSpTile;               // 2 x 16 KB output staging
  uint8_t* smem_w = smem_o + 2 * kSpTile;                          // wst x wn*128 B
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_w + (size_t)p.wst * w_stage_bytes);
  uint64_t* xfull = bars;            uint64_t* xempty = xfull + p.xs;
  uint64_t* wfull = xempty + p.xs;   uint64_t* wempty = wfull + p.wst;
  uint64_t* tfull = wempty + p.wst;  uint64_t* tempty = tfull + G;
  uint64_t* pfull = tempty + G;      uint64_t* pempty = pfull + G;
  uint64_t* sfull = pempty + G;      uint64_t* sempty = sfull + 2;
  uint64_t* big_bar = sempty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(big_bar + 1);
  float* b2_tab = reinterpret_cast<float*>(bars + 64);             // [V][Co+4] joint bias table
  const int warp = uniform_warp_idx(), lane = threadIdx.x % 32;
#ifdef STGCN_SP_TRACE
  const bool tr = p.trace != nullptr && blockIdx.x == 0; uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, last_ = 0;
#endif
  SPNS(10, threadIdx.x == 0);

  {  // padding rows of the activation tiles (never written by TMA) must be finite: they meet BigA's zero columns
    uint4* z = reinterpret_cast<uint4*>(smem_x);
    const int n16 = p.xs * kSpTile / 16;
    for (int i = threadIdx.x; i < n16; i += sp_ts_threads(G)) z[i] = make_uint4(0, 0, 0, 0);
  }
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&map_x); prefetch_tmap(&map_big); prefetch_tmap(&map_w); prefetch_tmap(&map_out);
    for (int s = 0; s < p.xs; ++s) { mbar_init(xfull + s, 1); mbar_init(xempty + s, 1); }
    for (int s = 0; s < p.wst; ++s) { mbar_init(wfull + s, 1); mbar_init(wempty + s, 1); }
    for (int a = 0; a < G; ++a) { mbar_init(tfull + a, 1); mbar_init(tempty + a, 128); mbar_init(pfull + a, 128); mbar_init(pempty + a, 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(sfull + a, 1); mbar_init(sempty + a, 128); }
    mbar_init(big_bar, BT ? 128 : 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, p.tmem_cols);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  SPNS(11, threadIdx.x == 0);
  pdl_wait();
  SPNS(12, threadIdx.x == 0);
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);      // warp-uniform for the compiler: the MMA operands stay in uniform registers (no R2UR per tcgen05.mma)
  const uint32_t t_tmem = tmem_base;                        // G x 64 columns: aggregation accumulators (fp32)
  const uint32_t p_tmem = tmem_base + 64 * G;               // G x 32 columns: the same chunks as packed bf16 pairs (A operand of MMA-2)
  const uint32_t s_tmem = tmem_base + 96 * G;               // ns x Co columns: output tiles
  const uint32_t big_tmem = s_tmem + (uint32_t)(p.ns * p.co);   // BT: k x 64 columns, the block-diagonal adjacency as packed bf16 pairs
  const int chunks = p.cblocks * p.k;                       // (cc, k) pairs per tile, cc outer
  int my_tiles = 0;
  for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) ++my_tiles;
  const int total = my_tiles * chunks;

  if (warp == 0) {
    // ===================== TMA producer: activation tiles (and the adjacency when it lives in shared memory) =====================
    if (!BT) {
      mbar_expect_tx_e(big_bar, (uint32_t)(p.k * 2 * kSpTile));
      for (int k = 0; k < p.k; ++k)
        for (int h = 0; h < 2; ++h) tma_load_2d_e(&map_big, smem_big + (size_t)(k * 2 + h) * kSpTile, big_bar, h * 64, k * 128);
    }
    int xs = 0; uint32_t xph = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      const int tl = p.rev ? p.total_tiles - 1 - tile : tile;
      const int j0 = (tl % p.m_tiles_per_b) * p.ft, b = tl / p.m_tiles_per_b;
      for (int cc = 0; cc < p.cblocks; ++cc) {
        mbar_wait(xempty + xs, xph ^ 1);
        mbar_expect_tx_e(xfull + xs, p.tile_bytes);
        tma_load_5d_e(&map_x, smem_x + (size_t)xs * kSpTile, xfull + xs, cc * 64, 0, 0, j0, b);
        if (++xs == p.xs) { xs = 0; xph ^= 1; }
      }
    }
  } else if (warp == 3) {
    // ===================== TMA producer: weight chunks (its own warp: never behind a full activation ring) =====================
    if (p.wres) {        // every (cc, k, part) chunk once, in the order issuer B walks them; wfull[0] carries all the bytes
      mbar_expect_tx_e(wfull, (uint32_t)(p.wst * w_stage_bytes));
      int slot = 0;
      for (int cc = 0; cc < p.cblocks; ++cc)
        for (int k = 0; k < p.k; ++k)
          for (int wp = 0; wp < p.wparts; ++wp, ++slot)
            tma_load_2d_e(&map_w, smem_w + (size_t)slot * w_stage_bytes, wfull, k * p.ci + cc * 64, wp * p.wn);
    } else {
      int ws = 0; uint32_t wph = 0;
      for (int t = 0; t < my_tiles; ++t)
        for (int cc = 0; cc < p.cblocks; ++cc)
          for (int k = 0; k < p.k; ++k)
            for (int wp = 0; wp < p.wparts; ++wp) {
              mbar_wait(wempty + ws, wph ^ 1);
              mbar_expect_tx_e(wfull + ws, (uint32_t)w_stage_bytes);
              tma_load_2d_e(&map_w, smem_w + (size_t)ws * w_stage_bytes, wfull + ws, k * p.ci + cc * 64, wp * p.wn);
              if (++ws == p.wst) { ws = 0; wph ^= 1; }
            }
    }
  } else if (warp == 1) {
    // ===================== issuer A: aggregation products T[tb] = BigA_k . x_cc (whole warp runs the loop, one elected lane issues) ====
    // Measured with ONE issuing warp for both products: every wait it passes costs ~50 cycles even when satisfied and the thread is
    // busy ~900 cycles per chunk for 384 cycles of tensor work; two warps halve that and decouple the two instruction streams.
    const uint32_t idesc1 = make_idesc_bf16(128, 64, 0, 1);        // A = BigA (K-major), B = x tile (MN-major)
    mbar_wait(big_bar, 0);
    const uint64_t big_d0 = make_smem_desc_sw128(smem_u32(smem_big), 16, 1024);
    const uint64_t x_d0 = make_smem_desc_sw128(smem_u32(smem_x), (uint32_t)kSpTile, 1024);
    int xs = 0; uint32_t xph = 0;
    int k1 = 0, tb = 0; uint32_t tpar = 0;                    // next partition, T buffer / round parity
    SPSTART();
    for (int g = 0; g < total; ++g) {
      mbar_wait(tempty + tb, tpar ^ 1);
      SPSEG(0);
      if (k1 == 0) { mbar_wait(xfull + xs, xph); }
      SPSEG(1);
      tc_fence_after();
      const uint64_t ad0 = big_d0 + (uint64_t)(k1 * (2 * kSpTile >> 4));
      const uint64_t bd0 = x_d0 + (uint64_t)(xs * (kSpTile >> 4));
      const uint32_t d = t_tmem + (uint32_t)(tb * 64);
      const uint32_t at0 = big_tmem + (uint32_t)(k1 * 64);
#pragma unroll
      for (int kk = 0; kk < 8; ++kk) {
        if (BT) umma_bf16_ts_elect(d, at0 + (uint32_t)(kk * 8), bd0 + (uint64_t)(kk * 128), idesc1, kk != 0 ? 1u : 0u);
        else umma_bf16_ss_elect(d, ad0 + (uint64_t)((kk / 4) * (kSpTile >> 4) + (kk % 4) * 2), bd0 + (uint64_t)(kk * 128), idesc1, kk != 0 ? 1u : 0u);
      }
      umma_commit_elect(tfull + tb);
      if (++k1 == p.k) { k1 = 0; umma_commit_elect(xempty + xs); if (++xs == p.xs) { xs = 0; xph ^= 1; } }
      if (++tb == G) { tb = 0; tpar ^= 1; }
      SPSEG(2);
    }
    SPDUMP(lane == 0, 0, 3, 0);
    SPNS(13, lane == 0);
  } else if (warp == 2) {
    // ===================== issuer B: mixing products S[sb] += P[pb] . W2'(k, cc) =====================
    const uint32_t idesc2 = make_idesc_bf16(128, p.wn, 0, 0);      // A = P (tensor memory), B = weights (K-major)
    const uint64_t w_d0 = make_smem_desc_sw128(smem_u32(smem_w), 16, 1024);
    const uint32_t w_step = (uint32_t)(w_stage_bytes >> 4);
    int ws = 0; uint32_t wph = 0;
    uint32_t sph = 0; int sb = 0;
    int c = 0, pb = 0; uint32_t ppar = 0;                     // chunk inside the tile, P buffer / round parity
    if (p.wres) { mbar_wait(wfull, 0); tc_fence_after(); }
    SPSTART();
    for (int g = 0; g < total; ++g) {
      if (c == 0) { mbar_wait(sempty + sb, ((sph >> sb) & 1u) ^ 1u); }   // the tile that last used this accumulator has left TMEM
      SPSEG(3);
      mbar_wait(pfull + pb, ppar);
      SPSEG(4);
      tc_fence_after();
      const uint32_t a_t = p_tmem + (uint32_t)(pb * 32);
      for (int wp = 0; wp < p.wparts; ++wp) {
        SPSEG(6);
        if (!p.wres) { mbar_wait(wfull + ws, wph); tc_fence_after(); }
        SPSEG(5);
        const uint64_t bd = w_d0 + (uint64_t)(ws * w_step);
        const uint32_t d = s_tmem + (uint32_t)(sb * p.co + wp * p.wn);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          umma_bf16_ts_elect(d, a_t + (uint32_t)(kk * 8), bd + (uint64_t)(kk * 2), idesc2, (c | kk) != 0 ? 1u : 0u);
        if (!p.wres) umma_commit_elect(wempty + ws);
        if (++ws == p.wst) { ws = 0; wph ^= 1; }
      }
      umma_commit_elect(pempty + pb);
      if (++pb == G) { pb = 0; ppar ^= 1; }
      if (++c == chunks) { c = 0; umma_commit_elect(sfull + sb); sph ^= (1u << sb); if (++sb == p.ns) sb = 0; }
      SPSEG(6);
    }
    SPDUMP(lane == 0, 3, 4, 3);
  } else if (warp < 8) {
    // ===================== output group: S -> + bias2d' -> bf16 -> staging -> TMA store =====================
    const int q = warp % 4, row = q * 32 + lane, et = threadIdx.x - 128;
    const int vv = row % p.v;
    auto gsync = [&](int id) { asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(128) : "memory"); };
    const int b2_ld = p.co + 4;
    for (int i = et; i < p.v * p.co; i += 128) b2_tab[(i / p.co) * b2_ld + (i % p.co)] = __ldg(p.b2d + i);
    gsync(1);
    const float* const my_b2 = b2_tab + (row < p.ft * p.v ? vv : 0) * b2_ld;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    uint32_t sph = 0; int sb = 0; int stage = 0;
    const int och = p.co / 64;
    SPSTART();
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      const int tl = p.rev ? p.total_tiles - 1 - tile : tile;
      const int j0 = (tl % p.m_tiles_per_b) * p.ft, b = tl / p.m_tiles_per_b;
      SPSEG(1);
      mbar_wait(sfull + sb, (sph >> sb) & 1u); sph ^= (1u << sb);
      SPSEG(0);
      tc_fence_after();
      const uint32_t s_cur = s_tmem + (uint32_t)(sb * p.co);
      for (int ch = 0; ch < och; ++ch) {
        uint32_t r2[2][32];
        tmem_ld_32x32(s_cur + lane_off + (uint32_t)(ch * 64), r2[0]);
        tmem_ld_32x32(s_cur + lane_off + (uint32_t)(ch * 64 + 32), r2[1]);
        if (et == 0) tma_store_wait_read<1>();                  // the store that last read this staging buffer (two chunks ago) is done with it
        tmem_ld_wait();
        if (ch == och - 1) { tc_fence_before(); mbar_arrive(sempty + sb); }     // the accumulator may be overwritten
        SPSEG(2);
        gsync(1);
        SPSEG(3);
        uint8_t* const my_o = smem_o + (size_t)stage * kSpTile;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const float4* b2 = reinterpret_cast<const float4*>(my_b2 + ch * 64 + half * 32);
          float f[32];
#pragma unroll
          for (int g4 = 0; g4 < 8; ++g4) {
            const float4 t4 = b2[g4];
            f[g4 * 4 + 0] = __uint_as_float(r2[half][g4 * 4 + 0]) + t4.x; f[g4 * 4 + 1] = __uint_as_float(r2[half][g4 * 4 + 1]) + t4.y;
            f[g4 * 4 + 2] = __uint_as_float(r2[half][g4 * 4 + 2]) + t4.z; f[g4 * 4 + 3] = __uint_as_float(r2[half][g4 * 4 + 3]) + t4.w;
          }
#pragma unroll
          for (int g4 = 0; g4 < 4; ++g4) {
            uint4 u;
            __nv_bfloat162 h0 = __floats2bfloat162_rn(f[g4 * 8 + 0], f[g4 * 8 + 1]);
            __nv_bfloat162 h1 = __floats2bfloat162_rn(f[g4 * 8 + 2], f[g4 * 8 + 3]);
            __nv_bfloat162 h2 = __floats2bfloat162_rn(f[g4 * 8 + 4], f[g4 * 8 + 5]);
            __nv_bfloat162 h3 = __floats2bfloat162_rn(f[g4 * 8 + 6], f[g4 * 8 + 7]);
            u.x = *reinterpret_cast<uint32_t*>(&h0); u.y = *reinterpret_cast<uint32_t*>(&h1);
            u.z = *reinterpret_cast<uint32_t*>(&h2); u.w = *reinterpret_cast<uint32_t*>(&h3);
            const int j = half * 4 + g4;
            *reinterpret_cast<uint4*>(my_o + row * 128 + ((j ^ (row & 7)) << 4)) = u;
          }
        }
        SPSEG(4);
        fence_proxy_async_smem();
        gsync(2);
        if (et == 0) { tma_store_5d(&map_out, my_o, ch * 64, 0, 0, j0, b); tma_store_commit(); }
        SPSEG(5);
        stage ^= 1;
      }
      if (++sb == p.ns) sb = 0;
    }
    SPDUMP(et == 0, 0, 6, 24);
    if (et == 0) tma_store_wait<0>();
    SPNS(14, et == 0);
  } else {
    // ===================== conversion groups: T (fp32) -> packed bf16 pairs -> P (tensor memory) =====================
    const int grp = (warp - 8) / 4, q = warp % 4;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    const uint32_t t_my = t_tmem + lane_off + (uint32_t)(grp * 64), p_my = p_tmem + lane_off + (uint32_t)(grp * 32);
    if (BT && grp == 0) {       // row `q*32+lane` of every BigA_k: 128 bf16 = 64 packed words, straight from global memory into this lane
      const int row = q * 32 + lane;
      for (int k = 0; k < p.k; ++k) {
        const uint4* src = reinterpret_cast<const uint4*>(p.big_g + ((size_t)k * 128 + row) * 128);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t w[32];
#pragma unroll
          for (int j = 0; j < 8; ++j) { const uint4 u = __ldg(src + h * 8 + j); w[4 * j] = u.x; w[4 * j + 1] = u.y; w[4 * j + 2] = u.z; w[4 * j + 3] = u.w; }
          tmem_st_32x32(big_tmem + lane_off + (uint32_t)(k * 64 + h * 32), w);
        }
      }
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(big_bar);
    }
    uint32_t par = 0;
    SPSTART();
    for (int g = grp; g < total; g += G, par ^= 1) {
      mbar_wait(tfull + grp, par);
      SPSEG(0);
      tc_fence_after();
      uint32_t r0[32], r1[32], pk[32];
      tmem_ld_32x32(t_my, r0);
      tmem_ld_32x32(t_my + 32, r1);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(tempty + grp);                                // the aggregation accumulator may be overwritten
      SPSEG(1);
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        __nv_bfloat162 h0 = __floats2bfloat162_rn(__uint_as_float(r0[2 * j]), __uint_as_float(r0[2 * j + 1]));
        __nv_bfloat162 h1 = __floats2bfloat162_rn(__uint_as_float(r1[2 * j]), __uint_as_float(r1[2 * j + 1]));
        pk[j] = *reinterpret_cast<uint32_t*>(&h0); pk[16 + j] = *reinterpret_cast<uint32_t*>(&h1);
      }
      if (p.xa_g) {      // training: this thread's row of the chunk (64 bf16 = 128 contiguous bytes of XA) straight from registers
        const int tix = g / chunks, c = g - tix * chunks, cc = c / p.k, kq = c - cc * p.k;
        const int tile = blockIdx.x + tix * (int)gridDim.x;
        const int tl = p.rev ? p.total_tiles - 1 - tile : tile;
        const int j0 = (tl % p.m_tiles_per_b) * p.ft, b = tl / p.m_tiles_per_b;
        const int row = q * 32 + lane, fr = j0 + row / p.v;
        if (row < p.ft * p.v && fr < p.t) {
          uint4* dst = reinterpret_cast<uint4*>(p.xa_g + ((size_t)(b * p.t + fr) * p.v + row % p.v) * ((size_t)p.k * p.ci) + (size_t)kq * p.ci + cc * 64);
#pragma unroll
          for (int j = 0; j < 4; ++j) {      // pk[0..15] = columns 0..31, pk[16..31] = columns 32..63
            dst[j] = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
            dst[4 + j] = make_uint4(pk[16 + 4 * j], pk[16 + 4 * j + 1], pk[16 + 4 * j + 2], pk[16 + 4 * j + 3]);
          }
        }
      }
      mbar_wait(pempty + grp, par ^ 1);                         // the mixing product that last read this operand has retired
      SPSEG(2);
      tc_fence_after();
      tmem_st_32x32(p_my, pk);
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(pfull + grp);
      SPSEG(3);
    }
    SPDUMP(warp == 8 && lane == 0, 0, 4, 16);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, p.tmem_cols); }
  SPNS(15, threadIdx.x == 0);
}

// ---------------------------------------------------------------------------
// host
// ---------------------------------------------------------------------------
static int32_t sp_act_map(CUtensorMap* m, const bf16* base, int cols, int v, int t, int nb, int ft) {
  uint64_t dims[5] = {(uint64_t)cols, (uint64_t)v, 1, (uint64_t)t, (uint64_t)nb};
  uint64_t str[4] = {(uint64_t)cols * 2, (uint64_t)v * cols * 2, (uint64_t)v * cols * 2, (uint64_t)t * v * cols * 2};
  uint32_t box[5] = {64, (uint32_t)v, 1, (uint32_t)ft, 1};
  return make_tmap_bf16(m, base, 5, dims, str, box);
}

struct SpPlan { int wn, wparts, xs, wst; size_t smem; bool ok; };
static SpPlan sp_plan(int v, int ci, int co, int k) {
  SpPlan pl; pl.ok = false;
  if (ci % 64 != 0 || co % 64 != 0 || co > 256 || v < 4 || v > 32 || k < 1 || k > 3) return pl;
  pl.wn = co % 128 == 0 ? 128 : 64; pl.wparts = co / pl.wn;
  const size_t fixed = (size_t)k * 2 * kSpTile + 2 * kSpTile + 1024 + 512 + (size_t)v * (co + 4) * sizeof(float);
  for (int xs = 3; xs >= 2 && !pl.ok; --xs)
    for (int wst = 4; wst >= 2 && !pl.ok; --wst) {
      if (wst < pl.wparts) continue;
      const size_t need = fixed + (size_t)xs * kSpTile + (size_t)wst * pl.wn * 128;
      if (need <= 226 * 1024) { pl.xs = xs; pl.wst = wst; pl.smem = need; pl.ok = true; }
    }
  return pl;
}
bool umma_spatial_supported(int v, int ci, int co, int k) { return sp_plan(v, ci, co, k).ok; }
// tensor-memory variant: no P tiles in shared memory, two output staging buffers instead; G conversion groups when the columns allow
struct SpTsPlan { int wn, wparts, xs, wst, g, ns; uint32_t cols; size_t smem; bool ok, bt, wres; };
static SpTsPlan sp_ts_plan(int v, int ci, int co, int k) {
  SpTsPlan pl; pl.ok = false;
  if (ci % 64 != 0 || co % 64 != 0 || co > 256 || v < 4 || v > 32 || k < 1 || k > 3) return pl;
  pl.wn = co % 128 == 0 ? 128 : 64; pl.wparts = co / pl.wn;
  // tensor-memory columns: G x (64 fp32 + 32 packed) + ns x Co <= 512; prefer three chunks in flight, then two output accumulators
  static const int force_g = [] { const char* e = getenv("STGCN_SPATIAL_G"); return e ? atoi(e) : 0; }();
  pl.g = 0;
  for (int g = force_g ? force_g : 2; g >= 2 && !pl.g; --g)
    for (int ns = 2; ns >= 1 && !pl.g; --ns) {
      if (96 * g + ns * co <= 512) { pl.g = g; pl.ns = ns; }
    }
  if (!pl.g) return pl;
  // the block-diagonal adjacency itself lives in tensor memory (A operand of MMA-1) when another k x 64 columns are free
  static const int bt_mode = [] { const char* e = getenv("STGCN_SPATIAL_BT"); return e ? atoi(e) : 1; }();
  pl.bt = bt_mode != 0 && 96 * pl.g + pl.ns * co + 64 * k <= 512;
  pl.cols = 32; while (pl.cols < (uint32_t)(96 * pl.g + pl.ns * co + (pl.bt ? 64 * k : 0))) pl.cols <<= 1;
  const size_t fixed = (pl.bt ? 0 : (size_t)k * 2 * kSpTile) + 2 * kSpTile + 1024 + 512 + (size_t)v * (co + 4) * sizeof(float);
  // all weight chunks resident (no ring, no per-chunk barrier traffic for issuer B) when they fit beside >= 3 activation tiles
  static const bool no_wres = getenv("STGCN_SPATIAL_NO_WRES") != nullptr;
  const size_t w_all = (size_t)k * ci * co * 2;
  pl.wres = false;
  for (int xs = 4; xs >= 3 && !pl.ok && !no_wres; --xs) {
    const size_t need = fixed + (size_t)xs * kSpTile + w_all;
    if (need <= 226 * 1024) { pl.xs = xs; pl.wst = (ci / 64) * k * pl.wparts; pl.smem = need; pl.ok = true; pl.wres = true; }
  }
  for (int xs = pl.bt ? 4 : 3; xs >= 2 && !pl.ok; --xs)
    for (int wst = 4; wst >= 2 && !pl.ok; --wst) {
      if (wst < pl.wparts) continue;
      const size_t need = fixed + (size_t)xs * kSpTile + (size_t)wst * pl.wn * 128;
      if (need <= 226 * 1024) { pl.xs = xs; pl.wst = wst; pl.smem = need; pl.ok = true; }
    }
  return pl;
}


static int num_sms_sp() {
  static int n = 0;
  if (!n) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev); if (n <= 0) n = 148; }
  return n;
}

// x (nb,t,v,ci) -> s (nb,t,v,co); xa (nb,t,v,k*ci) written too when non-null.  big = bf16 block-diagonal A (build_big_adjacency),
// w2p = bf16 [co][k*ci] folded weights, b2d = fp32 [v][co] joint bias table.
int32_t launch_spatial_fused(const bf16* x, int nb, int t, int v, int ci, int co, int k, const bf16* big, const bf16* w2p,
                             const float* b2d, bf16* xa, bf16* s_out, cudaStream_t st) {
  KernelTimer kt_(KC_UMMA_SPATIAL, 2.0 * nb * t * v * (double)k * ci * (co + v), st);
  const SpPlan pl = sp_plan(v, ci, co, k);
  if (!pl.ok) { set_error("fused spatial: unsupported shape V=%d Ci=%d Co=%d K=%d", v, ci, co, k); return STGCN_ERR_UNSUPPORTED; }
  SpParams p; memset(&p, 0, sizeof(p));
  p.rev = next_direction();
  p.nb = nb; p.t = t; p.v = v; p.ft = 128 / v; p.ci = ci; p.co = co; p.k = k;
  p.m_tiles_per_b = cdiv(t, p.ft); p.total_tiles = p.m_tiles_per_b * nb;
  p.cblocks = ci / 64; p.wn = pl.wn; p.wparts = pl.wparts; p.xs = pl.xs; p.wst = pl.wst;
  p.ns = (128 + 2 * co <= 512) ? 2 : 1;
  uint32_t cols = 32; while (cols < (uint32_t)(128 + p.ns * co)) cols <<= 1;
  p.tmem_cols = cols; p.tile_bytes = (uint32_t)(p.ft * v * 128);
  p.write_xa = xa != nullptr; p.b2d = b2d;
  CUtensorMap map_x, map_big, map_w, map_xa, map_out;
  STGCN_TRY(sp_act_map(&map_x, x, ci, v, t, nb, p.ft));
  STGCN_TRY(sp_act_map(&map_out, s_out, co, v, t, nb, p.ft));
  if (xa) { STGCN_TRY(sp_act_map(&map_xa, xa, k * ci, v, t, nb, p.ft)); } else map_xa = map_out;
  {
    uint64_t dims[2] = {128, (uint64_t)k * 128}; uint64_t str[1] = {256}; uint32_t box[2] = {64, 128};
    STGCN_TRY(make_tmap_bf16(&map_big, big, 2, dims, str, box));
  }
  {
    uint64_t dims[2] = {(uint64_t)k * ci, (uint64_t)co}; uint64_t str[1] = {(uint64_t)k * ci * 2}; uint32_t box[2] = {64, (uint32_t)pl.wn};
    STGCN_TRY(make_tmap_bf16(&map_w, w2p, 2, dims, str, box));
  }
  // inference (no XA store): the tensor-memory variant; STGCN_SPATIAL_TS=0 keeps the shared-memory kernel for A/B runs
  static const bool use_ts = [] { const char* e = getenv("STGCN_SPATIAL_TS"); return !(e && e[0] == '0'); }();
  const SpTsPlan tpl = sp_ts_plan(v, ci, co, k);
  static const bool ts_train = [] { const char* e = getenv("STGCN_SPATIAL_TS_TRAIN"); return !(e && e[0] == '0'); }();
  if (use_ts && (!xa || ts_train) && tpl.ok) {
#ifdef STGCN_SP_TRACE
    static long long* trace_buf = nullptr;
    if (getenv("STGCN_SPATIAL_TRACE")) { if (!trace_buf) cudaMalloc(&trace_buf, 32 * 8); cudaMemsetAsync(trace_buf, 0, 32 * 8, st); p.trace = trace_buf; }
#endif
    p.wn = tpl.wn; p.wparts = tpl.wparts; p.xs = tpl.xs; p.wst = tpl.wst; p.ns = tpl.ns; p.tmem_cols = tpl.cols; p.wres = tpl.wres ? 1 : 0;
    static bool ts_attr_set = false;
    if (!ts_attr_set) {
      cudaError_t e = cudaFuncSetAttribute(spatial_fused_ts_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(spatial_fused_ts_kernel<3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(spatial_fused_ts_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(spatial_fused_ts_kernel<3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
      if (e != cudaSuccess) { set_error("cudaFuncSetAttribute failed: %s", cudaGetErrorString(e)); return STGCN_ERR_CUDA; }
      ts_attr_set = true;
    }
    const int grid_ts = std::min(p.total_tiles, num_sms_sp());
    p.big_g = big; p.xa_g = xa;
    if (tpl.g == 3 && tpl.bt) launch_k(spatial_fused_ts_kernel<3, true>, grid_ts, sp_ts_threads(3), tpl.smem, st, map_x, map_big, map_w, map_out, p);
    else if (tpl.g == 3) launch_k(spatial_fused_ts_kernel<3, false>, grid_ts, sp_ts_threads(3), tpl.smem, st, map_x, map_big, map_w, map_out, p);
    else if (tpl.bt) launch_k(spatial_fused_ts_kernel<2, true>, grid_ts, sp_ts_threads(2), tpl.smem, st, map_x, map_big, map_w, map_out, p);
    else launch_k(spatial_fused_ts_kernel<2, false>, grid_ts, sp_ts_threads(2), tpl.smem, st, map_x, map_big, map_w, map_out, p);
    STGCN_CHECK_LAUNCH("spatial_fused_ts");
#ifdef STGCN_SP_TRACE
    if (p.trace) {
      cudaStreamSynchronize(st);
      long long h[32];
      cudaMemcpy(h, p.trace, sizeof(h), cudaMemcpyDeviceToHost);
      const double n = (double)(cdiv(p.total_tiles - 0, grid_ts)) * (ci / 64) * k;      // chunks of CTA 0
      fprintf(stderr, "[sp trace %d->%d G=%d bt=%d ns=%d xs=%d wst=%d] per chunk (cycles): MMA warp: tempty %.0f xfull %.0f issue1 %.0f | sempty %.0f pfull %.0f wfull %.0f mix %.0f other %.0f = %.0f ; conv0 (per own chunk): tfull %.0f ld %.0f pempty %.0f cvt+st %.0f ; O per tile: sfull %.0f loop %.0f ld %.0f bar1 %.0f bias+sts %.0f fence+bar2+store %.0f (wres %d) | ns: prologue %lld pdl_wait %lld issuerA-end %lld O-end %lld exit %lld\n",
              ci, co, tpl.g, (int)tpl.bt, tpl.ns, tpl.xs, tpl.wst, h[0] / n, h[1] / n, h[2] / n, h[3] / n, h[4] / n, h[5] / n, h[6] / n, h[7] / n,
              (h[0] + h[1] + h[2] + h[3] + h[4] + h[5] + h[6] + h[7]) / n, h[16] / (n / tpl.g), h[17] / (n / tpl.g), h[18] / (n / tpl.g), h[19] / (n / tpl.g),
              h[24] / (n / (ci / 64 * k)), h[25] / (n / (ci / 64 * k)), h[26] / (n / (ci / 64 * k)), h[27] / (n / (ci / 64 * k)), h[28] / (n / (ci / 64 * k)), h[29] / (n / (ci / 64 * k)), (int)tpl.wres,
              h[11] - h[10], h[12] - h[11], h[13] - h[12], h[14] - h[12], h[15] - h[12]);
    }
#endif
    return STGCN_OK;
  }
  // STGCN_SPATIAL_SUB=1 | 2: epilogue warps per 64-column chunk (see sp_threads); default chosen by same-box A/B
  static const int sub = [] { const char* v = getenv("STGCN_SPATIAL_SUB"); return (v && v[0] == '1') ? 1 : ((v && v[0] == '2') ? 2 : kSpDefaultSub); }();
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(spatial_fused_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(spatial_fused_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) { set_error("cudaFuncSetAttribute failed: %s", cudaGetErrorString(e)); return STGCN_ERR_CUDA; }
    attr_set = true;
  }
  const int grid = std::min(p.total_tiles, num_sms_sp());
  if (sub == 2) launch_k(spatial_fused_kernel<2>, grid, sp_threads(2), pl.smem, st, map_x, map_big, map_w, map_xa, map_out, p);
  else launch_k(spatial_fused_kernel<1>, grid, sp_threads(1), pl.smem, st, map_x, map_big, map_w, map_xa, map_out, p);
  STGCN_CHECK_LAUNCH("spatial_fused");
  return STGCN_OK;
}

}  // namespace stgcn
