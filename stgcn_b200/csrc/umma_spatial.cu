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
  const uint32_t tmem_base = *tmem_slot;
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
