// tcgen05 weight-gradient kernel (bf16 operands, fp32 accumulation in TMEM):
//
//     dw[tap][n][c] = sum_r dout[r][n] * in[row(r,tap)][c]
//
// for the temporal convolution (9 taps, stride 1|2; autograd of nn.Conv2d at
// st_graphconv.py:41-43), the 1x1 channel mixing (:136) and the residual 1x1 conv (:28).
// The reduction runs over rows = (sample, frame, joint), so BOTH operands are MN-major:
// their shared-memory tiles are [rows][64 channels] exactly as TMA delivers them.
//
// One work item = (n-tile of 128 output channels, c-tile, tap group, split of the row range).
// Per reduction step the CTA TMA-loads FJ whole frames of dout (FJ*V <= 128 rows, the
// unused rows stay zero from a one-time shared-memory clear) and ONE slab of
// FJ + span frames of the input; every tap of the group reads that slab through a
// descriptor whose start address is shifted by (tap offset)*V rows (validated on hardware
// by tests/test_gpu_umma.py: the 128B swizzle is a function of the absolute address).
// Each tap owns its own accumulator columns in TMEM; at the end of the item the epilogue
// warps add the tile into a packed fp32 buffer with red.global.add.
#include <stdlib.h>
#include <string.h>

#include "kernels.cuh"
#include "umma_common.cuh"

namespace stgcn {
using namespace umma;

constexpr int kWgThreads = 192;
constexpr int kWgMaxGroups = 8;
constexpr int kWgMaxTapsPerGroup = 8;
constexpr int kMaxSlabRows = 392;              // (taps - 1) * V + 128 rows, multiple of 8 (8*32 + 128 = 384 for V = 32)
constexpr int kATile = 128 * 128;              // 128 rows x 64 ch bf16

// stacked mode (n_out == 64): accumulator rows 0..63 hold tap w_tap[t], rows 64..127 tap w_tap2[t] (the next frame offset, or -1)
struct WgGroup { int ntaps; int parity; int h_min; int w_tap[kWgMaxTapsPerGroup]; int row_off[kWgMaxTapsPerGroup]; int w_tap2[kWgMaxTapsPerGroup]; };

struct WgParams {
  int nb, v, fj, jblocks;            // frames per reduction step, steps per sample
  int n_out, c, n_tiles, c_tiles, c_tile, n_slabs;   // n_slabs: 64-wide slabs of dout per tile (1 or 2)
  int ngroups, splits, kblocks_total, total_items;
  int stacked;                       // 64 output channels: the second dout slab is the SAME channels one frame earlier -> two taps per MMA
  int stages, rev;
  uint32_t tmem_cols, a_bytes, b_bytes;
  int b_slab;                        // bytes of one 64-channel input slab (rows * 128, rows % 8 == 0)
  WgGroup groups[kWgMaxGroups];
  float* packed;                     // [taps][n_out][c] fp32, zeroed by the host
  int terms;                         // 1, or 6 for split-bf16 fp32 operands: dout is [rows][3 n_out] = [hi | mid | lo], in is [rows][3 c];
                                     // every reduction step runs the six significant part products into the same accumulator
};
__device__ __constant__ int c_wg_term_d[6] = {1, 0, 2, 0, 1, 0};      // dout part of term i (smallest products first)
__device__ __constant__ int c_wg_term_i[6] = {1, 2, 0, 1, 0, 0};      // input part of term i

__global__ void __launch_bounds__(kWgThreads, 1)
wgrad_umma_kernel(const __grid_constant__ CUtensorMap map_dout, const __grid_constant__ CUtensorMap map_in,
                  const __grid_constant__ WgParams p) {
  pdl_trigger();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int a_stage = p.n_slabs * kATile;
  const int b_stage = (p.c_tile / 64) * p.b_slab;
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + (size_t)p.stages * a_stage;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_b + (size_t)p.stages * b_stage);
  uint64_t* empty_bar = full_bar + p.stages;
  uint64_t* tfull_bar = empty_bar + p.stages;
  uint64_t* tempty_bar = tfull_bar + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 1);
  float* tr_base = reinterpret_cast<float*>(full_bar + 32);      // 4 x [32][33] floats, after the barrier block
  const int warp = uniform_warp_idx(), lane = threadIdx.x % 32;

  // one-time clear: rows TMA never writes (tile padding) must read as zero, never as NaN bit patterns
  {
    uint4* z = reinterpret_cast<uint4*>(smem);
    const int n16 = (int)(((size_t)p.stages * (a_stage + b_stage)) / 16);
    for (int i = threadIdx.x; i < n16; i += kWgThreads) z[i] = make_uint4(0, 0, 0, 0);
  }
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&map_dout); prefetch_tmap(&map_in);
    for (int s = 0; s < p.stages; ++s) { mbar_init(full_bar + s, 1); mbar_init(empty_bar + s, 1); }
    mbar_init(tfull_bar, 1); mbar_init(tempty_bar, 128);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, p.tmem_cols);
  fence_proxy_async_smem();          // the generic-proxy zero fill must be ordered before TMA writes
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();                       // everything above (barriers, TMEM, descriptor prefetch) overlaps the previous kernel's tail
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);      // warp-uniform for the compiler: the MMA operands stay in uniform registers (no R2UR per tcgen05.mma)

  // item -> (n_t, c_t, group, split)
  auto decode = [&](int item, int& n_t, int& c_t, int& g, int& kb0, int& kb1) {
    if (p.rev) item = p.total_items - 1 - item;
    n_t = item % p.n_tiles; item /= p.n_tiles;
    c_t = item % p.c_tiles; item /= p.c_tiles;
    g = item % p.ngroups; item /= p.ngroups;
    const int per = (p.kblocks_total + p.splits - 1) / p.splits;
    kb0 = item * per; kb1 = min(p.kblocks_total, kb0 + per);
  };

  if (warp == 0) {
    {   // whole warp runs the loop, one elected lane arms barriers and issues copies
      int stage = 0; uint32_t phase = 0;
      for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
        int n_t, c_t, g, kb0, kb1; decode(item, n_t, c_t, g, kb0, kb1);
        const WgGroup& G = p.groups[g];
        for (int kb = kb0; kb < kb1; ++kb) {
          const int b = kb / p.jblocks, j0 = (kb % p.jblocks) * p.fj;
          for (int term = 0; term < p.terms; ++term) {
            const int dcol = p.terms > 1 ? c_wg_term_d[term] * p.n_out : 0, icol = p.terms > 1 ? c_wg_term_i[term] * p.c : 0;
            mbar_wait(empty_bar + stage, phase ^ 1);
            mbar_expect_tx_e(full_bar + stage, p.a_bytes + p.b_bytes);
            for (int s = 0; s < p.n_slabs; ++s)
              tma_load_5d_e(&map_dout, smem_a + (size_t)stage * a_stage + s * kATile, full_bar + stage,
                            dcol + (p.stacked ? 0 : n_t * 128 + s * 64), 0, 0, p.stacked ? j0 - s : j0, b);
            for (int s = 0; s < p.c_tile / 64; ++s)
              tma_load_5d_e(&map_in, smem_b + (size_t)stage * b_stage + s * p.b_slab, full_bar + stage, icol + c_t * p.c_tile + s * 64, 0,
                          G.parity, j0 + G.h_min, b);
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    {   // whole warp runs the issue loop, one elected lane issues (umma_common.cuh)
      const uint32_t idesc = make_idesc_bf16(128, p.c_tile, 1, 1);
      const uint32_t a_lbo = p.n_slabs == 2 ? (uint32_t)kATile : 0u;   // one slab: rows 64..127 alias rows 0..63 (ignored)
      int stage = 0; uint32_t phase = 0; uint32_t tphase = 0;
      for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
        int n_t, c_t, g, kb0, kb1; decode(item, n_t, c_t, g, kb0, kb1);
        const WgGroup& G = p.groups[g];
        mbar_wait(tempty_bar, tphase ^ 1);
        tc_fence_after();
        for (int kbt = kb0 * p.terms; kbt < kb1 * p.terms; ++kbt) {
          const int kb = kbt > kb0 * p.terms ? kb0 + 1 : kb0;      // only "first step of the item or not" matters below
          mbar_wait(full_bar + stage, phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem_a + (size_t)stage * a_stage);
          const uint32_t b_addr = smem_u32(smem_b + (size_t)stage * b_stage);
          for (int t = 0; t < G.ntaps; ++t) {
            const uint64_t ad = make_smem_desc_sw128(a_addr, a_lbo, 1024);
            const uint64_t bd = make_smem_desc_sw128(b_addr + (uint32_t)G.row_off[t] * 128u, (uint32_t)p.b_slab, 1024);
#pragma unroll
            for (int k = 0; k < 8; ++k)          // 8 x 16 rows of the 128-row reduction step (2048 B each)
              umma_bf16_ss_elect(tmem_base + (uint32_t)(t * p.c_tile), ad + (uint64_t)(k * 128), bd + (uint64_t)(k * 128), idesc,
                           (kb > kb0 || k > 0) ? 1u : 0u);
          }
          umma_commit_elect(empty_bar + stage);
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
        umma_commit_elect(tfull_bar);
        tphase ^= 1;
      }
    }
  } else {
    const int q = warp % 4;                      // TMEM lane quarter = accumulator rows q*32 .. q*32+31 (output channels)
    float* tr = tr_base + q * (32 * 33);         // this warp's 32x32 transpose tile (+1 padding: conflict-free both ways)
    uint32_t tphase = 0;
    for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
      int n_t, c_t, g, kb0, kb1; decode(item, n_t, c_t, g, kb0, kb1);
      const WgGroup& G = p.groups[g];
      mbar_wait(tfull_bar, tphase);
      tc_fence_after();
      const int n0 = p.stacked ? (q & 1) * 32 : n_t * 128 + q * 32;          // first output channel of this warp's rows
      const int rows_all = (kb1 > kb0) ? (p.stacked ? 32 : max(0, min(32, min(p.n_slabs * 64 - q * 32, p.n_out - n0)))) : 0;
      if (rows_all > 0) {
        for (int t = 0; t < G.ntaps; ++t) {
          const int tap = (p.stacked && q >= 2) ? G.w_tap2[t] : G.w_tap[t];
          const int rows_ok = tap >= 0 ? rows_all : 0;
          float* dst = p.packed + ((size_t)max(tap, 0) * p.n_out + n0) * p.c + (size_t)c_t * p.c_tile + lane;
          for (int cc = 0; cc < p.c_tile; cc += 32) {
            uint32_t r[32];
            tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(t * p.c_tile + cc), r);
            tmem_ld_wait();
            // transpose through shared memory so that one red instruction covers 32 CONSECUTIVE columns of one
            // output channel (a single 128-byte L2 atomic) instead of 32 scattered 4-byte atomics
#pragma unroll
            for (int i = 0; i < 32; ++i) tr[i * 33 + lane] = __uint_as_float(r[i]);
            __syncwarp();
            for (int rr = 0; rr < rows_ok; ++rr) atomicAdd(dst + (size_t)rr * p.c + cc, tr[lane * 33 + rr]);
            __syncwarp();
          }
        }
      }
      tc_fence_before();
      mbar_arrive(tempty_bar);
      tphase ^= 1;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, p.tmem_cols); }
}

// packed [taps][n][c] -> the reference's parameter layout
__global__ void __launch_bounds__(256) wgrad_unpack_kernel(const float* __restrict__ packed, WIndex wi, int taps, int n_out, int c,
                                                           int c_valid, float* __restrict__ dw) {
  pdl_enter();
  size_t total = (size_t)taps * n_out * c;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int cc = (int)(i % c); size_t r = i / c; int n = (int)(r % n_out); int tap = (int)(r / n_out);
    if (cc < c_valid) dw[wi.at(tap, n, cc)] = packed[i];       // columns beyond c_valid are zero padding of the reduction operand
  }
}

bool umma_wgrad_supported(const ConvGeom& g) {
  if (g.dgrad) return false;
  if (g.c % 64 != 0 || g.n_out % 64 != 0) return false;
  if (g.v < 1 || g.v > 32 || g.taps < 1 || g.taps > 15) return false;
  if (g.stride < 1 || g.stride > 2 || g.t_in % g.stride != 0) return false;
  return true;
}

size_t umma_wgrad_workspace_floats(const ConvGeom& g) { return (size_t)g.taps * g.n_out * g.c; }

static int num_sms_wg() {
  static int n = 0;
  if (!n) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev); if (n <= 0) n = 148; }
  return n;
}

int32_t launch_wgrad_umma(const ConvGeom& g, const bf16* dout, const bf16* in, const WIndex& wi, float* packed_ws, float* dw,
                          cudaStream_t st, int c_valid, bool packed_is_zero, cudaStream_t unpack_st, cudaEvent_t unpack_ev, int split) {
  KernelTimer kt_(KC_UMMA_WGRAD, 2.0 * g.nb * g.t_out * g.v * (double)(c_valid >= 0 ? c_valid : g.c) * g.n_out * g.taps, st);
  if (!umma_wgrad_supported(g)) { set_error("umma wgrad: unsupported geometry"); return STGCN_ERR_UNSUPPORTED; }
  WgParams p; memset(&p, 0, sizeof(p));
  p.rev = next_direction();
  const int s = g.stride;
  p.stacked = (g.n_out == 64 && g.taps > 1 && getenv("STGCN_NO_TAP_STACK") == nullptr) ? 1 : 0;
  // stacked: one extra reduction step per sample so that the frame-shifted slab also covers the last output frame
  p.nb = g.nb; p.v = g.v; p.fj = 128 / g.v; p.jblocks = cdiv(g.t_out + p.stacked, p.fj);
  p.n_out = g.n_out; p.c = g.c;
  p.n_tiles = cdiv(g.n_out, 128); p.n_slabs = (g.n_out >= 128 || p.stacked) ? 2 : 1;
  if (g.n_out % 128 != 0 && g.n_out > 64) { set_error("umma wgrad: n_out must be 64 or a multiple of 128"); return STGCN_ERR_UNSUPPORTED; }
  // tap groups: same input parity, consecutive frame offsets, at most `tg_max` taps (TMEM: taps*c_tile <= 512 columns)
  if (g.taps == 1) p.c_tile = g.c % 256 == 0 ? 256 : (g.c % 192 == 0 ? 192 : (g.c % 128 == 0 ? 128 : 64));
  else p.c_tile = g.c == 64 ? 64 : 128;
  p.c_tiles = g.c / p.c_tile;
  const int tg_max = std::min(kWgMaxTapsPerGroup, 512 / p.c_tile);
  int span_max = 0;
  p.ngroups = 0;
  for (int par = 0; par < s; ++par) {
    int taps_p[16], hoff_p[16], np = 0;
    for (int tap = 0; tap < g.taps; ++tap) {
      int d = tap - g.pad; int r = ((d % s) + s) % s;
      if (r == par) { taps_p[np] = tap; hoff_p[np] = (d - r) / s; ++np; }
    }
    if (!np) continue;
    if (p.stacked) {      // entries are PAIRS of consecutive frame offsets (taps_p is ordered by offset, step 1)
      const int npair = cdiv(np, 2);
      int ng = cdiv(npair, tg_max), per = cdiv(npair, ng);
      for (int i0 = 0; i0 < npair; i0 += per) {
        if (p.ngroups >= kWgMaxGroups) { set_error("umma wgrad: too many tap groups"); return STGCN_ERR_UNSUPPORTED; }
        WgGroup& G = p.groups[p.ngroups++];
        G.ntaps = std::min(per, npair - i0); G.parity = par; G.h_min = hoff_p[2 * i0];
        for (int i = 0; i < G.ntaps; ++i) {
          const int a = 2 * (i0 + i);
          G.w_tap[i] = taps_p[a]; G.row_off[i] = (hoff_p[a] - G.h_min) * g.v;
          G.w_tap2[i] = (a + 1 < np && hoff_p[a + 1] == hoff_p[a] + 1) ? taps_p[a + 1] : -1;
          if (a + 1 < np && G.w_tap2[i] < 0) { set_error("umma wgrad: taps of one parity are not consecutive"); return STGCN_ERR_UNSUPPORTED; }
        }
        span_max = std::max(span_max, hoff_p[2 * (i0 + G.ntaps - 1)] - G.h_min);
      }
      continue;
    }
    int ng = cdiv(np, tg_max), per = cdiv(np, ng);
    for (int i0 = 0; i0 < np; i0 += per) {
      if (p.ngroups >= kWgMaxGroups) { set_error("umma wgrad: too many tap groups"); return STGCN_ERR_UNSUPPORTED; }
      WgGroup& G = p.groups[p.ngroups++];
      G.ntaps = std::min(per, np - i0); G.parity = par; G.h_min = hoff_p[i0];
      for (int i = 0; i < G.ntaps; ++i) { G.w_tap[i] = taps_p[i0 + i]; G.row_off[i] = (hoff_p[i0 + i] - G.h_min) * g.v; }
      span_max = std::max(span_max, hoff_p[i0 + G.ntaps - 1] - G.h_min);
    }
  }
  const int slab_rows = (span_max * g.v + 128 + 7) / 8 * 8;      // every tap reads 128 rows starting at its row offset
  if (slab_rows > kMaxSlabRows) { set_error("umma wgrad: slab too large"); return STGCN_ERR_UNSUPPORTED; }
  p.b_slab = slab_rows * 128;
  p.kblocks_total = g.nb * p.jblocks;
  const int base_items = p.n_tiles * p.c_tiles * p.ngroups;
  p.splits = std::max(1, std::min(p.kblocks_total, num_sms_wg() / base_items));   // one item per CTA: a single epilogue each
  p.total_items = base_items * p.splits;
  p.a_bytes = (uint32_t)(p.n_slabs * p.fj * g.v * 128);
  p.b_bytes = (uint32_t)((p.c_tile / 64) * (p.fj + span_max) * g.v * 128);
  const size_t stage_bytes = (size_t)p.n_slabs * kATile + (size_t)(p.c_tile / 64) * p.b_slab;
  const size_t tail_bytes = 2048 + 4 * 32 * 33 * sizeof(float);     // align slack + barriers + transpose tiles
  int stages = (int)((220 * 1024 - tail_bytes) / stage_bytes);
  if (stages > 4) stages = 4;
  if (stages < 2) { set_error("umma wgrad: not enough shared memory"); return STGCN_ERR_UNSUPPORTED; }
  p.stages = stages;
  int maxt = 0; for (int i = 0; i < p.ngroups; ++i) maxt = std::max(maxt, p.groups[i].ntaps);
  uint32_t cols = 32; while (cols < (uint32_t)(maxt * p.c_tile)) cols <<= 1;
  p.tmem_cols = cols;
  p.packed = packed_ws;
  p.terms = split ? 6 : 1;
  if (!packed_is_zero) cudaMemsetAsync(packed_ws, 0, sizeof(float) * umma_wgrad_workspace_floats(g), st);

  CUtensorMap map_dout, map_in;
  const uint64_t C = (uint64_t)g.c * (split ? 3 : 1), V = g.v, N = (uint64_t)g.n_out * (split ? 3 : 1);   // split: [hi | mid | lo] column thirds
  {
    uint64_t dims[5] = {N, V, 1, (uint64_t)g.t_out, (uint64_t)g.nb};
    uint64_t str[4] = {N * 2, V * N * 2, V * N * 2, (uint64_t)g.t_out * V * N * 2};
    uint32_t box[5] = {64, (uint32_t)g.v, 1, (uint32_t)p.fj, 1};
    STGCN_TRY(make_tmap_bf16(&map_dout, dout, 5, dims, str, box));
  }
  {
    uint64_t dims[5] = {C, V, (uint64_t)s, (uint64_t)g.t_in / s, (uint64_t)g.nb};
    uint64_t str[4] = {C * 2, V * C * 2, (uint64_t)s * V * C * 2, (uint64_t)g.t_in * V * C * 2};
    uint32_t box[5] = {64, (uint32_t)g.v, 1, (uint32_t)(p.fj + span_max), 1};
    STGCN_TRY(make_tmap_bf16(&map_in, in, 5, dims, str, box));
  }
  const size_t smem = (size_t)stages * stage_bytes + tail_bytes;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(wgrad_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(smem) failed: %s", cudaGetErrorString(e)); return STGCN_ERR_CUDA; }
    attr_set = true;
  }
  const int grid = std::min(p.total_items, num_sms_wg());
  launch_k(wgrad_umma_kernel, grid, kWgThreads, smem, st, map_dout, map_in, p);
  STGCN_CHECK_LAUNCH("wgrad_umma");
  size_t total = umma_wgrad_workspace_floats(g);
  cudaStream_t ust = st;
  if (unpack_st && unpack_ev) {      // the layout change is off the data-gradient path: let it run beside the next kernels
    cudaEventRecord(unpack_ev, st); cudaStreamWaitEvent(unpack_st, unpack_ev, 0); ust = unpack_st;
  }
  launch_k(wgrad_unpack_kernel, (int)std::min<size_t>((total + 255) / 256, 148 * 4), 256, 0, ust, packed_ws, wi, g.taps, g.n_out, g.c, c_valid < 0 ? g.c : c_valid, dw);
  STGCN_CHECK_LAUNCH("wgrad_unpack");
  return STGCN_OK;
}

}  // namespace stgcn
