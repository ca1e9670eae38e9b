// tcgen05 implicit-GEMM engine (bf16 operands, fp32 accumulation in TMEM).
//
//   out[r][n] = sum_tap sum_c in[row(r,tap)][c] * w[tap][n][c]      (+ fused epilogue)
//
// Covers the two dense contractions of an ST-GCN block -- the 1x1 channel
// mixing (taps = 1; nn.Conv2d at st_graphconv.py:136 and :28) and the 9x1
// temporal convolution (nn.Conv2d(C,C,(9,1),stride=(s,1),padding=(4,0)) at
// st_graphconv.py:41-43) -- and their data gradients (transposed conv).
//
// Tiling: one output tile = FT whole frames x V joints (<=128 rows = the MMA M)
// by BN output channels.  The reduction runs over (tap, 64-channel block):
// every step TMA-loads the frame-shifted activation tile through a 5-D tensor
// map {C, V, stride-phase, T/stride, N} (zero fill outside [0,T) implements the
// conv padding; the phase dimension implements stride 2 without gathers) and
// the matching weight tile, both 128B-swizzled, and one elected thread issues
// four tcgen05.mma (K=16 each).  Accumulators are double-buffered in TMEM so
// the epilogue of tile i overlaps the MMAs of tile i+1.  Warp roles: 0 = TMA
// producer, 1 = MMA issuer (+TMEM alloc), 2..5 = epilogue (TMEM -> registers
// -> bias / folded BN / residual / ReLU / dropout -> bf16 -> swizzled smem ->
// TMA store).  Persistent CTAs, one per SM.
#include "kernels.cuh"
#include "umma_common.cuh"

namespace stgcn {

using namespace umma;

constexpr int kUmmaThreads = 320;          // TMA warp, MMA warp, two epilogue groups of four warps
constexpr int kMaxTaps = 16;
constexpr int kATileBytes = 128 * 128;      // 128 rows x 64 bf16
constexpr int kStagingBytes = 128 * 128;    // 128 rows x 64 bf16 (one TMA-store box)

struct TapEntry { int w_tap; int parity; int h_off; };

struct UmmaParams {
  int nb, v, ft, phases;
  int t_out_ph;          // output frames per phase
  int t_out;             // total output frames (row index for residual / dropout)
  int n_out, bn, n_tiles, kblocks_c, m_tiles_per_b, total_tiles;
  int rev;               // walk the tiles from the last to the first (ping-pong traversal, common.cuh)
  int stages;
  uint32_t tmem_cols;
  uint32_t a_bytes, b_bytes;
  int ntaps[2];
  TapEntry taps[2][kMaxTaps];
  // halo mode (stride 1, several taps): ONE activation tile of ft + (h_max - h_min) frames per 64-channel block serves every tap
  // through row-shifted descriptors; the weight tiles have their own ring, or stay resident when they all fit
  int halo, h_min, w_stages, w_resident, nstaging;
  uint32_t a_stage_bytes, a_halo_bytes;
  int a_off16[kMaxTaps];   // descriptor offset (16-byte units) of each tap's first row inside the halo tile
  // epilogue
  const float* bias; const float* bias2d; const float* scale; const float* shift;
  int b2_smem;           // bias2d staged in shared memory as [V][bn+4] floats (only when n_tiles == 1)
  float* stat_partials;  // [gridDim][2][n_out] per-CTA column sum / sum of squares of the STORED output (BatchNorm statistics)
  float* vsum;           // [V][n_out] += sum over frames of the stored output (per-joint column sums)
  const bf16* residual;
  int relu;
  float dropout_p, inv_keep;
  uint64_t seed, offset;
  const uint64_t* rng;   // device {seed, offset} (overrides the values above)
  // fp32 path on the tensor cores ("split bf16"): every fp32 operand value is carried as hi + mid + lo (three bf16, 24 significand
  // bits); the input tensor is [rows][3C] = [hi | mid | lo], the weights [tap][n][3C].  The reduction runs over SIX products per
  // logical 64-channel block -- (part of the activation, part of the weight) = mid.mid, hi.lo, lo.hi, hi.mid, mid.hi, hi.hi,
  // smallest first; the dropped ones are below 2^-24 relative -- all into the same fp32 accumulator in TMEM.
  int nterms, c_log;     // nterms: 0 (plain bf16) | 6; c_log = logical 64-channel blocks (kblocks_c = max(nterms,1) * c_log)
  float* out_f32;        // fp32 output [rows][n_out] written straight from the epilogue (no bf16 staging / TMA store)
  const float* residual_f32;
  int accumulate;        // out_f32 += (fp32 output only)
};
constexpr int kSplitParts = 3, kSplitTerms = 6;
__device__ __constant__ int c_term_a[kSplitTerms] = {1, 0, 2, 0, 1, 0};      // activation part of term i
__device__ __constant__ int c_term_w[kSplitTerms] = {1, 2, 0, 1, 0, 0};      // weight part of term i
// channel coordinate (elements) of reduction step kb in the activation / weight tensor
__device__ __forceinline__ int a_col_of(const UmmaParams& p, int kb) {
  if (!p.nterms) return kb * 64;
  const int term = kb / p.c_log, cb = kb - term * p.c_log;
  return (c_term_a[term] * p.c_log + cb) * 64;
}
__device__ __forceinline__ int w_col_of(const UmmaParams& p, int kb) {
  if (!p.nterms) return kb * 64;
  const int term = kb / p.c_log, cb = kb - term * p.c_log;
  return (c_term_w[term] * p.c_log + cb) * 64;
}

struct TileCoord { int n_t, ph, j0, b; };
__device__ __forceinline__ TileCoord decode_tile(const UmmaParams& p, int tile) {
  TileCoord t;
  if (p.rev) tile = p.total_tiles - 1 - tile;
  t.n_t = tile % p.n_tiles; tile /= p.n_tiles;
  t.ph = tile % p.phases; tile /= p.phases;
  t.j0 = (tile % p.m_tiles_per_b) * p.ft;
  t.b = tile / p.m_tiles_per_b;
  return t;
}

// EPM = compile-time mask of the epilogue options an instantiation keeps (the others fold away).  This matters because the
// epilogue warps are bound by dependent-issue latency (about 10 cycles per instruction with two warps per scheduler): the
// option-free instance used by every data-gradient GEMM needs 84 registers and a third of the instructions per chunk.
constexpr int EP_BIAS = 1, EP_BIAS2D = 2, EP_SCALE = 4, EP_RESID = 8, EP_RELU = 16, EP_DROPOUT = 32, EP_STATS = 64, EP_VSUM = 128, EP_ALL = 255;
constexpr int EP_F32OUT = 256;      // fp32 output straight from the epilogue (split-bf16 fp32 path); excludes the statistics options
template <int EPM>
__global__ void __launch_bounds__(kUmmaThreads, 1)
conv_gemm_umma_kernel(const __grid_constant__ CUtensorMap map_in, const __grid_constant__ CUtensorMap map_w,
                      const __grid_constant__ CUtensorMap map_out, const __grid_constant__ UmmaParams p) {
  const float* const o_bias = (EPM & EP_BIAS) ? p.bias : nullptr; const float* const o_bias2d = (EPM & EP_BIAS2D) ? p.bias2d : nullptr;
  const float* const o_scale = (EPM & EP_SCALE) ? p.scale : nullptr; const float* const o_shift = (EPM & EP_SCALE) ? p.shift : nullptr;
  const bf16* const o_residual = (EPM & EP_RESID) ? p.residual : nullptr;
  float* const o_stat_partials = (EPM & EP_STATS) ? p.stat_partials : nullptr; float* const o_vsum = (EPM & EP_VSUM) ? p.vsum : nullptr;
  const int o_b2_smem = (EPM & EP_BIAS2D) ? p.b2_smem : 0, o_relu = (EPM & EP_RELU) ? p.relu : 0;
  const float o_dropout_p = (EPM & EP_DROPOUT) ? p.dropout_p : 0.f;
  pdl_trigger();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte aligned carve-up (dynamic smem base is only guaranteed 16-byte aligned)
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* smem_a = smem;                                        // stages x a_stage_bytes (16 KB, or the halo tile)
  uint8_t* smem_b = smem_a + (size_t)p.stages * p.a_stage_bytes; // w_stages x bn*128 B (multiple of 1024)
  uint8_t* staging = smem_b + (size_t)p.w_stages * p.b_bytes;    // 2 groups x nstaging x 16 KB
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + (size_t)2 * p.nstaging * kStagingBytes);
  uint64_t* empty_bar = full_bar + 8;
  uint64_t* tfull_bar = full_bar + 16;
  uint64_t* tempty_bar = full_bar + 18;
  uint64_t* wfull_bar = full_bar + 20;                           // [16] halo mode: weight ring
  uint64_t* wempty_bar = full_bar + 36;                          // [16]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(full_bar + 52);
  float* b2_tab = reinterpret_cast<float*>(tmem_slot + 4);       // 16-byte aligned: barriers start on a 1024-byte boundary
  float* stat_acc = b2_tab + (o_b2_smem ? p.v * (p.bn + 4) : 0);  // [2][n_out]
  float* vsum_acc = stat_acc + (o_stat_partials ? 2 * p.n_out : 0);   // [V][n_out]

  const int warp = uniform_warp_idx(), lane = threadIdx.x % 32;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&map_in); prefetch_tmap(&map_w); prefetch_tmap(&map_out);
    for (int s = 0; s < p.stages; ++s) { mbar_init(full_bar + s, 1); mbar_init(empty_bar + s, 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar + a, 1); mbar_init(tempty_bar + a, 256); }
    for (int a = 0; a < 16; ++a) { mbar_init(wfull_bar + a, 1); mbar_init(wempty_bar + a, 1); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();                       // everything above (barriers, TMEM, descriptor prefetch) overlaps the previous kernel's tail
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);      // warp-uniform for the compiler: the MMA operands stay in uniform registers (no R2UR per tcgen05.mma)

  if (warp == 0) {
    // ===================== TMA producer =====================
    // the whole warp runs the loops (uniform registers for coordinates and addresses); one elected lane arms the barrier and issues
    if (p.halo) {
      const int nt = p.ntaps[0];
      if (p.w_resident) {                          // every (channel block, tap) weight tile, once
        if (elect_one()) mbar_expect_tx(wfull_bar, (uint32_t)(nt * p.kblocks_c) * p.b_bytes);
        __syncwarp();
        for (int kb = 0; kb < p.kblocks_c; ++kb)
          for (int ti = 0; ti < nt; ++ti)
            if (elect_one()) tma_load_3d(&map_w, smem_b + (size_t)(kb * nt + ti) * p.b_bytes, wfull_bar, w_col_of(p, kb), 0, p.taps[0][ti].w_tap);
      }
      // two cursors over the (tile, channel block) units: activation tiles run `lead` units ahead of the weight tiles
      int a_tile = blockIdx.x, a_kb = 0, w_tile = blockIdx.x, w_kb = 0;
      int stage = 0; uint32_t phase = 0; int ws = 0; uint32_t wph = 0;
      auto issue_a = [&]() {
        if (a_tile >= p.total_tiles) return;
        const TileCoord tc = decode_tile(p, a_tile);
        mbar_wait(empty_bar + stage, phase ^ 1);
        if (elect_one()) {
          mbar_expect_tx(full_bar + stage, p.a_halo_bytes);
          tma_load_5d(&map_in, smem_a + (size_t)stage * p.a_stage_bytes, full_bar + stage, a_col_of(p, a_kb), 0, 0, tc.j0 + p.h_min, tc.b);
        }
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
        if (++a_kb == p.kblocks_c) { a_kb = 0; a_tile += gridDim.x; }
      };
      const int lead = p.w_resident ? p.stages - 1 : p.stages - 2;
      for (int i = 0; i <= lead; ++i) issue_a();
      while (w_tile < p.total_tiles) {
        if (!p.w_resident) {
          const TileCoord tc = decode_tile(p, w_tile);
          for (int ti = 0; ti < nt; ++ti) {
            mbar_wait(wempty_bar + ws, wph ^ 1);
            if (elect_one()) {
              mbar_expect_tx(wfull_bar + ws, p.b_bytes);
              tma_load_3d(&map_w, smem_b + (size_t)ws * p.b_bytes, wfull_bar + ws, w_col_of(p, w_kb), tc.n_t * p.bn, p.taps[0][ti].w_tap);
            }
            if (++ws == p.w_stages) { ws = 0; wph ^= 1; }
          }
        }
        if (++w_kb == p.kblocks_c) { w_kb = 0; w_tile += gridDim.x; }
        issue_a();
      }
    } else {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        const TileCoord tc = decode_tile(p, tile);
        const int nt = p.ntaps[tc.ph];
        for (int ti = 0; ti < nt; ++ti) {
          const TapEntry te = p.taps[tc.ph][ti];
          for (int kb = 0; kb < p.kblocks_c; ++kb) {
            mbar_wait(empty_bar + stage, phase ^ 1);
            if (elect_one()) {
              mbar_expect_tx(full_bar + stage, p.a_bytes + p.b_bytes);
              tma_load_5d(&map_in, smem_a + (size_t)stage * p.a_stage_bytes, full_bar + stage, a_col_of(p, kb), 0, te.parity, tc.j0 + te.h_off, tc.b);
              tma_load_3d(&map_w, smem_b + (size_t)stage * p.b_bytes, full_bar + stage, w_col_of(p, kb), tc.n_t * p.bn, te.w_tap);
            }
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // the whole warp runs the loop, one elected lane issues (see umma_bf16_ss_elect)
    if (p.halo) {
      const uint32_t idesc = make_idesc_bf16(128, p.bn, 0, 0);
      const uint64_t a_d0 = make_smem_desc_sw128(smem_u32(smem_a), 16, 1024);
      const uint64_t w_d0 = make_smem_desc_sw128(smem_u32(smem_b), 16, 1024);
      const uint32_t a_step = p.a_stage_bytes >> 4, w_step = p.b_bytes >> 4;
      const int nt = p.ntaps[0];
      int stage = 0; uint32_t phase = 0; int ws = 0; uint32_t wph = 0;
      int as = 0; uint32_t aphase = 0;
      if (p.w_resident) { mbar_wait(wfull_bar, 0); tc_fence_after(); }
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        mbar_wait(tempty_bar + as, aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(as * p.bn);
        for (int kb = 0; kb < p.kblocks_c; ++kb) {
          mbar_wait(full_bar + stage, phase);
          tc_fence_after();
          const uint64_t a_st = a_d0 + (uint64_t)((uint32_t)stage * a_step);
          for (int ti = 0; ti < nt; ++ti) {
            uint64_t bdesc;
            if (p.w_resident) {
              bdesc = w_d0 + (uint64_t)((uint32_t)(kb * nt + ti) * w_step);
            } else {
              mbar_wait(wfull_bar + ws, wph);
              tc_fence_after();
              bdesc = w_d0 + (uint64_t)((uint32_t)ws * w_step);
            }
            const uint64_t adesc = a_st + (uint64_t)(uint32_t)p.a_off16[ti];      // this tap's rows: a row shift inside the halo tile
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_bf16_ss_elect(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | ti | k) != 0 ? 1u : 0u);
            if (!p.w_resident) { umma_commit_elect(wempty_bar + ws); if (++ws == p.w_stages) { ws = 0; wph ^= 1; } }
          }
          umma_commit_elect(empty_bar + stage);
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
        umma_commit_elect(tfull_bar + as);
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    } else {
      const uint32_t idesc = make_idesc_bf16(128, p.bn, 0, 0);
      int stage = 0; uint32_t phase = 0;
      int as = 0; uint32_t aphase = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        const TileCoord tc = decode_tile(p, tile);
        const int nkb = p.ntaps[tc.ph] * p.kblocks_c;
        mbar_wait(tempty_bar + as, aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(as * p.bn);
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(full_bar + stage, phase);
          tc_fence_after();
          const uint64_t adesc = make_smem_desc_sw128(smem_u32(smem_a + (size_t)stage * p.a_stage_bytes), 16, 1024);
          const uint64_t bdesc = make_smem_desc_sw128(smem_u32(smem_b + (size_t)stage * p.b_bytes), 16, 1024);
#pragma unroll
          for (int k = 0; k < 4; ++k)   // 4 x (K = 16 bf16 = 32 bytes) inside the 128-byte swizzle atom
            umma_bf16_ss_elect(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) != 0 ? 1u : 0u);
          umma_commit_elect(empty_bar + stage);          // frees the smem stage when these MMAs retire
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
        umma_commit_elect(tfull_bar + as);               // accumulator ready for the epilogue
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue warps: group 0 = warps 2..5, group 1 = warps 6..9 =====================
    // the groups take alternate 64-column chunks (alternate tiles when a tile is one chunk); each has its own staging buffers
    const int q = warp % 4;                        // TMEM lane quarter this warp may access
    const int row = q * 32 + lane;                 // tile row == TMEM lane
    const int grp = (warp - 2) >> 2;               // epilogue group
    const int et = threadIdx.x - 64 - grp * 128;   // 0..127 inside the group
    const int bar_a = 1 + 2 * grp, bar_b = 2 + 2 * grp;   // named barriers of this group
    const int et2 = threadIdx.x - 64;              // 0..255 over both groups
    const int fl = row / p.v, vv = row % p.v;
    int as = 0; uint32_t aphase = 0;
    int chunk_ctr = 0;                             // chunks this group has staged
    int gchunk = 0;                                // chunks of the CTA so far (both groups count them)
    uint8_t* const my_staging = staging + (size_t)grp * p.nstaging * kStagingBytes;
    const int b2_ld = p.bn + 4;                    // row pitch (floats): 16-byte rows, conflict-free across joints
    if (o_b2_smem)
      for (int i = et2; i < p.v * p.bn; i += 256) b2_tab[(i / p.bn) * b2_ld + (i % p.bn)] = __ldg(o_bias2d + i);
    if (o_stat_partials) for (int i = et2; i < 2 * p.n_out; i += 256) stat_acc[i] = 0.f;
    if (o_vsum) for (int i = et2; i < p.v * p.n_out; i += 256) vsum_acc[i] = 0.f;
    asm volatile("bar.sync 5, 256;" ::: "memory");
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      const TileCoord tc = decode_tile(p, tile);
      mbar_wait(tfull_bar + as, aphase);
      tc_fence_after();
      const int jj = tc.j0 + fl;
      const bool row_valid = row < p.ft * p.v && jj < p.t_out_ph;
      const long long grow = ((long long)tc.b * p.t_out + (long long)jj * p.phases + tc.ph) * p.v + vv;   // global output row
      for (int ch = 0; ch < p.bn / 64; ++ch) {
        if (((gchunk++) & 1) != grp) continue;
        uint8_t* stg = my_staging + (size_t)(chunk_ctr & (p.nstaging - 1)) * kStagingBytes;   // nstaging (per group) is 1 or 2
        if (!(EPM & EP_F32OUT) && chunk_ctr >= p.nstaging) {
          if (et == 0) { if (p.nstaging == 2) tma_store_wait_read<1>(); else tma_store_wait_read<0>(); }   // the store that last read this buffer has drained
          asm volatile("bar.sync %0, 128;" ::"r"(bar_a) : "memory");
        }
        ++chunk_ctr;
        uint32_t r2[2][32];                        // both 32-column halves in flight before the first use
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * p.bn + ch * 64), r2[0]);
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * p.bn + ch * 64 + 32), r2[1]);
        tmem_ld_wait();
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int col0 = ch * 64 + half * 32;    // column inside the bn-wide tile
          const int n0 = tc.n_t * p.bn + col0;     // global output channel of the first value
          float f[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(r2[half][i]);
          if (o_bias) {
            const float4* bp = reinterpret_cast<const float4*>(o_bias + n0);
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              const float4 t = __ldg(bp + g);
              f[g * 4 + 0] += t.x; f[g * 4 + 1] += t.y; f[g * 4 + 2] += t.z; f[g * 4 + 3] += t.w;
            }
          }
          if (o_b2_smem) {
            const float4* b2 = reinterpret_cast<const float4*>(b2_tab + vv * b2_ld + col0);
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              const float4 t = b2[g];
              f[g * 4 + 0] += t.x; f[g * 4 + 1] += t.y; f[g * 4 + 2] += t.z; f[g * 4 + 3] += t.w;
            }
          } else if (o_bias2d) {
            const float* b2 = o_bias2d + (size_t)vv * p.n_out + n0;
#pragma unroll
            for (int i = 0; i < 32; ++i) f[i] += __ldg(b2 + i);
          }
          if (o_scale) {
            const float4* sp = reinterpret_cast<const float4*>(o_scale + n0);
            const float4* hp = reinterpret_cast<const float4*>(o_shift + n0);
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              const float4 a4 = __ldg(sp + g), b4 = __ldg(hp + g);
              f[g * 4 + 0] = fmaf(a4.x, f[g * 4 + 0], b4.x); f[g * 4 + 1] = fmaf(a4.y, f[g * 4 + 1], b4.y);
              f[g * 4 + 2] = fmaf(a4.z, f[g * 4 + 2], b4.z); f[g * 4 + 3] = fmaf(a4.w, f[g * 4 + 3], b4.w);
            }
          }
          if ((EPM & EP_F32OUT) && p.residual_f32 && row_valid) {
            const float4* rp = reinterpret_cast<const float4*>(p.residual_f32 + (size_t)grow * p.n_out + n0);
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              const float4 t = __ldg(rp + g);
              f[g * 4 + 0] += t.x; f[g * 4 + 1] += t.y; f[g * 4 + 2] += t.z; f[g * 4 + 3] += t.w;
            }
          }
          if (!(EPM & EP_F32OUT) && o_residual && row_valid) {
            const uint4* rp = reinterpret_cast<const uint4*>(o_residual + (size_t)grow * p.n_out + n0);
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              uint4 u = __ldg(rp + g);
              const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
              for (int e = 0; e < 4; ++e) { f[g * 8 + 2 * e] += __low2float(h[e]); f[g * 8 + 2 * e + 1] += __high2float(h[e]); }
            }
          }
          if (o_relu) {
#pragma unroll
            for (int i = 0; i < 32; ++i) f[i] = fmaxf(f[i], 0.f);
          }
          if (o_dropout_p > 0.f && row_valid) {
            const uint64_t o = (uint64_t)grow * p.n_out + n0;
            const uint64_t d_seed = p.rng ? p.rng[0] : p.seed, d_off = p.rng ? p.rng[1] : p.offset;
#pragma unroll
            for (int i = 0; i < 32; ++i) f[i] *= dropout_scale(d_seed, d_off, o + i, o_dropout_p, p.inv_keep);
          }
          if (EPM & EP_F32OUT) {       // fp32 rows straight to global memory: 128 contiguous bytes per thread and half-chunk
            if (row_valid) {
              float4* op = reinterpret_cast<float4*>(p.out_f32 + (size_t)grow * p.n_out + n0);
#pragma unroll
              for (int g = 0; g < 8; ++g) {
                float4 t = make_float4(f[g * 4 + 0], f[g * 4 + 1], f[g * 4 + 2], f[g * 4 + 3]);
                if (p.accumulate) { const float4 o4 = op[g]; t.x += o4.x; t.y += o4.y; t.z += o4.z; t.w += o4.w; }
                op[g] = t;
              }
            }
            continue;
          }
          // bf16 pack and 128B-swizzled store: 16-byte chunk j of row lands at chunk (j ^ (row & 7))
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
        if (EPM & EP_F32OUT) continue;            // rows already stored; no staging tile, no statistics on this path
        fence_proxy_async_smem();                  // generic-proxy smem writes -> visible to the TMA engine
        asm volatile("bar.sync %0, 128;" ::"r"(bar_b) : "memory");
        if (et == 0) {
          tma_store_5d(&map_out, stg, tc.n_t * p.bn + ch * 64, 0, tc.ph, tc.j0, tc.b);
          tma_store_commit();
        }
        if (o_stat_partials || o_vsum) {
          // column reductions of the tile as stored (bf16), read back from the swizzled staging image
          const int vframes = min(p.ft, p.t_out_ph - tc.j0);
          const int nbase = tc.n_t * p.bn + ch * 64;
          if (o_stat_partials) {
            const int col = et & 63, r0 = (et >> 6) * 64, r1 = min(r0 + 64, vframes * p.v);
            const uint8_t* cp = stg + ((col & 7) << 1);
            const int cj = col >> 3;
            float s1 = 0.f, s2 = 0.f;
            for (int r = r0; r < r1; ++r) {
              const float vv2 = __bfloat162float(*reinterpret_cast<const bf16*>(cp + r * 128 + ((cj ^ (r & 7)) << 4)));
              s1 += vv2; s2 = fmaf(vv2, vv2, s2);
            }
            atomicAdd(stat_acc + nbase + col, s1);                 // two row-halves share a column
            atomicAdd(stat_acc + p.n_out + nbase + col, s2);
          }
          if (o_vsum) {
            for (int o = et; o < p.v * 64; o += 128) {
              const int v2 = o >> 6, col = o & 63, cj = col >> 3;
              float sacc = 0.f;
              for (int f2 = 0; f2 < vframes; ++f2) {
                const int r = f2 * p.v + v2;
                sacc += __bfloat162float(*reinterpret_cast<const bf16*>(stg + r * 128 + ((cj ^ (r & 7)) << 4) + ((col & 7) << 1)));
              }
              atomicAdd(vsum_acc + v2 * p.n_out + nbase + col, sacc);
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(tempty_bar + as);                // this thread is done reading accumulator `as`
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
    if (et == 0) tma_store_wait<0>();
    if (o_stat_partials || o_vsum) {
      asm volatile("bar.sync 5, 256;" ::: "memory");
      if (o_stat_partials)
        for (int i = et2; i < 2 * p.n_out; i += 256) o_stat_partials[(size_t)blockIdx.x * 2 * p.n_out + i] = stat_acc[i];
      if (o_vsum)
        for (int i = et2; i < p.v * p.n_out; i += 256) atomicAdd(o_vsum + i, vsum_acc[i]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, p.tmem_cols); }
}

constexpr int kEpSplit = EP_F32OUT | EP_BIAS | EP_BIAS2D | EP_SCALE | EP_RESID | EP_RELU | EP_DROPOUT;

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
PFN_encodeTiled get_encode_tiled() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess || !p) return nullptr;
    fn = reinterpret_cast<PFN_encodeTiled>(p);
  }
  return fn;
}

int32_t make_tmap_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                       const uint32_t* box) {
  PFN_encodeTiled enc = get_encode_tiled();
  if (!enc) { set_error("cuTensorMapEncodeTiled entry point unavailable (driver too old or no device)"); return STGCN_ERR_CUDA; }
  cuuint64_t gd[5]; cuuint64_t gs[4]; cuuint32_t bx[5]; cuuint32_t es[5];
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i < rank - 1; ++i) gs[i] = strides_bytes[i];
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d): rank %d dims %llu,%llu,%llu box %u,%u,%u", (int)r, rank,
              (unsigned long long)dims[0], (unsigned long long)dims[1], (unsigned long long)(rank > 2 ? dims[2] : 0), box[0], box[1],
              rank > 2 ? box[2] : 0);
    return STGCN_ERR_CUDA;
  }
  return STGCN_OK;
}

static int pick_bn(int n_out) {
  if (n_out <= 256) return n_out;
  if (n_out % 256 == 0) return 256;
  if (n_out % 192 == 0) return 192;
  if (n_out % 128 == 0) return 128;
  return 64;
}

static int num_sms() {
  static int n = 0;
  if (!n) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev); if (n <= 0) n = 148; }
  return n;
}

int umma_conv_gemm_grid(const ConvGeom& g) {       // CTAs the engine launches for this geometry (= number of stat partials)
  const int phases = g.dgrad ? g.stride : 1;
  const int ft = 128 / g.v, bn = pick_bn(g.n_out);
  const long long tiles = (long long)(g.n_out / bn) * phases * cdiv(g.t_out / phases, ft) * g.nb;
  return (int)std::min<long long>(tiles, num_sms());
}

bool umma_conv_gemm_supported(const ConvGeom& g) {
  if (g.c % 64 != 0 || g.n_out % 64 != 0) return false;
  if (g.v < 1 || g.v > 32 || g.taps < 1 || g.taps > kMaxTaps - 1) return false;
  if (g.stride < 1 || g.stride > 2) return false;
  if (g.dgrad && g.taps < g.stride) return false;   // every output phase needs at least one tap
  if (!g.dgrad && g.t_in % g.stride != 0) return false;
  if (g.dgrad && g.t_out % g.stride != 0) return false;
  return true;
}

static int32_t launch_conv_gemm_impl(const ConvGeom& g, const bf16* in, const bf16* w, const GemmEpilogue& ep, bf16* out, float* out_f32,
                                     cudaStream_t st) {
  KernelTimer kt_(KC_UMMA_GEMM, conv_algorithmic_flops(g), st);
  const int nterms = out_f32 ? kSplitTerms : 0;     // fp32 path: operands are [hi | mid | lo] bf16 triples, six products per channel block
  const int parts = out_f32 ? kSplitParts : 1;
  if (!umma_conv_gemm_supported(g)) { set_error("umma conv-gemm: unsupported geometry"); return STGCN_ERR_UNSUPPORTED; }
  if (ep.accumulate && !out_f32) { set_error("umma conv-gemm: accumulate epilogue unsupported"); return STGCN_ERR_UNSUPPORTED; }
  if (out_f32 && (ep.stat_partials || ep.vsum)) { set_error("umma conv-gemm: statistics epilogues need the bf16 output"); return STGCN_ERR_UNSUPPORTED; }
  if (((uintptr_t)ep.bias | (uintptr_t)ep.scale | (uintptr_t)ep.shift) & 15) {
    set_error("umma conv-gemm: bias / scale / shift must be 16-byte aligned"); return STGCN_ERR_INVALID;
  }
  UmmaParams p; memset(&p, 0, sizeof(p));
  p.rev = next_direction();
  const int s = g.stride;
  p.nb = g.nb; p.v = g.v; p.ft = 128 / g.v; p.t_out = g.t_out; p.n_out = g.n_out;
  p.bn = pick_bn(g.n_out); p.n_tiles = g.n_out / p.bn; p.c_log = g.c / 64; p.nterms = nterms; p.kblocks_c = p.c_log * (nterms ? nterms : 1);
  p.out_f32 = out_f32; p.accumulate = ep.accumulate; p.residual_f32 = out_f32 ? reinterpret_cast<const float*>(ep.residual) : nullptr;
  p.phases = g.dgrad ? s : 1;
  p.t_out_ph = g.t_out / p.phases;
  p.m_tiles_per_b = cdiv(p.t_out_ph, p.ft);
  p.total_tiles = p.n_tiles * p.phases * p.m_tiles_per_b * g.nb;
  // tap tables
  for (int ph = 0; ph < p.phases; ++ph) {
    int n = 0;
    for (int tap = 0; tap < g.taps; ++tap) {
      if (!g.dgrad) {
        int d = tap - g.pad;                              // input frame = j*s + d = s*(j + floor(d/s)) + (d mod s)
        int r = ((d % s) + s) % s;
        p.taps[ph][n].w_tap = tap; p.taps[ph][n].parity = r; p.taps[ph][n].h_off = (d - r) / s; ++n;
      } else {
        int d = ph + g.pad - tap;                         // output frame j = s*jj + ph ; input frame = (j + pad - tap)/s
        if (((d % s) + s) % s != 0) continue;
        p.taps[ph][n].w_tap = tap; p.taps[ph][n].parity = 0; p.taps[ph][n].h_off = d / s; ++n;
      }
    }
    p.ntaps[ph] = n;
    if (n == 0) { set_error("umma conv-gemm: empty tap phase"); return STGCN_ERR_UNSUPPORTED; }
  }
  p.a_bytes = (uint32_t)(p.ft * g.v * 128);
  p.b_bytes = (uint32_t)(p.bn * 128);
  p.b2_smem = (ep.bias2d && p.n_tiles == 1) ? 1 : 0;
  p.stat_partials = ep.stat_partials; p.vsum = ep.vsum;
  const size_t b2_bytes = (p.b2_smem ? (size_t)g.v * (p.bn + 4) * sizeof(float) : 0) +
                          (ep.stat_partials ? 2 * (size_t)g.n_out * sizeof(float) : 0) +
                          (ep.vsum ? (size_t)g.v * g.n_out * sizeof(float) : 0);
  const size_t budget = 220 * 1024 - (1024 /*align slack*/ + 512 /*barriers*/ + b2_bytes);
  p.nstaging = 1; p.a_stage_bytes = kATileBytes; p.halo = 0;        // staging buffers PER epilogue group
  static const bool no_halo = getenv("STGCN_NO_HALO") != nullptr;
  int hf = p.ft;
  if (!no_halo && s == 1 && p.ntaps[0] > 1) {
    // stride 1: all taps read frame-shifted windows of one (ft + span)-frame tile
    int hmin = p.taps[0][0].h_off, hmax = hmin;
    for (int i = 1; i < p.ntaps[0]; ++i) { hmin = std::min(hmin, p.taps[0][i].h_off); hmax = std::max(hmax, p.taps[0][i].h_off); }
    const int span = hmax - hmin;
    const size_t a_stage = (((size_t)span * g.v + 128) * 128 + 1023) / 1024 * 1024;   // the last tap still reads 128 rows
    const size_t w_all = (size_t)p.ntaps[0] * p.kblocks_c * p.b_bytes;
    bool ok = p.ft + span <= 256;
    int a_st = 0, w_st = 0, nstg = 1, resident = 0;
    if (ok && p.n_tiles == 1 && w_all <= 80 * 1024) {
      resident = 1; w_st = p.ntaps[0] * p.kblocks_c;
      a_st = (int)((budget - 2 * kStagingBytes - w_all) / a_stage);
      a_st = std::min(a_st, 4);
      ok = a_st >= 2;
    } else if (ok) {
      const size_t left = budget - 2 * kStagingBytes;
      a_st = 2;                                   // one halo tile feeds ntaps weight tiles: depth goes to the weight ring
      if (left < a_st * a_stage + 2 * (size_t)p.b_bytes) ok = false;
      else w_st = (int)std::min<size_t>(8, (left - a_st * a_stage) / p.b_bytes);
    }
    if (ok) {
      p.halo = 1; p.h_min = hmin; p.stages = a_st; p.w_stages = w_st; p.w_resident = resident; p.nstaging = nstg;
      p.a_stage_bytes = (uint32_t)a_stage; hf = p.ft + span; p.a_halo_bytes = (uint32_t)(hf * g.v * 128);
      for (int i = 0; i < p.ntaps[0]; ++i) p.a_off16[i] = (p.taps[0][i].h_off - hmin) * g.v * 8;
    }
  }
  size_t smem;
  if (!p.halo) {
    const size_t stage_bytes = kATileBytes + p.b_bytes;
    // several 64-column chunks per tile and a short reduction: the TMA-store drain latency, not the MMAs, paces the kernel,
    // so give the epilogue four staging buffers when the load ring stays at least four deep
    if ((budget - 4 * kStagingBytes) / stage_bytes >= 3) p.nstaging = 2;
    int stages = (int)((budget - (size_t)2 * p.nstaging * kStagingBytes) / stage_bytes);
    if (stages > 8) stages = 8;
    if (stages < 2) { set_error("umma conv-gemm: not enough shared memory for 2 stages"); return STGCN_ERR_UNSUPPORTED; }
    p.stages = stages; p.w_stages = stages;
  }
  smem = (size_t)p.stages * p.a_stage_bytes + (size_t)p.w_stages * p.b_bytes + (size_t)2 * p.nstaging * kStagingBytes + 1024 + 512 + b2_bytes;
  uint32_t cols = 32; while (cols < (uint32_t)(2 * p.bn)) cols <<= 1;
  p.tmem_cols = cols;
  p.bias = ep.bias; p.bias2d = ep.bias2d; p.scale = ep.scale; p.shift = ep.shift;
  p.residual = reinterpret_cast<const bf16*>(ep.residual); p.relu = ep.relu;
  p.dropout_p = ep.dropout_p; p.inv_keep = ep.dropout_p > 0.f ? 1.f / (1.f - ep.dropout_p) : 1.f; p.seed = ep.seed; p.offset = ep.offset; p.rng = ep.rng;

  CUtensorMap map_in, map_w, map_out;
  const uint64_t C = (uint64_t)g.c * parts, V = g.v, N = g.n_out;
  {
    // input viewed as {C, V, sp, T_in/sp, NB}; sp = conv stride in forward mode, 1 for dgrad
    const uint64_t sp = g.dgrad ? 1 : s;
    uint64_t dims[5] = {C, V, sp, (uint64_t)g.t_in / sp, (uint64_t)g.nb};
    uint64_t str[4] = {C * 2, V * C * 2, sp * V * C * 2, (uint64_t)g.t_in * V * C * 2};
    uint32_t box[5] = {64, (uint32_t)g.v, 1, (uint32_t)hf, 1};
    STGCN_TRY(make_tmap_bf16(&map_in, in, 5, dims, str, box));
  }
  {
    uint64_t dims[3] = {C, N, (uint64_t)g.taps};
    uint64_t str[2] = {C * 2, N * C * 2};
    uint32_t box[3] = {64, (uint32_t)p.bn, 1};
    STGCN_TRY(make_tmap_bf16(&map_w, w, 3, dims, str, box));
  }
  if (out_f32) map_out = map_in;       // unused: the epilogue stores fp32 rows itself
  else {
    const uint64_t ph = p.phases;
    uint64_t dims[5] = {N, V, ph, (uint64_t)g.t_out / ph, (uint64_t)g.nb};
    uint64_t str[4] = {N * 2, V * N * 2, ph * V * N * 2, (uint64_t)g.t_out * V * N * 2};
    uint32_t box[5] = {64, (uint32_t)g.v, 1, (uint32_t)p.ft, 1};
    STGCN_TRY(make_tmap_bf16(&map_out, out, 5, dims, str, box));
  }
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaSuccess;
#define STGCN_CG_ATTR(M) if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_gemm_umma_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    STGCN_CG_ATTR(0) STGCN_CG_ATTR(EP_BIAS2D) STGCN_CG_ATTR(EP_BIAS | EP_STATS) STGCN_CG_ATTR(EP_SCALE | EP_RELU)
    STGCN_CG_ATTR(EP_BIAS | EP_SCALE | EP_RELU) STGCN_CG_ATTR(EP_ALL) STGCN_CG_ATTR(kEpSplit)
#undef STGCN_CG_ATTR
    if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(smem) failed: %s", cudaGetErrorString(e)); return STGCN_ERR_CUDA; }
    attr_set = true;
  }
  const int grid = std::min(p.total_tiles, num_sms());
  const int need = (ep.bias ? EP_BIAS : 0) | (ep.bias2d ? EP_BIAS2D : 0) | (ep.scale ? EP_SCALE : 0) | (ep.residual ? EP_RESID : 0) |
                   (ep.relu ? EP_RELU : 0) | (ep.dropout_p > 0.f ? EP_DROPOUT : 0) | (ep.stat_partials ? EP_STATS : 0) | (ep.vsum ? EP_VSUM : 0);
  if (out_f32) {
    launch_k(conv_gemm_umma_kernel<kEpSplit>, grid, kUmmaThreads, smem, st, map_in, map_w, map_out, p);
    STGCN_CHECK_LAUNCH("conv_gemm_umma(split)");
    return STGCN_OK;
  }
  // the leanest instantiation that keeps every requested option
#define STGCN_CG_TRY(M) if ((need & ~(M)) == 0) { launch_k(conv_gemm_umma_kernel<M>, grid, kUmmaThreads, smem, st, map_in, map_w, map_out, p); } else
  STGCN_CG_TRY(0) STGCN_CG_TRY(EP_BIAS2D) STGCN_CG_TRY(EP_BIAS | EP_STATS) STGCN_CG_TRY(EP_SCALE | EP_RELU)
  STGCN_CG_TRY(EP_BIAS | EP_SCALE | EP_RELU) { launch_k(conv_gemm_umma_kernel<EP_ALL>, grid, kUmmaThreads, smem, st, map_in, map_w, map_out, p); }
#undef STGCN_CG_TRY
  STGCN_CHECK_LAUNCH("conv_gemm_umma");
  return STGCN_OK;
}

int32_t launch_conv_gemm_umma(const ConvGeom& g, const bf16* in, const bf16* w, const GemmEpilogue& ep, bf16* out, cudaStream_t st) {
  return launch_conv_gemm_impl(g, in, w, ep, out, nullptr, st);
}
// fp32 path: in3 [rows_in][3c] and w3 [taps][n_out][3c] hold (hi | mid | lo) bf16 splits of the fp32 operands; out is fp32 [rows_out][n_out]
int32_t launch_conv_gemm_umma_split(const ConvGeom& g, const bf16* in3, const bf16* w3, const GemmEpilogue& ep, float* out, cudaStream_t st) {
  if (!out) { set_error("umma conv-gemm (split): null output"); return STGCN_ERR_INVALID; }
  return launch_conv_gemm_impl(g, in3, w3, ep, nullptr, out, st);
}

}  // namespace stgcn
