// Host-side kernel chains of one ST-GCN block + the C-ABI entry points.
// Reference semantics: SpatialTemporalConv.forward (st_graphconv.py:85-109),
// residual_block (:60-82), SpatialConv.forward (:139-152); backward = what
// torch autograd derives for those.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include <nvtx3/nvToolsExt.h>      // header-only (dlopens the tools' injection library when a profiler is attached)

#include "kernels.cuh"

namespace stgcn {

// NVTX range around one library call ("stgcn fwd 64->128 s2 T300 N64"), on when STGCN_NVTX=1: nsys / ncu --nvtx then show the
// kernels grouped per ST-GCN block and direction.  Off by default (no call into the tools library at all).
struct NvtxRange {
  bool on;
  NvtxRange(const char* what, const StgcnBlockDesc* d) {
    static const bool enabled = getenv("STGCN_NVTX") != nullptr;
    on = enabled && d != nullptr;
    if (!on) return;
    char buf[96];
    snprintf(buf, sizeof(buf), "stgcn %s %d->%d s%d T%d V%d N%d %s", what, d->c_in, d->c_out, d->stride, d->t, d->v, d->n,
             d->dtype == STGCN_BF16 ? "bf16" : "f32");
    nvtxRangePushA(buf);
  }
  ~NvtxRange() { if (on) nvtxRangePop(); }
};

std::atomic<uint64_t> g_launches{0};
int next_direction() {
  static const bool on = getenv("STGCN_NO_PINGPONG") == nullptr;
  static thread_local int dir = 0;
  if (!on) return 0;
  dir ^= 1;
  return dir;
}
bool pdl_enabled() { static const bool on = getenv("STGCN_NO_PDL") == nullptr; return on; }
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
}

// ---- bump allocator over caller-provided device memory ---------------------
struct Arena {
  char* base; size_t cap; size_t off = 0; bool dry;
  Arena(void* b, size_t c) : base((char*)b), cap(c), dry(b == nullptr) {}
  template <class U> U* get(size_t count) {
    size_t bytes = align_up(count * sizeof(U), 256);
    size_t o = off; off += bytes;
    if (dry) return reinterpret_cast<U*>((uintptr_t)256);   // never dereferenced
    return reinterpret_cast<U*>(base + o);
  }
  bool ok() const { return dry || off <= cap; }
};

struct Dims {
  int n, ci, co, t, to, v, k, gamma, stride, pad;
  long long rows_in, rows_out, frames_in;
};
static Dims dims_of(const StgcnBlockDesc& d) {
  Dims m;
  m.n = d.n; m.ci = d.c_in; m.co = d.c_out; m.t = d.t; m.v = d.v; m.k = d.k; m.gamma = d.gamma; m.stride = d.stride; m.pad = d.pad;
  m.to = (d.t + 2 * d.pad - d.gamma) / d.stride + 1;
  m.rows_in = (long long)d.n * d.t * d.v; m.rows_out = (long long)d.n * m.to * d.v; m.frames_in = (long long)d.n * d.t;
  return m;
}

static int32_t validate(const StgcnBlockDesc* d) {
  if (!d) { set_error("null descriptor"); return STGCN_ERR_INVALID; }
  if (d->n < 1 || d->c_in < 1 || d->c_out < 1 || d->t < 1 || d->v < 1 || d->k < 1) { set_error("non-positive dimension"); return STGCN_ERR_INVALID; }
  if (d->v > 32 || d->k > 4) { set_error("V=%d K=%d unsupported (V<=32, K<=4)", d->v, d->k); return STGCN_ERR_UNSUPPORTED; }
  if (d->gamma < 1 || d->gamma > 15 || d->stride < 1 || d->pad < 0) { set_error("bad temporal conv geometry"); return STGCN_ERR_INVALID; }
  if (d->t + 2 * d->pad < d->gamma) { set_error("T=%d too short for gamma=%d pad=%d", d->t, d->gamma, d->pad); return STGCN_ERR_INVALID; }
  if (d->dtype != STGCN_F32 && d->dtype != STGCN_BF16) { set_error("dtype %d unsupported", d->dtype); return STGCN_ERR_UNSUPPORTED; }
  if (d->dropout_p < 0.f || d->dropout_p >= 1.f) { set_error("dropout_p out of range"); return STGCN_ERR_INVALID; }
  if (d->residual == STGCN_RES_IDENTITY && (d->c_in != d->c_out || d->stride != 1)) { set_error("identity residual needs C_in==C_out, stride 1"); return STGCN_ERR_INVALID; }
  if (d->residual == STGCN_RES_IDENTITY) {
    int to = (d->t + 2 * d->pad - d->gamma) / d->stride + 1;
    if (to != d->t) { set_error("identity residual needs T_out==T"); return STGCN_ERR_INVALID; }
  }
  if (d->residual < 0 || d->residual > 2) { set_error("bad residual mode"); return STGCN_ERR_INVALID; }
  return STGCN_OK;
}

// saved-for-backward layout
static bool fold_ok_desc(const StgcnBlockDesc& d, const Dims& m);
static bool smallc_ok_desc(const StgcnBlockDesc& d, const Dims& m);
constexpr int kSmallLd = 64;      // padded K*C_in of the small-C_in path
// operands the folded-BatchNorm spatial path prepares from (A, W, BN-1 coefficients): block-diagonal adjacency, W.diag(a) and its
// transpose, bias vectors.  Forward carves them from the saved buffer when there is one, so backward reuses them as they are.
struct FoldPrep { void* prep; bf16* w2p; bf16* w2pT; float* wb; float* wbb; bool valid; };
static FoldPrep carve_fold_prep(Arena& a, const Dims& m) {
  FoldPrep f;
  const int KCi = m.k * m.ci;
  f.prep = a.get<uint8_t>(umma_agg_prep_bytes(m.k));
  f.w2p = a.get<bf16>((size_t)m.co * KCi);
  f.w2pT = a.get<bf16>((size_t)m.co * KCi);
  f.wb = a.get<float>((size_t)m.k * m.co);
  f.wbb = a.get<float>((size_t)m.k * m.co);
  f.valid = true;
  return f;
}
template <class T> struct Saved { uint64_t* rng; float* stats1; float* stats2; T* S; T* U; T* XA; FoldPrep fp; };
// Inference with frozen parameters (stgcn_block_prepare / stgcn_block_forward_prepared): everything that depends only on the
// parameters -- eval BatchNorm coefficient tables, folded spatial operands, joint-bias table, packed temporal weights.
struct Prepared { float* stats1; float* stats2; FoldPrep fp; float* b2d; bf16* wt; };
static Prepared carve_prepared(Arena& a, const Dims& m) {
  Prepared q;
  q.stats1 = a.get<float>(4 * m.ci);
  q.stats2 = a.get<float>(4 * m.co);
  q.fp = carve_fold_prep(a, m);
  q.b2d = a.get<float>((size_t)m.v * m.co);
  q.wt = a.get<bf16>((size_t)m.gamma * m.co * m.co);
  return q;
}
template <class T> static Saved<T> carve_saved(Arena& a, const StgcnBlockDesc& d, const Dims& m) {
  Saved<T> s;
  s.rng = a.get<uint64_t>(2);          // dropout Philox {seed, offset} this call used (replayed by backward)
  s.stats1 = a.get<float>(4 * m.ci);
  s.stats2 = a.get<float>(4 * m.co);
  s.S = a.get<T>(m.rows_in * m.co);
  s.U = d.residual == STGCN_RES_NONE ? a.get<T>(m.rows_out * m.co) : nullptr;
  // folded-BatchNorm path: the aggregated raw input (rows x K*C_in) is kept so backward does not re-aggregate
  s.XA = fold_ok_desc(d, m) ? a.get<T>(m.rows_in * m.k * m.ci) : (smallc_ok_desc(d, m) ? a.get<T>(m.rows_in * kSmallLd) : nullptr);
  s.fp.valid = false;
  if (fold_ok_desc(d, m)) s.fp = carve_fold_prep(a, m);
  return s;
}

// weight-only operands produced ahead of the step by stgcn_train_prep_multi (train_prep.cu)
struct TrainPrep { bf16* wt_f; bf16* wt_d; void* big; bool has_big; };
static bool train_prep_supported(const StgcnBlockDesc& d, const Dims& m);
static TrainPrep carve_train_prep(Arena& a, const Dims& m) {
  TrainPrep t;
  t.wt_f = a.get<bf16>((size_t)m.gamma * m.co * m.co);
  t.wt_d = a.get<bf16>((size_t)m.gamma * m.co * m.co);
  t.has_big = umma_agg_supported(m.v, m.ci, m.k);
  t.big = t.has_big ? a.get<uint8_t>(umma_agg_prep_bytes(m.k)) : nullptr;
  return t;
}

// ---- side lane: a second stream per device for work that is off the data-gradient path of a chain (per-joint sums, weight-gradient
// layout change).  Forked from / joined to the caller's stream with events inside one library call, so the caller never sees it;
// under CUDA-graph capture the fork/join become graph edges.  STGCN_NO_SIDE=1 keeps everything on the caller's stream.
constexpr int kLaneEvents = 6;
struct SideLane { cudaStream_t s = nullptr; cudaEvent_t ev[kLaneEvents] = {}; bool ok = false; };
static SideLane* side_lane() {
  static const bool off = getenv("STGCN_NO_SIDE") != nullptr;
  if (off) return nullptr;
  static thread_local SideLane lanes[64];      // per host thread and device: calls from different threads never share events
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  SideLane& l = lanes[dev];
  if (!l.s) {
    if (cudaStreamCreateWithFlags(&l.s, cudaStreamNonBlocking) != cudaSuccess) { l.s = nullptr; cudaGetLastError(); return nullptr; }
    l.ok = true;
    for (int i = 0; i < kLaneEvents; ++i)
      if (cudaEventCreateWithFlags(&l.ev[i], cudaEventDisableTiming) != cudaSuccess) { l.ok = false; cudaGetLastError(); }
  }
  return l.ok ? &l : nullptr;
}

// ---- SyncBN: global-batch BatchNorm sums through the caller's all-reduce (StgcnBlockParams.sync_*) --------------------------------
struct SyncCtx {
  float* buf = nullptr; void (*fn)(void*, void*) = nullptr; void* user = nullptr; int world = 1; cudaStream_t st = nullptr;
  bool on() const { return fn != nullptr && buf != nullptr && world > 1; }
  void allreduce() const { fn(user, (void*)st); }          // enqueues on st, does not synchronise
};
// a half-filled link is a caller bug that would silently fall back to per-replica statistics on SOME ranks: refuse it
static int32_t sync_check(const StgcnBlockParams& P) {
  if (P.sync_allreduce && P.sync_world > 1 && !P.sync_buf) { set_error("SyncBN link: sync_allreduce is set but sync_buf is null"); return STGCN_ERR_INVALID; }
  if (P.sync_world < 0) { set_error("SyncBN link: negative sync_world"); return STGCN_ERR_INVALID; }
  return STGCN_OK;
}
static SyncCtx sync_of(const StgcnBlockDesc& d, const StgcnBlockParams& P, cudaStream_t st, bool dry) {
  SyncCtx s;
  if (!dry && d.training && P.sync_allreduce && P.sync_buf && P.sync_world > 1) {
    s.buf = P.sync_buf; s.fn = P.sync_allreduce; s.user = P.sync_user; s.world = (int)P.sync_world; s.st = st;
  }
  return s;
}
// training-mode statistics -> coefficient table; over all ranks when sync is on
static int32_t finalize_stats(const SyncCtx& sy, const float* partials, int nblk, int C, double rows, const float* gamma, const float* beta,
                              float* rm, float* rv, long long* nbt, float eps, float momentum, float* stats, cudaStream_t st) {
  if (!sy.on()) return launch_bn_finalize(partials, nblk, C, rows, gamma, beta, rm, rv, nbt, eps, momentum, stats, st);
  STGCN_TRY(launch_reduce_partials(partials, nblk, C, sy.buf, st));
  sy.allreduce();
  return launch_bn_finalize(sy.buf, 1, C, rows * sy.world, gamma, beta, rm, rv, nbt, eps, momentum, stats, st);
}
// BatchNorm-backward sums: parameter gradients from the LOCAL sums, the sums the apply pass uses over all ranks when sync is on
static int32_t finalize_bwd_sums(const SyncCtx& sy, const float* partials, int nblk, int C, float* sums_local, float* dgamma, float* dbeta,
                                 cudaStream_t st, const float** sums_use, double* count_scale) {
  float* dst = sy.on() ? sy.buf : sums_local;
  STGCN_TRY(launch_bn_bwd_finalize(partials, nblk, C, dst, dgamma, dbeta, st));
  if (sy.on()) sy.allreduce();
  *sums_use = dst; *count_scale = sy.on() ? (double)sy.world : 1.0;
  return STGCN_OK;
}

// ---- GEMM dispatch: tcgen05 engine for bf16 when the shape fits, else CUDA cores ----
template <class T> struct Engine {
  const StgcnBlockDesc& d; Arena& ws; cudaStream_t st;
  bool want_umma() const { return std::is_same<T, bf16>::value && !(d.flags & STGCN_FLAG_FORCE_SIMT); }
  // fp32 path on the tensor cores: both GEMM operands are split into (hi | mid | lo) bf16 triples by an element-wise pre-pass and
  // the tcgen05 engines run six products per reduction step with fp32 accumulation in TMEM (error ~1e-6 of the operand scale, an
  // order inside the fp32 path's 1e-4).  STGCN_FLAG_FORCE_SIMT or STGCN_FP32_SIMT=1 keep the CUDA-core GEMMs (A/B, debugging).
  bool want_tc32() const {
    static const bool off = getenv("STGCN_FP32_SIMT") != nullptr;
    return std::is_same<T, float>::value && !off && !(d.flags & STGCN_FLAG_FORCE_SIMT);
  }
  int32_t gemm(const ConvGeom& g, const T* in, const float* wsrc, const WIndex& wi, const GemmEpilogue& ep, T* out) {
    if constexpr (std::is_same<T, float>::value) {
      if (want_tc32() && umma_conv_gemm_supported(g) && !ep.stat_partials && !ep.vsum) {
        const long long rows_in = (long long)g.nb * g.t_in * g.v;
        bf16* in3 = ws.template get<bf16>((size_t)rows_in * 3 * g.c);
        float* wpf = ws.template get<float>((size_t)g.taps * g.n_out * g.c);
        bf16* wp3 = ws.template get<bf16>((size_t)g.taps * g.n_out * 3 * g.c);
        if (ws.dry) return STGCN_OK;
        if (!ws.ok()) { set_error("workspace too small"); return STGCN_ERR_WORKSPACE; }
        STGCN_TRY(launch_split_bf16(in, rows_in, g.c, in3, st));
        STGCN_TRY(launch_pack_weights<float>(wsrc, wi, g.taps, g.n_out, g.c, wpf, st, 0));
        STGCN_TRY(launch_split_bf16(wpf, (long long)g.taps * g.n_out, g.c, wp3, st));
        return launch_conv_gemm_umma_split(g, in3, wp3, ep, out, st);
      }
    }
    if (want_umma() && umma_conv_gemm_supported(g)) {
      bf16* wp = ws.template get<bf16>((size_t)g.taps * g.n_out * g.c);
      if (ws.dry) return STGCN_OK;
      if (!ws.ok()) { set_error("workspace too small"); return STGCN_ERR_WORKSPACE; }
      STGCN_TRY(launch_pack_weights<bf16>(wsrc, wi, g.taps, g.n_out, g.c, wp, st));
      return launch_conv_gemm_umma(g, reinterpret_cast<const bf16*>(in), wp, ep, reinterpret_cast<bf16*>(out), st);
    }
    float* wp = ws.template get<float>((size_t)g.taps * g.n_out * g.c);
    if (ws.dry) return STGCN_OK;
    if (!ws.ok()) { set_error("workspace too small"); return STGCN_ERR_WORKSPACE; }
    STGCN_TRY(launch_pack_weights<float>(wsrc, wi, g.taps, g.n_out, g.c, wp, st, std::is_same<T, bf16>::value ? 1 : 0));
    return launch_conv_gemm_simt<T>(g, in, wp, ep, out, st);
  }
  // Weights packed ahead of their GEMM (they depend on parameters only): on the side lane, beside whatever the caller's stream runs
  // first, or -- without a lane -- right here.  The buffer is carved in every mode so the dry run plans the same arena.
  // Use with gemm_packed() after waiting on `done` (recorded on the side lane).
  bf16* prepack(const ConvGeom& g, const float* wsrc, const WIndex& wi, SideLane* lane, cudaEvent_t fork, cudaEvent_t done, int32_t* rc) {
    *rc = STGCN_OK;
    bf16* wp = ws.template get<bf16>((size_t)g.taps * g.n_out * g.c);
    if (ws.dry || !ws.ok()) return wp;
    if (lane) {
      cudaEventRecord(fork, st); cudaStreamWaitEvent(lane->s, fork, 0);
      *rc = launch_pack_weights<bf16>(wsrc, wi, g.taps, g.n_out, g.c, wp, lane->s);
      cudaEventRecord(done, lane->s);
    } else {
      *rc = launch_pack_weights<bf16>(wsrc, wi, g.taps, g.n_out, g.c, wp, st);
    }
    return wp;
  }
  // true when gemm() runs on the tcgen05 engine, whose epilogue can also produce BatchNorm statistics / per-joint sums
  bool gemm_on_umma(const ConvGeom& g) const { return want_umma() && umma_conv_gemm_supported(g); }
  // same, with the weights already packed as bf16 [taps][n_out][c] by the caller (folded BatchNorm path)
  int32_t gemm_packed(const ConvGeom& g, const T* in, const bf16* wp, const GemmEpilogue& ep, T* out) {
    if (ws.dry) return STGCN_OK;
    if constexpr (std::is_same<T, bf16>::value) {
      if (umma_conv_gemm_supported(g)) return launch_conv_gemm_umma(g, in, wp, ep, out, st);
    }
    set_error("gemm_packed: geometry not supported by the tcgen05 engine"); return STGCN_ERR_UNSUPPORTED;
  }
  // dw (reference layout, overwritten) = sum over rows of dout (x) in
  bool wgrad_on_umma(const ConvGeom& g) const { return want_umma() && umma_wgrad_supported(g) && (g.n_out == 64 || g.n_out % 128 == 0); }
  // packed_pre: split-K accumulator carved by the caller (packed_zero: and already cleared by an earlier kernel of the chain)
  int32_t wgrad(const ConvGeom& g, const T* dout, const T* in, const WIndex& wi, float* dw, size_t dw_floats, int c_valid = -1,
                float* packed_pre = nullptr, bool packed_zero = false, cudaStream_t unpack_st = nullptr, cudaEvent_t unpack_ev = nullptr) {
    if constexpr (std::is_same<T, float>::value) {
      if (want_tc32() && umma_wgrad_supported(g) && (g.n_out == 64 || g.n_out % 128 == 0) && c_valid < 0) {
        const long long rows_out = (long long)g.nb * g.t_out * g.v, rows_in = (long long)g.nb * g.t_in * g.v;
        bf16* d3 = ws.template get<bf16>((size_t)rows_out * 3 * g.n_out);
        bf16* i3 = ws.template get<bf16>((size_t)rows_in * 3 * g.c);
        float* packed = ws.template get<float>(umma_wgrad_workspace_floats(g));
        if (ws.dry) return STGCN_OK;
        if (!ws.ok()) { set_error("workspace too small"); return STGCN_ERR_WORKSPACE; }
        STGCN_TRY(launch_split_bf16(dout, rows_out, g.n_out, d3, st));
        STGCN_TRY(launch_split_bf16(in, rows_in, g.c, i3, st));
        return launch_wgrad_umma(g, d3, i3, wi, packed, dw, st, -1, false, nullptr, nullptr, 1);
      }
    }
    if (wgrad_on_umma(g)) {
      float* packed = packed_pre ? packed_pre : ws.template get<float>(umma_wgrad_workspace_floats(g));
      if (ws.dry) return STGCN_OK;
      if (!ws.ok()) { set_error("workspace too small"); return STGCN_ERR_WORKSPACE; }
      return launch_wgrad_umma(g, reinterpret_cast<const bf16*>(dout), reinterpret_cast<const bf16*>(in), wi, packed, dw, st, c_valid,
                               packed_pre && packed_zero, unpack_st, unpack_ev);
    }
    if (c_valid >= 0) { set_error("wgrad: padded reduction operand needs the tcgen05 engine"); return STGCN_ERR_UNSUPPORTED; }
    if (ws.dry) return STGCN_OK;
    cudaMemsetAsync(dw, 0, sizeof(float) * dw_floats, st);
    return launch_wgrad_simt<T>(g, dout, in, wi, dw, st);
  }
};

static ConvGeom geom(int nb, int t_in, int t_out, int v, int c, int n_out, int taps, int stride, int pad, int dgrad) {
  ConvGeom g; g.nb = nb; g.t_in = t_in; g.t_out = t_out; g.v = v; g.c = c; g.n_out = n_out; g.taps = taps; g.stride = stride; g.pad = pad; g.dgrad = dgrad;
  return g;
}
static WIndex widx(int ndiv, long long snh, long long snl, int cdiv_, long long sch, long long scl, long long st) {
  WIndex w; w.ndiv = ndiv; w.snh = snh; w.snl = snl; w.cdiv = cdiv_; w.sch = sch; w.scl = scl; w.st = st; return w;
}
constexpr int kNoDiv = 1 << 30;
static bool train_prep_supported(const StgcnBlockDesc& d, const Dims& m) {
  if (d.dtype != STGCN_BF16 || (d.flags & STGCN_FLAG_FORCE_SIMT) || d.residual != STGCN_RES_NONE) return false;
  return umma_conv_gemm_supported(geom(m.n, m.t, m.to, m.v, m.co, m.co, m.gamma, m.stride, m.pad, 0)) &&
         umma_conv_gemm_supported(geom(m.n, m.to, m.t, m.v, m.co, m.co, m.gamma, m.stride, m.pad, 1));
}

// ---------------------------------------------------------------------------
// spatial graph convolution:  S = sum_k (A_k h) W_k^T + bias through A
// ---------------------------------------------------------------------------
// BatchNorm-1 can be folded into the 1x1 weights when nothing non-linear sits between it and the aggregation
// (default block, st_graphconv.py:98-99) and every contraction of the spatial stage runs on the tcgen05 engine.
static bool fold_shapes_ok(const Dims& m) {
  if (!umma_agg_supported(m.v, m.ci, m.k) || m.k * m.co > 1024) return false;
  const int KCi = m.k * m.ci;
  return umma_conv_gemm_supported(geom(m.n, m.t, m.t, m.v, KCi, m.co, 1, 1, 0, 0)) &&
         umma_conv_gemm_supported(geom(m.n, m.t, m.t, m.v, m.co, KCi, 1, 1, 0, 0));
}
static bool fold_ok_desc(const StgcnBlockDesc& d, const Dims& m) {       // same decision from the descriptor alone (block chains)
  return d.dtype == STGCN_BF16 && !(d.flags & STGCN_FLAG_FORCE_SIMT) && d.residual == STGCN_RES_NONE && fold_shapes_ok(m);
}
// small C_in (block 0): K*C_in <= 64 is zero-padded to one 64-channel reduction block so GEMM-1, its weight gradient and
// dXA run on the tcgen05 engine; the tiny aggregation itself stays on CUDA cores.
static bool smallc_shapes_ok(const Dims& m) {
  if (m.ci % 64 == 0 || m.k * m.ci > kSmallLd || m.co % 64 != 0 || m.k * m.co > 1024) return false;
  const ConvGeom g1 = geom(m.n, m.t, m.t, m.v, kSmallLd, m.co, 1, 1, 0, 0);
  return umma_conv_gemm_supported(g1) && umma_conv_gemm_supported(geom(m.n, m.t, m.t, m.v, m.co, kSmallLd, 1, 1, 0, 0)) &&
         umma_wgrad_supported(g1) && (m.co == 64 || m.co % 128 == 0);
}
static bool smallc_ok_desc(const StgcnBlockDesc& d, const Dims& m) {
  return d.dtype == STGCN_BF16 && !(d.flags & STGCN_FLAG_FORCE_SIMT) && d.residual == STGCN_RES_NONE && smallc_shapes_ok(m);
}
template <class T> static bool smallc_ok(const Engine<T>& e, const Dims& m, const float* aff_a, int relu1) {
  if (!std::is_same<T, bf16>::value || !aff_a || relu1 || !e.want_umma()) return false;
  return smallc_shapes_ok(m);
}
template <class T> static bool fold_ok(const Engine<T>& e, const Dims& m, const float* aff_a, int relu1) {
  if (!std::is_same<T, bf16>::value || !aff_a || relu1 || !e.want_umma()) return false;
  return fold_shapes_ok(m);
}
template <class T> struct Bn1Fold {      // backward of the folded BatchNorm-1 (spatial_bwd fills dgamma / dbeta / dx itself)
  const float* stats; int training; float* dgamma; float* dbeta; T* dx; bool done; const T* xa_saved;
  const FoldPrep* prep_saved;   // operands prepared (and kept) by forward, or null
  float* t1_ready;      // per-joint column sums of dS already produced by the temporal dgrad epilogue (or null)
  float* t1_buf = nullptr;      // per-joint sum buffer carved by the caller (folded path)
  float* packed_m = nullptr;    // split-K accumulator of the mixing weight gradient carved by the caller (folded path)
  bool zeroed = false;          // dA, t1_buf and packed_m were already cleared by an earlier kernel of the chain
  bool bias_grads_done = false; // dW_bias and the rank-1 part of dA were already issued by the caller (side lane)
  const SyncCtx* sync = nullptr; // SyncBN: the dx-correction coefficients come from sums over all ranks
};
template <class T>
static int32_t spatial_fwd(Engine<T>& e, const Dims& m, const T* x, const float* aff_a, const float* aff_b, int relu1,
                           const float* A, const float* W, const float* Wb, T* S, T* xa_keep = nullptr,
                           const FoldPrep* keep_prep = nullptr, const float* b2d_ready = nullptr, void* big_ext = nullptr) {
  // big_ext: block-diagonal adjacency operand already built for this step (stgcn_train_prep_multi); folded path only
  Arena& ws = e.ws;
  T* XA = xa_keep ? xa_keep : ws.template get<T>(m.rows_in * m.k * m.ci);
  float* b2d = ws.template get<float>((size_t)m.v * m.co);
  if (b2d_ready) b2d = const_cast<float*>(b2d_ready);      // operands prepared ahead of time (keep_prep holds the rest)
  bool tc_agg = false;
  if constexpr (std::is_same<T, bf16>::value) tc_agg = e.want_umma() && umma_agg_supported(m.v, m.ci, m.k);
  if constexpr (std::is_same<T, bf16>::value) {
    if (smallc_ok<T>(e, m, aff_a, relu1)) {
      // small C_in: CUDA-core aggregation of the raw input into zero-padded 64-channel rows, then the folded-BN GEMM-1
      T* XAp = xa_keep ? xa_keep : ws.template get<T>(m.rows_in * kSmallLd);
      bf16* w2p = ws.template get<bf16>((size_t)m.co * kSmallLd);
      bf16* w2pT = ws.template get<bf16>((size_t)m.co * kSmallLd);
      float* wb = ws.template get<float>((size_t)m.k * m.co);
      float* wbb = ws.template get<float>((size_t)m.k * m.co);
      float* rs = ws.template get<float>((size_t)m.k * m.v);
      if (ws.dry) return STGCN_OK;
      if (!ws.ok()) return STGCN_ERR_WORKSPACE;
      STGCN_TRY(launch_adjacency_rowsum(A, m.k, m.v, rs, e.st));
      STGCN_TRY(launch_bn1_fold(W, Wb, A, rs, aff_a, aff_b, m.k, m.v, m.co, m.ci, w2p, w2pT, wb, wbb, b2d, e.st, kSmallLd));
      STGCN_TRY(launch_agg_small_fwd<T>(x, m.frames_in, m.v, m.ci, m.k, kSmallLd, A, XAp, e.st));
      GemmEpilogue ep; ep.bias2d = b2d;
      ConvGeom g1 = geom(m.n, m.t, m.t, m.v, kSmallLd, m.co, 1, 1, 0, 0); g1.c_valid = m.k * m.ci;
      return e.gemm_packed(g1, XAp, w2p, ep, S);
    }
    if (fold_ok<T>(e, m, aff_a, relu1)) {
      // BatchNorm-1 folded into the 1x1 weights: aggregate the RAW input, GEMM-1 against W.diag(a), shift through bias2d
      const int KCi = m.k * m.ci;
      const FoldPrep fp = (keep_prep && keep_prep->valid) ? *keep_prep : carve_fold_prep(ws, m);
      void* prep = big_ext ? big_ext : fp.prep; bf16* w2p = fp.w2p; bf16* w2pT = fp.w2pT; float* wb = fp.wb; float* wbb = fp.wbb;
      float* rsum2 = ws.template get<float>((size_t)2 * m.k * m.v);
      float* bvec2 = ws.template get<float>((size_t)2 * m.k * m.co);
      if (ws.dry) return STGCN_OK;
      if (!ws.ok()) return STGCN_ERR_WORKSPACE;
      // Single-kernel aggregation + mixing: default when XA is not wanted by a backward pass (inference: no XA round trip through
      // HBM, measured 0.78 vs 0.92 ms per 32 sequences); with the XA store it only ties the two-kernel pair, which then stays the
      // default.  STGCN_FUSED_SPATIAL=1 / =0 forces it on / off.
      static const char* fused_env = getenv("STGCN_FUSED_SPATIAL");
      const bool want_fused = fused_env ? (fused_env[0] != '0') : (xa_keep == nullptr);
      const bool fused = want_fused && umma_spatial_supported(m.v, m.ci, m.co, m.k);
      if (!b2d_ready) {
        if (!big_ext) STGCN_TRY(launch_build_big_adjacency(A, m.k, m.v, prep, e.st));
        STGCN_TRY(launch_bn1_fold(W, Wb, A, umma_agg_rowsum_q(prep, m.k), aff_a, aff_b, m.k, m.v, m.co, m.ci, w2p, w2pT, wb, wbb, b2d, e.st, 0,
                                  rsum2, bvec2));
      }
      if (fused)     // one kernel: aggregation -> (TMEM -> bf16 smem) -> channel mixing; XA goes to HBM only when backward wants it
        return launch_spatial_fused(x, m.n, m.t, m.v, m.ci, m.co, m.k, umma_agg_big(prep, m.k), w2p, b2d, xa_keep, S, e.st);
      STGCN_TRY(launch_agg_fwd_umma(x, m.n, m.t, m.v, m.ci, m.k, prep, nullptr, nullptr, XA, e.st));
      GemmEpilogue ep; ep.bias2d = b2d;
      return e.gemm_packed(geom(m.n, m.t, m.t, m.v, KCi, m.co, 1, 1, 0, 0), XA, w2p, ep, S);
    }
  }
  if (tc_agg) {
    // tensor-core aggregation: BN-1 affine rides in the epilogue; a ReLU in between (residual block) needs h stored
    void* prep = ws.template get<uint8_t>(umma_agg_prep_bytes(m.k));
    T* P1 = relu1 ? ws.template get<T>(m.rows_in * m.ci) : nullptr;
    if (!ws.dry) {
      if (!ws.ok()) return STGCN_ERR_WORKSPACE;
      if constexpr (std::is_same<T, bf16>::value) {
        STGCN_TRY(launch_build_big_adjacency(A, m.k, m.v, prep, e.st));
        if (relu1) {
          STGCN_TRY(launch_affine_act<T>(x, m.rows_in, m.ci, aff_a, aff_b, 1, 0.f, 0, 0, P1, e.st));
          STGCN_TRY(launch_agg_fwd_umma(P1, m.n, m.t, m.v, m.ci, m.k, prep, nullptr, nullptr, XA, e.st));
        } else {
          STGCN_TRY(launch_agg_fwd_umma(x, m.n, m.t, m.v, m.ci, m.k, prep, aff_a, aff_b, XA, e.st));
        }
      }
      STGCN_TRY(launch_bias2d(A, Wb, m.k, m.v, m.co, b2d, e.st));
    }
  } else if (!ws.dry) {
    if (!ws.ok()) return STGCN_ERR_WORKSPACE;
    STGCN_TRY(launch_agg_fwd<T>(x, m.frames_in, m.v, m.ci, m.k, A, aff_a, aff_b, relu1, XA, e.st));
    STGCN_TRY(launch_bias2d(A, Wb, m.k, m.v, m.co, b2d, e.st));
  }
  GemmEpilogue ep; ep.bias2d = b2d;
  // W2[n=co][c=k*Ci+ci] = W[(k*Co+co)*Ci + ci]
  WIndex wi = widx(kNoDiv, 0, m.ci, m.ci, (long long)m.co * m.ci, 1, 0);
  return e.gemm(geom(m.n, m.t, m.t, m.v, m.k * m.ci, m.co, 1, 1, 0, 0), XA, W, wi, ep, S);
}

// gradients of the above given dS; dx_pre = d/d(h) masked by relu1 (NOT yet through the BN)
template <class T>
static int32_t spatial_bwd(Engine<T>& e, const Dims& m, const T* x, const float* aff_a, const float* aff_b, int relu1,
                           const float* A, const float* W, const float* Wb, const T* dS, T* dh, float* dA, float* dW, float* dWb,
                           Bn1Fold<T>* bf = nullptr) {
  Arena& ws = e.ws;
  const int KCi = m.k * m.ci;
  T* XA = ws.template get<T>(m.rows_in * KCi);
  T* dXA = ws.template get<T>(m.rows_in * KCi);
  float* t1 = (bf && bf->t1_ready) ? bf->t1_ready : ((bf && bf->t1_buf) ? bf->t1_buf : ws.template get<float>((size_t)m.v * m.co));
  const bool have_t1 = bf && bf->t1_ready;
  if constexpr (std::is_same<T, bf16>::value) {
    if (bf && !bf->dx && bf->xa_saved && smallc_ok<T>(e, m, aff_a, relu1)) {
      // small C_in, no input gradient wanted (first block): everything through the padded 64-channel operand kept by forward
      bf->done = true;
      T* dXAp = ws.template get<T>(m.rows_in * kSmallLd);
      bf16* w2p = ws.template get<bf16>((size_t)m.co * kSmallLd);
      bf16* w2pT = ws.template get<bf16>((size_t)m.co * kSmallLd);
      float* wb = ws.template get<float>((size_t)m.k * m.co);
      float* wbb = ws.template get<float>((size_t)m.k * m.co);
      float* b2d = ws.template get<float>((size_t)m.v * m.co);
      float* rs = ws.template get<float>((size_t)m.k * m.v);
      float* coef = ws.template get<float>((size_t)3 * m.ci);
      if (!ws.dry) {
        if (!ws.ok()) return STGCN_ERR_WORKSPACE;
        STGCN_TRY(launch_adjacency_rowsum(A, m.k, m.v, rs, e.st));
        STGCN_TRY(launch_bn1_fold(W, Wb, A, rs, aff_a, aff_b, m.k, m.v, m.co, m.ci, w2p, w2pT, wb, wbb, b2d, e.st, kSmallLd));
        cudaMemsetAsync(dA, 0, sizeof(float) * (size_t)m.k * m.v * m.v, e.st);
        if (!have_t1) {
          cudaMemsetAsync(t1, 0, sizeof(float) * (size_t)m.v * m.co, e.st);
          STGCN_TRY(launch_vcolsum<T>(dS, m.frames_in, m.v, m.co, t1, e.st));
        }
      }
      STGCN_TRY(e.wgrad(geom(m.n, m.t, m.t, m.v, kSmallLd, m.co, 1, 1, 0, 0), dS, bf->xa_saved,
                        widx(kNoDiv, 0, m.ci, m.ci, (long long)m.co * m.ci, 1, 0), dW, (size_t)m.k * m.co * m.ci, KCi));
      GemmEpilogue ep;
      ConvGeom gx = geom(m.n, m.t, m.t, m.v, m.co, kSmallLd, 1, 1, 0, 0); gx.n_valid = KCi;
      STGCN_TRY(e.gemm_packed(gx, dS, w2pT, ep, dXAp));
      if (ws.dry) return STGCN_OK;
      STGCN_TRY(launch_agg_dA<T>(dXAp, x, m.frames_in, m.v, m.ci, m.k, nullptr, nullptr, 0, dA, e.st, kSmallLd));
      STGCN_TRY(launch_bn1_fold_grads(W, dW, bf->stats, t1, rs, m.k, m.v, m.co, m.ci, bf->training, (double)m.rows_in, bf->dgamma,
                                      bf->dbeta, coef, e.st));
      return launch_spatial_bias_grads(t1, A, Wb, m.k, m.v, m.co, dWb, dA, e.st, wbb);
    }
    if (bf && fold_ok<T>(e, m, aff_a, relu1)) {
      bf->done = true;
      const bool have_prep = bf->prep_saved && bf->prep_saved->valid;
      const FoldPrep fp = have_prep ? *bf->prep_saved : carve_fold_prep(ws, m);
      void* prep = fp.prep; bf16* w2p = fp.w2p; bf16* w2pT = fp.w2pT; float* wb = fp.wb; float* wbb = fp.wbb;
      float* b2d = ws.template get<float>((size_t)m.v * m.co);
      float* coef = ws.template get<float>((size_t)3 * m.ci);
      if (!ws.dry) {
        if (!ws.ok()) return STGCN_ERR_WORKSPACE;
        if (!have_prep) {
          STGCN_TRY(launch_build_big_adjacency(A, m.k, m.v, prep, e.st));
          STGCN_TRY(launch_bn1_fold(W, Wb, A, umma_agg_rowsum_q(prep, m.k), aff_a, aff_b, m.k, m.v, m.co, m.ci, w2p, w2pT, wb, wbb, b2d, e.st));
        }
        if (!bf->xa_saved) STGCN_TRY(launch_agg_fwd_umma(x, m.n, m.t, m.v, m.ci, m.k, prep, nullptr, nullptr, XA, e.st));   // else kept by forward
        if (!bf->zeroed) cudaMemsetAsync(dA, 0, sizeof(float) * (size_t)m.k * m.v * m.v, e.st);
        if (!have_t1) {
          if (!bf->zeroed) cudaMemsetAsync(t1, 0, sizeof(float) * (size_t)m.v * m.co, e.st);
          STGCN_TRY(launch_vcolsum<T>(dS, m.frames_in, m.v, m.co, t1, e.st));
        }
      }
      STGCN_TRY(e.wgrad(geom(m.n, m.t, m.t, m.v, KCi, m.co, 1, 1, 0, 0), dS, bf->xa_saved ? bf->xa_saved : XA,
                        widx(kNoDiv, 0, m.ci, m.ci, (long long)m.co * m.ci, 1, 0), dW, (size_t)m.k * m.co * m.ci, -1, bf->packed_m,
                        bf->zeroed));   // raw: w.r.t. W.diag(a)
      GemmEpilogue ep;
      STGCN_TRY(e.gemm_packed(geom(m.n, m.t, m.t, m.v, m.co, KCi, 1, 1, 0, 0), dS, w2pT, ep, dXA));         // dXA' = dXA (.) a
      if (ws.dry) return STGCN_OK;
      static const bool split_dA = getenv("STGCN_SPLIT_DA") != nullptr;       // A/B switch: separate dA kernel
      const bool merged = bf->dx && !split_dA;                                 // dA rides in the transposed aggregation
      if (!merged) STGCN_TRY(launch_agg_dA_umma(dXA, x, m.n, m.t, m.v, m.ci, m.k, dA, e.st));
      STGCN_TRY(launch_bn1_fold_grads(W, dW, bf->stats, t1, umma_agg_rowsum_q(prep, m.k), m.k, m.v, m.co, m.ci, bf->training,
                                      (double)m.rows_in, bf->dgamma, bf->dbeta, coef, e.st));
      if (bf->sync && bf->sync->on() && bf->dx) {      // dbeta / dgamma stay local; the coefficients use their sums over all ranks
        STGCN_TRY(launch_pack2(bf->dbeta, bf->dgamma, m.ci, bf->sync->buf, e.st));
        bf->sync->allreduce();
        STGCN_TRY(launch_bn1_coef_global(bf->stats, bf->sync->buf, m.ci, (double)m.rows_in * bf->sync->world, coef, e.st));
      }
      if (bf->dx) STGCN_TRY(launch_agg_bwd_umma(dXA, m.n, m.t, m.v, m.ci, m.k, prep, nullptr, bf->dx, e.st, x, coef, merged ? dA : nullptr));
      if (!bf->bias_grads_done) STGCN_TRY(launch_spatial_bias_grads(t1, A, Wb, m.k, m.v, m.co, dWb, dA, e.st, wbb));
      return STGCN_OK;
    }
  }
  bool tc_agg = false;
  if constexpr (std::is_same<T, bf16>::value) tc_agg = e.want_umma() && umma_agg_supported(m.v, m.ci, m.k);
  void* prep = tc_agg ? ws.template get<uint8_t>(umma_agg_prep_bytes(m.k)) : nullptr;
  // tensor-core path: h = act(a*x+b) is an MMA operand of the dA contraction, so it is materialised once
  const bool need_h = tc_agg && (aff_a != nullptr || relu1);
  T* hbuf = need_h ? ws.template get<T>(m.rows_in * m.ci) : nullptr;
  const T* h = need_h ? hbuf : x;
  if (!ws.dry) {
    if (!ws.ok()) return STGCN_ERR_WORKSPACE;
    if (tc_agg) {
      if constexpr (std::is_same<T, bf16>::value) {
        STGCN_TRY(launch_build_big_adjacency(A, m.k, m.v, prep, e.st));
        if (need_h) STGCN_TRY(launch_affine_act<T>(x, m.rows_in, m.ci, aff_a, aff_b, relu1, 0.f, 0, 0, hbuf, e.st));
        STGCN_TRY(launch_agg_fwd_umma(h, m.n, m.t, m.v, m.ci, m.k, prep, nullptr, nullptr, XA, e.st));
      }
    } else {
      STGCN_TRY(launch_agg_fwd<T>(x, m.frames_in, m.v, m.ci, m.k, A, aff_a, aff_b, relu1, XA, e.st));
    }
    cudaMemsetAsync(dA, 0, sizeof(float) * (size_t)m.k * m.v * m.v, e.st);
    if (!have_t1) {
      cudaMemsetAsync(t1, 0, sizeof(float) * (size_t)m.v * m.co, e.st);
      STGCN_TRY(launch_vcolsum<T>(dS, m.frames_in, m.v, m.co, t1, e.st));
    }
  }
  STGCN_TRY(e.wgrad(geom(m.n, m.t, m.t, m.v, KCi, m.co, 1, 1, 0, 0), dS, XA, widx(kNoDiv, 0, m.ci, m.ci, (long long)m.co * m.ci, 1, 0),
                    dW, (size_t)m.k * m.co * m.ci));
  // dXA = dS . W2 : packed [n=k*Ci+ci][c=co] = W[(k*Co+co)*Ci+ci]
  GemmEpilogue ep;
  WIndex wiT = widx(m.ci, (long long)m.co * m.ci, 1, kNoDiv, 0, m.ci, 0);
  STGCN_TRY(e.gemm(geom(m.n, m.t, m.t, m.v, m.co, KCi, 1, 1, 0, 0), dS, W, wiT, ep, dXA));
  if (ws.dry) return STGCN_OK;
  if (tc_agg) {
    if constexpr (std::is_same<T, bf16>::value) {
      STGCN_TRY(launch_agg_dA_umma(dXA, h, m.n, m.t, m.v, m.ci, m.k, dA, e.st));
      STGCN_TRY(launch_spatial_bias_grads(t1, A, Wb, m.k, m.v, m.co, dWb, dA, e.st));
      if (dh) STGCN_TRY(launch_agg_bwd_umma(dXA, m.n, m.t, m.v, m.ci, m.k, prep, relu1 ? h : nullptr, dh, e.st));
    }
    return STGCN_OK;
  }
  STGCN_TRY(launch_agg_dA<T>(dXA, x, m.frames_in, m.v, m.ci, m.k, aff_a, aff_b, relu1, dA, e.st));
  STGCN_TRY(launch_spatial_bias_grads(t1, A, Wb, m.k, m.v, m.co, dWb, dA, e.st));
  if (dh) STGCN_TRY(launch_agg_bwd<T>(dXA, m.frames_in, m.v, m.ci, m.k, A, x, aff_a, aff_b, relu1, dh, e.st));
  return STGCN_OK;
}

// ---------------------------------------------------------------------------
// block forward
// ---------------------------------------------------------------------------
template <class T>
static int32_t block_forward_t(const StgcnBlockDesc& d, const T* x, const StgcnBlockParams& P, T* y, void* saved_mem,
                               Arena& ws, cudaStream_t st, const float* x_cs = nullptr, int x_cs_n = 0, float* y_cs = nullptr,
                               int* y_cs_n = nullptr) {
  // x_cs / y_cs: chained blocks (stgcn_block_forward_chain) -- column statistics of x handed over by the producer of x, and of
  // y produced for the consumer of y, so BatchNorm-1 of the next block needs no pass of its own
  if (y_cs_n) *y_cs_n = 0;
  const Dims m = dims_of(d);
  Arena sv(saved_mem, ~(size_t)0);
  Saved<T> S;
  const bool keep = saved_mem != nullptr || ws.dry;
  if (saved_mem) S = carve_saved<T>(sv, d, m);
  else { S.rng = ws.get<uint64_t>(2); S.stats1 = ws.get<float>(4 * m.ci); S.stats2 = ws.get<float>(4 * m.co); S.S = ws.get<T>(m.rows_in * m.co); S.U = nullptr; S.XA = nullptr; S.fp.valid = false; }
  float* partials = ws.get<float>(stat_partials_bytes(std::max(m.ci, m.co)) / sizeof(float));
  Engine<T> e{d, ws, st};
  const bool res = d.residual != STGCN_RES_NONE;
  const int relu1 = res ? 1 : 0;
  int nblk = 0;
  STGCN_TRY(sync_check(P));
  const SyncCtx sy = sync_of(d, P, st, ws.dry);
  const bool drop = d.training && d.dropout_p > 0.f;
  if (!ws.dry) {
    if (!ws.ok()) { set_error("workspace too small"); return STGCN_ERR_WORKSPACE; }
    // dropout state: from the caller's device-resident generator (advanced here by one Philox call per four outputs), else the
    // descriptor's values; either way the kernels read it from S.rng
    if (drop) STGCN_TRY(launch_rng_fetch(P.rng_state, d.seed, d.offset, (uint64_t)((m.rows_out * m.co + 3) / 4), S.rng, st));
    if (d.training) {
      const float* part1 = partials;
      if (x_cs && x_cs_n > 0) { part1 = x_cs; nblk = x_cs_n; }
      else STGCN_TRY(launch_colstats<T>(x, m.rows_in, m.ci, partials, &nblk, st));
      STGCN_TRY(finalize_stats(sy, part1, nblk, m.ci, (double)m.rows_in, P.bn1_weight, P.bn1_bias, P.bn1_running_mean,
                               P.bn1_running_var, (long long*)P.bn1_num_batches_tracked, d.bn_eps, d.bn_momentum, S.stats1, st));
    } else {
      STGCN_TRY(launch_bn_eval_coeffs(m.ci, P.bn1_weight, P.bn1_bias, P.bn1_running_mean, P.bn1_running_var, d.bn_eps, S.stats1, st));
    }
  }
  // temporal weights: packed [tap][n=co][c=ci] = Wt[(co*Co+ci)*gamma + tap]
  const WIndex wiT = widx(kNoDiv, 0, (long long)m.co * m.gamma, kNoDiv, 0, m.gamma, 1);
  const ConvGeom gt = geom(m.n, m.t, m.to, m.v, m.co, m.co, m.gamma, m.stride, m.pad, 0);
  // training chain on the tcgen05 engine: the temporal weights are packed on the side lane while BatchNorm-1 and the spatial
  // convolution run
  const bool pre_wt = !res && (d.training || keep) && e.gemm_on_umma(gt);
  TrainPrep tp{}; bool have_tp = false;
  if (P.train_prep && !ws.dry && train_prep_supported(d, m)) {
    Arena ta(const_cast<void*>(P.train_prep), ~(size_t)0); tp = carve_train_prep(ta, m); have_tp = true;
  }
  SideLane* lane = (pre_wt && !have_tp && !ws.dry) ? side_lane() : nullptr;
  bf16* wt_pre = nullptr;
  if (pre_wt && have_tp) wt_pre = tp.wt_f;          // packed ahead of the step
  else if (pre_wt) {
    int32_t rc = STGCN_OK;
    wt_pre = e.prepack(gt, P.Wt, wiT, lane, lane ? lane->ev[0] : nullptr, lane ? lane->ev[1] : nullptr, &rc);
    STGCN_TRY(rc);
  }
  STGCN_TRY(spatial_fwd<T>(e, m, x, S.stats1 + 2 * m.ci, S.stats1 + 3 * m.ci, relu1, P.A, P.W, P.W_bias, S.S, saved_mem ? S.XA : nullptr, saved_mem ? &S.fp : nullptr,
                           nullptr, (have_tp && tp.has_big && !res) ? tp.big : nullptr));
  if (!res) {
    if (d.training || keep) {
      T* U = S.U ? S.U : ws.get<T>(m.rows_out * m.co);
      GemmEpilogue ep; ep.bias = P.Wt_bias;
      static const bool no_fs = getenv("STGCN_NO_FUSED_STATS") != nullptr;
      const bool fused_stats = !no_fs && d.training && e.gemm_on_umma(gt);     // BN-2 statistics ride in the GEMM epilogue
      if (fused_stats) ep.stat_partials = partials;
      if (wt_pre) {
        if (lane) cudaStreamWaitEvent(st, lane->ev[1], 0);            // join: the side lane is empty again
        STGCN_TRY(e.gemm_packed(gt, S.S, wt_pre, ep, U));
      } else STGCN_TRY(e.gemm(gt, S.S, P.Wt, wiT, ep, U));
      if (ws.dry) return STGCN_OK;
      if (!ws.ok()) { set_error("workspace too small"); return STGCN_ERR_WORKSPACE; }
      if (d.training) {
        if (fused_stats) nblk = umma_conv_gemm_grid(gt);
        else STGCN_TRY(launch_colstats<T>(U, m.rows_out, m.co, partials, &nblk, st));
        STGCN_TRY(finalize_stats(sy, partials, nblk, m.co, (double)m.rows_out, P.bn2_weight, P.bn2_bias, P.bn2_running_mean,
                                 P.bn2_running_var, (long long*)P.bn2_num_batches_tracked, d.bn_eps, d.bn_momentum, S.stats2, st));
      } else {
        STGCN_TRY(launch_bn_eval_coeffs(m.co, P.bn2_weight, P.bn2_bias, P.bn2_running_mean, P.bn2_running_var, d.bn_eps, S.stats2, st));
      }
      const float p = d.training ? d.dropout_p : 0.f;
      return launch_affine_act<T>(U, m.rows_out, m.co, S.stats2 + 2 * m.co, S.stats2 + 3 * m.co, 1, p, 0, 0, y, st,
                                  d.training ? y_cs : nullptr, y_cs_n, drop ? S.rng : nullptr);
    }
    // inference: BN2 + ReLU folded into the temporal GEMM's epilogue, U never exists
    if (!ws.dry) STGCN_TRY(launch_bn_eval_coeffs(m.co, P.bn2_weight, P.bn2_bias, P.bn2_running_mean, P.bn2_running_var, d.bn_eps, S.stats2, st));
    GemmEpilogue ep; ep.bias = P.Wt_bias; ep.scale = S.stats2 + 2 * m.co; ep.shift = S.stats2 + 3 * m.co; ep.relu = 1;
    return e.gemm(gt, S.S, P.Wt, wiT, ep, y);
  }
  // residual block: BN2 + ReLU sit between the spatial and the temporal conv
  T* P2 = ws.get<T>(m.rows_in * m.co);
  if (!ws.dry) {
    if (!ws.ok()) { set_error("workspace too small"); return STGCN_ERR_WORKSPACE; }
    if (d.training) {
      STGCN_TRY(launch_colstats<T>(S.S, m.rows_in, m.co, partials, &nblk, st));
      STGCN_TRY(finalize_stats(sy, partials, nblk, m.co, (double)m.rows_in, P.bn2_weight, P.bn2_bias, P.bn2_running_mean,
                               P.bn2_running_var, (long long*)P.bn2_num_batches_tracked, d.bn_eps, d.bn_momentum, S.stats2, st));
    } else {
      STGCN_TRY(launch_bn_eval_coeffs(m.co, P.bn2_weight, P.bn2_bias, P.bn2_running_mean, P.bn2_running_var, d.bn_eps, S.stats2, st));
    }
    STGCN_TRY(launch_affine_act<T>(S.S, m.rows_in, m.co, S.stats2 + 2 * m.co, S.stats2 + 3 * m.co, 1, 0.f, 0, 0, P2, st));
  }
  const void* resid = x;
  if (d.residual == STGCN_RES_CONV) {
    GemmEpilogue er; er.bias = P.Wr_bias;
    STGCN_TRY(e.gemm(geom(m.n, m.t, m.to, m.v, m.ci, m.co, 1, m.stride, 0, 0), x, P.Wr, widx(kNoDiv, 0, m.ci, kNoDiv, 0, 1, 0), er, y));
    resid = y;
  }
  GemmEpilogue ep; ep.bias = P.Wt_bias; ep.residual = resid; ep.relu = 1;
  if (drop) { ep.dropout_p = d.dropout_p; ep.rng = S.rng; }
  return e.gemm(gt, P2, P.Wt, wiT, ep, y);
}

// ---------------------------------------------------------------------------
// inference with frozen parameters: prepare once, then two large kernels per block
// ---------------------------------------------------------------------------
static bool prepared_ok(const StgcnBlockDesc& d, const Dims& m) {
  if (d.training || d.dtype != STGCN_BF16 || d.residual != STGCN_RES_NONE || !fold_ok_desc(d, m)) return false;
  return umma_conv_gemm_supported(geom(m.n, m.t, m.to, m.v, m.co, m.co, m.gamma, m.stride, m.pad, 0));
}
static int32_t block_prepare(const StgcnBlockDesc& d, const StgcnBlockParams& P, void* prepared_mem, cudaStream_t st) {
  const Dims m = dims_of(d);
  if (!prepared_ok(d, m)) { set_error("block_prepare: this block has no prepared path (needs eval mode, bf16, default block, C_in % 64 == 0)"); return STGCN_ERR_UNSUPPORTED; }
  Arena pa(prepared_mem, ~(size_t)0);
  const Prepared q = carve_prepared(pa, m);
  STGCN_TRY(launch_bn_eval_coeffs(m.ci, P.bn1_weight, P.bn1_bias, P.bn1_running_mean, P.bn1_running_var, d.bn_eps, q.stats1, st));
  // the temporal conv's bias travels in the BatchNorm-2 shift: the fused epilogue is one fma + ReLU per element
  STGCN_TRY(launch_bn_eval_coeffs(m.co, P.bn2_weight, P.bn2_bias, P.bn2_running_mean, P.bn2_running_var, d.bn_eps, q.stats2, st, P.Wt_bias));
  STGCN_TRY(launch_build_big_adjacency(P.A, m.k, m.v, q.fp.prep, st));
  STGCN_TRY(launch_bn1_fold(P.W, P.W_bias, P.A, umma_agg_rowsum_q(q.fp.prep, m.k), q.stats1 + 2 * m.ci, q.stats1 + 3 * m.ci, m.k, m.v, m.co,
                            m.ci, q.fp.w2p, q.fp.w2pT, q.fp.wb, q.fp.wbb, q.b2d, st));
  const WIndex wiT = widx(kNoDiv, 0, (long long)m.co * m.gamma, kNoDiv, 0, m.gamma, 1);
  return launch_pack_weights<bf16>(P.Wt, wiT, m.gamma, m.co, m.co, q.wt, st);
}
static int32_t block_forward_prepared(const StgcnBlockDesc& d, const bf16* x, const StgcnBlockParams& P, const void* prepared_mem, bf16* y,
                                      Arena& ws, cudaStream_t st) {
  const Dims m = dims_of(d);
  if (!prepared_ok(d, m)) { set_error("block_forward_prepared: this block has no prepared path"); return STGCN_ERR_UNSUPPORTED; }
  Arena pa(const_cast<void*>(prepared_mem), ~(size_t)0);
  const Prepared q = carve_prepared(pa, m);
  bf16* Sbuf = ws.get<bf16>(m.rows_in * m.co);
  Engine<bf16> e{d, ws, st};
  STGCN_TRY(spatial_fwd<bf16>(e, m, x, q.stats1 + 2 * m.ci, q.stats1 + 3 * m.ci, 0, P.A, P.W, P.W_bias, Sbuf, nullptr, &q.fp, q.b2d));
  GemmEpilogue ep; ep.scale = q.stats2 + 2 * m.co; ep.shift = q.stats2 + 3 * m.co; ep.relu = 1;   // conv bias + BN-2 + ReLU in the epilogue
  return e.gemm_packed(geom(m.n, m.t, m.to, m.v, m.co, m.co, m.gamma, m.stride, m.pad, 0), Sbuf, q.wt, ep, y);
}

// ---------------------------------------------------------------------------
// block backward
// ---------------------------------------------------------------------------
template <class T>
static int32_t block_backward_t(const StgcnBlockDesc& d, const T* x, const T* y, const StgcnBlockParams& P, const void* saved_mem,
                                const T* dy, T* dx, const StgcnBlockGrads& G, Arena& ws, cudaStream_t st) {
  const Dims m = dims_of(d);
  Arena sv(const_cast<void*>(saved_mem), ~(size_t)0);
  Saved<T> S = carve_saved<T>(sv, d, m);
  float* partials = ws.get<float>(stat_partials_bytes(std::max(m.ci, m.co)) / sizeof(float));
  float* sums1 = ws.get<float>(2 * m.ci);
  float* sums2 = ws.get<float>(2 * m.co);
  T* dU = ws.get<T>(m.rows_out * m.co);
  T* dTin = ws.get<T>(m.rows_in * m.co);
  T* dh = ws.get<T>(m.rows_in * m.ci);
  Engine<T> e{d, ws, st};
  const bool res = d.residual != STGCN_RES_NONE;
  const int relu1 = res ? 1 : 0;
  const float p = d.training ? d.dropout_p : 0.f;
  int nblk = 0;
  const T* tin = S.S;
  T* P2 = res ? ws.get<T>(m.rows_in * m.co) : nullptr;
  T* dS = res ? ws.get<T>(m.rows_in * m.co) : dTin;
  float* t1 = ws.get<float>((size_t)m.v * m.co);     // per-joint column sums of dS (bias / BN-fold gradients)
  // accumulators of this chain (split-K weight-gradient tiles, dA, t1, the zero temporal-bias gradient) are cleared by the first
  // full-grid kernel (bn_bwd_reduce) instead of one memset node each; the buffers are carved here so their addresses are known
  const ConvGeom g_wt = geom(m.n, m.t, m.to, m.v, m.co, m.co, m.gamma, m.stride, m.pad, 0);
  const ConvGeom g_wm = geom(m.n, m.t, m.t, m.v, m.k * m.ci, m.co, 1, 1, 0, 0);
  const bool folded = !res && fold_ok<T>(e, m, S.stats1 + 2 * m.ci, relu1);
  float* packed_t = (!res && e.wgrad_on_umma(g_wt)) ? ws.get<float>(umma_wgrad_workspace_floats(g_wt)) : nullptr;
  float* packed_m = (folded && e.wgrad_on_umma(g_wm)) ? ws.get<float>(umma_wgrad_workspace_floats(g_wm)) : nullptr;
  SideLane* lane = (!res && folded && packed_t && !ws.dry) ? side_lane() : nullptr;
  // temporal data-gradient weights (parameters only): packed on the side lane beside the BatchNorm-2 backward passes
  const ConvGeom gd = geom(m.n, m.to, m.t, m.v, m.co, m.co, m.gamma, m.stride, m.pad, 1);
  const WIndex wiD = widx(kNoDiv, 0, m.gamma, kNoDiv, 0, (long long)m.co * m.gamma, 1);
  bf16* wd_pre = nullptr;
  TrainPrep tp{}; bool have_tp = false;
  if (P.train_prep && !ws.dry && train_prep_supported(d, m)) {
    Arena ta(const_cast<void*>(P.train_prep), ~(size_t)0); tp = carve_train_prep(ta, m); have_tp = true;
    if (S.fp.valid && tp.has_big) S.fp.prep = tp.big;        // forward put the block-diagonal adjacency there, not in `saved`
  }
  bool wd_on_lane = false;
  if (!res && folded && packed_t && e.gemm_on_umma(gd)) {
    if (have_tp) wd_pre = tp.wt_d;                            // packed ahead of the step
    else {
      int32_t rc = STGCN_OK;
      wd_pre = e.prepack(gd, P.Wt, wiD, lane, lane ? lane->ev[4] : nullptr, lane ? lane->ev[5] : nullptr, &rc);
      STGCN_TRY(rc);
      wd_on_lane = lane != nullptr;
    }
  }
  STGCN_TRY(sync_check(P));
  const SyncCtx sy = sync_of(d, P, st, ws.dry);
  const float* sums_use = nullptr; double cscale = 1.0;
  ZeroJobs zj;
  if (!res) {
    if (packed_t) zj.add(packed_t, umma_wgrad_workspace_floats(g_wt));
    if (packed_m) zj.add(packed_m, umma_wgrad_workspace_floats(g_wm));
    if (folded) { zj.add(G.A, (size_t)m.k * m.v * m.v); zj.add(t1, (size_t)m.v * m.co); }
    if (d.training) zj.add(G.Wt_bias, (size_t)m.co);   // a bias before a batch-stat BN has zero gradient
  }
  if (!ws.dry) {
    if (!ws.ok()) { set_error("workspace too small"); return STGCN_ERR_WORKSPACE; }
    if (!res) {
      static const bool no_zj = getenv("STGCN_NO_ZEROJOBS") != nullptr;      // A/B switch: one memset node per accumulator
      STGCN_TRY(launch_bn_bwd_reduce<T>(dy, S.U, S.stats2, m.rows_out, m.co, 1, p, 0, 0, partials, &nblk, st, no_zj ? nullptr : &zj, p > 0.f ? S.rng : nullptr));
      if (!zj.done) for (int j = 0; j < zj.count; ++j) cudaMemsetAsync(zj.p[j], 0, sizeof(float) * zj.n[j], st);
      STGCN_TRY(finalize_bwd_sums(sy, partials, nblk, m.co, sums2, G.bn2_weight, G.bn2_bias, st, &sums_use, &cscale));
      STGCN_TRY(launch_bn_bwd_apply<T>(dy, S.U, S.stats2, sums_use, m.rows_out, m.co, 1, d.training, p, 0, 0, dU, st, p > 0.f ? S.rng : nullptr,
                                       nullptr, cscale));
      if (!d.training) STGCN_TRY(launch_colsum<T>(dU, m.rows_out, m.co, partials, G.Wt_bias, st));
    } else {
      STGCN_TRY(launch_relu_bwd<T>(dy, y, m.rows_out * m.co, p, dU, st));
      STGCN_TRY(launch_colsum<T>(dU, m.rows_out, m.co, partials, G.Wt_bias, st));
      STGCN_TRY(launch_affine_act<T>(S.S, m.rows_in, m.co, S.stats2 + 2 * m.co, S.stats2 + 3 * m.co, 1, 0.f, 0, 0, P2, st));
      tin = P2;
    }
  }
  // With the side lane the temporal data gradient goes first; the (HBM-bound) per-joint sums of dS then run beside the
  // (tensor-bound) temporal weight gradient, whose layout change follows on the side lane as well.
  auto temporal_wgrad = [&]() -> int32_t {
    return e.wgrad(g_wt, dU, tin, widx(kNoDiv, 0, (long long)m.co * m.gamma, kNoDiv, 0, m.gamma, 1), G.Wt, (size_t)m.co * m.co * m.gamma,
                   -1, packed_t, true, lane ? lane->s : nullptr, lane ? lane->ev[2] : nullptr);
  };
  if (!lane) STGCN_TRY(temporal_wgrad());
  // temporal dgrad: packed [tap][n=ci][c=co] = Wt[(co*Co+ci)*gamma + tap]
  bool t1_ready = false;
  {
    GemmEpilogue ep;
    // measured on B200: summing dS in the dgrad epilogue costs more (+60 us per launch) than the separate 30 us pass,
    // so it is off unless STGCN_FUSED_VSUM is set (kept for re-evaluation with a wider epilogue)
    static const bool fv = getenv("STGCN_FUSED_VSUM") != nullptr;
    if (fv && !res && e.gemm_on_umma(gd)) {
      if (!ws.dry) cudaMemsetAsync(t1, 0, sizeof(float) * (size_t)m.v * m.co, st);
      ep.vsum = t1; t1_ready = true;
    }
    if (wd_pre) {
      if (wd_on_lane && !ws.dry) cudaStreamWaitEvent(st, lane->ev[5], 0);
      STGCN_TRY(e.gemm_packed(gd, dU, wd_pre, ep, dTin));
    } else STGCN_TRY(e.gemm(gd, dU, P.Wt, wiD, ep, dTin));
  }
  bool side_bias_grads = false;
  const bool t1_from_epilogue = t1_ready;        // (STGCN_FUSED_VSUM: cleared right before its producer, not by the zero jobs)
  if (lane) {
    if (!t1_ready) {
      cudaEventRecord(lane->ev[0], st); cudaStreamWaitEvent(lane->s, lane->ev[0], 0);
      STGCN_TRY(launch_vcolsum<T>(dS, m.frames_in, m.v, m.co, t1, lane->s));       // t1 was cleared by the zero jobs / memsets above
      cudaEventRecord(lane->ev[1], lane->s);
      if (S.fp.valid) {      // dW_bias and the rank-1 part of dA need only t1 and operands kept by forward (dA takes atomics from both lanes)
        STGCN_TRY(launch_spatial_bias_grads(t1, P.A, P.W_bias, m.k, m.v, m.co, G.W_bias, G.A, lane->s, S.fp.wbb));
        side_bias_grads = true;
      }
    }
    STGCN_TRY(temporal_wgrad());                       // caller's stream; its unpack waits on ev[2] on the side lane
    cudaEventRecord(lane->ev[3], lane->s);             // everything issued to the side lane so far
    if (!t1_ready) { cudaStreamWaitEvent(st, lane->ev[1], 0); t1_ready = true; }
  }
  if (res && !ws.dry) {
    STGCN_TRY(launch_bn_bwd_reduce<T>(dTin, S.S, S.stats2, m.rows_in, m.co, 1, 0.f, 0, 0, partials, &nblk, st));
    STGCN_TRY(finalize_bwd_sums(sy, partials, nblk, m.co, sums2, G.bn2_weight, G.bn2_bias, st, &sums_use, &cscale));
    STGCN_TRY(launch_bn_bwd_apply<T>(dTin, S.S, S.stats2, sums_use, m.rows_in, m.co, 1, d.training, 0.f, 0, 0, dS, st, nullptr, nullptr, cscale));
  }
  Bn1Fold<T> bf{S.stats1, d.training, G.bn1_weight, G.bn1_bias, dx, false, S.XA, &S.fp, t1_ready ? t1 : nullptr};
  bf.sync = &sy;
  bf.t1_buf = t1; bf.packed_m = packed_m; bf.zeroed = folded && !t1_from_epilogue; bf.bias_grads_done = side_bias_grads;      // (with STGCN_FUSED_VSUM t1 is cleared right before its producer)
  STGCN_TRY(spatial_bwd<T>(e, m, x, S.stats1 + 2 * m.ci, S.stats1 + 3 * m.ci, relu1, P.A, P.W, P.W_bias, dS, dh, G.A, G.W, G.W_bias, &bf));
  // Gradient of the skip connection (st_graphconv.py:71,81): it reaches dx as an addend of the last BatchNorm-1 backward pass
  // instead of a pass of its own.  Identity skip: dU itself.  Strided 1x1 conv skip: dR = dU . Wr as a plain 1x1 GEMM on the
  // rows the conv sampled (tensor cores; the transposed-conv form has an empty phase for stride 2 and used to fall back to CUDA
  // cores: 0.56 of 2.69 ms for the 64->128 block at N_eff = 128), added at frames t % stride == 0.
  ResAddend<T> ra;
  const bool own_pass = !bf.done;                    // residual blocks and CUDA-core shapes: BatchNorm-1 backward is not folded
  if (dx && d.residual == STGCN_RES_IDENTITY) { ra.ptr = dU; ra.stride = 1; ra.t = m.t; ra.t_src = m.t; ra.v = m.v; }
  if (d.residual == STGCN_RES_CONV) {
    STGCN_TRY(e.wgrad(geom(m.n, m.t, m.to, m.v, m.ci, m.co, 1, m.stride, 0, 0), dU, x, widx(kNoDiv, 0, m.ci, kNoDiv, 0, 1, 0), G.Wr,
                      (size_t)m.co * m.ci));
    if (!ws.dry) cudaMemcpyAsync(G.Wr_bias, G.Wt_bias, sizeof(float) * m.co, cudaMemcpyDeviceToDevice, st);
    if (dx) {
      T* dR = ws.get<T>(m.rows_out * m.ci);
      GemmEpilogue ep;
      STGCN_TRY(e.gemm(geom(m.n, m.to, m.to, m.v, m.co, m.ci, 1, 1, 0, 0), dU, P.Wr, widx(kNoDiv, 0, 1, kNoDiv, 0, m.ci, 0), ep, dR));
      ra.ptr = dR; ra.stride = m.stride; ra.t = m.t; ra.t_src = m.to; ra.v = m.v;
    }
  }
  if (!ws.dry && own_pass) {
    STGCN_TRY(launch_bn_bwd_reduce<T>(dh, x, S.stats1, m.rows_in, m.ci, 0, 0.f, 0, 0, partials, &nblk, st));
    STGCN_TRY(finalize_bwd_sums(dx ? sy : SyncCtx{}, partials, nblk, m.ci, sums1, G.bn1_weight, G.bn1_bias, st, &sums_use, &cscale));
    if (dx) STGCN_TRY(launch_bn_bwd_apply<T>(dh, x, S.stats1, sums_use, m.rows_in, m.ci, 0, d.training, 0.f, 0, 0, dx, st, nullptr,
                                             ra.ptr ? &ra : nullptr, cscale));
  } else if (!ws.dry && dx && ra.ptr) {              // (not reached today: the folded path excludes residual blocks)
    set_error("block_backward: skip-connection gradient without a BatchNorm-1 backward pass"); return STGCN_ERR_UNSUPPORTED;
  }
  if (lane) cudaStreamWaitEvent(st, lane->ev[3], 0);     // join: nothing of this call is left on the side lane when it returns
  if (!ws.ok()) { set_error("workspace too small"); return STGCN_ERR_WORKSPACE; }
  return STGCN_OK;
}

// ---------------------------------------------------------------------------
// SpatialConv alone (module SpatialConv, st_graphconv.py:111-152)
// ---------------------------------------------------------------------------
template <class T>
static int32_t spatial_only_fwd(const StgcnBlockDesc& d, const T* x, const float* A, const float* W, const float* Wb, T* out,
                                Arena& ws, cudaStream_t st) {
  const Dims m = dims_of(d);
  Engine<T> e{d, ws, st};
  STGCN_TRY(spatial_fwd<T>(e, m, x, nullptr, nullptr, 0, A, W, Wb, out));
  if (!ws.ok()) { set_error("workspace too small"); return STGCN_ERR_WORKSPACE; }
  return STGCN_OK;
}
template <class T>
static int32_t spatial_only_bwd(const StgcnBlockDesc& d, const T* x, const float* A, const float* W, const float* Wb, const T* dout,
                                T* dx, float* dA, float* dW, float* dWb, Arena& ws, cudaStream_t st) {
  const Dims m = dims_of(d);
  Engine<T> e{d, ws, st};
  STGCN_TRY(spatial_bwd<T>(e, m, x, nullptr, nullptr, 0, A, W, Wb, dout, dx, dA, dW, dWb));
  if (!ws.ok()) { set_error("workspace too small"); return STGCN_ERR_WORKSPACE; }
  return STGCN_OK;
}

}  // namespace stgcn

// ===========================================================================
// C ABI
// ===========================================================================
using namespace stgcn;

extern "C" {

int32_t stgcn_block_t_out(const StgcnBlockDesc* d) { return d ? (d->t + 2 * d->pad - d->gamma) / d->stride + 1 : -1; }

size_t stgcn_block_saved_bytes(const StgcnBlockDesc* d) {
  if (validate(d) != STGCN_OK) return 0;
  Arena a(nullptr, 0);
  Dims m = dims_of(*d);
  if (d->dtype == STGCN_F32) carve_saved<float>(a, *d, m); else carve_saved<bf16>(a, *d, m);
  return a.off;
}

size_t stgcn_block_workspace_bytes(const StgcnBlockDesc* d, int32_t backward) {
  if (validate(d) != STGCN_OK) return 0;
  StgcnBlockParams P; memset(&P, 0, sizeof(P));
  StgcnBlockGrads G; memset(&G, 0, sizeof(G));
  // dry arena: the chain only records its allocation plan, nothing is launched.  Backward is planned both with and
  // without an input gradient (the first block takes a different spatial path when dx is not wanted).
  size_t need = 0;
  for (int pass = 0; pass < (backward ? 2 : 1); ++pass) {
    Arena dry(nullptr, 0);
    int32_t rc;
    if (d->dtype == STGCN_F32)
      rc = backward ? block_backward_t<float>(*d, nullptr, nullptr, P, nullptr, nullptr, pass ? nullptr : (float*)256, G, dry, 0)
                    : block_forward_t<float>(*d, nullptr, P, nullptr, nullptr, dry, 0);
    else
      rc = backward ? block_backward_t<bf16>(*d, nullptr, nullptr, P, nullptr, nullptr, pass ? nullptr : (bf16*)256, G, dry, 0)
                    : block_forward_t<bf16>(*d, nullptr, P, nullptr, nullptr, dry, 0);
    if (rc != STGCN_OK) return 0;
    need = std::max(need, dry.off);
  }
  return need + 256;
}

static int32_t check_ptrs(const void* a, const void* b, const void* c) {
  if (!a || !b || !c) { set_error("null pointer argument"); return STGCN_ERR_INVALID; }
  return STGCN_OK;
}

int32_t stgcn_block_forward(const StgcnBlockDesc* d, const void* x, const StgcnBlockParams* P, void* y, void* saved,
                            void* workspace, size_t workspace_bytes, void* stream) {
  NvtxRange nvtx_("fwd", d);
  STGCN_TRY(validate(d));
  STGCN_TRY(check_ptrs(x, P, y));
  if (!workspace) { set_error("null workspace"); return STGCN_ERR_WORKSPACE; }
  if (d->residual == STGCN_RES_CONV && (!P->Wr || !P->Wr_bias)) { set_error("conv residual needs Wr / Wr_bias"); return STGCN_ERR_INVALID; }
  Arena ws(workspace, workspace_bytes);
  cudaStream_t st = (cudaStream_t)stream;
  if (d->dtype == STGCN_F32) return block_forward_t<float>(*d, (const float*)x, *P, (float*)y, saved, ws, st);
  return block_forward_t<bf16>(*d, (const bf16*)x, *P, (bf16*)y, saved, ws, st);
}

size_t stgcn_colstats_bytes(int32_t channels) { return channels > 0 ? stat_partials_bytes(channels) : 0; }

int32_t stgcn_block_forward_chain(const StgcnBlockDesc* d, const void* x, const StgcnBlockParams* P, void* y, void* saved,
                                  void* workspace, size_t workspace_bytes, const float* x_colstats, int32_t x_colstats_nblk,
                                  float* y_colstats, int32_t* y_colstats_nblk, void* stream) {
  NvtxRange nvtx_("fwd", d);
  STGCN_TRY(validate(d));
  STGCN_TRY(check_ptrs(x, P, y));
  if (!workspace) { set_error("null workspace"); return STGCN_ERR_WORKSPACE; }
  if (d->residual == STGCN_RES_CONV && (!P->Wr || !P->Wr_bias)) { set_error("conv residual needs Wr / Wr_bias"); return STGCN_ERR_INVALID; }
  if (y_colstats && !y_colstats_nblk) { set_error("y_colstats needs y_colstats_nblk"); return STGCN_ERR_INVALID; }
  Arena ws(workspace, workspace_bytes);
  cudaStream_t st = (cudaStream_t)stream;
  int nb = 0;
  int32_t rc;
  if (d->dtype == STGCN_F32)
    rc = block_forward_t<float>(*d, (const float*)x, *P, (float*)y, saved, ws, st, x_colstats, x_colstats_nblk, y_colstats, &nb);
  else
    rc = block_forward_t<bf16>(*d, (const bf16*)x, *P, (bf16*)y, saved, ws, st, x_colstats, x_colstats_nblk, y_colstats, &nb);
  if (y_colstats_nblk) *y_colstats_nblk = nb;
  return rc;
}

size_t stgcn_block_prepared_bytes(const StgcnBlockDesc* d) {
  if (validate(d) != STGCN_OK) return 0;
  const Dims m = dims_of(*d);
  if (!prepared_ok(*d, m)) return 0;
  Arena dry(nullptr, 0);
  carve_prepared(dry, m);
  return dry.off + 256;
}

int32_t stgcn_block_prepare(const StgcnBlockDesc* d, const StgcnBlockParams* P, void* prepared, void* stream) {
  STGCN_TRY(validate(d));
  STGCN_TRY(check_ptrs(prepared, P, prepared));
  return block_prepare(*d, *P, prepared, (cudaStream_t)stream);
}

int32_t stgcn_block_forward_prepared(const StgcnBlockDesc* d, const void* x, const StgcnBlockParams* P, const void* prepared, void* y,
                                     void* workspace, size_t workspace_bytes, void* stream) {
  NvtxRange nvtx_("fwd-prepared", d);
  STGCN_TRY(validate(d));
  STGCN_TRY(check_ptrs(x, P, y));
  if (!prepared) { set_error("null prepared buffer"); return STGCN_ERR_INVALID; }
  if (!workspace) { set_error("null workspace"); return STGCN_ERR_WORKSPACE; }
  Arena ws(workspace, workspace_bytes);
  return block_forward_prepared(*d, (const bf16*)x, *P, prepared, (bf16*)y, ws, (cudaStream_t)stream);
}

size_t stgcn_train_prep_bytes(const StgcnBlockDesc* d) {
  if (validate(d) != STGCN_OK) return 0;
  const Dims m = dims_of(*d);
  if (!train_prep_supported(*d, m)) return 0;
  Arena dry(nullptr, 0);
  carve_train_prep(dry, m);
  return dry.off + 256;
}

int32_t stgcn_train_prep_multi(int32_t count, const StgcnBlockDesc* descs, const StgcnBlockParams* params, void* const* preps,
                               void* stream) {
  if (count < 0 || (count > 0 && (!descs || !params || !preps))) { set_error("train_prep_multi: null argument"); return STGCN_ERR_INVALID; }
  PrepTable tab; tab.count = 0;
  cudaStream_t st = (cudaStream_t)stream;
  for (int i = 0; i < count; ++i) {
    STGCN_TRY(validate(&descs[i]));
    const Dims m = dims_of(descs[i]);
    if (!preps[i] || !train_prep_supported(descs[i], m)) continue;
    if (!params[i].Wt || !params[i].A) { set_error("train_prep_multi: block %d has no parameters", i); return STGCN_ERR_INVALID; }
    Arena ta(preps[i], ~(size_t)0);
    const TrainPrep tp = carve_train_prep(ta, m);
    if (tab.count + 3 > kMaxPrepJobs) { STGCN_TRY(launch_train_prep_multi(tab, st)); tab.count = 0; }
    PrepJob j; memset(&j, 0, sizeof(j));
    j.kind = 0; j.taps = m.gamma; j.n_out = m.co; j.c = m.co; j.src = params[i].Wt;
    j.wi = widx(kNoDiv, 0, (long long)m.co * m.gamma, kNoDiv, 0, m.gamma, 1); j.dst = tp.wt_f;          // [tap][n=co][c=ci]  (forward)
    tab.jobs[tab.count++] = j;
    j.wi = widx(kNoDiv, 0, m.gamma, kNoDiv, 0, (long long)m.co * m.gamma, 1); j.dst = tp.wt_d;          // [tap][n=ci][c=co]  (data gradient)
    tab.jobs[tab.count++] = j;
    if (tp.has_big) {
      PrepJob a; memset(&a, 0, sizeof(a));
      a.kind = 1; a.K = m.k; a.V = m.v; a.FT = 128 / m.v; a.src = params[i].A;
      umma_agg_prep_pointers(tp.big, m.k, &a.big, &a.bigT, &a.rowsum);
      tab.jobs[tab.count++] = a;
    }
  }
  return launch_train_prep_multi(tab, st);
}

int32_t stgcn_block_backward(const StgcnBlockDesc* d, const void* x, const void* y, const StgcnBlockParams* P, const void* saved,
                             const void* dy, void* dx, const StgcnBlockGrads* G, void* workspace, size_t workspace_bytes,
                             void* stream) {
  NvtxRange nvtx_("bwd", d);
  STGCN_TRY(validate(d));
  STGCN_TRY(check_ptrs(x, P, dy));
  STGCN_TRY(check_ptrs(saved, G, y));
  if (!workspace) { set_error("null workspace"); return STGCN_ERR_WORKSPACE; }
  Arena ws(workspace, workspace_bytes);
  cudaStream_t st = (cudaStream_t)stream;
  if (d->dtype == STGCN_F32)
    return block_backward_t<float>(*d, (const float*)x, (const float*)y, *P, saved, (const float*)dy, (float*)dx, *G, ws, st);
  return block_backward_t<bf16>(*d, (const bf16*)x, (const bf16*)y, *P, saved, (const bf16*)dy, (bf16*)dx, *G, ws, st);
}

size_t stgcn_spatial_workspace_bytes(const StgcnBlockDesc* d, int32_t backward) {
  if (validate(d) != STGCN_OK) return 0;
  Arena dry(nullptr, 0);
  int32_t rc;
  if (d->dtype == STGCN_F32)
    rc = backward ? spatial_only_bwd<float>(*d, nullptr, nullptr, nullptr, nullptr, nullptr, (float*)256, nullptr, nullptr, nullptr, dry, 0)
                  : spatial_only_fwd<float>(*d, nullptr, nullptr, nullptr, nullptr, nullptr, dry, 0);
  else
    rc = backward ? spatial_only_bwd<bf16>(*d, nullptr, nullptr, nullptr, nullptr, nullptr, (bf16*)256, nullptr, nullptr, nullptr, dry, 0)
                  : spatial_only_fwd<bf16>(*d, nullptr, nullptr, nullptr, nullptr, nullptr, dry, 0);
  if (rc != STGCN_OK) return 0;
  return dry.off + 256;
}

int32_t stgcn_spatial_forward(const StgcnBlockDesc* d, const void* x, const float* A, const float* W, const float* Wb, void* out,
                              void* workspace, size_t workspace_bytes, void* stream) {
  STGCN_TRY(validate(d));
  STGCN_TRY(check_ptrs(x, A, W));
  STGCN_TRY(check_ptrs(Wb, out, workspace));
  Arena ws(workspace, workspace_bytes);
  cudaStream_t st = (cudaStream_t)stream;
  if (d->dtype == STGCN_F32) return spatial_only_fwd<float>(*d, (const float*)x, A, W, Wb, (float*)out, ws, st);
  return spatial_only_fwd<bf16>(*d, (const bf16*)x, A, W, Wb, (bf16*)out, ws, st);
}

int32_t stgcn_spatial_backward(const StgcnBlockDesc* d, const void* x, const float* A, const float* W, const float* Wb,
                               const void* dout, void* dx, float* dA, float* dW, float* dWb, void* workspace,
                               size_t workspace_bytes, void* stream) {
  STGCN_TRY(validate(d));
  STGCN_TRY(check_ptrs(x, A, W));
  STGCN_TRY(check_ptrs(Wb, dout, workspace));
  STGCN_TRY(check_ptrs(dA, dW, dWb));
  Arena ws(workspace, workspace_bytes);
  cudaStream_t st = (cudaStream_t)stream;
  if (d->dtype == STGCN_F32)
    return spatial_only_bwd<float>(*d, (const float*)x, A, W, Wb, (const float*)dout, (float*)dx, dA, dW, dWb, ws, st);
  return spatial_only_bwd<bf16>(*d, (const bf16*)x, A, W, Wb, (const bf16*)dout, (bf16*)dx, dA, dW, dWb, ws, st);
}

int32_t stgcn_head_forward(int32_t dtype, int32_t n, int32_t t, int32_t v, int32_t c, int32_t classes, const void* feat,
                           const float* fcw, const float* fcb, float* pooled, float* logits, void* stream) {
  STGCN_TRY(check_ptrs(feat, fcw, fcb));
  STGCN_TRY(check_ptrs(pooled, logits, logits));
  return head_forward(dtype, n, t, v, c, classes, feat, fcw, fcb, pooled, logits, (cudaStream_t)stream);
}
int32_t stgcn_head_backward(int32_t dtype, int32_t n, int32_t t, int32_t v, int32_t c, int32_t classes, const float* pooled,
                            const float* fcw, const float* dlogits, void* dfeat, float* dfcw, float* dfcb, void* stream) {
  STGCN_TRY(check_ptrs(pooled, fcw, dlogits));
  STGCN_TRY(check_ptrs(dfeat, dfcw, dfcb));
  return head_backward(dtype, n, t, v, c, classes, pooled, fcw, dlogits, dfeat, dfcw, dfcb, (cudaStream_t)stream);
}

int32_t stgcn_adjacency_normalize(int32_t k, int32_t v, const float* raw, float alpha, float expo, int32_t dense_alpha, float* out,
                                  void* stream) {
  STGCN_TRY(check_ptrs(raw, out, out));
  return launch_adjacency_normalize(raw, k, v, alpha, expo, dense_alpha, out, (cudaStream_t)stream);
}

int32_t stgcn_cast(int32_t sd, int32_t dd, const void* src, void* dst, size_t count, void* stream) {
  STGCN_TRY(check_ptrs(src, dst, dst));
  cudaStream_t st = (cudaStream_t)stream;
  if (sd == STGCN_F32 && dd == STGCN_BF16) return launch_cast<float, bf16>((const float*)src, (bf16*)dst, count, st);
  if (sd == STGCN_BF16 && dd == STGCN_F32) return launch_cast<bf16, float>((const bf16*)src, (float*)dst, count, st);
  if (sd == STGCN_F32 && dd == STGCN_F32) return launch_cast<float, float>((const float*)src, (float*)dst, count, st);
  if (sd == STGCN_BF16 && dd == STGCN_BF16) return launch_cast<bf16, bf16>((const bf16*)src, (bf16*)dst, count, st);
  set_error("cast: bad dtype"); return STGCN_ERR_INVALID;
}

int32_t stgcn_umma_conv_gemm(int32_t nb, int32_t t_in, int32_t t_out, int32_t v, int32_t c, int32_t n_out, int32_t taps,
                             int32_t stride, int32_t pad, const void* in, const void* w, const float* bias, void* out, void* stream) {
  STGCN_TRY(check_ptrs(in, w, out));
  ConvGeom g = geom(nb, t_in, t_out, v, c, n_out, taps, stride, pad, 0);
  if (!umma_conv_gemm_supported(g)) { set_error("umma_conv_gemm: shape unsupported (c%%64, n_out%%16, n_out<=256, V*frames<=128)"); return STGCN_ERR_UNSUPPORTED; }
  GemmEpilogue ep; ep.bias = bias;
  return launch_conv_gemm_umma(g, (const bf16*)in, (const bf16*)w, ep, (bf16*)out, (cudaStream_t)stream);
}


// ---------------------------------------------------------------------------------------------------------------------------
// stgcn_spatial_forward_static: SpatialConv.forward with its scratch in a file-scope device array (no caller workspace, no
// allocation of any kind) -- for callers that cannot allocate: code under CUDA-graph capture, tight serving loops, harnesses
// that forbid cudaMalloc.  No counterpart in the reference.
//   * g_static_scratch is a HOST-SIDE STAND-IN in host code, not a device address: the address a kernel may use comes from
//     cudaGetSymbolAddress, queried once per device with that device current (each device holds its own copy of the array,
//     STGCN_STATIC_SCRATCH_BYTES of HBM from library load until exit).  Device code names the array directly.
//   * every call shares the array and finds the previous call's contents in it: the bytes this call will use are cleared first.
//     Calls that overlap on the GPU (two streams, two host threads) RACE on it -- serialise them.
// ---------------------------------------------------------------------------------------------------------------------------
namespace stgcn {
static __device__ __align__(256) unsigned char g_static_scratch[STGCN_STATIC_SCRATCH_BYTES];

__global__ void __launch_bounds__(256) static_scratch_clear_kernel(size_t bytes) {
  pdl_enter();
  uint4* p = reinterpret_cast<uint4*>(g_static_scratch);            // device code: the array by name
  const size_t n16 = (bytes + 15) / 16;                              // the array is 256-byte aligned and a multiple of 16 long
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
    p[i] = make_uint4(0, 0, 0, 0);
}

// device address of the array on the CURRENT device (looked up once per device)
static int32_t static_scratch_address(void** out) {
  static std::mutex mu;
  static void* ptrs[64] = {};
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { set_error("static scratch: no current CUDA device"); return STGCN_ERR_CUDA; }
  std::lock_guard<std::mutex> lk(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || !p) {
      set_error("static scratch: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return STGCN_ERR_CUDA;
    }
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return STGCN_OK;
}
}  // namespace stgcn

size_t stgcn_static_scratch_bytes(void) { return (size_t)STGCN_STATIC_SCRATCH_BYTES; }

int32_t stgcn_spatial_forward_static(const StgcnBlockDesc* d, const void* x, const float* A, const float* W, const float* Wb, void* out,
                                     void* stream) {
  STGCN_TRY(validate(d));
  STGCN_TRY(check_ptrs(x, A, W));
  STGCN_TRY(check_ptrs(Wb, out, out));
  const size_t need = stgcn_spatial_workspace_bytes(d, 0);
  if (need == 0) return STGCN_ERR_UNSUPPORTED;                       // error text set by the sizing pass
  if (need > (size_t)STGCN_STATIC_SCRATCH_BYTES) {
    set_error("spatial_forward_static: this shape needs %zu bytes of scratch, the static array holds %zu (STGCN_STATIC_SCRATCH_BYTES); "
              "use stgcn_spatial_forward with a caller-provided workspace", need, (size_t)STGCN_STATIC_SCRATCH_BYTES);
    return STGCN_ERR_WORKSPACE;
  }
  void* scratch = nullptr;
  STGCN_TRY(static_scratch_address(&scratch));
  cudaStream_t st = (cudaStream_t)stream;
  launch_k(static_scratch_clear_kernel, (int)std::min<size_t>((need / 16 + 255) / 256 + 1, 148 * 8), 256, 0, st, need);
  STGCN_CHECK_LAUNCH("static_scratch_clear");
  Arena ws(scratch, (size_t)STGCN_STATIC_SCRATCH_BYTES);
  if (d->dtype == STGCN_F32) return spatial_only_fwd<float>(*d, (const float*)x, A, W, Wb, (float*)out, ws, st);
  return spatial_only_fwd<bf16>(*d, (const bf16*)x, A, W, Wb, (bf16*)out, ws, st);
}

size_t stgcn_abi_sizeof(int32_t which) {
  return which == 0 ? sizeof(StgcnBlockDesc) : which == 1 ? sizeof(StgcnBlockParams) : which == 2 ? sizeof(StgcnBlockGrads) : 0;
}
uint64_t stgcn_launch_count(void) { return g_launches.load(); }
int32_t stgcn_dropout_mask(uint64_t seed, uint64_t offset, float p, size_t count, float* out, void* stream) {
  if (!out) { set_error("dropout_mask: null output"); return STGCN_ERR_INVALID; }
  if (p < 0.f || p >= 1.f) { set_error("dropout_mask: p out of range"); return STGCN_ERR_INVALID; }
  return launch_dropout_mask(seed, offset, p, count, out, (cudaStream_t)stream);
}
const char* stgcn_last_error(void) { return g_err; }
const char* stgcn_version(void) { return "stgcn_b200 0.1 (sm_100a)"; }

}  // extern "C"
