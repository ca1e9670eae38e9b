// Internal launch API shared by the .cu files of libstgcn_b200.
#pragma once
#include <algorithm>
#include <type_traits>

#include "common.cuh"

namespace stgcn {

// epilogue of the implicit-GEMM engines, applied in this order:
//   t = acc (+bias[n]) (+bias2d[v][n]); t = scale[n]*t + shift[n]; t += residual[r][n];
//   relu; dropout; (+= out[r][n])
struct GemmEpilogue {
  const float* bias = nullptr;
  const float* bias2d = nullptr;     // [V][n_out]
  const float* scale = nullptr;      // folded BatchNorm (eval) a,b
  const float* shift = nullptr;
  const void* residual = nullptr;    // same type / shape as out
  int relu = 0;
  int accumulate = 0;
  float dropout_p = 0.f;
  uint64_t seed = 0, offset = 0;
  const uint64_t* rng = nullptr;     // device {seed, offset}: overrides the two values above when set
  float* stat_partials = nullptr;    // tcgen05 engine only: per-CTA column sum / sumsq of the stored output, [grid][2][n_out]
  float* vsum = nullptr;             // tcgen05 engine only: [V][n_out] += per-joint column sums of the stored output
};

// element (tap, n, c) of a weight tensor kept in the reference's layout
struct WIndex {
  int ndiv = 1 << 30; long long snh = 0, snl = 0;
  int cdiv = 1 << 30; long long sch = 0, scl = 0;
  long long st = 0;
  __host__ __device__ __forceinline__ long long at(int tap, int n, int c) const {
    return (long long)(n / ndiv) * snh + (long long)(n % ndiv) * snl + (long long)(c / cdiv) * sch + (long long)(c % cdiv) * scl +
           (long long)tap * st;
  }
};

// timing.cu: kernel classes for the optional device timing
enum KernelClass { KC_UMMA_GEMM = 0, KC_SIMT_GEMM, KC_WGRAD, KC_AGG_FWD, KC_AGG_BWD, KC_AGG_DA, KC_BN_STATS, KC_ELEMWISE,
                   KC_PACK, KC_HEAD, KC_UMMA_WGRAD, KC_UMMA_SPATIAL, KC_COUNT };
struct KernelTimer {     // records an event pair around the launches issued in its scope (when timing is enabled)
  KernelTimer(int cls, double work, cudaStream_t st);
  ~KernelTimer();
  int idx_; cudaStream_t st_;
};

// Small fp32 buffers that must be zero before later kernels of a chain accumulate into them (split-K weight-gradient tiles, dA,
// per-joint sums).  Instead of one memset node each -- which also breaks the programmatic launch chain -- the first full-grid
// kernel of the chain clears them on the side (bn_bwd_reduce's fast path).
struct ZeroJobs {
  float* p[6]; unsigned n[6]; int count = 0; bool done = false;
  void add(float* ptr, size_t floats) { if (ptr && floats && count < 6) { p[count] = ptr; n[count] = (unsigned)floats; ++count; } }
};

// Gradient of a skip connection added inside the last element-wise backward pass of a residual block: element e of the
// (n, t, v, C) tensor being written receives ptr[((n*t_src + t/stride)*v + v_)*C + c] when t % stride == 0 (stride 1, t_src = t:
// the identity skip; stride 2: the strided 1x1 conv skip, whose data gradient only reaches the frames it sampled).
template <class T> struct ResAddend {
  const T* ptr = nullptr; int stride = 1, t = 1, t_src = 1, v = 1;
  __device__ __forceinline__ long long source(long long e, int C) const {      // element index in ptr, or -1: nothing to add
    const long long row = e / C; const int c = (int)(e - row * C);
    const long long frame = row / v; const int vv = (int)(row - frame * v);
    const long long n = frame / t; const int tt = (int)(frame - n * t);
    if (tt % stride != 0) return -1;
    return ((n * t_src + tt / stride) * v + vv) * (long long)C + c;
  }
};

// bn_kernels.cu
size_t stat_partials_bytes(int C);
template <class T> int32_t launch_colstats(const T* x, long long rows, int C, float* partials, int* nblk, cudaStream_t st);
int32_t launch_bn_finalize(const float* partials, int nblk, int C, double count, const float* gamma, const float* beta,
                           float* rm, float* rv, long long* nbt, float eps, float momentum, float* stats, cudaStream_t st);
int32_t launch_bn_eval_coeffs(int C, const float* gamma, const float* beta, const float* rm, const float* rv, float eps,
                              float* stats, cudaStream_t st, const float* pre_bias = nullptr);
// out_partials / out_nblk (optional): also produce per-CTA column statistics of the stored output (see affine_act_stats_kernel)
template <class T> int32_t launch_affine_act(const T* x, long long rows, int C, const float* a, const float* b, int relu, float p,
                                            uint64_t seed, uint64_t offset, T* y, cudaStream_t st, float* out_partials = nullptr,
                                            int* out_nblk = nullptr, const uint64_t* rng = nullptr);
template <class T> int32_t launch_bn_bwd_reduce(const T* dy, const T* z, const float* stats, long long rows, int C, int relu, float p,
                                               uint64_t seed, uint64_t offset, float* partials, int* nblk, cudaStream_t st,
                                               ZeroJobs* zero = nullptr, const uint64_t* rng = nullptr);
int32_t launch_bn_bwd_finalize(const float* partials, int nblk, int C, float* sums, float* dgamma, float* dbeta, cudaStream_t st);
template <class T> int32_t launch_bn_bwd_apply(const T* dy, const T* z, const float* stats, const float* sums, long long rows, int C,
                                              int relu, int training, float p, uint64_t seed, uint64_t offset, T* dz, cudaStream_t st,
                                              const uint64_t* rng = nullptr, const ResAddend<T>* addend = nullptr, double count_scale = 1.0);
int32_t launch_reduce_partials(const float* partials, int nblk, int C, float* out, cudaStream_t st);
int32_t launch_pack2(const float* a, const float* b, int C, float* out, cudaStream_t st);
int32_t launch_bn1_coef_global(const float* stats, const float* g, int C, double rows_global, float* coef, cudaStream_t st);
int32_t launch_rng_fetch(uint64_t* state, uint64_t seed, uint64_t offset, uint64_t advance, uint64_t* slot, cudaStream_t st);
int32_t launch_dropout_mask(uint64_t seed, uint64_t offset, float p, size_t n, float* out, cudaStream_t st);
template <class T> int32_t launch_relu_bwd(const T* dy, const T* out, long long total, float p, T* dz, cudaStream_t st);
template <class TS, class TD> int32_t launch_cast(const TS* s, TD* d, size_t n, cudaStream_t st);
template <class T> int32_t launch_add_inplace(T* d, const T* s, size_t n, cudaStream_t st);

// agg_kernels.cu
template <class T> int32_t launch_agg_fwd(const T* x, long long frames, int V, int C, int K, const float* A, const float* a,
                                         const float* b, int relu, T* xa, cudaStream_t st);
template <class T> int32_t launch_agg_bwd(const T* dxa, long long frames, int V, int C, int K, const float* A, const T* x,
                                         const float* a, const float* b, int relu, T* dh, cudaStream_t st);
template <class T> int32_t launch_agg_dA(const T* dxa, const T* x, long long frames, int V, int C, int K, const float* a,
                                        const float* b, int relu, float* dA, cudaStream_t st, int ld = 0);
template <class T> int32_t launch_agg_small_fwd(const T* x, long long frames, int V, int C, int K, int ld, const float* A, T* xa,
                                               cudaStream_t st);
int32_t launch_adjacency_rowsum(const float* A, int K, int V, float* rs, cudaStream_t st);
int32_t launch_adjacency_normalize(const float* raw, int K, int V, float alpha, float expo, int dense_alpha, float* out, cudaStream_t st);
template <class T> int32_t launch_vcolsum(const T* ds, long long frames, int V, int C, float* t1, cudaStream_t st);
int32_t launch_bias2d(const float* A, const float* bias, int K, int V, int Co, float* out, cudaStream_t st);
int32_t launch_spatial_bias_grads(const float* t1, const float* A, const float* bias, int K, int V, int Co, float* dbias, float* dA,
                                  cudaStream_t st, const float* bias_dA = nullptr);
int32_t launch_bn1_fold(const float* W, const float* Wb, const float* A, const float* rowsum_q, const float* a, const float* b, int K,
                        int V, int Co, int Ci, bf16* w2p, bf16* w2pT, float* wb, float* wbb, float* bias2d, cudaStream_t st,
                        int ld = 0, float* rsum2 = nullptr, float* bvec2 = nullptr);
int32_t launch_bn1_fold_grads(const float* W, float* dW, const float* stats, const float* t1, const float* rowsum_q, int K, int V,
                              int Co, int Ci, int training, double rows, float* dgamma, float* dbeta, float* coef, cudaStream_t st);

// gemm_simt.cu
template <class T> int32_t launch_conv_gemm_simt(const ConvGeom& g, const T* in, const float* w, const GemmEpilogue& ep, T* out,
                                                cudaStream_t st);
template <class T> int32_t launch_wgrad_simt(const ConvGeom& g, const T* dout, const T* in, const WIndex& wi, float* dw, cudaStream_t st);
template <class TD> int32_t launch_pack_weights(const float* src, const WIndex& wi, int taps, int n_out, int c, TD* dst, cudaStream_t st,
                                                int round_bf16 = 0);
template <class T> int32_t launch_colsum(const T* x, long long rows, int C, float* partials, float* out, cudaStream_t st);

// umma_gemm.cu  (tcgen05 / TMA engine, bf16)
bool umma_conv_gemm_supported(const ConvGeom& g);
int umma_conv_gemm_grid(const ConvGeom& g);
int32_t launch_conv_gemm_umma(const ConvGeom& g, const bf16* in, const bf16* w, const GemmEpilogue& ep, bf16* out, cudaStream_t st);
// fp32 operands carried as (hi | mid | lo) bf16 triples: in3 [rows][3c], w3 [taps][n_out][3c]; fp32 output (UmmaParams::nterms)
int32_t launch_conv_gemm_umma_split(const ConvGeom& g, const bf16* in3, const bf16* w3, const GemmEpilogue& ep, float* out, cudaStream_t st);
// dst[r] = [bf16(x) | bf16(x - hi) | bf16(x - hi - mid)] over the C channels of src[r]   (3C bf16 per row)
int32_t launch_split_bf16(const float* src, long long rows, int C, bf16* dst, cudaStream_t st);

// umma_wgrad.cu  (tcgen05 weight gradient, bf16, MN-major operands)
bool umma_wgrad_supported(const ConvGeom& g);
size_t umma_wgrad_workspace_floats(const ConvGeom& g);
int32_t launch_wgrad_umma(const ConvGeom& g, const bf16* dout, const bf16* in, const WIndex& wi, float* packed_ws, float* dw,
                          cudaStream_t st, int c_valid = -1, bool packed_is_zero = false, cudaStream_t unpack_st = nullptr,
                          cudaEvent_t unpack_ev = nullptr, int split = 0);

// umma_agg.cu  (tensor-core adjacency aggregation, bf16)
bool umma_agg_supported(int V, int C, int K);
size_t umma_agg_prep_bytes(int K);
int32_t launch_build_big_adjacency(const float* A, int K, int V, void* prep_mem, cudaStream_t st);
int32_t launch_agg_fwd_umma(const bf16* x, int nb, int t, int v, int c, int k, const void* prep, const float* aff_a,
                            const float* aff_b, bf16* xa, cudaStream_t st);
int32_t launch_agg_bwd_umma(const bf16* dxa, int nb, int t, int v, int c, int k, const void* prep, const bf16* mask_h, bf16* dh,
                            cudaStream_t st, const bf16* bn_x = nullptr, const float* bn_coef = nullptr, float* dA = nullptr);
const float* umma_agg_rowsum_q(const void* prep, int k);
int32_t launch_agg_dA_umma(const bf16* dxa, const bf16* h, int nb, int t, int v, int c, int k, float* dA, cudaStream_t st);

// umma_spatial.cu  (fused aggregation + 1x1 channel mixing, forward)
bool umma_spatial_supported(int v, int ci, int co, int k);
int32_t launch_spatial_fused(const bf16* x, int nb, int t, int v, int ci, int co, int k, const bf16* big, const bf16* w2p,
                             const float* b2d, bf16* xa, bf16* s_out, cudaStream_t st);
const bf16* umma_agg_big(const void* prep, int k);

// train_prep.cu  (weight-only operands of every block of a stack in one launch)
struct PrepJob {
  int kind;                                  // 0: weight pack, 1: block-diagonal adjacency
  int taps, n_out, c; WIndex wi;             // kind 0
  int K, V, FT;                              // kind 1
  const float* src; bf16* dst;               // kind 0: weights -> packed; kind 1: src = A
  bf16* big; bf16* bigT; float* rowsum;      // kind 1
};
constexpr int kMaxPrepJobs = 48;
struct PrepTable { int count; PrepJob jobs[kMaxPrepJobs]; };
int32_t launch_train_prep_multi(const PrepTable& tab, cudaStream_t st);
// where the three parts of an aggregation prep buffer (umma_agg_prep_bytes) live
void umma_agg_prep_pointers(void* prep, int K, bf16** big, bf16** bigT, float** rowsum);

// head_kernels.cu
int32_t head_forward(int dtype, int n, int t, int v, int c, int classes, const void* feat, const float* fcw, const float* fcb,
                     float* pooled, float* logits, cudaStream_t st);
int32_t head_backward(int dtype, int n, int t, int v, int c, int classes, const float* pooled, const float* fcw,
                      const float* dlogits, void* dfeat, float* dfcw, float* dfcb, cudaStream_t st);

}  // namespace stgcn
