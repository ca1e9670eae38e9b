/*
 * stgcn_b200 -- C-ABI of the B200-native ST-GCN block path.
 *
 * The reference (nagyrajmund/st-gcn) has no FFI layer: its hot path is the
 * Python class SpatialTemporalConv (src/network/st_graphconv.py:4-109) calling
 * stock torch modules.  This header is the boundary a maintainer binds instead
 * (ctypes stub in INTEGRATION.md): plain pointers and sizes, no torch types.
 *
 * Conventions
 *   - every data pointer is a DEVICE pointer owned by the caller (PyTorch);
 *     the library never allocates or frees caller-visible memory;
 *   - activations are channels-last: x[n][t][v][c] (logical (N,C,T,V) with
 *     torch.channels_last strides), element type given by StgcnBlockDesc.dtype;
 *   - parameters and parameter gradients are fp32 in the reference's
 *     state_dict layouts (st_graphconv.py:28,34,41-46,129,136);
 *   - all launches go to the cudaStream_t passed in (as void*), no host sync,
 *     no allocation: CUDA-graph capturable;
 *   - return value 0 = ok, negative = error; stgcn_last_error() gives the text.
 *     There is no CPU fallback: unsupported input is an error.
 */
#ifndef STGCN_B200_H_
#define STGCN_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define STGCN_OK 0
#define STGCN_ERR_INVALID (-1)
#define STGCN_ERR_UNSUPPORTED (-2)
#define STGCN_ERR_CUDA (-3)
#define STGCN_ERR_WORKSPACE (-4)

enum { STGCN_F32 = 0, STGCN_BF16 = 1 };
enum { STGCN_RES_NONE = 0, STGCN_RES_IDENTITY = 1, STGCN_RES_CONV = 2 };
/* StgcnBlockDesc.flags */
enum {
  STGCN_FLAG_FORCE_SIMT = 1,     /* bf16 storage but CUDA-core GEMMs (debug / A-B) */
};

/* One SpatialTemporalConv call: replaces SpatialTemporalConv.__init__ arguments
 * (st_graphconv.py:9) + the runtime shape of forward's f_in (st_graphconv.py:85). */
typedef struct StgcnBlockDesc {
  int32_t n;          /* batch (persons folded in), f_in.shape[0]            */
  int32_t c_in;       /* C_in                                                 */
  int32_t c_out;      /* C_out                                                */
  int32_t t;          /* input frames T                                       */
  int32_t v;          /* joints V = A.shape[1]  (<= 32)                       */
  int32_t k;          /* partitions K = A.shape[0] (<= 4)                     */
  int32_t gamma;      /* temporal kernel size (odd, <= 15)                    */
  int32_t stride;     /* temporal stride (1 or 2)                             */
  int32_t pad;        /* temporal zero padding                                */
  int32_t dtype;      /* STGCN_F32 | STGCN_BF16 : activation storage type     */
  int32_t training;   /* module.training: batch statistics + running update   */
  int32_t residual;   /* STGCN_RES_*  (st_graphconv.py:23-28)                 */
  int32_t flags;
  float dropout_p;    /* nn.Dropout p, 0 = none (st_graphconv.py:53-58)       */
  float bn_eps;       /* 1e-5                                                 */
  float bn_momentum;  /* 0.1                                                  */
  uint64_t seed;      /* Philox seed / offset for dropout                     */
  uint64_t offset;
} StgcnBlockDesc;

/* Parameters and buffers, named like the reference's state_dict. */
typedef struct StgcnBlockParams {
  const float* bn1_weight;  const float* bn1_bias;      /* batch_n.*    (C_in)  */
  float* bn1_running_mean;  float* bn1_running_var;     /* updated in training  */
  int64_t* bn1_num_batches_tracked;
  const float* A;                                       /* spatialConv.A (K,V,V) */
  const float* W;                                       /* spatialConv.W.weight (K*C_out, C_in) */
  const float* W_bias;                                  /* spatialConv.W.bias (K*C_out) */
  const float* Wt;                                      /* temporalConv.weight (C_out,C_out,gamma) */
  const float* Wt_bias;                                 /* temporalConv.bias (C_out) */
  const float* bn2_weight;  const float* bn2_bias;      /* batch_n_2.*  (C_out) */
  float* bn2_running_mean;  float* bn2_running_var;
  int64_t* bn2_num_batches_tracked;
  const float* Wr;                                      /* apply_residual.weight (C_out,C_in) or NULL */
  const float* Wr_bias;                                 /* apply_residual.bias (C_out) or NULL */
  /* nn.Dropout state (st_graphconv.py:53-58), DEVICE resident: rng_state[0] = Philox seed, rng_state[1] = offset.  When non-NULL
   * (and training with dropout_p > 0) forward takes its mask from this state and advances the offset by ceil(outputs / 4), so a
   * CUDA graph of the step draws a fresh mask per replay; the state used is kept in `saved` and replayed by backward.  NULL: the
   * descriptor's seed / offset values are used. */
  uint64_t* rng_state;
  /* Optional (NULL = derive inside the call): the block's weight-only training operands as written by stgcn_train_prep_multi --
   * valid for as long as A and temporalConv.weight are unchanged.  Pass the same pointer to forward and backward. */
  const void* train_prep;
  /* Optional SyncBN (global-batch BatchNorm statistics under data parallel; no reference counterpart -- the reference is single
   * process -- SURVEY.md 8(f)-4).  When sync_allreduce != NULL and sync_world > 1, every training-mode BatchNorm of the block uses
   * the statistics of the batch over ALL ranks: at each of its four exchange points (BatchNorm-1 / -2, forward and backward) the
   * library writes the 2*C local sums to sync_buf (device, >= 2 * max(C_in, C_out) floats), calls sync_allreduce(sync_user, stream)
   * -- which must enqueue an in-place SUM all-reduce of sync_buf over the ranks on `stream` (e.g. ncclAllReduce, or
   * torch.distributed.all_reduce with that stream current) and return without synchronising -- and continues from the global
   * sums with sync_world times the local row count.  Parameter gradients stay local sums (the data-parallel gradient all-reduce
   * completes them).  Every rank must make the same calls in the same order. */
  float* sync_buf;
  void (*sync_allreduce)(void* user, void* stream);
  void* sync_user;
  int64_t sync_world;
} StgcnBlockParams;

/* Parameter gradients (fp32, overwritten).  Same shapes as the parameters. */
typedef struct StgcnBlockGrads {
  float* bn1_weight; float* bn1_bias;
  float* A;
  float* W; float* W_bias;
  float* Wt; float* Wt_bias;
  float* bn2_weight; float* bn2_bias;
  float* Wr; float* Wr_bias;
} StgcnBlockGrads;

/* T_out = (T + 2*pad - gamma)/stride + 1 */
int32_t stgcn_block_t_out(const StgcnBlockDesc* desc);

/* Bytes of the `saved` buffer forward fills for backward (0 when not training
 * is still valid: eval-mode backward is supported and saves the same tensors). */
size_t stgcn_block_saved_bytes(const StgcnBlockDesc* desc);

/* Scratch bytes needed by forward (backward=0) or backward (backward=1). */
size_t stgcn_block_workspace_bytes(const StgcnBlockDesc* desc, int32_t backward);

/* SpatialTemporalConv.forward (st_graphconv.py:85-109), both the default path
 * (:97-100) and residual_block (:60-82), trailing ReLU (:105) and dropout (:108).
 *   x (n,t,v,c_in) -> y (n,t_out,v,c_out); saved may be NULL when no backward
 *   will follow. */
int32_t stgcn_block_forward(const StgcnBlockDesc* desc, const void* x, const StgcnBlockParams* params,
                            void* y, void* saved, void* workspace, size_t workspace_bytes, void* stream);

/* Chained blocks (training): the BatchNorm-1 batch statistics of block i+1 (st_graphconv.py:34, :98) are column statistics of
 * block i's output, which block i's last pass can produce for free.  Same as stgcn_block_forward, plus
 *   x_colstats / x_colstats_nblk : per-CTA partial column sums of x written by the call that produced x (NULL / 0: none)
 *   y_colstats / y_colstats_nblk : buffer of stgcn_colstats_bytes(c_out) bytes to receive the partials of y; *nblk is set to
 *                                  the number of partial rows written, 0 when this path cannot produce them (dropout, residual
 *                                  block, eval mode) -- the consumer then computes its own statistics.
 * The statistics describe y as stored; they are stale once y is modified. */
size_t stgcn_colstats_bytes(int32_t channels);
int32_t stgcn_block_forward_chain(const StgcnBlockDesc* desc, const void* x, const StgcnBlockParams* params, void* y, void* saved,
                                  void* workspace, size_t workspace_bytes, const float* x_colstats, int32_t x_colstats_nblk,
                                  float* y_colstats, int32_t* y_colstats_nblk, void* stream);

/* Inference with frozen parameters (no reference counterpart: the reference re-derives nothing, it calls cuDNN per layer).
 * Everything that depends only on the parameters -- eval-mode BatchNorm tables (st_graphconv.py:34,46), the BatchNorm-1-folded
 * 1x1 weights, the joint-bias table, the block-diagonal adjacency operand and the packed temporal weights -- is computed once by
 * stgcn_block_prepare into a caller-owned device buffer of stgcn_block_prepared_bytes(desc) bytes; stgcn_block_forward_prepared
 * then runs SpatialTemporalConv.forward (eval mode, default block) as two kernels.  The buffer does not depend on n or t.
 * stgcn_block_prepared_bytes returns 0 when the block has no prepared path (training mode, fp32, residual block, C_in % 64 != 0);
 * callers then use stgcn_block_forward.  The caller must prepare again whenever a parameter or a running statistic changes. */
size_t stgcn_block_prepared_bytes(const StgcnBlockDesc* desc);
int32_t stgcn_block_prepare(const StgcnBlockDesc* desc, const StgcnBlockParams* params, void* prepared, void* stream);
int32_t stgcn_block_forward_prepared(const StgcnBlockDesc* desc, const void* x, const StgcnBlockParams* params, const void* prepared,
                                     void* y, void* workspace, size_t workspace_bytes, void* stream);

/* Weight-only operands of the training chain for a whole stack of blocks in ONE launch (no reference counterpart: the reference
 * calls cuDNN per layer).  For each block i: the temporal weights (st_graphconv.py:41-43) packed as bf16 for the forward GEMM and
 * for its data gradient, and the block-diagonal bf16 expansions of A (st_graphconv.py:129) the tensor-core aggregation uses, written
 * to preps[i] (a device buffer of stgcn_train_prep_bytes(&descs[i]) bytes; 0 = this block has no such operands: fp32, CUDA-core
 * shapes).  Call it once after the parameters changed (after the optimizer step / at the start of a training forward) and hand
 * preps[i] to the block calls through StgcnBlockParams.train_prep; the ~40 small launches those operands otherwise cost per step
 * (measured: 0.33 ms of a 6.09 ms cfg4 step) collapse into one.  descs / params / preps are HOST arrays. */
size_t stgcn_train_prep_bytes(const StgcnBlockDesc* desc);
int32_t stgcn_train_prep_multi(int32_t count, const StgcnBlockDesc* descs, const StgcnBlockParams* params, void* const* preps,
                               void* stream);

/* Autograd of the above (reference: torch autograd over st_graphconv.py:85-109).
 *   dy (n,t_out,v,c_out) -> dx (n,t,v,c_in) (may be NULL: first block) + grads. */
int32_t stgcn_block_backward(const StgcnBlockDesc* desc, const void* x, const void* y,
                             const StgcnBlockParams* params, const void* saved, const void* dy,
                             void* dx, const StgcnBlockGrads* grads,
                             void* workspace, size_t workspace_bytes, void* stream);

/* SpatialConv.forward alone (st_graphconv.py:139-152), for the SpatialConv
 * module: x (n,t,v,c_in) -> out (n,t,v,c_out); no BN, no activation. */
size_t stgcn_spatial_workspace_bytes(const StgcnBlockDesc* desc, int32_t backward);
int32_t stgcn_spatial_forward(const StgcnBlockDesc* desc, const void* x, const float* A, const float* W,
                              const float* W_bias, void* out, void* workspace, size_t workspace_bytes,
                              void* stream);
int32_t stgcn_spatial_backward(const StgcnBlockDesc* desc, const void* x, const float* A, const float* W,
                               const float* W_bias, const void* dout, void* dx, float* dA, float* dW,
                               float* dW_bias, void* workspace, size_t workspace_bytes, void* stream);

/* SpatialConv.forward WITHOUT a caller workspace and without any allocation.  No counterpart in the reference: added
 * for callers that cannot allocate (code under CUDA-graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  Same result as stgcn_spatial_forward; the scratch (aggregated input, joint-bias table, packed weights)
 * lives in a static device array of STGCN_STATIC_SCRATCH_BYTES inside the library, which every GPU that loads the library holds
 * until exit.  Shapes whose stgcn_spatial_workspace_bytes(desc, 0) exceeds that limit are rejected with STGCN_ERR_WORKSPACE and a
 * message naming both sizes.  The call launches on `stream`, never calls cudaMalloc / cudaFree or an allocating library routine.
 * ALL calls share the one array (it keeps the last call's contents; the part a call uses is cleared first): calls that overlap on
 * the GPU -- two streams, two threads -- race; serialise them. */
#define STGCN_STATIC_SCRATCH_BYTES (16u << 20)
size_t stgcn_static_scratch_bytes(void);
int32_t stgcn_spatial_forward_static(const StgcnBlockDesc* desc, const void* x, const float* A, const float* W,
                                     const float* W_bias, void* out, void* stream);

/* Head of L_STGCN.forward (lightning_model.py:105-107): global average pool
 * over (T,V) + Linear(C, classes).  feat (n,t,v,c) -> logits (n,classes) fp32. */
int32_t stgcn_head_forward(int32_t dtype, int32_t n, int32_t t, int32_t v, int32_t c, int32_t classes,
                           const void* feat, const float* fc_weight, const float* fc_bias,
                           float* pooled /* (n,c) */, float* logits, void* stream);
int32_t stgcn_head_backward(int32_t dtype, int32_t n, int32_t t, int32_t v, int32_t c, int32_t classes,
                            const float* pooled, const float* fc_weight, const float* dlogits,
                            void* dfeat, float* dfc_weight, float* dfc_bias, void* stream);

/* Loss of the reference's training_step (src/lightning_model.py:202: F.cross_entropy(output, y), mean reduction).
 * One launch: loss[0] = mean over rows with labels[r] != ignore_index of (logsumexp(logits[r]) - logits[r][labels[r]]) and
 * dlogits_unit = d loss / d logits (n,classes) for an incoming gradient of 1.  _backward scales it by grad_loss[0] (device). */
int32_t stgcn_cross_entropy_forward(int32_t n, int32_t classes, const float* logits, const int64_t* labels,
                                    int64_t ignore_index, float* loss, float* dlogits_unit, void* stream);
int32_t stgcn_cross_entropy_backward(int32_t n, int32_t classes, const float* dlogits_unit, const float* grad_loss,
                                     float* dlogits, void* stream);

/* Optimizer of the reference (src/lightning_model.py:196-197: optim.Adam(self.parameters(), lr)): one step of torch.optim.Adam
 * (no amsgrad, L2 weight decay) over `count` fp32 tensors.  The four tables are HOST arrays of DEVICE pointers, sizes[i] =
 * elements of tensor i.  The step number lives on the DEVICE: step_state points to 16 bytes {int64 step; float[2] scratch}
 * (zero-initialised by the caller).  A one-thread kernel advances it by one (step == 0) or sets it to `step` (step >= 1) and forms
 * the bias corrections in double before the update runs, so a CUDA graph of the training step replays correctly.
 * One update launch per 64 tensors. */
int32_t stgcn_adam_step(int32_t count, float* const* params, const float* const* grads, float* const* exp_avg,
                        float* const* exp_avg_sq, const int64_t* sizes, double lr, double beta1, double beta2, double eps,
                        double weight_decay, int64_t step, void* step_state, void* stream);

/* Input pipeline in front of block 0: loop-padding collate (loopy_pad_collate_fn / pad_array_with_loops, src/data/util.py:12-47),
 * the 2-D affine augmentation (augment_data, src/data/augmentation.py:8-69) and the cast to the storage type, one launch.
 * `packed` holds the n ragged sequences back to back in fp32: sequence i is lengths[i] frames of v*c values starting at element
 * offsets[i] (offsets / lengths are DEVICE arrays).  out (n, t_out, v, c) of `dtype`: out[i][t] = seq_i[t mod lengths[i]]
 * (np.pad mode='wrap'); with xform != NULL (device, n x 4 floats = the T[:2,:2] block of the reference's 3x3 matrix, row major)
 * the first two channels of every joint become (x T00 + y T10, x T01 + y T11) -- np.dot(frames, T) with a zero homogeneous
 * coordinate, so T's translation entries never act, exactly as in the reference. */
int32_t stgcn_collate_pad_augment(int32_t dtype, int32_t n, int32_t t_out, int32_t v, int32_t c, const float* packed,
                                  const int64_t* offsets, const int32_t* lengths, const float* xform, void* out, void* stream);

/* Keep-scale of nn.Dropout(p) as the library draws it (Philox4x32-10): out[i] = 1/(1-p) if element i of the block output
 * (flat index over (n,t_out,v,c_out)) is kept under state {seed, offset}, else 0.  For replaying a mask in tests / oracles. */
int32_t stgcn_dropout_mask(uint64_t seed, uint64_t offset, float p, size_t count, float* out, void* stream);

/* Adjacency normalisation of the model constructor on the device (normalize, src/data/adjacency.py:161-183):
 * out[k] = L . raw[k] . L with L = (diag(rowsum(raw[k])) + alpha) ** expo, alpha added to EVERY entry of the diagonal matrix when
 * dense_alpha != 0 (what the reference computes: a dense result with entries ~1e4..1e5) or to the diagonal only (dense_alpha = 0:
 * the D^-1/2 A D^-1/2 its comment intends).  raw, out: (k, v, v) fp32 device tensors of 0/1 partition matrices, v <= 32. */
int32_t stgcn_adjacency_normalize(int32_t k, int32_t v, const float* raw, float alpha, float expo, int32_t dense_alpha, float* out,
                                  void* stream);

/* Layout helpers: fp32 (n,t,v,c) <-> dtype (n,t,v,c) casts on device. */
int32_t stgcn_cast(int32_t src_dtype, int32_t dst_dtype, const void* src, void* dst, size_t count, void* stream);

/* tcgen05 implicit-GEMM engine exposed for unit tests:
 *   out[r][n] = sum_tap sum_c in[row(r,tap)][c] * w[tap][n][c]   (bf16 in, fp32 accumulate, bf16 out)
 * in (nb,t_in,v,c) ; w (taps,n_out,c) bf16 ; out (nb,t_out,v,n_out) bf16;
 * input frame for output frame j and tap i:  j*stride - pad + i  (zero outside). */
int32_t stgcn_umma_conv_gemm(int32_t nb, int32_t t_in, int32_t t_out, int32_t v, int32_t c, int32_t n_out,
                             int32_t taps, int32_t stride, int32_t pad, const void* in, const void* w,
                             const float* bias, void* out, void* stream);

/* tcgen05 descriptor probe (tests only): mode 0 = K-major A with a start address shifted by row_off rows,
 * mode 1 = MN-major A (a_rows x 128) and B (64 x 64): D[m][n] = sum_k A[k+row_off][m]*B[k][n]. d: 128x64 fp32.
 * base_off < 0 leaves the descriptor's base-offset field 0. */
int32_t stgcn_umma_probe(int32_t mode, int32_t row_off, int32_t base_off, int32_t a_rows, const void* a,
                         const void* b, float* d, void* stream);

/* Optional device timing per kernel class (cudaEvent pairs on the launching stream; off by default). */
void stgcn_timing_enable(int32_t on);
void stgcn_timing_reset(void);
int32_t stgcn_timing_classes(void);
const char* stgcn_timing_class_name(int32_t cls);
int32_t stgcn_timing_read(int32_t cls, double* ms, double* work, uint64_t* launches);

/* sizeof of the library's own view of a boundary struct: which = 0 StgcnBlockDesc, 1 StgcnBlockParams, 2 StgcnBlockGrads (else 0).
 * A binding compares it with its mirror of the struct at load time. */
size_t stgcn_abi_sizeof(int32_t which);

/* Number of CUDA kernels this library has launched in this process (bench's gpu_launches). */
uint64_t stgcn_launch_count(void);
const char* stgcn_last_error(void);
const char* stgcn_version(void);

#ifdef __cplusplus
}
#endif
#endif  /* STGCN_B200_H_ */
