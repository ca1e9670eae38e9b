"""Drop-in modules for the reference's ST-GCN unit (src/network/st_graphconv.py).

Same constructor signatures, attribute names and state_dict keys/shapes as the
reference, so checkpoints load both ways and `optim.Adam(model.parameters())`
works unchanged.  The stock nn.Conv2d / nn.BatchNorm2d sub-modules are kept only
as PARAMETER CONTAINERS (and so that `torch.manual_seed` gives the reference's
initial weights); their forward is never called -- all arithmetic runs in
libstgcn_b200.so on the current CUDA stream.  No CPU fallback: a CPU tensor is an
error.
"""
import torch
import torch.nn as nn

from . import _lib as L
from . import functional as Fn


def _momentum(bn):
    if bn.momentum is None:
        # torch switches to a cumulative moving average then; the library implements the exponential one only
        raise RuntimeError('stgcn_b200: BatchNorm2d(momentum=None) (cumulative average) is not supported')
    return float(bn.momentum)


def prepare_training_operands(blocks, n, t, device):
    """Weight-only operands of a chain of SpatialTemporalConv blocks (packed temporal weights in both layouts, block-diagonal
    adjacency) in ONE launch (stgcn_train_prep_multi) instead of ~4 small launches per block inside the step.  `n`, `t`: batch
    and frame count of the input of blocks[0]; later blocks see the frame counts the chain produces.  Marks each block so that
    its next forward uses the operands; called by the models at the start of a training forward and by a stand-alone block for
    itself.  Must run after the last change of the parameters and before the forward that uses them (both are per forward)."""
    import ctypes as C
    import os
    # Off unless STGCN_TRAIN_PREP=1: measured on B200 (cfg4, same box, alternating runs) the step takes 6.12 ms with the batched
    # launch and 6.08-6.10 ms without it although 27 launches disappear -- the per-block packs already hide behind programmatic
    # dependent launch and the side lane, while the batched kernel sits at the head of the forward chain.
    if os.environ.get('STGCN_TRAIN_PREP', '0') in ('', '0'):
        for b in blocks:
            b._train_prep = None
            b._tp_fresh = True
        return
    lib = L.lib()
    k = len(blocks)
    descs = (L.BlockDesc * k)()
    params = (L.BlockParams * k)()
    preps = (C.c_void_p * k)()
    any_prep = False
    for i, b in enumerate(blocks):
        sc = b.spatialConv
        d = L.BlockDesc(n=n, c_in=b.C_in, c_out=b.C_out, t=t, v=sc.V, k=sc.K, gamma=b.gamma, stride=b.temporal_stride,
                        pad=b.temporal_padding, dtype=L.dtype_code(b.compute_dtype), training=1, residual=b._res_mode,
                        flags=b.flags, dropout_p=0.0, bn_eps=1e-5, bn_momentum=0.1, seed=0, offset=0)
        descs[i] = d
        nbytes = lib.stgcn_train_prep_bytes(C.byref(d)) if (sc.A.is_cuda and sc.A.device == device) else 0
        if nbytes:
            if b._train_prep is None or b._train_prep.numel() != nbytes or b._train_prep.device != device:
                b._train_prep = torch.empty(nbytes, dtype=torch.uint8, device=device)
            params[i].A = (b._A_override if b._A_override is not None else sc.A).data_ptr()
            params[i].Wt = b.temporalConv.weight.data_ptr()
            preps[i] = b._train_prep.data_ptr()
            b._tp_fresh = True
            any_prep = True
        else:
            b._train_prep = None
            b._tp_fresh = True          # nothing to prepare for this block (fp32 / CUDA-core shapes): do not ask again this forward
        t = (t + 2 * b.temporal_padding - b.gamma) // b.temporal_stride + 1
        if t < 1:
            break
    if any_prep:
        L.check(lib.stgcn_train_prep_multi(k, descs, params, preps, L.stream_ptr(device)), 'train_prep_multi')


class SpatialConv(nn.Module):
    """Spatial graph convolution, reference st_graphconv.py:111-152.

    out[n,c,t,v] = sum_k sum_w A[k,v,w] * (W f_in + b)[n, k*C_out + c, t, w]
    """

    def __init__(self, C_in, C_out, A, compute_dtype=torch.float32):
        super().__init__()
        self.C_in = C_in
        self.C_out = C_out
        self.A = nn.Parameter(A.float())       # detached trainable leaf, like st_graphconv.py:129
        self.K = self.A.shape[0]
        self.V = self.A.shape[1]
        self.W = nn.Conv2d(C_in, self.K * C_out, (1, 1))
        self.compute_dtype = compute_dtype
        self.flags = 0

    def forward(self, f_in):
        cfg = {'dtype': self.compute_dtype, 'flags': self.flags}
        w = self.W.weight.view(self.K * self.C_out, self.C_in)
        out = Fn.spatial_forward(f_in, cfg, self.A, w, self.W.bias)
        return out


class SpatialTemporalConv(nn.Module):
    """ST-GCN unit, reference st_graphconv.py:4-109 (default path :97-100, residual_block :60-82)."""

    def __init__(self, C_in, C_out, A, gamma, temporal_stride, temporal_padding, dropout_rate=0.5, residual=False,
                 compute_dtype=torch.float32):
        super().__init__()
        self._res_mode = L.RES_NONE
        if residual:
            if C_in == C_out and temporal_stride == 1:
                self.apply_residual = lambda x: x
                self._res_mode = L.RES_IDENTITY
            else:
                self.apply_residual = nn.Conv2d(C_in, C_out, kernel_size=1, stride=(temporal_stride, 1))
                self._res_mode = L.RES_CONV
        self.residual = residual
        self.batch_n = nn.BatchNorm2d(C_in)
        self.spatialConv = SpatialConv(C_in, C_out, A, compute_dtype)
        self.temporalConv = nn.Conv2d(C_out, C_out, kernel_size=(gamma, 1), stride=(temporal_stride, 1),
                                      padding=(temporal_padding, 0))
        self.batch_n_2 = nn.BatchNorm2d(C_out)
        self.relu = nn.ReLU(inplace=True)
        self.dropout = nn.Dropout(dropout_rate, inplace=True) if dropout_rate != 0 else None
        self.C_in, self.C_out = C_in, C_out
        self.gamma, self.temporal_stride, self.temporal_padding = gamma, temporal_stride, temporal_padding
        self.compute_dtype = compute_dtype
        self.flags = 0
        self._calls = 0
        # dropout generator {seed, offset} on the device (advanced by the library, so a captured step draws fresh masks);
        # non-persistent: the state_dict keeps the reference's keys
        self.register_buffer('_rng_state', None, persistent=False)
        self._prep_cache = {}      # frozen-parameter inference operands (functional._BlockFn), keyed by tensor versions
        self._A_override = None    # effective adjacency of this forward when the model learns edge-importance masks
        self._train_prep = None    # weight-only training operands (prepare_training_operands), refreshed every training forward
        self._tp_fresh = False
        self._sync = None          # SyncBN link (parallel.enable_sync_batchnorm): training statistics over all ranks

    def _ensure_rng(self, device):
        """Create the device-resident dropout generator on first use: seeded once from torch's CPU generator (so
        torch.manual_seed makes runs repeatable), never read back by the host."""
        if self.dropout is not None and (self._rng_state is None or self._rng_state.device != device):
            self._rng_state = torch.tensor([int(torch.randint(0, 2 ** 62, (1,))), 0], dtype=torch.int64).to(device)

    def forward(self, f_in):
        sc, tc, b1, b2 = self.spatialConv, self.temporalConv, self.batch_n, self.batch_n_2
        # learnable edge importance (model.py, beyond the reference): the model hands in A_base * Mask for this forward
        A_eff = self._A_override if self._A_override is not None else sc.A
        params = [b1.weight, b1.bias, A_eff, sc.W.weight.view(sc.K * sc.C_out, sc.C_in), sc.W.bias,
                  tc.weight.view(self.C_out, self.C_out, self.gamma), tc.bias, b2.weight, b2.bias]
        raw = [b1.weight, b1.bias, sc.A if self._A_override is None else None, sc.W.weight, sc.W.bias, tc.weight, tc.bias,
               b2.weight, b2.bias]
        if self._res_mode == L.RES_CONV:
            params += [self.apply_residual.weight.view(self.C_out, self.C_in), self.apply_residual.bias]
            raw += [self.apply_residual.weight, self.apply_residual.bias]
        else:
            params += [None, None]
            raw += [None, None]
        need_grad = torch.is_grad_enabled() and (f_in.requires_grad or any(p is not None and p.requires_grad for p in params))
        p_drop = float(self.dropout.p) if (self.dropout is not None and self.training) else 0.0
        if p_drop > 0:
            self._ensure_rng(f_in.device)
        self._calls += 1
        if self.training:
            self._prep_cache.clear()       # running statistics are about to change behind torch's version counters
        cfg = {'dtype': self.compute_dtype, 'c_in': self.C_in, 'c_out': self.C_out, 'gamma': self.gamma,
               'stride': self.temporal_stride, 'pad': self.temporal_padding, 'training': self.training,
               'residual': self._res_mode, 'dropout_p': p_drop, 'eps': b1.eps,
               'momentum': _momentum(b1), 'seed': 0, 'need_grad': need_grad,
               'rng_state': self._rng_state if p_drop > 0 else None,
               'grad_bufs': [Fn.grad_buf_of(q) for q in raw],
               'flags': self.flags, 'prep_cache': self._prep_cache}
        if self._sync is not None and self.training:
            cfg['sync'] = self._sync.on(f_in.device, max(self.C_in, self.C_out))
        if need_grad and f_in.is_cuda:
            if not self._tp_fresh:          # stand-alone use: prepare for this block alone (models do it for the whole stack)
                prepare_training_operands([self], f_in.shape[0], f_in.shape[2], f_in.device)
            self._tp_fresh = False
            cfg['train_prep'] = self._train_prep
        buffers = (b1.running_mean, b1.running_var, b1.num_batches_tracked,
                   b2.running_mean, b2.running_var, b2.num_batches_tracked)
        # chained blocks: the block that produced f_in may have left its column statistics on the tensor (valid only for this
        # very tensor, unmodified since, and only used for training-mode BatchNorm-1)
        cs = getattr(f_in, '_stgcn_colstats', None)
        if (cs is not None and self.training and cs[2] == f_in._version and cs[3] == self.compute_dtype
                and f_in.dtype == self.compute_dtype and cs[4] == self.C_in):
            cfg['x_stats'] = (cs[0], cs[1])
        y = Fn.block_forward(f_in, cfg, buffers, params)
        ys = cfg.get('y_stats')
        if ys is not None:
            y._stgcn_colstats = (ys[0], ys[1], y._version, self.compute_dtype, self.C_out)
        return y
