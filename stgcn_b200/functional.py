"""torch.autograd bridges over the C ABI.

Tensors cross the boundary as raw device pointers of channels-last buffers
(N,T,V,C); logical shapes seen by callers stay (N,C,T,V) like the reference
(st_graphconv.py:85-109).
"""
import ctypes as C

import torch

from . import _lib as L


def _to_ntvc(x, dtype):
    """logical (N,C,T,V) -> contiguous physical (N,T,V,C) of `dtype` (no copy when already so).  A dtype change of an already
    channels-last CUDA tensor (block 0's fp32 input on the bf16 path) is done by the library's own cast kernel, so that no torch
    kernel runs on the block path; anything that needs a real transpose falls back to torch (outside the benchmarked chain)."""
    xp = x.permute(0, 2, 3, 1)
    if xp.dtype == dtype:
        return xp.contiguous()
    if xp.is_cuda and xp.is_contiguous() and not (torch.is_grad_enabled() and x.requires_grad) \
            and xp.dtype in (torch.float32, torch.bfloat16) and dtype in (torch.float32, torch.bfloat16):
        out = torch.empty(xp.shape, dtype=dtype, device=xp.device)
        L.check(L.lib().stgcn_cast(L.dtype_code(xp.dtype), L.dtype_code(dtype), L.ptr(xp), L.ptr(out), xp.numel(),
                                   L.stream_ptr(xp.device)), 'cast')
        return out
    return xp.to(dtype).contiguous()


def _from_ntvc(y):
    return y.permute(0, 3, 1, 2)


def _f32(t):
    if t is None:
        return None
    if t.dtype != torch.float32 or not t.is_contiguous():
        raise RuntimeError('stgcn_b200: parameters must be contiguous float32 tensors')
    return t


def _grad_out(param, buf):
    """Tensor a parameter gradient is written to: the caller-provided slice (same element count, fp32, same device) or a new one."""
    if buf is not None and buf.numel() == param.numel() and buf.dtype == param.dtype and buf.device == param.device:
        return buf.view(param.shape)
    return torch.empty_like(param)


def grad_buf_of(p):
    """The data-parallel reducer's bucket slice for parameter `p`, when backward may write the gradient straight into it: only
    while `p.grad is None` (zero_grad(set_to_none=True), torch's default).  If `p.grad` exists it already aliases that slice, and
    autograd would add the freshly written gradient to ITSELF; a separate tensor is used then and accumulates correctly."""
    if p is None or p.grad is not None:
        return None
    return getattr(p, '_stgcn_grad_buf', None)


class _BlockFn(torch.autograd.Function):
    """SpatialTemporalConv.forward / its autograd (st_graphconv.py:85-109)."""

    @staticmethod
    def forward(ctx, x, cfg, buffers, bn1_w, bn1_b, A, W, Wb, Wt, bt, bn2_w, bn2_b, Wr, Wrb):
        # cfg['rng_state']: int64[2] device tensor {seed, offset} of the block's dropout generator (None without dropout)
        L.require_cuda(x, 'f_in')
        lib = L.lib()
        dt = cfg['dtype']
        xp = _to_ntvc(x, dt)
        n, t, v, ci = xp.shape
        if ci != cfg['c_in'] or v != A.shape[1]:
            raise RuntimeError('stgcn_b200: f_in has shape (N,%d,T,%d), block expects C_in=%d V=%d'
                               % (ci, v, cfg['c_in'], A.shape[1]))
        d = L.BlockDesc(n=n, c_in=ci, c_out=cfg['c_out'], t=t, v=v, k=A.shape[0], gamma=cfg['gamma'],
                        stride=cfg['stride'], pad=cfg['pad'], dtype=L.dtype_code(dt), training=int(cfg['training']),
                        residual=cfg['residual'], flags=cfg.get('flags', 0), dropout_p=cfg['dropout_p'],
                        bn_eps=cfg['eps'], bn_momentum=cfg['momentum'], seed=cfg['seed'], offset=0)
        t_out = lib.stgcn_block_t_out(C.byref(d))
        if t_out < 1:
            raise RuntimeError('stgcn_b200: sequence too short for the temporal kernel')
        need_grad = cfg['need_grad']
        rm1, rv1, nbt1, rm2, rv2, nbt2 = buffers
        P = L.BlockParams(L.ptr(_f32(bn1_w)), L.ptr(_f32(bn1_b)), L.ptr(rm1), L.ptr(rv1), L.ptr(nbt1),
                          L.ptr(_f32(A)), L.ptr(_f32(W)), L.ptr(_f32(Wb)), L.ptr(_f32(Wt)), L.ptr(_f32(bt)),
                          L.ptr(_f32(bn2_w)), L.ptr(_f32(bn2_b)), L.ptr(rm2), L.ptr(rv2), L.ptr(nbt2),
                          L.ptr(_f32(Wr)), L.ptr(_f32(Wrb)), L.ptr(cfg.get('rng_state')), L.ptr(cfg.get('train_prep'))).set_sync(cfg.get('sync'))
        y = torch.empty((n, t_out, v, cfg['c_out']), dtype=dt, device=x.device)
        saved = None
        if need_grad:
            sb = lib.stgcn_block_saved_bytes(C.byref(d))
            saved = torch.empty(sb, dtype=torch.uint8, device=x.device)
        wsb = lib.stgcn_block_workspace_bytes(C.byref(d), 0)
        if wsb == 0:
            raise RuntimeError('stgcn_b200: ' + lib.stgcn_last_error().decode())
        ws = L.workspace(x.device, wsb)
        cache = cfg.get('prep_cache')
        blob = None
        if cache is not None and not need_grad and not cfg['training']:
            # frozen-parameter inference: parameter-only operands are prepared once and reused until a parameter or buffer
            # changes (tensor version counters; the owning module also clears the cache after every training-mode forward,
            # because the library updates the running statistics through raw pointers)
            tensors = (bn1_w, bn1_b, A, W, Wb, Wt, bt, bn2_w, bn2_b, rm1, rv1, rm2, rv2)
            key = (dt, str(x.device), ci, v, tuple((t_.data_ptr(), t_._version) for t_ in tensors))
            if cache.get('key') != key:
                pb = lib.stgcn_block_prepared_bytes(C.byref(d))
                cache['blob'] = None
                if pb > 0:
                    buf = torch.empty(pb, dtype=torch.uint8, device=x.device)
                    L.check(lib.stgcn_block_prepare(C.byref(d), C.byref(P), L.ptr(buf), L.stream_ptr(x.device)), 'block_prepare')
                    cache['blob'] = buf
                cache['key'] = key
            blob = cache['blob']
        if blob is not None:
            L.check(lib.stgcn_block_forward_prepared(C.byref(d), L.ptr(xp), C.byref(P), L.ptr(blob), L.ptr(y), L.ptr(ws),
                                                     ws.numel(), L.stream_ptr(x.device)), 'block_forward_prepared')
        else:
            # chained blocks: column statistics of x handed over by the block that produced it, and those of y for the next one
            xs = cfg.get('x_stats')
            ys = None
            if cfg['training'] and cfg['residual'] == L.RES_NONE and cfg['dropout_p'] == 0.0:
                ys = torch.empty(lib.stgcn_colstats_bytes(cfg['c_out']) // 4, dtype=torch.float32, device=x.device)
            nb = C.c_int32(0)
            L.check(lib.stgcn_block_forward_chain(C.byref(d), L.ptr(xp), C.byref(P), L.ptr(y), L.ptr(saved), L.ptr(ws), ws.numel(),
                                                  L.ptr(xs[0]) if xs else None, xs[1] if xs else 0, L.ptr(ys), C.byref(nb),
                                                  L.stream_ptr(x.device)), 'block_forward')
            cfg['y_stats'] = (ys, nb.value) if (ys is not None and nb.value > 0) else None
        if need_grad:
            ctx.d = d
            ctx.grad_bufs = cfg.get('grad_bufs')
            ctx.sync = cfg.get('sync')                    # SyncBN link (parallel.enable_sync_batchnorm): backward exchanges sums too
            ctx.train_prep = cfg.get('train_prep')        # the same weight-only operands must reach backward (and stay alive)
            ctx.buffers = buffers
            ctx.x_dtype = x.dtype
            ctx.has_res_conv = Wr is not None
            ctx.save_for_backward(xp, y, saved, bn1_w, bn1_b, A, W, Wb, Wt, bt, bn2_w, bn2_b, Wr, Wrb)
        return _from_ntvc(y)

    @staticmethod
    def backward(ctx, gy):
        lib = L.lib()
        xp, y, saved, bn1_w, bn1_b, A, W, Wb, Wt, bt, bn2_w, bn2_b, Wr, Wrb = ctx.saved_tensors
        d = ctx.d
        rm1, rv1, nbt1, rm2, rv2, nbt2 = ctx.buffers
        dyp = _to_ntvc(gy, y.dtype)
        P = L.BlockParams(L.ptr(bn1_w), L.ptr(bn1_b), L.ptr(rm1), L.ptr(rv1), L.ptr(nbt1), L.ptr(A), L.ptr(W),
                          L.ptr(Wb), L.ptr(Wt), L.ptr(bt), L.ptr(bn2_w), L.ptr(bn2_b), L.ptr(rm2), L.ptr(rv2),
                          L.ptr(nbt2), L.ptr(Wr), L.ptr(Wrb), None, L.ptr(ctx.train_prep)).set_sync(ctx.sync)
        # gradients go straight into the data-parallel reducer's bucket slices when there are any (parallel.GradBucketReducer)
        gb = ctx.grad_bufs or [None] * 11
        g = {k: _grad_out(v, gb[i]) for i, (k, v) in enumerate((('bn1_w', bn1_w), ('bn1_b', bn1_b), ('A', A), ('W', W), ('Wb', Wb),
                                                                 ('Wt', Wt), ('bt', bt), ('bn2_w', bn2_w), ('bn2_b', bn2_b)))}
        gWr = _grad_out(Wr, gb[9]) if Wr is not None else None
        gWrb = _grad_out(Wrb, gb[10]) if Wrb is not None else None
        G = L.BlockGrads(L.ptr(g['bn1_w']), L.ptr(g['bn1_b']), L.ptr(g['A']), L.ptr(g['W']), L.ptr(g['Wb']),
                         L.ptr(g['Wt']), L.ptr(g['bt']), L.ptr(g['bn2_w']), L.ptr(g['bn2_b']), L.ptr(gWr), L.ptr(gWrb))
        dx = torch.empty_like(xp) if ctx.needs_input_grad[0] else None
        wsb = lib.stgcn_block_workspace_bytes(C.byref(d), 1)
        ws = L.workspace(xp.device, wsb)
        L.check(lib.stgcn_block_backward(C.byref(d), L.ptr(xp), L.ptr(y), C.byref(P), L.ptr(saved), L.ptr(dyp),
                                         L.ptr(dx), C.byref(G), L.ptr(ws), ws.numel(), L.stream_ptr(xp.device)),
                'block_backward')
        gx = _from_ntvc(dx).to(ctx.x_dtype) if dx is not None else None
        return (gx, None, None, g['bn1_w'], g['bn1_b'], g['A'], g['W'], g['Wb'], g['Wt'], g['bt'], g['bn2_w'],
                g['bn2_b'], gWr, gWrb)


def block_forward(x, cfg, buffers, params):
    return _BlockFn.apply(x, cfg, buffers, *params)


class _SpatialFn(torch.autograd.Function):
    """SpatialConv.forward / its autograd (st_graphconv.py:139-152)."""

    @staticmethod
    def forward(ctx, x, cfg, A, W, Wb):
        L.require_cuda(x, 'f_in')
        lib = L.lib()
        dt = cfg['dtype']
        xp = _to_ntvc(x, dt)
        n, t, v, ci = xp.shape
        co = W.shape[0] // A.shape[0]
        d = L.BlockDesc(n=n, c_in=ci, c_out=co, t=t, v=v, k=A.shape[0], gamma=1, stride=1, pad=0,
                        dtype=L.dtype_code(dt), training=0, residual=0, flags=cfg.get('flags', 0), dropout_p=0.0,
                        bn_eps=1e-5, bn_momentum=0.1, seed=0, offset=0)
        out = torch.empty((n, t, v, co), dtype=dt, device=x.device)
        wsb = lib.stgcn_spatial_workspace_bytes(C.byref(d), 0)
        if wsb == 0:
            raise RuntimeError('stgcn_b200: ' + lib.stgcn_last_error().decode())
        ws = L.workspace(x.device, wsb)
        L.check(lib.stgcn_spatial_forward(C.byref(d), L.ptr(xp), L.ptr(_f32(A)), L.ptr(_f32(W)), L.ptr(_f32(Wb)),
                                          L.ptr(out), L.ptr(ws), ws.numel(), L.stream_ptr(x.device)), 'spatial_forward')
        ctx.d = d
        ctx.x_dtype = x.dtype
        ctx.save_for_backward(xp, A, W, Wb)
        return _from_ntvc(out)

    @staticmethod
    def backward(ctx, gout):
        lib = L.lib()
        xp, A, W, Wb = ctx.saved_tensors
        d = ctx.d
        gp = _to_ntvc(gout, xp.dtype)
        dx = torch.empty_like(xp) if ctx.needs_input_grad[0] else None
        dA, dW, dWb = torch.empty_like(A), torch.empty_like(W), torch.empty_like(Wb)
        wsb = lib.stgcn_spatial_workspace_bytes(C.byref(d), 1)
        ws = L.workspace(xp.device, wsb)
        L.check(lib.stgcn_spatial_backward(C.byref(d), L.ptr(xp), L.ptr(A), L.ptr(W), L.ptr(Wb), L.ptr(gp), L.ptr(dx),
                                           L.ptr(dA), L.ptr(dW), L.ptr(dWb), L.ptr(ws), ws.numel(),
                                           L.stream_ptr(xp.device)), 'spatial_backward')
        gx = _from_ntvc(dx).to(ctx.x_dtype) if dx is not None else None
        return gx, None, dA, dW, dWb


def spatial_forward(x, cfg, A, W, Wb):
    return _SpatialFn.apply(x, cfg, A, W, Wb)


class _HeadFn(torch.autograd.Function):
    """avg_pool2d over (T,V) + fc_layer (lightning_model.py:105-107). feat is logical (N,C,T,V)."""

    @staticmethod
    def forward(ctx, feat, fcw, fcb):
        L.require_cuda(feat, 'features')
        lib = L.lib()
        fp = feat.permute(0, 2, 3, 1)
        if not fp.is_contiguous():
            fp = fp.contiguous()
        n, t, v, c = fp.shape
        classes = fcw.shape[0]
        pooled = torch.empty((n, c), dtype=torch.float32, device=feat.device)
        logits = torch.empty((n, classes), dtype=torch.float32, device=feat.device)
        L.check(lib.stgcn_head_forward(L.dtype_code(fp.dtype), n, t, v, c, classes, L.ptr(fp), L.ptr(_f32(fcw)),
                                       L.ptr(_f32(fcb)), L.ptr(pooled), L.ptr(logits), L.stream_ptr(feat.device)),
                'head_forward')
        ctx.shape = (n, t, v, c, classes)
        ctx.dt = fp.dtype
        ctx.grad_bufs = (grad_buf_of(fcw), grad_buf_of(fcb))
        ctx.save_for_backward(pooled, fcw)
        return logits

    @staticmethod
    def backward(ctx, gl):
        lib = L.lib()
        pooled, fcw = ctx.saved_tensors
        n, t, v, c, classes = ctx.shape
        gl = gl.contiguous().float()
        dfeat = torch.empty((n, t, v, c), dtype=ctx.dt, device=gl.device)
        dW = _grad_out(fcw, ctx.grad_bufs[0])
        db = ctx.grad_bufs[1].view(classes) if (ctx.grad_bufs[1] is not None and ctx.grad_bufs[1].numel() == classes) \
            else torch.empty((classes,), dtype=torch.float32, device=gl.device)
        scratch = pooled.clone()      # head_backward reuses the pooled buffer as scratch
        L.check(lib.stgcn_head_backward(L.dtype_code(ctx.dt), n, t, v, c, classes, L.ptr(scratch), L.ptr(fcw), L.ptr(gl),
                                        L.ptr(dfeat), L.ptr(dW), L.ptr(db), L.stream_ptr(gl.device)), 'head_backward')
        return _from_ntvc(dfeat), dW, db


def head_forward(feat, fcw, fcb):
    return _HeadFn.apply(feat, fcw, fcb)


class _CrossEntropyFn(torch.autograd.Function):
    """F.cross_entropy(output, y), mean reduction (lightning_model.py:202): loss and d loss / d logits in one launch."""

    @staticmethod
    def forward(ctx, logits, labels, ignore_index):
        L.require_cuda(logits, 'logits')
        L.require_cuda(labels, 'labels')
        if logits.dim() != 2 or labels.dim() != 1 or labels.shape[0] != logits.shape[0]:
            raise RuntimeError('stgcn_b200.cross_entropy: logits (N,classes) and labels (N,) expected')
        if labels.dtype != torch.int64:
            raise RuntimeError('stgcn_b200.cross_entropy: labels must be int64 class indices')
        lib = L.lib()
        lg = _f32(logits)
        lb = labels.contiguous()
        n, classes = lg.shape
        loss = torch.empty((), dtype=torch.float32, device=lg.device)
        dl = torch.empty_like(lg)
        L.check(lib.stgcn_cross_entropy_forward(n, classes, L.ptr(lg), L.ptr(lb), int(ignore_index), L.ptr(loss), L.ptr(dl),
                                                L.stream_ptr(lg.device)), 'cross_entropy_forward')
        ctx.save_for_backward(dl)
        ctx.in_dtype = logits.dtype
        return loss

    @staticmethod
    def backward(ctx, gloss):
        (dl,) = ctx.saved_tensors
        lib = L.lib()
        g = gloss.contiguous().float()
        out = torch.empty_like(dl)
        L.check(lib.stgcn_cross_entropy_backward(dl.shape[0], dl.shape[1], L.ptr(dl), L.ptr(g), L.ptr(out),
                                                 L.stream_ptr(dl.device)), 'cross_entropy_backward')
        return out.to(ctx.in_dtype), None, None


def dropout_mask(seed, offset, p, count, device):
    """Keep-scales (1/(1-p) or 0) of `count` consecutive block-output elements under Philox state (seed, offset): what
    nn.Dropout(p) at st_graphconv.py:108 multiplies by in this library.  Test / oracle helper."""
    out = torch.empty(int(count), dtype=torch.float32, device=device)
    L.check(L.lib().stgcn_dropout_mask(int(seed), int(offset), float(p), int(count), L.ptr(out), L.stream_ptr(out.device)),
            'dropout_mask')
    return out


def cross_entropy(logits, labels, ignore_index=-100):
    return _CrossEntropyFn.apply(logits, labels, ignore_index)
