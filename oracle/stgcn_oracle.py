"""CPU oracle for the ST-GCN block stack  --  TEST INFRASTRUCTURE ONLY.

This file is a plain restatement (explicit tensor formulas, torch CPU ops, no
nn.Module, no conv2d / batch_norm library calls) of the algorithm the reference
implements in

    /root/reference/src/network/st_graphconv.py   (SpatialTemporalConv, SpatialConv)
    /root/reference/src/lightning_model.py:91-112 (L_STGCN.forward)
    /root/reference/src/network/stgcn.py:56-80    (STGCN.forward)

The reference's arithmetic lives in a third-party dependency that is not under
/root/reference: PyTorch ATen (`conv2d`, `batch_norm`, `einsum`, `relu`,
`dropout`, `avg_pool2d`, `linear`).  The reference pins no version; the oracle
restates the published semantics of those operators (torch 2.11.0 here) at the
reference's call sites, which are cited per function below.

Pinning: the reference ships no golden vectors, so parity is
pinned against outputs of the reference itself, imported unmodified in the
build container by `tests/golden/make_golden.py`; the resulting fixtures are
committed under `tests/golden/*.npz` and `tests/test_oracle_golden.py` checks
this oracle (forward, every gradient, running statistics) against them.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline /
`--impl reference` legs may import this module.  The product path
(`stgcn_b200/`) never does.

Conventions: activations are logical (N, C, T, V) like the reference; all maths
in the dtype of the inputs (float32 by default, float64 for tight checks).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import torch

BN_EPS = 1e-5          # torch.nn.BatchNorm2d default, st_graphconv.py:34,46
BN_MOMENTUM = 0.1      # torch.nn.BatchNorm2d default


# ----------------------------------------------------------------------------
# elementary operators (restated, differentiable through autograd)
# ----------------------------------------------------------------------------

def batch_norm(x, gamma, beta, running_mean, running_var, training, eps=BN_EPS,
               momentum=BN_MOMENTUM, update_running=True):
    """nn.BatchNorm2d as used at st_graphconv.py:73,77,98,100.

    training: normalise with the biased batch variance over (N,T,V); update
    running_mean/var in place with momentum 0.1 and the *unbiased* variance.
    eval: normalise with the running statistics.
    """
    if training:
        mean = x.mean(dim=(0, 2, 3))
        var = ((x - mean[None, :, None, None]) ** 2).mean(dim=(0, 2, 3))
        if update_running and running_mean is not None:
            n = x.numel() // x.shape[1]
            with torch.no_grad():
                running_mean.mul_(1 - momentum).add_(momentum * mean.detach())
                unbiased = var.detach() * (n / max(n - 1, 1))
                running_var.mul_(1 - momentum).add_(momentum * unbiased)
    else:
        mean, var = running_mean, running_var
    inv = torch.rsqrt(var + eps)
    return (x - mean[None, :, None, None]) * (inv * gamma)[None, :, None, None] \
        + beta[None, :, None, None]


def conv1x1(x, weight, bias, stride_t=1):
    """nn.Conv2d(C_in, C_o, 1, stride=(s,1)) -- st_graphconv.py:28,136.
    weight (C_o, C_in, 1, 1)."""
    if stride_t != 1:
        x = x[:, :, ::stride_t, :]
    w = weight.reshape(weight.shape[0], weight.shape[1])
    y = torch.einsum('oc,nctv->notv', w, x)
    if bias is not None:
        y = y + bias[None, :, None, None]
    return y


def temporal_conv(x, weight, bias, stride_t, pad_t):
    """nn.Conv2d(C, C, (gamma,1), stride=(s,1), padding=(p,0)) -- st_graphconv.py:41-43.
    weight (C_o, C_i, gamma, 1); zero padding of p frames on both ends of T."""
    n, c, t, v = x.shape
    gamma = weight.shape[2]
    t_out = (t + 2 * pad_t - gamma) // stride_t + 1
    zpad = torch.zeros(n, c, pad_t, v, dtype=x.dtype, device=x.device)
    xp = torch.cat([zpad, x, zpad], dim=2)
    y = None
    for tap in range(gamma):
        # frames  tap + s*j  of the padded input, j = 0..t_out-1
        sl = xp[:, :, tap: tap + stride_t * (t_out - 1) + 1: stride_t, :]
        term = torch.einsum('oc,nctv->notv', weight[:, :, tap, 0], sl)
        y = term if y is None else y + term
    if bias is not None:
        y = y + bias[None, :, None, None]
    return y


def spatial_conv(x, A, W, Wb):
    """SpatialConv.forward -- st_graphconv.py:139-152.

    Z = W(x) (1x1 conv to K*C_out channels, channel = k*C_out + c), viewed as
    (N,K,C_out,T,V); out[n,c,t,v] = sum_k sum_w A[k,v,w] * Z[n,k,c,t,w]
    (the einsum 'kvw,nkctw->nctv' at :150 contracts A's SECOND joint index).
    """
    k = A.shape[0]
    c_out = W.shape[0] // k
    z = conv1x1(x, W, Wb)
    n, _, t, v = z.shape
    z = z.reshape(n, k, c_out, t, v)
    return torch.einsum('kvw,nkctw->nctv', A, z)


def dropout(x, p, training, generator=None, mask=None):
    """nn.Dropout(p) -- st_graphconv.py:55,108.  `mask` (0/1 keep mask) may be
    supplied so that a different RNG (the CUDA Philox stream) can be replayed."""
    if not training or p == 0:
        return x
    if mask is None:
        mask = (torch.rand(x.shape, generator=generator) >= p).to(x.dtype)
    return x * mask / (1.0 - p)


# ----------------------------------------------------------------------------
# block / model
# ----------------------------------------------------------------------------

@dataclass
class BlockCfg:
    c_in: int
    c_out: int
    gamma: int = 9
    stride: int = 1
    pad: int = 4
    dropout_rate: float = 0.0
    residual: bool = False


def block_forward(x, p: Dict[str, torch.Tensor], cfg: BlockCfg, training: bool,
                  drop_mask=None, update_running=True):
    """SpatialTemporalConv.forward -- st_graphconv.py:85-109.

    p uses the reference's state_dict names: batch_n.{weight,bias,running_mean,
    running_var}, spatialConv.{A,W.weight,W.bias}, temporalConv.{weight,bias},
    batch_n_2.*, apply_residual.{weight,bias} (conv residual only).
    """
    x = x.to(p['spatialConv.A'].dtype)          # the hard .float() casts, :73,:98
    if cfg.residual:
        # residual_block, :60-82 (full pre-activation) + trailing ReLU :105
        h = batch_norm(x, p['batch_n.weight'], p['batch_n.bias'],
                       p.get('batch_n.running_mean'), p.get('batch_n.running_var'),
                       training, update_running=update_running)
        h = torch.relu(h)
        s = spatial_conv(h, p['spatialConv.A'], p['spatialConv.W.weight'], p['spatialConv.W.bias'])
        h2 = batch_norm(s, p['batch_n_2.weight'], p['batch_n_2.bias'],
                        p.get('batch_n_2.running_mean'), p.get('batch_n_2.running_var'),
                        training, update_running=update_running)
        h2 = torch.relu(h2)
        u = temporal_conv(h2, p['temporalConv.weight'], p['temporalConv.bias'], cfg.stride, cfg.pad)
        if 'apply_residual.weight' in p:
            r = conv1x1(x, p['apply_residual.weight'], p['apply_residual.bias'], cfg.stride)
        else:
            r = x
        out = u + r
    else:
        h = batch_norm(x, p['batch_n.weight'], p['batch_n.bias'],
                       p.get('batch_n.running_mean'), p.get('batch_n.running_var'),
                       training, update_running=update_running)
        s = spatial_conv(h, p['spatialConv.A'], p['spatialConv.W.weight'], p['spatialConv.W.bias'])
        u = temporal_conv(s, p['temporalConv.weight'], p['temporalConv.bias'], cfg.stride, cfg.pad)
        out = batch_norm(u, p['batch_n_2.weight'], p['batch_n_2.bias'],
                         p.get('batch_n_2.running_mean'), p.get('batch_n_2.running_var'),
                         training, update_running=update_running)
    y = torch.relu(out)
    return dropout(y, cfg.dropout_rate, training, mask=drop_mask)


# the layer table of lightning_model.py:65-86 / stgcn.py:40-51
LAYER_TABLE = [(None, 64, 1), (64, 64, 1), (64, 64, 1), (64, 64, 1), (64, 128, 2),
               (128, 128, 1), (128, 128, 1), (128, 256, 2), (256, 256, 1), (256, 256, 1)]


def stack_cfgs(c_in, gamma=9, dropout_rate=0.0, residual=False) -> List[BlockCfg]:
    pad = (gamma - 1) // 2
    return [BlockCfg(c_in if ci is None else ci, co, gamma, s, pad, dropout_rate, residual)
            for ci, co, s in LAYER_TABLE]


def split_state(state: Dict[str, torch.Tensor], i: int) -> Dict[str, torch.Tensor]:
    pre = f'conv.{i}.'
    return {k[len(pre):]: v for k, v in state.items() if k.startswith(pre)}


def model_forward(x, state: Dict[str, torch.Tensor], cfgs: List[BlockCfg], training: bool,
                  softmax=False, update_running=True):
    """L_STGCN.forward -- lightning_model.py:91-112 (softmax=False) and
    STGCN.forward -- stgcn.py:56-80 (softmax=True).  x is (N,T,V,C_in)."""
    h = x.permute(0, 3, 1, 2)
    for i, cfg in enumerate(cfgs):
        h = block_forward(h, split_state(state, i), cfg, training, update_running=update_running)
    feat = h.mean(dim=(2, 3))                                    # avg_pool2d over (T,V), :105
    logits = feat @ state['fc_layer.weight'].t() + state['fc_layer.bias']   # :107
    if softmax:
        logits = torch.softmax(logits, dim=1)                   # stgcn.py:77
    return logits


def cross_entropy(logits, labels):
    """F.cross_entropy, mean reduction -- lightning_model.py:202."""
    lse = torch.logsumexp(logits, dim=1)
    picked = logits.gather(1, labels[:, None])[:, 0]
    return (lse - picked).mean()


# ----------------------------------------------------------------------------
# input pipeline in front of block 0 (numpy, like the reference's own code)
# ----------------------------------------------------------------------------

def pad_with_loops(x, target_len):
    """pad_array_with_loops, src/data/util.py:12-30: np.pad(mode='wrap') on the time axis of (N,T,V,C) = frame t repeats
    frame t mod T."""
    import numpy as np
    t = x.shape[1]
    if t >= target_len:
        return x
    idx = np.arange(target_len) % t
    return x[:, idx]


def loopy_pad_collate(batch):
    """loopy_pad_collate_fn, src/data/util.py:33-47: batch = [(f_in (N_i,T_i,V,C), labels (N_i,)), ...] -> (sum N_i, T*, V, C), labels."""
    import numpy as np
    t_max = max(x.shape[1] for x, _ in batch)
    return (np.concatenate([pad_with_loops(x, t_max) for x, _ in batch]), np.concatenate([np.asarray(l) for _, l in batch]))


AUG_ROTATIONS = [15, -15, 5, -5, 10, -10]           # src/data/augmentation.py:15-17
AUG_TRANSLATIONS = [[5, 5], [0, 5], [5, 0]]
AUG_SCALES = [1.05, 1.1, 0.95]


def augmentation_matrix(rng):
    """The 3x3 matrix augment_data draws (src/data/augmentation.py:19-48), same sequence of `rng` calls: two of
    {rotation, translation, scaling, flip} with replacement, composed in that fixed order by left multiplication."""
    import numpy as np
    chosen = rng.choice(['rotation', 'translation', 'scaling', 'flip'], 2)
    T = np.eye(3)
    if 'rotation' in chosen:
        th = np.radians(rng.choice(AUG_ROTATIONS))
        c, s = np.cos(th), np.sin(th)
        T = np.array([[c, s, 0], [-s, c, 0], [0, 0, 1]]).dot(T)
    if 'translation' in chosen:
        tx, ty = AUG_TRANSLATIONS[rng.choice(range(3))]
        T = np.array([[1, 0, tx], [0, 1, ty], [0, 0, 1]]).dot(T)
    if 'scaling' in chosen:
        f = rng.choice(AUG_SCALES)
        T = np.array([[f, 0, 0], [0, f, 0], [0, 0, 1]]).dot(T)
    if 'flip' in chosen:
        T = np.array([[-1, 0, 0], [0, 1, 0], [0, 0, 1]]).dot(T)
    return T


def augment_apply(seqs, T):
    """_transform_sequence, src/data/augmentation.py:53-66: (x, y, 0) . T, first two components kept -- the homogeneous
    coordinate is 0, so only T[:2,:2] acts."""
    import numpy as np
    seqs = np.asarray(seqs)
    return seqs[..., :2] @ np.asarray(T)[:2, :2]


def cross_entropy_grad(logits, labels):
    """d cross_entropy / d logits for an incoming gradient of 1: (softmax - onehot) / N."""
    p = torch.softmax(logits, dim=1)
    oh = torch.zeros_like(p)
    oh[torch.arange(logits.shape[0]), labels] = 1.0
    return (p - oh) / logits.shape[0]


def adam_step(p, g, m, v, step, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
    """One step of torch.optim.Adam (no amsgrad) -- the optimizer of lightning_model.py:196-197 -- written out.
    Returns the new (p, m, v); `step` is the 1-based step number.  Pinned against torch.optim.Adam in
    tests/test_oracle_train.py."""
    b1, b2 = betas
    g = g + weight_decay * p
    m = m + (1.0 - b1) * (g - m)
    v = b2 * v + (1.0 - b2) * g * g
    bc1 = 1.0 - b1 ** step
    bc2 = 1.0 - b2 ** step
    denom = v.sqrt() / (bc2 ** 0.5) + eps
    return p - (lr / bc1) * (m / denom), m, v


# ----------------------------------------------------------------------------
# explicit backward of the default (non-residual) block -- restates what autograd
# does for the reference, in the factorisation the CUDA kernels use
# ----------------------------------------------------------------------------

def block_backward_explicit(x, p, cfg: BlockCfg, training: bool, dy):
    """Manual gradient of `block_forward` (residual=False, dropout 0), used to
    validate the derivation the CUDA backward chain follows.  Returns
    (dx, grads) with grads keyed like `p`."""
    assert not cfg.residual and cfg.dropout_rate == 0
    A, W, Wb = p['spatialConv.A'], p['spatialConv.W.weight'], p['spatialConv.W.bias']
    Wt, bt = p['temporalConv.weight'], p['temporalConv.bias']
    K, V = A.shape[0], A.shape[1]
    Co, Ci = cfg.c_out, cfg.c_in
    N, _, T, _ = x.shape
    g = {}

    def bn_fwd(z, name):
        if training:
            mu = z.mean(dim=(0, 2, 3)); var = z.var(dim=(0, 2, 3), unbiased=False)
        else:
            mu, var = p[name + '.running_mean'], p[name + '.running_var']
        inv = torch.rsqrt(var + BN_EPS)
        zh = (z - mu[None, :, None, None]) * inv[None, :, None, None]
        return zh, inv

    def bn_bwd(dout, zh, inv, name):
        gam = p[name + '.weight']
        g[name + '.weight'] = (dout * zh).sum(dim=(0, 2, 3))
        g[name + '.bias'] = dout.sum(dim=(0, 2, 3))
        if training:
            m = dout.numel() // dout.shape[1]
            dz = (gam * inv)[None, :, None, None] * (
                dout - (g[name + '.bias'] / m)[None, :, None, None]
                - zh * (g[name + '.weight'] / m)[None, :, None, None])
        else:
            dz = (gam * inv)[None, :, None, None] * dout
        return dz

    # forward recompute
    xh, inv1 = bn_fwd(x, 'batch_n')
    h = xh * p['batch_n.weight'][None, :, None, None] + p['batch_n.bias'][None, :, None, None]
    # aggregate-first factorisation: XA[n,k,ci,t,v] = sum_w A[k,v,w] h[n,ci,t,w]
    xa = torch.einsum('kvw,nctw->nkctv', A, h)
    Wk = W.reshape(K, Co, Ci)
    rowsum = A.sum(dim=2)                                              # (K,V)
    s = torch.einsum('koc,nkctv->notv', Wk, xa) + torch.einsum('ko,kv->ov', Wb.reshape(K, Co), rowsum)[None, :, None, :]
    u = temporal_conv(s, Wt, bt, cfg.stride, cfg.pad)
    uh, inv2 = bn_fwd(u, 'batch_n_2')
    yhat = uh * p['batch_n_2.weight'][None, :, None, None] + p['batch_n_2.bias'][None, :, None, None]

    # backward
    dyh = dy * (yhat > 0).to(dy.dtype)
    du = bn_bwd(dyh, uh, inv2, 'batch_n_2')
    g['temporalConv.bias'] = du.sum(dim=(0, 2, 3))
    T_out = u.shape[2]
    ds = torch.zeros_like(s)
    dWt = torch.zeros_like(Wt)
    sp = torch.cat([torch.zeros(N, Co, cfg.pad, V, dtype=s.dtype), s, torch.zeros(N, Co, cfg.pad, V, dtype=s.dtype)], dim=2)
    dsp = torch.zeros_like(sp)
    for tap in range(cfg.gamma):
        sl = slice(tap, tap + cfg.stride * (T_out - 1) + 1, cfg.stride)
        dWt[:, :, tap, 0] = torch.einsum('notv,nctv->oc', du, sp[:, :, sl, :])
        dsp[:, :, sl, :] += torch.einsum('oc,notv->nctv', Wt[:, :, tap, 0], du)
    ds = dsp[:, :, cfg.pad: cfg.pad + T, :]
    g['temporalConv.weight'] = dWt
    # spatial
    g['spatialConv.W.weight'] = torch.einsum('notv,nkctv->koc', ds, xa).reshape(K * Co, Ci, 1, 1)
    g['spatialConv.W.bias'] = torch.einsum('notv,kv->ko', ds, rowsum).reshape(K * Co)
    dxa = torch.einsum('koc,notv->nkctv', Wk, ds)
    dA = torch.einsum('nkctv,nctw->kvw', dxa, h)
    dA = dA + torch.einsum('notv,ko->kv', ds, Wb.reshape(K, Co))[:, :, None]
    g['spatialConv.A'] = dA
    dh = torch.einsum('kvw,nkctv->nctw', A, dxa)
    dx = bn_bwd(dh, xh, inv1, 'batch_n')
    return dx, g


# ----------------------------------------------------------------------------
# adjacency (data/adjacency.py) -- restated for input generation in tests
# ----------------------------------------------------------------------------

BODY25_ADJ = {0: [1, 15, 16], 1: [0, 2, 5, 8], 2: [1, 3], 3: [2, 4], 4: [3], 5: [1, 6], 6: [5, 7],
              7: [6], 8: [1, 9, 12], 9: [8, 10], 10: [9, 11], 11: [10, 22, 24], 12: [8, 13],
              13: [12, 14], 14: [13, 19, 21], 15: [0, 17], 16: [0, 18], 17: [15], 18: [16],
              19: [14, 20], 20: [19], 21: [14], 22: [11, 23], 23: [22], 24: [11]}   # util.py:156-180


def _neighbourhood(adj, root, d):
    """Limited graph search, adjacency.py:13-32: returns (all, rings) where
    rings[i] are the nodes first reached at hop i+1."""
    seen, frontier, rings = [root], [root], []
    for _ in range(d):
        new = []
        while frontier:
            cur = frontier.pop(0)
            fresh = [j for j in adj[cur] if j not in seen]
            seen.extend(fresh); new.extend(fresh)
        frontier = list(new); rings.append(list(new))
    return seen, rings


def adjacency_matrices(strategy: int, d: int = 1, distances=None, adj=BODY25_ADJ, opposite=None):
    """create_adjacency_matrices -- adjacency.py:34-158 (0 uni, 1 distance,
    2 spatial configuration, 3 symmetrical)."""
    V = len(adj)
    if strategy == 0:
        A = torch.zeros(V, V)
        for i in range(V):
            for j in _neighbourhood(adj, i, d)[0]:
                A[i, j] = 1
        return [A]
    if strategy == 1 or strategy == 3:
        mats = [torch.eye(V)] + [torch.zeros(V, V) for _ in range(d)]
        for i in range(V):
            _, rings = _neighbourhood(adj, i, d)
            for dist, ring in enumerate(rings):
                for j in ring:
                    mats[dist + 1][i, j] = 1
                    if strategy == 3 and j in opposite:
                        mats[dist + 1][i, opposite[j]] = 1
            if strategy == 3 and i in opposite:
                mats[d][i, opposite[i]] = 1        # adjacency.py:155-156 (uses last `dist`)
        return mats
    if strategy == 2:
        mats = [torch.zeros(V, V) for _ in range(3)]
        for i in range(V):
            for j in _neighbourhood(adj, i, d)[0]:
                if distances[j] == distances[i]:
                    lab = 0
                elif distances[j] < distances[i]:
                    lab = 1
                else:
                    lab = 2
                mats[lab][i, j] = 1
        return mats
    raise ValueError(strategy)


def normalize_adjacency(mats, expo=-0.5, alpha=0.001):
    """normalize -- adjacency.py:161-183.  NB: alpha is added to EVERY entry of
    diag(rowsum) before the element-wise power, so Lambda is dense."""
    out = torch.empty(len(mats), mats[0].shape[0], mats[0].shape[1])
    for i, A in enumerate(mats):
        lam = (torch.diag(A.sum(dim=1)) + alpha) ** expo
        out[i] = lam @ A @ lam
    return out


# ----------------------------------------------------------------------------
# bf16 storage emulation: the same maths with every tensor the CUDA chain stores
# in bf16 (activations, GEMM weight operands, activation gradients) rounded to
# bf16 at the same points, accumulation in fp32.  Needed because a ReLU network's
# gradients cannot be compared across precisions in the max norm: an activation
# that is +eps in fp32 and -eps after bf16 rounding flips its ReLU mask and
# changes that element's gradient by O(1) (measured: ~5-10 % of max|grad| between
# bf16 and fp32 runs of the same block, identical for CUDA-core and tensor-core
# GEMMs).  With matched rounding points the masks agree and the 2e-2 bound is
# meaningful.
# ----------------------------------------------------------------------------

def _rb(t):
    return t.to(torch.bfloat16).to(t.dtype)


def _q(t):
    """round to bf16 in forward (straight-through) and round the incoming gradient in backward"""
    out = t + (_rb(t) - t).detach()
    if out.requires_grad:
        out.register_hook(_rb)
    return out


def _bn_coeffs(x, gamma, beta, rm, rv, training, update_running, eps=BN_EPS, momentum=BN_MOMENTUM):
    if training:
        mean = x.mean(dim=(0, 2, 3))
        var = ((x - mean[None, :, None, None]) ** 2).mean(dim=(0, 2, 3))
        if update_running and rm is not None:
            n = x.numel() // x.shape[1]
            with torch.no_grad():
                rm.mul_(1 - momentum).add_(momentum * mean.detach())
                rv.mul_(1 - momentum).add_(momentum * var.detach() * (n / max(n - 1, 1)))
    else:
        mean, var = rm, rv
    a = gamma * torch.rsqrt(var + eps)
    return a, beta - a * mean


def block_forward_bf16(x, p, cfg: BlockCfg, training: bool, update_running=True, keep_u=True, round_A=None, drop_mask=None):
    """`block_forward` with bf16 storage emulation at the CUDA chain's tensor boundaries
    (x, XA, S, [P2], U, R, y and their gradients; GEMM weights rounded to bf16).
    keep_u=False models the fused inference epilogue where U is never stored."""
    A = p['spatialConv.A']
    K = A.shape[0]
    Co, Ci = cfg.c_out, cfg.c_in
    x = _q(x.float())
    a1, b1 = _bn_coeffs(x, p['batch_n.weight'], p['batch_n.bias'], p.get('batch_n.running_mean'),
                        p.get('batch_n.running_var'), training, update_running)
    h = x * a1[None, :, None, None] + b1[None, :, None, None]
    if cfg.residual:
        h = torch.relu(h)
    if round_A is None:
        round_A = cfg.c_in % 64 == 0             # the tensor-core aggregation (C_in % 64 == 0) takes A as a bf16 operand
    W3 = p['spatialConv.W.weight'].reshape(K, Co, Ci)
    Wb = p['spatialConv.W.bias'].reshape(K, Co)
    small_c = (Ci % 64 != 0) and (K * Ci <= 64) and (Co % 64 == 0) and not round_A
    if (round_A or small_c) and not cfg.residual:
        # tensor-core default block: BatchNorm-1 is folded into the 1x1 weights.  The RAW input is aggregated
        # (bf16 A operand, XA stored in bf16), GEMM-1 runs against bf16(W.diag(a1)), and the BN shift travels
        # through the weights into the per-joint bias:  sum_k rowsum(Aq)_k[v] * (W_k b1)[co].
        # small C_in (block 0): same algebra, but the aggregation runs on CUDA cores with the fp32 A.
        Aq = A + (_rb(A) - A).detach() if round_A else A   # forward rounding only: dA stays an fp32 accumulation
        xa = _q(torch.einsum('kvw,nctw->nkctv', Aq, x))
        Wk = _q(W3 * a1[None, None, :])
        wb = torch.einsum('koc,c->ko', W3, b1)
        bias2d = torch.einsum('ko,kv->ov', Wb, A.sum(dim=2)) + torch.einsum('ko,kv->ov', wb, Aq.sum(dim=2))
        s = _q(torch.einsum('koc,nkctv->notv', Wk, xa) + bias2d[None, :, None, :])
    else:
        if round_A:
            Aq = A + (_rb(A) - A).detach()
            h = _q(h)                            # a ReLU between BN-1 and the aggregation: h is stored in bf16
            xa = _q(torch.einsum('kvw,nctw->nkctv', Aq, h))
        else:
            xa = _q(torch.einsum('kvw,nctw->nkctv', A, h))                  # aggregate on the C_in side
        Wk = _q(W3)
        bias2d = torch.einsum('ko,kv->ov', Wb, A.sum(dim=2))
        s = _q(torch.einsum('koc,nkctv->notv', Wk, xa) + bias2d[None, :, None, :])
    Wt = _q(p['temporalConv.weight'])
    if not cfg.residual:
        u = temporal_conv(s, Wt, p['temporalConv.bias'], cfg.stride, cfg.pad)
        if keep_u:
            u = _q(u)
        a2, b2 = _bn_coeffs(u, p['batch_n_2.weight'], p['batch_n_2.bias'], p.get('batch_n_2.running_mean'),
                            p.get('batch_n_2.running_var'), training, update_running)
        return _q(dropout(torch.relu(u * a2[None, :, None, None] + b2[None, :, None, None]), cfg.dropout_rate, training,
                          mask=drop_mask))
    a2, b2 = _bn_coeffs(s, p['batch_n_2.weight'], p['batch_n_2.bias'], p.get('batch_n_2.running_mean'),
                        p.get('batch_n_2.running_var'), training, update_running)
    p2 = _q(torch.relu(s * a2[None, :, None, None] + b2[None, :, None, None]))
    u = temporal_conv(p2, Wt, p['temporalConv.bias'], cfg.stride, cfg.pad)
    if 'apply_residual.weight' in p:
        r = _q(conv1x1(x, _q(p['apply_residual.weight']), p['apply_residual.bias'], cfg.stride))
    else:
        r = x
    return _q(dropout(torch.relu(u + r), cfg.dropout_rate, training, mask=drop_mask))


def model_forward_bf16(x, state, cfgs, training, update_running=True, keep_u=True):
    h = x.permute(0, 3, 1, 2)
    for i, cfg in enumerate(cfgs):
        h = block_forward_bf16(h, split_state(state, i), cfg, training, update_running, keep_u)
    feat = h.mean(dim=(2, 3))
    return feat @ state['fc_layer.weight'].t() + state['fc_layer.bias']
