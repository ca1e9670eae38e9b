"""GPU parity, second set: the reference's features that round 1 left thinly covered.

  * nn.Dropout (st_graphconv.py:53-58,106-109) with the library's Philox mask replayed in the oracle: y, dx, every gradient;
    a captured training step must draw a fresh mask at every replay;
  * the STGCN twin (stgcn.py:56-80: softmax output, dropout 0.5, edge-importance masks);
  * L_STGCN(use_edge_importance=True) (lightning_model.py:53-59): jittered masks scale A at construction, get no gradient;
  * SpatialConv alone on the bf16 / tcgen05 path (st_graphconv.py:139-152);
  * residual blocks (st_graphconv.py:60-82) at full frame counts on the tcgen05 path;
  * a captured graph must survive growth of the library's scratch arena.
Tolerances: fp32 1e-4, bf16 2e-2 (max|got-ref| / max|ref| per tensor), structurally-zero gradients absolutely.
"""
import argparse
import contextlib
import io
import os

import numpy as np
import pytest
import torch

from _helpers import assert_argmax_consistent, bf16_grad_bound, check_dx_kink_tolerant, rel_err
from oracle import stgcn_oracle as O
from test_gpu_block import build_block, oracle_block, run_block

pytestmark = pytest.mark.gpu
TOL = {torch.float32: 1e-4, torch.bfloat16: 2e-2}


def _rand_bn(m):
    with torch.no_grad():
        for bn in (m.batch_n, m.batch_n_2):
            bn.weight.uniform_(0.5, 1.5); bn.bias.uniform_(-0.3, 0.3)
            bn.running_mean.uniform_(-0.2, 0.2); bn.running_var.uniform_(0.6, 1.4)


# ---------------------------------------------------------------------------------------------------------------------
# dropout with a replayed mask
# ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize('residual', [False, True], ids=['default', 'residual'])
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16], ids=['float32', 'bfloat16'])
def test_dropout_replayed_mask_matches_oracle(dtype, residual):
    """C % 64 == 0, so bf16 runs the tcgen05 chain: dropout in the element-wise pass (default block) and in the temporal GEMM's
    EP_DROPOUT epilogue (residual block).  The mask the library drew is rebuilt from its (seed, offset) through the C ABI
    (stgcn_dropout_mask) and handed to the oracle (oracle.dropout(mask=...))."""
    from stgcn_b200.functional import dropout_mask
    ci = co = 64; T, V, K, N, p_drop = 24, 25, 3, 3, 0.5
    torch.manual_seed(4321)
    A = torch.randn(K, V, V) * 0.2
    with contextlib.redirect_stdout(io.StringIO()):
        m = build_block(ci, co, A, 1, residual, dtype=dtype, dropout=p_drop)
    _rand_bn(m)
    p = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    x = torch.randn(N, ci, T, V); gy = torch.randn(N, co, T, V)
    m.train()
    m._ensure_rng(torch.device('cuda'))
    seed, off0 = [int(v) for v in m._rng_state.cpu()]
    y, dx, grads = run_block(m, x, gy, True)
    seed1, off1 = [int(v) for v in m._rng_state.cpu()]
    assert seed1 == seed and off1 == off0 + (N * T * V * co + 3) // 4           # advanced by one Philox call per 4 outputs
    scale = dropout_mask(seed, off0, p_drop, N * T * V * co, 'cuda').cpu().view(N, T, V, co).permute(0, 3, 1, 2)
    keep = (scale > 0).float()
    assert set(torch.unique(scale).tolist()) <= {0.0, 1.0 / (1.0 - p_drop)}
    assert 0.45 < float(keep.mean()) < 0.55
    cfg = O.BlockCfg(ci, co, 9, 1, 4, p_drop, residual)

    def oracle(bf16):
        q = {k: v.clone() for k, v in p.items()}
        leaves = {}
        for k, v in q.items():
            if v.dtype.is_floating_point and 'running' not in k:
                v.requires_grad_(True); leaves[k] = v
        xr = x.clone().requires_grad_(True)
        f = O.block_forward_bf16 if bf16 else O.block_forward
        yr = f(xr, q, cfg, True, drop_mask=keep)
        (yr * gy).sum().backward()
        return yr.detach(), xr.grad, {k: v.grad for k, v in leaves.items()}

    tol = TOL[dtype]
    y_ref, dx_ref, g_ref = oracle(False)
    g32, dx32 = g_ref, dx_ref
    assert rel_err(y, y_ref) < tol
    dropped_only = float(((y == 0) & (keep == 0)).float().sum() / (keep == 0).float().sum())
    assert dropped_only == 1.0
    if dtype == torch.bfloat16:
        y_ref, dx_ref, g_ref = oracle(True)
        assert rel_err(y, y_ref) < tol
    kinks = check_dx_kink_tolerant(dx, dx_ref, tol if dtype == torch.float32 else bf16_grad_bound(tol, dx_ref, dx32))
    flips = int((((y > 0) != (y_ref > 0)) & (keep > 0)).sum())
    assert flips <= max(2, int(1e-4 * y.numel())), flips
    for k, v in grads.items():
        if not residual and k == 'temporalConv.bias':
            assert float(v.abs().max()) <= 1e-5 * float(gy.abs().sum())
            continue
        bound = (tol if (kinks == 0 and flips == 0) else 5e-2) if dtype == torch.float32 else bf16_grad_bound(tol, g_ref[k], g32[k])
        assert rel_err(v.reshape(g_ref[k].shape), g_ref[k]) < bound, (k, bound)
    # a second call draws a different mask; eval mode drops nothing
    y2, _, _ = run_block(m, x, gy, True)
    assert float(((y2 == 0) != (y == 0)).float().mean()) > 0.2
    m.eval()
    with torch.no_grad():
        ye = m(x.cuda()).float().cpu()
    assert float((ye == 0).float().mean()) < 0.75


def test_captured_train_step_with_dropout_draws_fresh_masks():
    """STGCN's default dropout 0.5 (stgcn.py:41-50) inside CapturedTrainStep: every replay advances the device-resident
    generator, so two replays on the same batch give different losses, and the generator offsets advance by the same amount
    per step as in eager mode."""
    from stgcn_b200 import CapturedTrainStep, STGCN
    from stgcn_b200.optim import Adam
    torch.manual_seed(0)
    A = torch.randn(3, 25, 25) * 0.15
    model = STGCN(3, 9, 10, edge_importance=False, compute_dtype=torch.bfloat16, A=A).cuda().train()
    opt = Adam(model.parameters(), lr=0.0)                 # lr 0: the only thing that changes between replays is the mask
    x = torch.randn(4, 40, 25, 3).cuda(); y = torch.tensor([1, 2, 3, 4]).cuda()

    class NLL:
        def __call__(self, probs, labels):
            from stgcn_b200 import cross_entropy
            return cross_entropy(torch.log(probs.clamp_min(1e-12)), labels)
    step = CapturedTrainStep(model, opt, x, y, loss_fn=NLL())
    offs0 = [int(b._rng_state[1]) for b in model.conv]
    l1 = float(step(x, y)); offs1 = [int(b._rng_state[1]) for b in model.conv]
    l2 = float(step(x, y)); offs2 = [int(b._rng_state[1]) for b in model.conv]
    assert np.isfinite(l1) and np.isfinite(l2) and l1 != l2
    assert all(b - a == c - b and b > a for a, b, c in zip(offs0, offs1, offs2))
    step.close()
    with pytest.raises(RuntimeError):
        step(x, y)


# ---------------------------------------------------------------------------------------------------------------------
# model variants
# ---------------------------------------------------------------------------------------------------------------------
def _hp(**kw):
    d = dict(C_in=3, nr_classes=60, gamma=9, dropout_rate=0.0, residual=False, use_edge_importance=False,
             max_mask_jitter=0.001, partitioning=2, d=1, distance_file=None, distances=np.random.RandomState(0).rand(25))
    d.update(kw)
    return argparse.Namespace(**d)


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16], ids=['float32', 'bfloat16'])
def test_stgcn_twin_softmax_output_matches_oracle(dtype):
    """STGCN (stgcn.py:7-80): default dropout 0.5 is inactive in eval mode, output = softmax probabilities."""
    from stgcn_b200 import STGCN
    torch.manual_seed(3)
    A = torch.randn(2, 25, 25) * 0.15
    model = STGCN(3, 9, 60, edge_importance=True, compute_dtype=dtype, A=A).cuda()
    assert all(b.dropout is not None and b.dropout.p == 0.5 for b in model.conv)
    assert len(list(model.Masks)) == 10
    x = torch.randn(4, 40, 25, 3)
    model.train()
    with torch.no_grad():
        for _ in range(2):
            model(x.cuda())                       # calibrate the running statistics (dropout active here)
    model.eval()
    state = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    with torch.no_grad():
        got = model(x.cuda()).float().cpu()
    ref = O.model_forward(x, state, O.stack_cfgs(3, 9, 0.5, False), False, softmax=True)
    assert torch.allclose(got.sum(1), torch.ones(4), atol=1e-3)
    assert rel_err(got, ref) < TOL[dtype]
    assert_argmax_consistent(got, ref, exact=dtype == torch.float32)


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16], ids=['float32', 'bfloat16'])
def test_edge_importance_masks_scale_A_and_get_no_gradient(dtype):
    """use_edge_importance=True (lightning_model.py:53-59): each block's A is A * Masks[i] at construction, re-wrapped as a
    detached leaf (st_graphconv.py:129), so Masks never receive a gradient while spatialConv.A does."""
    from stgcn_b200 import L_STGCN
    torch.manual_seed(5)
    A = torch.randn(3, 25, 25) * 0.15
    model = L_STGCN(_hp(use_edge_importance=True, max_mask_jitter=0.2), compute_dtype=dtype, A=A)
    for i, b in enumerate(model.conv):
        assert torch.allclose(b.spatialConv.A.detach(), A * model.Masks[i].detach())
        assert not torch.allclose(model.Masks[i].detach(), torch.ones_like(A))
    state = {k: v.detach().clone() for k, v in model.state_dict().items()}
    model = model.cuda().train()
    x = torch.randn(4, 40, 25, 3); labels = torch.tensor([1, 5, 30, 59])
    logits = model(x.cuda())
    torch.nn.functional.cross_entropy(logits, labels.cuda()).backward()
    assert all(mk.grad is None for mk in model.Masks)
    assert all(b.spatialConv.A.grad is not None and float(b.spatialConv.A.grad.abs().max()) > 0 for b in model.conv)
    ref = O.model_forward(x, state, O.stack_cfgs(3, 9, 0.0, False), True)
    assert rel_err(logits, ref) < TOL[dtype]


def test_learnable_edge_importance_gives_the_masks_their_gradient():
    """edge_importance_mode='learnable' (beyond the reference): every block uses A_base * Masks[i]; Masks[i].grad
    must equal the oracle's gradient of the same function, spatialConv.A stays frozen.  fp32, 1e-4."""
    from stgcn_b200 import L_STGCN
    torch.manual_seed(6)
    A = torch.randn(3, 25, 25) * 0.15
    model = L_STGCN(_hp(use_edge_importance=True, max_mask_jitter=0.2, edge_importance_mode='learnable'), A=A)
    assert all(torch.equal(b.spatialConv.A.detach(), A) and not b.spatialConv.A.requires_grad for b in model.conv)
    state = {k: v.detach().clone() for k, v in model.state_dict().items()}
    masks = [state['Masks.%d' % i].clone().requires_grad_(True) for i in range(10)]
    for i in range(10):
        state['conv.%d.spatialConv.A' % i] = A * masks[i]
    x = torch.randn(4, 24, 25, 3); labels = torch.tensor([1, 5, 30, 59])
    ref = O.model_forward(x, state, O.stack_cfgs(3, 9, 0.0, False), True)
    O.cross_entropy(ref, labels).backward()
    model = model.cuda().train()
    logits = model(x.cuda())
    eff = [b._A_override for b in model.conv]            # the adjacency each block used in this forward
    for e in eff:
        e.retain_grad()
    torch.nn.functional.cross_entropy(logits, labels.cuda()).backward()
    assert rel_err(logits, ref) < 1e-4
    for i, mk in enumerate(model.Masks):
        assert mk.grad is not None and model.conv[i].spatialConv.A.grad is None
        assert torch.allclose(eff[i].detach().cpu(), A * state['Masks.%d' % i], rtol=1e-6, atol=1e-7)
        # the plumbing, exactly: d/dMask = A_base (.) d/dA_eff, with d/dA_eff the gradient the library computed for this block
        assert rel_err(mk.grad, A.cuda() * eff[i].grad) < 1e-6, i
        # and against the oracle: gradients of a 10-block network on a 4-sample batch are reproducible to ~5e-3 in fp32 (ReLU
        # kinks; measured 4.5e-3 here, tests/test_gpu_model.py allows 0.12 for the golden models for the same reason)
        assert rel_err(mk.grad, masks[i].grad) < 2e-2, i


def test_spatial_conv_module_bf16_tcgen05_matches_oracle():
    """SpatialConv alone with C_in % 64 == 0 in bf16: tensor-core aggregation + tcgen05 1x1 GEMM, and their gradients."""
    from stgcn_b200 import SpatialConv
    torch.manual_seed(12)
    ci, co, T, V, K, N = 64, 128, 20, 25, 3, 3
    A = torch.randn(K, V, V) * 0.3
    m = SpatialConv(ci, co, A, compute_dtype=torch.bfloat16).cuda()
    x = torch.randn(N, ci, T, V)
    xg = x.cuda().requires_grad_(True)
    out = m(xg)
    go = torch.randn(N, co, T, V)
    (out.float() * go.cuda()).sum().backward()
    rb = lambda t: t.to(torch.bfloat16).float()
    # oracle on the operands as the bf16 chain sees them (x, A, W rounded to bf16), fp32 arithmetic
    Ar = A.clone().requires_grad_(True); W = m.W.weight.detach().cpu().clone().requires_grad_(True)
    b = m.W.bias.detach().cpu().clone().requires_grad_(True); xr = rb(x).requires_grad_(True)
    Aq = Ar + (rb(A) - A); Wq = W + (rb(W.detach()) - W.detach())
    ref = O.spatial_conv(xr, Aq, Wq, b)
    (ref * rb(go)).sum().backward()
    tol = TOL[torch.bfloat16]
    assert rel_err(out, ref) < tol
    assert rel_err(xg.grad, xr.grad) < tol
    assert rel_err(m.A.grad, Ar.grad) < tol
    assert rel_err(m.W.weight.grad, W.grad) < tol
    assert rel_err(m.W.bias.grad, b.grad) < tol
    # and against the plain fp32 oracle on the unrounded operands
    ref32 = O.spatial_conv(x, A, m.W.weight.detach().cpu(), m.W.bias.detach().cpu())
    assert rel_err(out, ref32) < tol


RES_FULL = [
    # ci, co, T, V, K, stride   (the reference's best configuration is the residual one, report section 5.7)
    (64, 64, 300, 25, 3, 1),
    (64, 128, 300, 25, 3, 2),
]


@pytest.mark.parametrize('case', RES_FULL, ids=lambda c: '%d-%d-T%d-V%d-K%d-s%d' % c)
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16], ids=['float32', 'bfloat16'])
def test_residual_block_full_frames_matches_oracle(case, dtype):
    ci, co, T, V, K, s = case
    torch.manual_seed(321 + ci + co)
    A = torch.randn(K, V, V) * 0.2
    with contextlib.redirect_stdout(io.StringIO()):
        m = build_block(ci, co, A, s, True, dtype=dtype)
    with torch.no_grad():
        for bn in (m.batch_n, m.batch_n_2):
            bn.weight.uniform_(0.5, 1.5); bn.bias.uniform_(-0.3, 0.3)
    p = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    N = 2
    x = torch.randn(N, ci, T, V)
    T_out = (T - 1) // s + 1
    gy = torch.randn(N, co, T_out, V)
    cfg = O.BlockCfg(ci, co, 9, s, 4, 0.0, True)
    tol = TOL[dtype]
    y_ref, dx_ref, g_ref, _ = oracle_block(x, gy, p, cfg, True)
    g32, dx32 = g_ref, dx_ref
    y, dx, grads = run_block(m, x, gy, True)
    assert rel_err(y, y_ref) < tol
    if dtype == torch.bfloat16:
        y_ref, dx_ref, g_ref, _ = oracle_block(x, gy, p, cfg, True, bf16=True)
        assert rel_err(y, y_ref) < tol
    kinks = check_dx_kink_tolerant(dx, dx_ref, tol if dtype == torch.float32 else bf16_grad_bound(tol, dx_ref, dx32), max_frames=20)
    flips = int(((y > 0) != (y_ref > 0)).sum())
    assert flips <= max(2, int(1e-4 * y.numel())), flips
    for k, v in grads.items():
        bound = (tol if (kinks == 0 and flips == 0) else 2e-2) if dtype == torch.float32 else bf16_grad_bound(tol, g_ref[k], g32[k])
        assert rel_err(v.reshape(g_ref[k].shape), g_ref[k]) < bound, (k, bound)


# ---------------------------------------------------------------------------------------------------------------------
# scratch-arena growth under a live graph
# ---------------------------------------------------------------------------------------------------------------------
def test_captured_step_survives_workspace_growth():
    """Capture a training step at a small shape, then run a LARGER eager forward (the arena grows), then replay: the replayed
    step must equal an eager step on a twin model (the outgrown arena stays alive for the graph)."""
    from stgcn_b200 import CapturedTrainStep, L_STGCN, _lib, cross_entropy
    from stgcn_b200.optim import Adam
    torch.manual_seed(0)
    A = torch.randn(3, 25, 25) * 0.15
    m1 = L_STGCN(_hp(), compute_dtype=torch.bfloat16, A=A).cuda().train()
    m2 = L_STGCN(_hp(), compute_dtype=torch.bfloat16, A=A).cuda().train()
    m2.load_state_dict(m1.state_dict())
    o1, o2 = Adam(m1.parameters(), lr=1e-4), Adam(m2.parameters(), lr=1e-4)    # Adam's first step is lr * sign(g)
    x = torch.randn(4, 32, 25, 3).cuda(); y = torch.tensor([1, 2, 3, 4]).cuda()
    step = CapturedTrainStep(m1, o1, x, y)
    before = _lib.current_workspace(x.device)
    junk = torch.full((256 << 20,), 7, dtype=torch.uint8, device='cuda')      # whatever the allocator hands out next
    # a later, larger call outgrows the arena (requested directly: earlier tests of this process may already have grown the
    # process-wide arena beyond anything a model call here would need)
    grown = _lib.workspace(x.device, before.numel() + (64 << 20))
    assert grown is not before and _lib.current_workspace(x.device) is grown
    grown.fill_(0x5a)
    with torch.no_grad():
        m2.eval(); m2(torch.randn(8, 64, 25, 3).cuda()); m2.train()              # and is used by other calls
    l1 = float(step(x, y))
    o2.zero_grad(set_to_none=True)
    loss2 = cross_entropy(m2(x), y); loss2.backward(); o2.step()
    assert abs(l1 - float(loss2)) < 2e-2 * max(1.0, abs(float(loss2)))
    for (k, a), (_, b) in zip(m1.named_parameters(), m2.named_parameters()):
        # Adam's first step moves every weight by lr * sign(g): two correct runs differ by at most 2 lr (a near-zero gradient of
        # the other sign); a replay that wrote its scratch into foreign memory shows up as garbage here and in the canary below
        assert float((a - b).abs().max()) <= 2.5e-4, k
    assert int(junk[:1024].sum()) == 7 * 1024 and int(junk[-1024:].sum()) == 7 * 1024
    step.close()


def test_captured_step_refuses_when_an_older_autograd_graph_is_alive():
    """A `loss` kept from an eager step pins the parameters' AccumulateGrad nodes to the default stream; capture of backward
    would fail with cudaErrorStreamCaptureImplicit and leave torch's capture state poisoned.  The constructor must notice during
    its warm-up and raise BEFORE capturing; after `del loss` construction works."""
    from stgcn_b200 import CapturedTrainStep, L_STGCN, cross_entropy
    from stgcn_b200.optim import Adam
    torch.manual_seed(0)
    m = L_STGCN(_hp(), compute_dtype=torch.bfloat16, A=torch.randn(3, 25, 25) * 0.15).cuda().train()
    opt = Adam(m.parameters(), lr=1e-4)
    x = torch.randn(4, 32, 25, 3).cuda(); y = torch.tensor([1, 2, 3, 4]).cuda()
    loss = cross_entropy(m(x), y)
    loss.backward()
    for _ in range(2):            # detected every time, not only the first time in a process (torch warns once by default)
        with pytest.raises(RuntimeError, match='autograd graph of an earlier step'):
            CapturedTrainStep(m, opt, x, y)
    del loss
    step = CapturedTrainStep(m, opt, x, y)
    assert np.isfinite(float(step(x, y)))
    step.close()
    torch.empty(4, device='cuda').uniform_()          # torch's CUDA generator is not stuck in capture mode


def test_batched_training_operands_give_the_same_step(monkeypatch):
    """stgcn_train_prep_multi (opt-in, STGCN_TRAIN_PREP=1): the weight-only operands of all ten blocks from one launch must give
    the same logits and gradients as every block deriving them inside its own call, also after the weights changed."""
    from stgcn_b200 import L_STGCN, cross_entropy
    monkeypatch.setenv('STGCN_NO_PINGPONG', '1')       # same tile order in both runs (bf16 run-to-run noise, tools/gpu_determinism.py)
    torch.manual_seed(0)
    A = torch.randn(3, 25, 25) * 0.15
    model = L_STGCN(_hp(), compute_dtype=torch.bfloat16, A=A).cuda().train()
    state = {k: v.detach().clone() for k, v in model.state_dict().items()}
    x = torch.randn(4, 40, 25, 3).cuda(); y = torch.tensor([1, 5, 30, 59]).cuda()

    def run(flag):
        monkeypatch.setenv('STGCN_TRAIN_PREP', flag)
        model.load_state_dict(state)
        outs = []
        for it in range(2):                              # the second pass runs on modified weights: operands must follow
            model.zero_grad(set_to_none=True)
            lg = model(x)
            cross_entropy(lg, y).backward()
            outs.append((lg.detach().float().cpu(), torch.cat([p.grad.flatten().float().cpu() for p in model.parameters()])))
            del lg
            with torch.no_grad():
                for b in model.conv:
                    b.temporalConv.weight.mul_(1.25); b.spatialConv.A.mul_(0.8)
        return outs
    off, on = run('0'), run('1')
    assert all(b._train_prep is not None for b in model.conv)            # the batched path really ran
    for (l0, g0), (l1, g1) in zip(off, on):
        assert rel_err(l1, l0) < 2e-2
        assert float((g1 - g0).abs().max() / g0.abs().max()) < 2e-2
    assert rel_err(off[1][0], off[0][0]) > 1e-3                          # the weight change was visible


@pytest.mark.parametrize('skeleton,V', [('body25', 25), ('openpose18', 18)])
@pytest.mark.parametrize('dense_alpha', [True, False], ids=['reference-dense-alpha', 'diagonal-alpha'])
def test_device_adjacency_normalisation_matches_host(skeleton, V, dense_alpha):
    """stgcn_adjacency_normalize against graph.normalize (which reproduces the reference's adjacency.py bit for bit on the golden
    fixture): every strategy, d = 1 and 2, both skeletons."""
    from stgcn_b200 import graph
    dist = np.random.RandomState(0).rand(V)
    for strat in range(4):
        for d in (1, 2):
            host = graph.get_normalized_adjacency_matrices(strat, d, distances=dist, skeleton=skeleton, dense_alpha=dense_alpha)
            dev = graph.get_normalized_adjacency_matrices(strat, d, distances=dist, skeleton=skeleton, dense_alpha=dense_alpha,
                                                          device=torch.device('cuda'))
            assert dev.is_cuda and dev.shape == host.shape
            assert rel_err(dev, host) < 1e-5, (strat, d)


def test_cross_entropy_out_of_range_label_poisons_loss_without_oob():
    import stgcn_b200
    logits = torch.randn(6, 10).cuda().requires_grad_(True)
    labels = torch.tensor([1, 2, 10, 3, -1, 4]).cuda()
    loss = stgcn_b200.cross_entropy(logits, labels)
    loss.backward()
    assert torch.isnan(loss)
    assert torch.isfinite(logits.grad).all()
    assert float(logits.grad[2].abs().max()) == 0.0 and float(logits.grad[4].abs().max()) == 0.0


def test_batchnorm_momentum_none_is_rejected():
    A = torch.randn(3, 25, 25) * 0.2
    m = build_block(8, 8, A, 1, False)
    m.batch_n.momentum = None
    with pytest.raises(RuntimeError):
        m(torch.randn(2, 8, 12, 25).cuda())


# ---- SyncBN plumbing on one GPU: a world of two IDENTICAL ranks ----------------------------------------------------------------
SYNC_CASES = [
    # ci, co, T, stride, residual, dtype
    (16, 24, 20, 1, False, torch.float32),       # CUDA-core / split path, BatchNorm-1 backward as its own pass
    (64, 64, 20, 1, False, torch.bfloat16),      # tcgen05 path: BatchNorm-1 backward folded into the spatial weights
    (64, 128, 20, 2, False, torch.bfloat16),
    (64, 64, 14, 1, True, torch.bfloat16),       # residual block: BatchNorm-2 sits between the convolutions
    (16, 32, 20, 2, True, torch.float32),
]


@pytest.mark.parametrize('case', SYNC_CASES, ids=lambda c: '%d-%d-T%d-s%d-%s-%s' % (c[0], c[1], c[2], c[3], 'res' if c[4] else 'def', str(c[5])[6:]))
def test_sync_batchnorm_with_two_identical_ranks_equals_local_statistics(case):
    """StgcnBlockParams.sync_* (include/stgcn_b200.h): with an exchange step that doubles the sums (= SUM over two ranks holding
    the same shard) and world = 2, every mean / variance / backward correction is that of the local batch, so outputs, dx and
    all gradients must equal the local-statistics run; only the running variance differs, by the unbiased-count factor
    (2m/(2m-1)) / (m/(m-1)).  All exchange points must have fired: 2 forward + 2 backward per block (+1 when the folded
    BatchNorm-1 backward is its own exchange)."""
    from stgcn_b200.parallel import SyncBatchNormLink
    ci, co, T, s, res, dtype = case
    V, K, N = 25, 3, 4
    torch.manual_seed(99)
    A = torch.randn(K, V, V) * 0.2
    with contextlib.redirect_stdout(io.StringIO()):
        m = build_block(ci, co, A, s, res, dtype=dtype)
    _rand_bn(m)
    state = {k: v.detach().clone() for k, v in m.state_dict().items()}
    x = torch.randn(N, ci, T, V)
    gy = torch.randn(N, co, (T - 1) // s + 1, V)
    y0, dx0, g0 = run_block(m, x, gy, True)
    sd0 = {k: v.detach().clone() for k, v in m.state_dict().items()}
    m.load_state_dict(state)
    m.zero_grad(set_to_none=True)
    link = SyncBatchNormLink(world=2, allreduce=lambda buf: buf.mul_(2.0))
    m._sync = link
    y1, dx1, g1 = run_block(m, x, gy, True)
    sd1 = m.state_dict()
    assert link.calls in (4, 5), link.calls
    tol = 1e-5 if dtype == torch.float32 else 2e-2            # fp32: same kernels, sums doubled and halved; bf16: rounding of y re-drawn
    assert rel_err(y1, y0) < tol
    assert rel_err(dx1, dx0) < (1e-4 if dtype == torch.float32 else 2e-2)
    for k in g0:
        assert rel_err(g1[k], g0[k]) < (1e-4 if dtype == torch.float32 else 2e-2), k
    rows = {'batch_n': N * T * V, 'batch_n_2': N * (T if res else (T - 1) // s + 1) * V}
    for bn, mrows in rows.items():
        assert rel_err(sd1[bn + '.running_mean'], sd0[bn + '.running_mean']) < 1e-5
        # running_var = 0.9 * start + 0.1 * unbiased batch variance: undo the blend, compare the batch parts
        b0 = (sd0[bn + '.running_var'] - 0.9 * state[bn + '.running_var']) / 0.1
        b1 = (sd1[bn + '.running_var'] - 0.9 * state[bn + '.running_var']) / 0.1
        factor = (2 * mrows / (2 * mrows - 1)) / (mrows / (mrows - 1))
        assert rel_err(b1, b0 * factor) < 1e-4, bn
        assert int(sd1[bn + '.num_batches_tracked']) == int(sd0[bn + '.num_batches_tracked'])
    # eval mode never exchanges
    m.eval(); n_before = link.calls
    with torch.no_grad():
        m(x.cuda())
    assert link.calls == n_before


def test_fused_spatial_kernel_with_aggregate_store_in_training():
    """STGCN_FUSED_SPATIAL=1 routes the training forward through the tensor-memory spatial kernel, which then also stores the
    aggregated rows for backward's weight gradient straight from its conversion registers (umma_spatial.cu).  The switch is read
    once per process, so the block parity cases run in a child process with it set."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, STGCN_FUSED_SPATIAL='1')
    r = subprocess.run([sys.executable, '-m', 'pytest', os.path.join(root, 'tests', 'test_gpu_block.py'), '-q', '-x', '-m', 'gpu',
                        '-k', 'test_block_matches_oracle and bfloat16'], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-1000:]
    assert ' passed' in r.stdout and 'failed' not in r.stdout


# ---- tensor-memory spatial kernel (inference default): shapes beyond the model's own ----------------------------------------------
TS_SHAPES = [
    # ci, co, V, K, T, stride   (T chosen so that the last tile of frames is ragged; every tensor-memory plan is visited:
    #                            adjacency in tensor memory + resident weights, weight ring, one / two output accumulators, 64- and
    #                            128-wide weight stages, 1..3 partitions, 4..7 frames per tile)
    (64, 64, 25, 3, 23, 1),
    (64, 192, 25, 3, 13, 1),
    (128, 64, 18, 2, 20, 1),
    (192, 256, 20, 1, 9, 1),
    (256, 128, 32, 3, 11, 2),
    (64, 128, 16, 3, 30, 2),
    (256, 256, 25, 3, 7, 1),
]


@pytest.mark.parametrize('shape', TS_SHAPES, ids=lambda s: '%d-%d-V%d-K%d-T%d-s%d' % s)
def test_inference_block_on_the_tensor_memory_spatial_kernel_matches_oracle(shape):
    """Eval-mode, no-grad bf16 block (the path CapturedForward replays): spatial_fused_ts_kernel (aggregation and channel mixing with
    the intermediate in tensor memory, umma_spatial.cu) followed by the temporal GEMM with BatchNorm-2 + ReLU in its epilogue,
    against the fp32 oracle block (st_graphconv.py:85-109) at the bf16 tolerance; a second call must give the same bits."""
    ci, co, V, K, T, s = shape
    torch.manual_seed(ci + co + V + K)
    A = torch.randn(K, V, V) * 0.2
    with contextlib.redirect_stdout(io.StringIO()):
        m = build_block(ci, co, A, s, False, dtype=torch.bfloat16).eval()
    _rand_bn(m)
    p = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    x = torch.randn(3, ci, T, V)
    from stgcn_b200 import _lib
    with torch.no_grad():
        y = m(x.cuda())
        _lib.timing_reset(); _lib.timing_enable(True)
        y2 = m(x.cuda())
        torch.cuda.synchronize(); _lib.timing_enable(False)
        y_ref = O.block_forward(x, p, O.BlockCfg(ci, co, 9, s, 4, 0.0, False), False)
    assert _lib.timing_read().get('umma_spatial', (0, 0, 0))[2] == 1, 'the fused spatial kernel did not run for this shape'
    assert rel_err(y.float(), y_ref) < TOL[torch.bfloat16]
    assert float(y_ref.abs().max()) > 0.1
    assert rel_err(y2.float(), y.float()) < 1e-2          # (bitwise only with the traversal direction fixed)


def test_sync_batchnorm_link_without_a_buffer_is_refused():
    """A half-filled SyncBN link (callback but no exchange buffer) must fail loudly, not fall back to per-replica statistics."""
    from stgcn_b200.parallel import SyncBatchNormLink
    torch.manual_seed(0)
    with contextlib.redirect_stdout(io.StringIO()):
        m = build_block(16, 16, torch.randn(2, 25, 25) * 0.2, 1, False).train()
    link = SyncBatchNormLink(world=2, allreduce=lambda buf: None)
    link.buf = torch.empty(0, device='cuda')            # data_ptr() == 0
    link.on = lambda device, channels: link
    m._sync = link
    with pytest.raises(RuntimeError, match='sync_buf is null'):
        m(torch.randn(2, 16, 12, 25).cuda())
    assert link.calls == 0
