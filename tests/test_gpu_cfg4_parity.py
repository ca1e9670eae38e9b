"""Model-level parity of the path bench.py times: bf16, the full 10-block stack, one TRAINING step at BASELINE cfg4's
per-GPU shape (32 sequences x 2 persons = N_eff 64, T = 300, V = 25, K = 3, 60 classes), with a sane-magnitude adjacency and
with the reference's own normalised adjacency (entries ~6e4, adjacency.py:161-183), and at cfg5's per-GPU shape (128 sequences =
N_eff 256, V = 18, 400 classes, the OpenPose-18 graph bench.py uses); eager and through CapturedTrainStep.

Train-mode BatchNorm couples every row of the batch, so the oracle has to see the whole batch.  It is the same CPU
restatement (oracle/stgcn_oracle.py: plain torch tensor formulas, device-agnostic), evaluated here on the GPU box's device
in fp32 with TF32 off; nothing of the product is involved in it.

What is asserted, and why the model-level gradient bound is not "every tensor within 2e-2 in the max norm":
  * logits within 2e-2 of the fp32 oracle, loss within 2e-2, argmax identical on every row whose reference margin exceeds
    the measured error (measured on B200: logits 2.6e-3 / 6e-3, loss 5e-6, 64 of 64 rows decisive and identical);
  * gradients of a 10-block bf16 ReLU network on a random-label batch have an irreducible noise floor that NO bf16
    implementation gets under: ~1e-3 of all pre-activations lie within one bf16 ulp of zero and take the other ReLU branch, each
    moving its gradient by O(1), and the BatchNorm / bias gradients are sums of ~5e5 such terms with heavy cancellation.  The
    test MEASURES that floor with the checker alone -- oracle with bf16 storage emulation (oracle.block_forward_bf16, pinned to
    the fp32 oracle with rounding off by tests/test_oracle_golden.py) against the fp32 oracle: 0.2 .. 0.9 per tensor in the max
    norm on B200, where single blocks give ~5e-2 -- and requires the CUDA path to be AS CLOSE TO FP32 AS THE EMULATION IS:
    per tensor  ||got - fp32||_2 / ||fp32||_2  <=  3 x the emulation's same ratio + 2e-2, and over all gradient elements
    together  <= 1.5 x the emulation's + 2e-2, with a cosine to the fp32 gradient no worse than the emulation's - 0.02;
  * the strict 2e-2 per-tensor max-norm bound on gradients is asserted where it is meaningful: per block, at the full frame
    count, against the emulation (tests/test_gpu_fullsize.py, tests/test_gpu_round2.py, tests/test_gpu_block.py);
  * tensors whose gradient is structurally zero (a bias in front of a batch-statistics BatchNorm) are checked absolutely.
The measured numbers are written to gpurun_out/cfg4_parity_<adjacency>.json for DESIGN.md.
"""
import argparse
import json
import os

import numpy as np
import pytest
import torch

from _helpers import assert_argmax_consistent, rel_err
from oracle import stgcn_oracle as O

pytestmark = pytest.mark.gpu
M, T, K = 2, 300, 3
TOL = 2e-2
# (sequences per GPU, V, classes): BASELINE cfg4 (NTU shape, 32 per GPU) and cfg5 (Kinetics-skeleton shape, V = 18, 400 classes,
# 128 per GPU = N 1024 at 8 GPUs, with the OpenPose-18 graph bench.py uses)
WORKLOADS = {'cfg4': (32, 25, 60), 'cfg5': (128, 18, 400)}


def _hp(V, CLASSES):
    return argparse.Namespace(C_in=3, nr_classes=CLASSES, gamma=9, dropout_rate=0.0, residual=False, use_edge_importance=False,
                              max_mask_jitter=0.001, partitioning=2, d=1, distance_file=None,
                              distances=np.random.RandomState(0).rand(V))


def _adjacency(kind, V):
    from stgcn_b200 import graph
    if kind == 'sane':
        return torch.randn(K, V, V, generator=torch.Generator().manual_seed(0)) * 0.2
    if kind == 'openpose18':
        A = graph.get_normalized_adjacency_matrices(graph.Strategy.SPATIAL_CONFIGURATION, 1, distances=np.random.RandomState(0).rand(V),
                                                    skeleton='openpose18', dense_alpha=False)
        return A / A.abs().max()
    A = graph.get_normalized_adjacency_matrices(2, 1, distances=np.random.RandomState(0).rand(V))     # SPATIAL, d=1
    assert A.shape == (K, V, V) and float(A.abs().max()) > 1e4            # the reference's dense-alpha normalisation
    return A


def _oracle_step(x, labels, state, bf16):
    """loss, logits and every gradient of one training-mode forward+backward of the oracle on x's device (fp32, TF32 off)."""
    st = {k: v.detach().clone() for k, v in state.items()}
    leaves = {}
    for k, v in st.items():
        if v.dtype.is_floating_point and 'running' not in k and not k.startswith('Masks'):
            v.requires_grad_(True); leaves[k] = v
    cfgs = O.stack_cfgs(3, 9, 0.0, False)
    f = O.model_forward_bf16 if bf16 else O.model_forward
    logits = f(x, st, cfgs, True)
    loss = O.cross_entropy(logits, labels)
    loss.backward()
    return float(loss), logits.detach().cpu(), {k: v.grad.detach().cpu() for k, v in leaves.items() if v.grad is not None}


@pytest.mark.parametrize('workload,adjacency', [('cfg4', 'sane'), ('cfg4', 'reference'), ('cfg5', 'openpose18')])
def test_bf16_training_step_matches_oracle_at_full_shape(workload, adjacency):
    from stgcn_b200 import CapturedTrainStep, L_STGCN, cross_entropy
    from stgcn_b200.optim import Adam
    N_SEQ, V, CLASSES = WORKLOADS[workload]
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(0)
    A = _adjacency(adjacency, V)
    model = L_STGCN(_hp(V, CLASSES), compute_dtype=torch.bfloat16, A=A).cuda().train()
    g = torch.Generator().manual_seed(100)
    x = torch.randn(N_SEQ * M, T, V, 3, generator=g).cuda()
    labels = torch.randint(0, CLASSES, (N_SEQ * M,), generator=g).cuda()
    with torch.no_grad():
        for bn in [m for m in model.modules() if isinstance(m, torch.nn.BatchNorm2d)]:
            bn.weight.uniform_(0.7, 1.3); bn.bias.uniform_(-0.2, 0.2)
    state = {k: v.detach().clone() for k, v in model.state_dict().items()}

    # ---- the product: eager step -----------------------------------------------------------------------------------
    model.zero_grad(set_to_none=True)
    logits = model(x)
    loss = cross_entropy(logits, labels)
    loss.backward()
    got_logits = logits.detach().float().cpu()
    got = {k: p.grad.detach().float().cpu() for k, p in model.named_parameters() if p.grad is not None}
    loss_eager = float(loss)
    # drop the autograd graph of the eager step: while it lives, the parameters' AccumulateGrad nodes stay bound to the stream
    # it ran on (the legacy default stream), and a later capture of backward on another stream would make the legacy stream
    # wait on a capturing one (cudaErrorStreamCaptureImplicit)
    del logits, loss
    model.zero_grad(set_to_none=True)
    assert torch.isfinite(got_logits).all()

    # ---- the checker -------------------------------------------------------------------------------------------------
    loss32, logits32, g32 = _oracle_step(x.float(), labels, state, bf16=False)
    loss16, logits16, g16 = _oracle_step(x.float(), labels, state, bf16=True)
    report = {'workload': workload, 'n_eff': N_SEQ * M, 'V': V, 'classes': CLASSES, 'adjacency': adjacency, 'max_abs_A': float(A.abs().max()),
              'logits_vs_fp32': rel_err(got_logits, logits32), 'logits_vs_emulation': rel_err(got_logits, logits16),
              'emulation_logits_vs_fp32': rel_err(logits16, logits32),
              'loss': loss_eager, 'loss_fp32': loss32, 'loss_emulation': loss16, 'grads': {}}
    assert set(got) == set(g32), set(got) ^ set(g32)
    worst = {'vs_emulation': 0.0, 'vs_fp32': 0.0, 'floor': 0.0}
    failures = []

    def l2(a, b):
        return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))
    cat = {'got': [], 'g32': [], 'g16': []}
    for k in sorted(g32):
        ref32, ref16, mine = g32[k], g16[k], got[k].reshape(g32[k].shape)
        if k.endswith('temporalConv.bias'):
            # in front of a batch-statistics BatchNorm: mathematically zero; the oracle's own value is
            # round-off of the size printed here
            report['grads'][k] = {'abs_max': float(mine.abs().max()), 'oracle_abs_max': float(ref32.abs().max())}
            if float(mine.abs().max()) > 1e-4:
                failures.append((k, 'structurally zero', float(mine.abs().max())))
            continue
        e16, e32, floor = rel_err(mine, ref16), rel_err(mine, ref32), rel_err(ref16, ref32)
        l32, lfloor = l2(mine, ref32), l2(ref16, ref32)
        report['grads'][k] = {'vs_emulation': e16, 'vs_fp32': e32, 'emulation_vs_fp32': floor, 'l2_vs_fp32': l32,
                              'l2_emulation_vs_fp32': lfloor}
        worst = {'vs_emulation': max(worst['vs_emulation'], e16), 'vs_fp32': max(worst['vs_fp32'], e32),
                 'floor': max(worst['floor'], floor)}
        if l32 >= 3 * lfloor + TOL:
            failures.append((k, 'l2 vs fp32 %.3g, emulation %.3g' % (l32, lfloor)))
        for name, t in (('got', mine), ('g32', ref32), ('g16', ref16)):
            cat[name].append(t.flatten().double())
    G = {k: torch.cat(v) for k, v in cat.items()}
    cos = lambda a, b: float((a @ b) / (a.norm() * b.norm()))
    report['global'] = {'l2_vs_fp32': l2(G['got'], G['g32']), 'l2_emulation_vs_fp32': l2(G['g16'], G['g32']),
                        'l2_vs_emulation': l2(G['got'], G['g16']), 'cos_vs_fp32': cos(G['got'], G['g32']),
                        'cos_emulation_vs_fp32': cos(G['g16'], G['g32']), 'elements': int(G['g32'].numel())}
    report['worst'] = worst
    n_dec, n_rows = assert_argmax_consistent(got_logits, logits32)
    report['argmax_rows_decisive'] = [n_dec, n_rows]

    # ---- the same step through the captured graph (what bench.py replays) ------------------------------------------
    model.load_state_dict(state)
    opt = Adam(model.parameters(), lr=1e-4)
    step = CapturedTrainStep(model, opt, x, labels)
    loss_c = float(step(x, labels))
    report['loss_captured'] = loss_c
    after = {k: v.detach().clone() for k, v in model.state_dict().items()}
    step.close()
    model.load_state_dict(state)
    opt2 = Adam(model.parameters(), lr=1e-4)
    opt2.zero_grad(set_to_none=True)
    loss_e = cross_entropy(model(x), labels); loss_e.backward(); opt2.step()
    loss_e = float(loss_e)
    cap_err = max(rel_err(after[k], v) for k, v in model.state_dict().items() if v.dtype.is_floating_point)
    report['captured_vs_eager_state'] = cap_err

    os.makedirs('gpurun_out', exist_ok=True)
    with open(os.path.join('gpurun_out', '%s_parity_%s.json' % (workload, adjacency)), 'w') as f:
        json.dump(report, f, indent=1)
    print(json.dumps({k: v for k, v in report.items() if k != 'grads'}))
    assert report['logits_vs_fp32'] < TOL, report['logits_vs_fp32']
    assert abs(loss_eager - loss32) < TOL * max(1.0, abs(loss32))
    assert abs(loss_c - loss_e) < 1e-2 * max(1.0, abs(loss_e))
    assert cap_err < 2e-2, cap_err          # Adam's first step is lr * sign(g): a near-zero gradient of the other sign moves a weight by 2 lr
    assert not failures, failures[:8]
    gl = report['global']
    assert gl['l2_vs_fp32'] < 1.5 * gl['l2_emulation_vs_fp32'] + TOL, gl
    assert gl['cos_vs_fp32'] > gl['cos_emulation_vs_fp32'] - 0.02, gl
    assert n_dec >= n_rows // 2, (n_dec, n_rows)
