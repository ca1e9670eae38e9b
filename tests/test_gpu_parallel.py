"""Data-parallel parity on hardware (needs 2 GPUs; skipped otherwise): two NCCL ranks, one process per GPU, the real
10-block model in bf16.  SURVEY.md section 7.3-3: BatchNorm statistics are per replica, so the averaged gradients must equal
the MEAN OF THE PER-SHARD ORACLE GRADIENTS (oracle with bf16 storage emulation, evaluated on rank 0's device), and each
rank's logits must equal the oracle's on that rank's shard.  Run both eagerly and through CapturedTrainStep (the all-reduce
is captured in the graph), and with the reference's edge-importance Masks in the parameter list (they never get a gradient).
"""
import argparse
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
TOL = 2e-2


def _free_port():
    s = socket.socket(); s.bind(('127.0.0.1', 0)); p = s.getsockname()[1]; s.close(); return p


def _hp(edge):
    return argparse.Namespace(C_in=3, nr_classes=60, gamma=9, dropout_rate=0.0, residual=False, use_edge_importance=edge,
                              max_mask_jitter=0.01, partitioning=2, d=1, distance_file=None, distances=None)


def _worker(rank, world, port, edge, out):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, 'tests')):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device('cuda', rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)
    from _helpers import rel_err
    from oracle import stgcn_oracle as O
    from stgcn_b200 import CapturedTrainStep, L_STGCN, cross_entropy
    from stgcn_b200.optim import Adam
    from stgcn_b200.parallel import GradBucketReducer, shard_batch
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(0)
    A = torch.randn(3, 25, 25, generator=torch.Generator().manual_seed(0)) * 0.2
    model = L_STGCN(_hp(edge), compute_dtype=torch.bfloat16, A=A).to(dev).train()
    state = {k: v.detach().clone() for k, v in model.state_dict().items()}
    g = torch.Generator().manual_seed(5)
    x = torch.randn(16, 60, 25, 3, generator=g); y = torch.randint(0, 60, (16,), generator=g)
    xs, ys = shard_batch(x, y, rank, world)
    xs, ys = xs.to(dev), ys.to(dev)
    reducer = GradBucketReducer(model.parameters(), 3)
    cfgs = O.stack_cfgs(3, 9, 0.0, False)

    def oracle_shard(r):
        st = {k: v.detach().clone() for k, v in state.items()}
        leaves = {}
        for k, v in st.items():
            if v.dtype.is_floating_point and 'running' not in k and not k.startswith('Masks'):
                v.requires_grad_(True); leaves[k] = v
        xr, yr = shard_batch(x, y, r, world)
        logits = O.model_forward_bf16(xr.to(dev), st, cfgs, True)
        O.cross_entropy(logits, yr.to(dev)).backward()
        return logits.detach(), {k: v.grad for k, v in leaves.items()}

    # eager step
    model.zero_grad(set_to_none=True)
    logits = model(xs)
    cross_entropy(logits, ys).backward()
    reducer.finish()
    res = {}
    ref_logits, _ = oracle_shard(rank)
    res['logits'] = rel_err(logits.float(), ref_logits)
    aliased = sum(int(p.grad.data_ptr() == p._stgcn_grad_buf.data_ptr()) for p in model.parameters() if p.grad is not None)
    res['aliased'] = aliased / max(1, sum(1 for p in model.parameters() if p.grad is not None))
    if rank == 0:
        per = [oracle_shard(r)[1] for r in range(world)]
        worst = 0.0
        for k, p in model.named_parameters():
            if k.startswith('Masks'):
                assert p.grad is None
                continue
            if k.endswith('temporalConv.bias'):
                continue
            ref = sum(q[k] for q in per) / world
            worst = max(worst, rel_err(p.grad.float(), ref))
        res['grads_vs_mean_of_shard_oracles'] = worst
    # replicas must hold identical averaged gradients
    flat = torch.cat([p.grad.flatten() for p in model.parameters() if p.grad is not None])
    other = flat.clone()
    dist.broadcast(other, src=0)
    res['replica_grad_diff'] = float((flat - other).abs().max())
    # captured step (all-reduce inside the graph) == eager step on a fresh copy of the same state
    model.load_state_dict(state)
    opt = Adam(model.parameters(), lr=1e-4)
    step = CapturedTrainStep(model, opt, xs, ys, reducer=reducer)
    l_cap = float(step(xs, ys))
    cap_state = {k: v.detach().clone() for k, v in model.state_dict().items()}
    step.close()
    model.load_state_dict(state)
    opt2 = Adam(model.parameters(), lr=1e-4)
    opt2.zero_grad(set_to_none=True)
    l_e = cross_entropy(model(xs), ys); l_e.backward(); reducer.finish(); opt2.step()
    res['captured_loss_diff'] = abs(l_cap - float(l_e))
    res['captured_state_diff'] = max(rel_err(cap_state[k], v) for k, v in model.state_dict().items() if v.dtype.is_floating_point)
    # replicas stay in sync after the step
    w = torch.cat([p.detach().flatten() for p in model.parameters()])
    w0 = w.clone(); dist.broadcast(w0, src=0)
    res['replica_weight_diff'] = float((w - w0).abs().max())
    out.put((rank, res))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize('edge', [False, True], ids=['plain', 'edge_importance'])
def test_two_rank_nccl_gradients_match_mean_of_shard_oracles(edge):
    if torch.cuda.device_count() < 2:
        pytest.skip('needs 2 GPUs (gpurun --gpus 2)')
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, edge, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = dict(q.get(timeout=600) for _ in range(2))
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    print(results)
    for r in range(2):
        res = results[r]
        assert res['logits'] < TOL, res
        assert res['replica_grad_diff'] == 0.0, res
        assert res['replica_weight_diff'] == 0.0, res
        assert res['captured_loss_diff'] < 1e-2, res
        assert res['captured_state_diff'] < 2e-2, res
        assert res['aliased'] > 0.9, res                      # gradients were written straight into the bucket slices
    assert results[0]['grads_vs_mean_of_shard_oracles'] < TOL * 2, results[0]
