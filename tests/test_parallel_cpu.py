"""world_size-2 gloo test of the data-parallel gradient reducer (host logic of the N>1 path)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket(); s.bind(('127.0.0.1', 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from stgcn_b200.parallel import GradBucketReducer, shard_batch
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.ReLU(), torch.nn.Linear(16, 16), torch.nn.ReLU(),
                                torch.nn.Linear(16, 4))
    red = GradBucketReducer(model.parameters(), num_buckets=3)
    assert 2 <= len(red.buckets) <= 3 and sum(len(b) for b in red.buckets) == 6
    assert red.buckets[0][0] is list(model.parameters())[-1]           # backward order: last layer first
    x = torch.randn(8, 8, generator=torch.Generator().manual_seed(1)); y = torch.randint(0, 4, (8,), generator=torch.Generator().manual_seed(2))
    xs, ys = shard_batch(x, y, rank, world)
    for step in range(2):                                               # reducer is reusable across steps
        model.zero_grad()
        torch.nn.functional.cross_entropy(model(xs), ys).backward()
        red.finish()
    g = torch.cat([p.grad.flatten() for p in model.parameters()])
    if rank == 0:
        model.zero_grad()
        # mean over ranks of per-shard mean losses == full-batch mean loss (equal shard sizes)
        torch.nn.functional.cross_entropy(model(x), y).backward()
        ref = torch.cat([p.grad.flatten() for p in model.parameters()])
        out.put(float((g - ref).abs().max() / ref.abs().max()))
    dist.destroy_process_group()


def test_bucket_reducer_matches_full_batch_gradients():
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    err = q.get(timeout=120)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert err < 1e-6


def _model_worker(rank, world, port, out):
    """The real drop-in model's parameter set (incl. the edge-importance Masks that never get a gradient) through the
    reducer; arithmetic by the CPU oracle (the product has no CPU path), which is all the reducer needs: autograd gradients
    arriving at the model's Parameters."""
    import argparse
    import numpy as np
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from oracle import stgcn_oracle as O
    from stgcn_b200 import L_STGCN
    from stgcn_b200.parallel import GradBucketReducer, shard_batch
    torch.manual_seed(0)
    hp = argparse.Namespace(C_in=3, nr_classes=7, gamma=9, dropout_rate=0.0, residual=False, use_edge_importance=True,
                            max_mask_jitter=0.01, partitioning=0, d=1, distance_file=None, distances=None)
    model = L_STGCN(hp, A=torch.randn(2, 25, 25, generator=torch.Generator().manual_seed(3)) * 0.2)
    names = [n for n, _ in model.named_parameters()]
    assert any(n.startswith('Masks') for n in names)
    state = dict(model.state_dict(keep_vars=True))
    cfgs = O.stack_cfgs(3, 9, 0.0, False)
    x = torch.randn(4, 12, 25, 3, generator=torch.Generator().manual_seed(1))
    y = torch.randint(0, 7, (4,), generator=torch.Generator().manual_seed(2))

    def grads_of(xs, ys):
        model.zero_grad(set_to_none=True)
        O.cross_entropy(O.model_forward(xs, state, cfgs, True, update_running=False), ys).backward()
        return {n: (p.grad.clone() if p.grad is not None else None) for n, p in model.named_parameters()}

    ref = None
    if rank == 0:     # mean over shards of the per-shard gradients (per-replica BatchNorm statistics)
        per = [grads_of(*shard_batch(x, y, r, world)) for r in range(world)]
        ref = {n: (sum(g[n] for g in per) / world if per[0][n] is not None else None) for n in names}
    red = GradBucketReducer(model.parameters(), num_buckets=3)
    assert len(red.buckets) == 3
    xs, ys = shard_batch(x, y, rank, world)
    errs = []
    for step in range(3):      # step 0 learns which gradients arrive; later steps launch buckets as soon as those are in
        model.zero_grad(set_to_none=True)
        O.cross_entropy(O.model_forward(xs, state, cfgs, True, update_running=False), ys).backward()
        early = sum(red.launched)
        red.finish()
        if step > 0:
            assert early >= 2, early                       # the Masks' bucket no longer waits for them
        if rank == 0:
            for n, p in model.named_parameters():
                if n.startswith('Masks'):
                    assert p.grad is None
                else:
                    errs.append(float((p.grad - ref[n]).abs().max() / ref[n].abs().max().clamp_min(1e-12)))
    if rank == 0:
        out.put(max(errs))
    dist.destroy_process_group()


def test_bucket_reducer_on_the_real_model_with_edge_importance_masks():
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_model_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    err = q.get(timeout=300)
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert err < 1e-5


def test_bucket_slices_are_not_handed_out_while_a_gradient_exists():
    """functional.grad_buf_of: with zero_grad(set_to_none=False) p.grad aliases the bucket slice; writing the new gradient into
    that slice and letting autograd accumulate would double it, so no slice is handed out then."""
    from stgcn_b200.functional import grad_buf_of
    from stgcn_b200.parallel import GradBucketReducer
    lin = torch.nn.Linear(4, 3)
    red = GradBucketReducer(lin.parameters(), num_buckets=1)
    assert all(grad_buf_of(p) is p._stgcn_grad_buf for p in lin.parameters())
    lin(torch.randn(2, 4)).sum().backward()
    red.finish()
    g1 = [p.grad.clone() for p in lin.parameters()]
    assert all(p.grad.data_ptr() == p._stgcn_grad_buf.data_ptr() for p in lin.parameters())        # adopted into the bucket
    assert all(grad_buf_of(p) is None for p in lin.parameters())                                    # a gradient exists now
    lin.zero_grad(set_to_none=False)
    assert all(grad_buf_of(p) is None for p in lin.parameters())
    lin(torch.randn(2, 4)).sum().backward()
    red.finish()
    assert all(torch.isfinite(p.grad).all() for p in lin.parameters())
    lin.zero_grad(set_to_none=True)
    assert all(grad_buf_of(p) is p._stgcn_grad_buf for p in lin.parameters())
    red.remove()
    assert all(grad_buf_of(p) is None for p in lin.parameters()) and g1 is not None
