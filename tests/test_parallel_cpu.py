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
