"""Data-parallel training support: one process per GPU, batch sharded over ranks, weights replicated.

The reference has no distributed code; what the path needs is one exchange step: the
all-reduce of the parameter gradients (3.06 M fp32 values).  Gradients are packed into a few flat buckets in
REVERSE block order (blocks 7-9 hold 74 % of the bytes and finish first in backward); as soon as every
gradient of a bucket has been produced, the bucket is all-reduced with NCCL on a side stream while the
remaining backward kernels keep running.  BatchNorm statistics stay per replica (PyTorch DDP default).
Inference needs no collective.  Works with backend 'gloo' on CPU tensors for tests.
"""
import torch
import torch.distributed as dist


class GradBucketReducer:
    def __init__(self, params, num_buckets=3, process_group=None, average=True):
        self.params = [p for p in params if p.requires_grad]
        self.pg = process_group
        self.average = average
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        # buckets over the reversed parameter list (backward order), split by cumulative size
        rev = list(reversed(self.params))
        total = sum(p.numel() for p in rev)
        target = total / max(1, num_buckets)
        self.buckets, cur, acc = [], [], 0
        for p in rev:
            cur.append(p); acc += p.numel()
            if acc >= target and len(self.buckets) < num_buckets - 1:
                self.buckets.append(cur); cur, acc = [], 0
        if cur:
            self.buckets.append(cur)
        self.flat, self.views, self.owner = [], [], {}
        for bi, b in enumerate(self.buckets):
            flat = torch.zeros(sum(p.numel() for p in b), dtype=b[0].dtype, device=b[0].device)
            off, vs = 0, []
            for p in b:
                vs.append(flat[off:off + p.numel()].view_as(p)); off += p.numel()
                self.owner[p] = bi
            self.flat.append(flat); self.views.append(vs)
        self.pending = [0] * len(self.buckets)
        self.works = []
        self.side = torch.cuda.Stream() if self.params and self.params[0].is_cuda else None
        self.hooks = [p.register_post_accumulate_grad_hook(self._hook) for p in self.params]
        self.reset()

    def reset(self):
        self.pending = [len(b) for b in self.buckets]
        self.works = []

    def _hook(self, p):
        bi = self.owner[p]
        self.pending[bi] -= 1
        if self.pending[bi] == 0:
            self._launch(bi)

    def _launch(self, bi):
        if self.world == 1:
            return
        flat, vs = self.flat[bi], self.views[bi]
        torch._foreach_copy_(vs, [p.grad for p in self.buckets[bi]])
        if self.side is not None:
            self.side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.side):
                w = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.pg, async_op=True)
        else:
            w = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.pg, async_op=True)
        self.works.append((bi, w))

    def finish(self):
        """Wait for the bucket all-reduces and scatter the (averaged) sums back into .grad."""
        if self.world > 1:
            for bi, w in self.works:
                w.wait()
            if self.side is not None:
                torch.cuda.current_stream().wait_stream(self.side)
            for bi, _ in self.works:
                if self.average:
                    self.flat[bi].div_(self.world)
                torch._foreach_copy_([p.grad for p in self.buckets[bi]], self.views[bi])
        self.reset()

    def remove(self):
        for h in self.hooks:
            h.remove()


def shard_batch(x, y, rank, world):
    """Contiguous shard of the leading (sequence) axis; the persons of one sequence stay on one rank."""
    n = x.shape[0]
    per = n // world
    return x[rank * per:(rank + 1) * per], y[rank * per:(rank + 1) * per]
