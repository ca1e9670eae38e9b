"""Data-parallel training support: one process per GPU, batch sharded over ranks, weights replicated.

The reference has no distributed code; what the path needs is one exchange step: the
all-reduce of the parameter gradients (3.06 M fp32 values).  Gradients live in a few flat buckets laid out in
REVERSE parameter order (backward order): the first bucket closes once ~70 % of the bytes are in (the head and
blocks 9-7, which finish first), the last one holds only the early blocks (~7 %), so what is still in flight
when backward ends is small.  The library's backward writes each gradient straight into its slice of a bucket
(`p.grad` IS the slice -- no pack / unpack copies); as soon as every gradient of a bucket has been produced the
bucket is all-reduced on a side stream (NCCL `AVG`, so no separate divide) while the remaining backward
kernels keep running.  BatchNorm statistics stay per replica by default (PyTorch DDP default);
enable_sync_batchnorm() switches the blocks to global-batch statistics (SyncBN).  Inference needs no
collective.  Works with backend 'gloo' on CPU tensors for tests (SUM + divide there).

Constraint: a parameter must be used once per step (no weight sharing between calls inside one backward).
"""
import ctypes

import torch
import torch.distributed as dist


class GradBucketReducer:
    def __init__(self, params, num_buckets=3, process_group=None, average=True, fractions=(0.70, 0.93)):
        self.params = [p for p in params if p.requires_grad]
        self.pg = process_group
        self.average = average
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.use_avg = bool(average and dist.is_initialized() and dist.get_backend(process_group) == 'nccl')
        rev = list(reversed(self.params))
        total = max(1, sum(p.numel() for p in rev))
        if fractions is None or num_buckets != len(fractions) + 1:
            fractions = tuple((i + 1) / num_buckets for i in range(num_buckets - 1))
        self.buckets, cur, acc = [], [], 0
        for p in rev:
            cur.append(p); acc += p.numel()
            if len(self.buckets) < len(fractions) and acc >= fractions[len(self.buckets)] * total:
                self.buckets.append(cur); cur = []
        if cur:
            self.buckets.append(cur)
        self.flat, self.views, self.owner = [], [], {}
        for bi, b in enumerate(self.buckets):
            flat = torch.zeros(sum(p.numel() for p in b), dtype=b[0].dtype, device=b[0].device)
            off, vs = 0, []
            for i, p in enumerate(b):
                v = flat[off:off + p.numel()].view_as(p); off += p.numel()
                vs.append(v)
                self.owner[p] = (bi, i)
                p._stgcn_grad_buf = v          # the library's backward writes the gradient here (functional._grad_out)
            self.flat.append(flat); self.views.append(vs)
        self.expect = [set(range(len(b))) for b in self.buckets]   # gradients a bucket waits for (learned from the first step)
        self.learned = False
        self.side = torch.cuda.Stream() if self.params and self.params[0].is_cuda else None
        self.hooks = [p.register_post_accumulate_grad_hook(self._hook) for p in self.params]
        self.reset()

    def reset(self):
        self.pending = [len(e) for e in self.expect]
        self.fired = [set() for _ in self.buckets]
        self.launched = [False] * len(self.buckets)
        self.works = []

    def _hook(self, p):
        bi, i = self.owner[p]
        if i in self.fired[bi]:
            raise RuntimeError('stgcn_b200.GradBucketReducer: a parameter received two gradients in one step '
                               '(weight sharing inside one backward is not supported)')
        self.fired[bi].add(i)
        v = self.views[bi][i]
        if p.grad.data_ptr() != v.data_ptr():        # produced elsewhere (stock autograd op) or cloned by autograd: adopt it
            v.copy_(p.grad)
            p.grad = v
        if i not in self.expect[bi]:                 # a gradient nobody expected (first seen after the learning step): this
            self.learned = False                     # step it is reduced only if its bucket is still open; learn again
            return
        self.pending[bi] -= 1
        if self.pending[bi] == 0:
            self._launch(bi)

    def _launch(self, bi):
        self.launched[bi] = True
        if self.world == 1:
            return
        flat = self.flat[bi]
        op = dist.ReduceOp.AVG if self.use_avg else dist.ReduceOp.SUM
        if self.side is not None:
            self.side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.side):
                w = dist.all_reduce(flat, op=op, group=self.pg, async_op=True)
        else:
            w = dist.all_reduce(flat, op=op, group=self.pg, async_op=True)
        self.works.append((bi, w))

    def finish(self):
        """Flush buckets still waiting (parameters that got no gradient this step, e.g. the reference's `Masks`, count as
        zeros -- the same on every rank), wait for the all-reduces; `.grad` then holds the rank average."""
        for bi in range(len(self.buckets)):
            if not self.launched[bi]:
                for i in range(len(self.buckets[bi])):
                    if i not in self.fired[bi]:
                        self.views[bi][i].zero_()
                self._launch(bi)
        if self.world > 1:
            for bi, w in self.works:
                w.wait()
            if self.side is not None:
                torch.cuda.current_stream().wait_stream(self.side)
            if self.average and not self.use_avg:
                for bi, _ in self.works:
                    self.flat[bi].div_(self.world)
        if not self.learned:        # from now on a bucket fires as soon as the gradients that really arrive are in
            self.expect = [set(f) if f else {0} for f in self.fired]
            self.learned = True
        self.reset()

    def remove(self):
        for h in self.hooks:
            h.remove()
        for p in self.params:
            if hasattr(p, '_stgcn_grad_buf'):
                del p._stgcn_grad_buf


def shard_batch(x, y, rank, world):
    """Contiguous shard of the leading (sequence) axis; the persons of one sequence stay on one rank."""
    n = x.shape[0]
    per = n // world
    return x[rank * per:(rank + 1) * per], y[rank * per:(rank + 1) * per]


class SyncBatchNormLink:
    """Global-batch BatchNorm statistics for the blocks of a data-parallel replica (SyncBN; the reference is single-process and
    has no counterpart).

    The library does the arithmetic; this object is the exchange step it calls back into (StgcnBlockParams.sync_*,
    include/stgcn_b200.h): at each of a block's BatchNorm exchange points the library writes 2*C local sums into `buf` and calls
    `allreduce`, which must enqueue an in-place SUM over the ranks on the CURRENT stream (the library is always called on
    torch's current stream) and return without a host synchronisation.  Default: torch.distributed.all_reduce on `group`
    (NCCL: stream-ordered).  All ranks must run the same blocks in the same order (they do: one model, replicated).
    """

    def __init__(self, world=None, group=None, allreduce=None):
        self.world = int(world if world is not None else dist.get_world_size(group))
        self.group = group
        self.buf = None
        self._allreduce = allreduce or (lambda buf: dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=self.group))
        self.calls = 0

        def _cb(user, stream):
            self.calls += 1
            self._allreduce(self.buf)
        self._cfn = ctypes.CFUNCTYPE(None, ctypes.c_void_p, ctypes.c_void_p)(_cb)      # kept alive with the link
        self.fn_ptr = ctypes.cast(self._cfn, ctypes.c_void_p).value

    def on(self, device, channels):
        """The link with its exchange buffer on `device`, at least 2*channels floats (called by the block every forward)."""
        if self.buf is None or self.buf.device != device or self.buf.numel() < 2 * channels:
            self.buf = torch.zeros(2 * max(channels, 256), dtype=torch.float32, device=device)
        return self


def enable_sync_batchnorm(model, group=None, world=None, allreduce=None):
    """Switch every SpatialTemporalConv of `model` to global-batch BatchNorm statistics in training mode; returns the link
    (keep it, or call disable_sync_batchnorm).  With world == 1 the blocks keep their local path."""
    from .modules import SpatialTemporalConv
    link = SyncBatchNormLink(world, group, allreduce)
    for m in model.modules():
        if isinstance(m, SpatialTemporalConv):
            m._sync = link if link.world > 1 else None
    return link


def disable_sync_batchnorm(model):
    from .modules import SpatialTemporalConv
    for m in model.modules():
        if isinstance(m, SpatialTemporalConv):
            m._sync = None
