"""One training step of the drop-in model as ONE CUDA graph.

The reference's step is Lightning's `training_step` + `optimizer.step()` (src/lightning_model.py:196-205): forward,
F.cross_entropy, backward, Adam.  Here that is ~260 kernel launches of 10-60 us each; issued one by one from Python the
host needs about as long to enqueue them as the B200 needs to run them.  `CapturedTrainStep` records the whole step once
(zero_grad -> forward -> loss -> backward -> gradient all-reduce -> optimizer) with `torch.cuda.graph` and replays it per
batch: one `cudaGraphLaunch` per step, the programmatic-dependent-launch edges between the library's kernels included.

    step = CapturedTrainStep(model, optimizer, example_x, example_y)       # model.train(), stgcn_b200.optim.Adam
    for x, y in loader:                                                     # device tensors of the example's shape / dtype
        loss = step(x, y)                                                   # 0-dim device tensor, valid until the next call

What makes the step capturable: the library never synchronises or allocates, BatchNorm's running statistics and
`num_batches_tracked` are advanced by kernels, and the optimizer keeps its step number on the device.  The two warm-up
steps that precede capture run on the example batch and are UNDONE afterwards (parameters, buffers and optimizer state are
restored in place), so constructing the object does not train the model.

Drop every autograd graph of earlier EAGER steps on the same parameters before constructing the object (`del loss, logits`):
while such a graph lives, the parameters' AccumulateGrad nodes stay bound to the stream that step ran on (normally the legacy
default stream), and capturing backward on the capture stream then fails with cudaErrorStreamCaptureImplicit.
"""
import torch

from . import _lib
from .functional import cross_entropy


class CapturedTrainStep:
    def __init__(self, model, optimizer, example_x, example_y, loss_fn=cross_entropy, reducer=None, warmup=2):
        if not example_x.is_cuda:
            raise RuntimeError('stgcn_b200.CapturedTrainStep: CUDA tensors required (there is no CPU path)')
        if not model.training:
            raise RuntimeError('stgcn_b200.CapturedTrainStep: put the model in train() mode first')
        self.model, self.opt, self.loss_fn, self.reducer = model, optimizer, loss_fn, reducer
        self._blocks = [m for m in model.modules() if hasattr(m, '_prep_cache')]
        for m in self._blocks:            # dropout state lives on the device and is advanced by the library: every replay draws
            m._ensure_rng(example_x.device)   # a fresh mask; create it now so the warm-up below can be undone exactly
        self.x = torch.empty_like(example_x)
        self.y = torch.empty_like(example_y)
        self.x.copy_(example_x); self.y.copy_(example_y)
        dev = example_x.device
        # ---- warm-up (sizes the scratch arena, creates optimizer state), then undo it in place ---------------------------
        saved = [t.detach().clone() for t in self._model_tensors()]
        opt_saved = {p: {k: v.clone() for k, v in st.items() if torch.is_tensor(v)} for p, st in optimizer.state.items()}
        dev_saved = {k: t.clone() for k, t in getattr(optimizer, '_dev_step', {}).items()}
        n_warm = max(1, warmup)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        import warnings
        warn_always = torch.is_warn_always_enabled()
        torch.set_warn_always(True)        # the stream-mismatch warning is a once-per-process C++ warning otherwise
        try:
            with torch.cuda.stream(side), warnings.catch_warnings(record=True) as caught:
                warnings.simplefilter('always')
                for _ in range(n_warm):
                    self._eager_step()
        finally:
            torch.set_warn_always(warn_always)
        torch.cuda.current_stream(dev).wait_stream(side)
        stale = any('AccumulateGrad' in str(w.message) and 'stream' in str(w.message) for w in caught)
        for w in caught:
            if not ('AccumulateGrad' in str(w.message) and 'stream' in str(w.message)):
                warnings.warn_explicit(w.message, w.category, w.filename, w.lineno)
        with torch.no_grad():
            for t, s in zip(self._model_tensors(), saved):
                t.copy_(s)
            for p, st in optimizer.state.items():           # state created by the warm-up starts from zero, like a fresh optimizer
                for k, v in st.items():
                    if torch.is_tensor(v):
                        old = opt_saved.get(p, {}).get(k)
                        v.copy_(old) if old is not None else v.zero_()
            for k, t in getattr(optimizer, '_dev_step', {}).items():     # device-resident step numbers (stgcn_b200.optim.Adam)
                if k in dev_saved:
                    t.copy_(dev_saved[k])
                else:
                    t[0] -= n_warm
        if stale:      # found by the warm-up (it runs on a side stream), BEFORE a capture that would fail and poison the process
            raise RuntimeError('stgcn_b200.CapturedTrainStep: an autograd graph of an earlier step on these parameters is still '
                               'alive (e.g. a `loss` or `logits` tensor you kept), which pins their AccumulateGrad nodes to the '
                               'stream that step ran on; backward cannot be captured on another stream.  Delete those tensors '
                               '(`del loss, logits`) and construct the object again.')
        # ---- capture ----------------------------------------------------------------------------------------------------------
        optimizer.zero_grad(set_to_none=True)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss = self._eager_step()
        # the graph holds raw addresses inside the library's scratch arena: keep that arena alive for as long as the graph
        self._workspace = _lib.current_workspace(dev)

    def close(self):
        """Destroy the captured graph (and with it every reference it holds: NCCL work of the gradient all-reduce, the scratch
        arena).  Call before `torch.distributed.destroy_process_group()`; the object cannot be called afterwards."""
        if getattr(self, 'graph', None) is not None:
            torch.cuda.synchronize()
            self.graph.reset()
            self.graph = None
            self.loss = None
            self._workspace = None

    def _model_tensors(self):
        return list(self.model.parameters()) + list(self.model.buffers())

    def _eager_step(self):
        self.opt.zero_grad(set_to_none=True)
        loss = self.loss_fn(self.model(self.x), self.y)
        loss.backward()
        if self.reducer is not None:
            self.reducer.finish()
        self.opt.step()
        return loss.detach()

    def __call__(self, x, y):
        """Copy the batch into the captured input buffers (device or pinned-host source, stream-ordered) and replay."""
        if x.shape != self.x.shape or y.shape != self.y.shape:
            raise RuntimeError('stgcn_b200.CapturedTrainStep: batch shape %s differs from the captured %s' %
                               (tuple(x.shape), tuple(self.x.shape)))
        if self.graph is None:
            raise RuntimeError('stgcn_b200.CapturedTrainStep: called after close()')
        self.x.copy_(x, non_blocking=True)
        self.y.copy_(y, non_blocking=True)
        self.graph.replay()
        for m in self._blocks:            # parameters and running statistics changed behind torch's version counters
            m._prep_cache.clear()
        return self.loss


class CapturedForward:
    """Eval-mode forward of the drop-in model as one CUDA graph (frozen parameters: the prepared two-kernels-per-block path).

        fwd = CapturedForward(model.eval(), example_x)
        logits = fwd(x)              # (N, classes) device tensor, valid until the next call

    Re-capture (construct a new object) after the parameters or the running statistics change: the graph replays the operands
    prepared at capture time."""

    def __init__(self, model, example_x, warmup=2):
        if not example_x.is_cuda:
            raise RuntimeError('stgcn_b200.CapturedForward: CUDA tensors required (there is no CPU path)')
        if model.training:
            raise RuntimeError('stgcn_b200.CapturedForward: put the model in eval() mode first')
        self.model = model
        self.x = torch.empty_like(example_x)
        self.x.copy_(example_x)
        dev = example_x.device
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(max(1, warmup)):
                model(self.x)
        torch.cuda.current_stream(dev).wait_stream(side)
        self.graph = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self.out = model(self.x)
        # the graph reads the operands prepared during the warm-up: keep them alive even if the modules drop their caches
        self._keep = [m._prep_cache.get('blob') for m in model.modules() if hasattr(m, '_prep_cache')]
        self._workspace = _lib.current_workspace(dev)      # and the scratch arena whose addresses it recorded

    def __call__(self, x):
        if x.shape != self.x.shape:
            raise RuntimeError('stgcn_b200.CapturedForward: batch shape %s differs from the captured %s' %
                               (tuple(x.shape), tuple(self.x.shape)))
        self.x.copy_(x, non_blocking=True)
        self.graph.replay()
        return self.out

    def stream(self, host_batches, out_host=None):
        """Serving loop over an iterable of PINNED host batches: the host->device copy of batch i+1 runs on a copy stream while the
        graph computes batch i (two device staging buffers), and each batch's logits go to a pinned host buffer.  Generator: yields
        the host logits of batch i once they have arrived, while batch i+1 is already running; the yielded buffer is valid until
        the generator is advanced again (it is one of two that alternate) unless `out_host`, a list of pinned buffers, provides
        one per batch."""
        dev = self.x.device
        main = torch.cuda.current_stream(dev)
        if not hasattr(self, '_copy_stream'):
            self._copy_stream = torch.cuda.Stream(device=dev)
            self._stage = [torch.empty_like(self.x) for _ in range(2)]
            self._ev_in = [torch.cuda.Event() for _ in range(2)]
            self._ev_free = [torch.cuda.Event() for _ in range(2)]
            self._ev_out = [torch.cuda.Event() for _ in range(2)]
            self._out_host = [torch.empty(self.out.shape, dtype=torch.float32).pin_memory() for _ in range(2)]
            self._used = [False, False]

        def prefetch(hx, j):
            if hx.shape != self.x.shape or not hx.is_pinned():
                raise RuntimeError('stgcn_b200.CapturedForward.stream: batches must be pinned host tensors of the captured shape')
            with torch.cuda.stream(self._copy_stream):
                if self._used[j]:
                    self._copy_stream.wait_event(self._ev_free[j])       # the graph input took over this staging buffer's last batch
                self._stage[j].copy_(hx, non_blocking=True)
                self._ev_in[j].record(self._copy_stream)
            self._used[j] = True

        it = iter(host_batches)
        nxt = next(it, None)
        if nxt is not None:
            prefetch(nxt, 0)
        i = 0
        pending = None
        while nxt is not None:
            j = i & 1
            cur, nxt = nxt, next(it, None)
            if nxt is not None:
                prefetch(nxt, j ^ 1)
            main.wait_event(self._ev_in[j])
            self.x.copy_(self._stage[j], non_blocking=True)
            self._ev_free[j].record(main)
            self.graph.replay()
            dst = out_host[i] if out_host is not None else self._out_host[j]
            dst.copy_(self.out, non_blocking=True)
            self._ev_out[j].record(main)
            if pending is not None:                    # hand out batch i-1 while batch i computes
                pending[0].synchronize()
                yield pending[1]
            pending = (self._ev_out[j], dst)
            i += 1
        if pending is not None:
            pending[0].synchronize()
            yield pending[1]
