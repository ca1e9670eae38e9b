"""Host -> device input pipeline for the block stack (plumbing around the hot path, no reference counterpart).

The reference feeds the model from a torch DataLoader and lets Lightning move each batch to the GPU right before the step
(`src/lightning_model.py:199-253`).  `HostBatchPrefetcher` keeps that contract (pinned host batches in, device batches
out) but issues the copy of batch i+1 on a side stream while step i computes, so the PCIe transfer leaves the critical path.
"""
import torch


class HostBatchPrefetcher:
    """Ring of device buffers filled by pinned-host -> device copies on a dedicated stream.

        pf = HostBatchPrefetcher('cuda:0')
        pf.push(batch0)                 # async copy
        for i in range(n):
            pf.push(batch[i + 1])       # overlaps step i
            x, y = pf.pop()             # current stream waits for the copy of batch i only
            step(x, y)
            pf.done()                   # everything queued so far may still read x, y; their slot is reusable after it

    `push` never blocks the host; reuse of a slot is ordered after the `done` that released it by stream events.
    """

    def __init__(self, device, depth=2):
        self.device = torch.device(device)
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.depth = depth
        self.slots = [None] * depth          # device tensors of the slot
        self.ready = [None] * depth          # event: copy into the slot finished
        self.freed = [None] * depth          # event: consumer work on the slot finished
        self.head = self.tail = self.count = 0
        self.last = None

    def push(self, host_batch):
        if self.count == self.depth:
            raise RuntimeError('HostBatchPrefetcher: ring full (pop/done before pushing more than `depth` batches)')
        s = self.head
        if self.slots[s] is None or any(d.shape != h.shape or d.dtype != h.dtype for d, h in zip(self.slots[s], host_batch)):
            self.slots[s] = tuple(torch.empty(h.shape, dtype=h.dtype, device=self.device) for h in host_batch)
            self.freed[s] = None
        with torch.cuda.stream(self.copy_stream):
            if self.freed[s] is not None:
                self.copy_stream.wait_event(self.freed[s])       # never overwrite a batch the compute stream still reads
            for d, h in zip(self.slots[s], host_batch):
                d.copy_(h, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        self.ready[s] = ev
        self.head = (s + 1) % self.depth
        self.count += 1

    def pop(self):
        if self.count == 0:
            raise RuntimeError('HostBatchPrefetcher: nothing pushed')
        s = self.tail
        torch.cuda.current_stream(self.device).wait_event(self.ready[s])
        self.tail = (s + 1) % self.depth
        self.count -= 1
        self.last = s
        return self.slots[s]

    def done(self):
        if self.last is None:
            return
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.freed[self.last] = ev
        self.last = None
