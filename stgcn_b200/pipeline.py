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


# ---- the reference's collate + augmentation, on the device ------------------------------------------------------------------------

AUG_ROTATIONS = (15, -15, 5, -5, 10, -10)          # augment_data's tables, src/data/augmentation.py:15-17
AUG_TRANSLATIONS = ((5, 5), (0, 5), (5, 0))
AUG_SCALES = (1.05, 1.1, 0.95)


def sample_augmentation(rng=None):
    """Draw the 2x2 block of the matrix `augment_data` would draw (src/data/augmentation.py:19-48) from `rng` (numpy's global
    generator by default, the one the reference uses), consuming the same random numbers in the same order.  The reference
    multiplies (x, y, 0) by its 3x3 matrix, so the translation it draws never acts; it is drawn here too (to stay in step with
    the generator) and dropped."""
    import numpy as np
    rng = np.random if rng is None else rng
    chosen = rng.choice(['rotation', 'translation', 'scaling', 'flip'], 2)
    m = np.eye(2)
    if 'rotation' in chosen:
        th = np.radians(rng.choice(AUG_ROTATIONS))
        c, s = np.cos(th), np.sin(th)
        m = np.array([[c, s], [-s, c]]).dot(m)
    if 'translation' in chosen:
        rng.choice(range(3))
    if 'scaling' in chosen:
        m = rng.choice(AUG_SCALES) * m
    if 'flip' in chosen:
        m = np.array([[-1.0, 0.0], [0.0, 1.0]]).dot(m)
    return m


class DeviceCollate:
    """`loopy_pad_collate_fn` (src/data/util.py:33-47) with the augmentation and the storage cast folded in, as one kernel.

        collate = DeviceCollate('cuda:0', dtype=torch.bfloat16)
        x, labels = collate(batch)                       # batch = [(f_in (N_i,T_i,V,C) array, labels (N_i,) array), ...]
        x, labels = collate(batch, xforms=[sample_augmentation() or None for _ in batch])

    The ragged sequences are packed into one pinned host buffer, copied once, and expanded on the device to (sum N_i, T*, V, C)
    -- T* = the longest sequence, shorter ones wrap around -- in the layout and dtype block 0 reads.  `xforms[i]` (2x2, or the
    reference's 3x3) applies to every sequence of batch element i, like `KTHDataset.__getitem__` applying one `augment_data`
    call per item (src/data/datasets.py:153-155)."""

    def __init__(self, device, dtype=torch.float32):
        self.device = torch.device(device)
        self.dtype = dtype
        self._pinned = None

    def __call__(self, batch, xforms=None):
        import ctypes as C
        import numpy as np
        from . import _lib as L
        if self.device.type != 'cuda':
            raise RuntimeError('stgcn_b200.DeviceCollate: a CUDA device is required (there is no CPU path)')
        arrays = [np.ascontiguousarray(np.asarray(x), dtype=np.float32) for x, _ in batch]
        labels = [np.asarray(l).reshape(-1) for _, l in batch]
        if not arrays:
            raise RuntimeError('stgcn_b200.DeviceCollate: empty batch')
        v, c = arrays[0].shape[2], arrays[0].shape[3]
        if any(a.ndim != 4 or a.shape[2] != v or a.shape[3] != c for a in arrays):
            raise RuntimeError('stgcn_b200.DeviceCollate: every element must be (N_i, T_i, V, C) with one V and C')
        t_out = max(a.shape[1] for a in arrays)
        offs, lens, mats, total = [], [], [], 0
        for i, a in enumerate(arrays):
            for p in range(a.shape[0]):
                offs.append(total + p * a.shape[1] * v * c); lens.append(a.shape[1])
                if xforms is not None:
                    m = np.eye(2) if xforms[i] is None else np.asarray(xforms[i], dtype=np.float64)[:2, :2]
                    mats.append(m.reshape(4))
            total += a.size
        n = len(offs)
        if self._pinned is None or self._pinned.numel() < total:
            self._pinned = torch.empty(max(total, 1), dtype=torch.float32).pin_memory()
        host = self._pinned[:total].numpy()
        pos = 0
        for a in arrays:
            host[pos:pos + a.size] = a.reshape(-1); pos += a.size
        packed = self._pinned[:total].to(self.device, non_blocking=True)
        meta = torch.tensor(offs + lens, dtype=torch.int64)
        d_off = meta[:n].to(self.device)
        d_len = meta[n:].to(torch.int32).to(self.device)
        d_mat = torch.tensor(np.stack(mats), dtype=torch.float32).to(self.device) if xforms is not None else None
        out = torch.empty((n, t_out, v, c), dtype=self.dtype, device=self.device)
        with torch.cuda.device(self.device):
            L.check(L.lib().stgcn_collate_pad_augment(L.dtype_code(self.dtype), n, t_out, v, c, L.ptr(packed), L.ptr(d_off), L.ptr(d_len),
                                                      L.ptr(d_mat), L.ptr(out), L.stream_ptr(self.device)), 'collate_pad_augment')
        return out, torch.from_numpy(np.concatenate(labels)).to(self.device)
