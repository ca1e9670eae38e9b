"""Adam for the drop-in model: torch.optim.Adam semantics (the reference's `configure_optimizers`,
src/lightning_model.py:196-197), executed by ONE update launch of the library per 64 parameter tensors
(`stgcn_adam_step`) instead of torch's multi-tensor kernels.  Same constructor arguments and the same
`state_dict` layout (`step`, `exp_avg`, `exp_avg_sq` per parameter) as torch.optim.Adam, so optimizer
checkpoints move both ways.  The step number lives on the device (a one-thread kernel advances it and forms
the bias corrections), so a CUDA graph of the whole training step replays correctly
(`stgcn_b200.training.CapturedTrainStep`).  fp32 CUDA parameters only; anything else raises (no fallback).
"""
import ctypes as C

import torch

from . import _lib as L


class Adam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, amsgrad=False):
        if amsgrad:
            raise RuntimeError('stgcn_b200.optim.Adam: amsgrad is not implemented')
        if lr < 0 or eps < 0 or not (0 <= betas[0] < 1) or not (0 <= betas[1] < 1) or weight_decay < 0:
            raise ValueError('stgcn_b200.optim.Adam: invalid hyper-parameter')
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=False))
        self._tables = {}
        self._dev_step = {}            # (group index, device) -> int64[2] device tensor: {step, two float corrections}

    # ---- step state -----------------------------------------------------------------------------------------------
    def _moments(self, p):
        st = self.state[p]
        if 'exp_avg' not in st:
            st['exp_avg'] = torch.zeros_like(p, memory_format=torch.preserve_format)
            st['exp_avg_sq'] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return st

    def _step_tensor(self, gi, dev, plist):
        key = (gi, dev)
        t = self._dev_step.get(key)
        if t is None:
            start = 0
            for p in plist:                       # resume from a loaded state_dict
                s = self.state[p].get('step')
                if s is not None:
                    start = max(start, int(s))
            t = torch.zeros(2, dtype=torch.int64, device=dev)
            t[0] = start
            self._dev_step[key] = t
        return t

    def steps_taken(self, group=0):
        """Step number of a parameter group, read from the device (synchronises)."""
        vals = [int(t[0]) for (gi, _), t in self._dev_step.items() if gi == group]
        return max(vals) if vals else 0

    def state_dict(self):
        for (gi, dev), t in self._dev_step.items():
            n = float(int(t[0]))
            for p in self.param_groups[gi]['params']:
                if p in self.state and p.device == dev:
                    self.state[p]['step'] = torch.tensor(n, dtype=torch.float32)
        return super().state_dict()

    def load_state_dict(self, sd):
        super().load_state_dict(sd)
        self._tables.clear()
        self._dev_step.clear()                    # re-seeded from the loaded `step` entries on the next step()

    def reset_state(self):
        """Zero the moments and the step IN PLACE (pointers captured by a CUDA graph stay valid)."""
        for st in self.state.values():
            for k in ('exp_avg', 'exp_avg_sq'):
                if k in st:
                    st[k].zero_()
            st.pop('step', None)
        for t in self._dev_step.values():
            t.zero_()

    # ---- the update -------------------------------------------------------------------------------------------------
    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = L.lib()
        for gi, group in enumerate(self.param_groups):
            by_dev = {}
            for p in group['params']:
                if p.grad is None:
                    continue
                if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError('stgcn_b200.optim.Adam: parameters must be contiguous fp32 CUDA tensors')
                g = p.grad
                if g.is_sparse or g.dtype != torch.float32 or g.device != p.device:
                    raise RuntimeError('stgcn_b200.optim.Adam: gradients must be dense fp32 on the parameter\'s device')
                by_dev.setdefault(p.device, []).append(p)
            for dev, plist in by_dev.items():
                n = len(plist)
                key = (gi, dev, tuple(id(p) for p in plist))
                tb = self._tables.get(key)
                if tb is None:      # parameter / moment pointers never change; only the gradient table is refreshed per step
                    arr = C.c_void_p * n
                    moms = [self._moments(p) for p in plist]
                    tb = (arr(*[p.data_ptr() for p in plist]), arr(), arr(*[m['exp_avg'].data_ptr() for m in moms]),
                          arr(*[m['exp_avg_sq'].data_ptr() for m in moms]), (C.c_int64 * n)(*[p.numel() for p in plist]))
                    if len(self._tables) > 8:
                        self._tables.clear()
                    self._tables[key] = tb
                P, G, M, V, S = tb
                keep = []
                for i, p in enumerate(plist):
                    g = p.grad
                    if not g.is_contiguous():
                        g = g.contiguous(); keep.append(g)
                    G[i] = g.data_ptr()
                    if P[i] != p.data_ptr():                      # storage was swapped (.to(), load_state_dict of the model)
                        st = self._moments(p)
                        P[i], M[i], V[i] = p.data_ptr(), st['exp_avg'].data_ptr(), st['exp_avg_sq'].data_ptr()
                step_t = self._step_tensor(gi, dev, plist)
                with torch.cuda.device(dev):
                    L.check(lib.stgcn_adam_step(n, P, G, M, V, S, float(group['lr']), float(group['betas'][0]),
                                                float(group['betas'][1]), float(group['eps']), float(group['weight_decay']),
                                                0, L.ptr(step_t), L.stream_ptr(dev)), 'adam_step')
        return loss
