"""Deterministic RNG-free parameter fill shared by make_golden.py and the tests."""
import numpy as np
import torch


def det_fill(t: torch.Tensor, tag: int, scale: float, offset: float = 0.0):
    """Deterministic, RNG-free parameter fill that any machine can reproduce:
    value_i = offset + scale * sin(0.37*i*(1+tag%7) + 1.3*tag)."""
    n = t.numel()
    i = torch.arange(n, dtype=torch.float64)
    v = offset + scale * torch.sin(0.37 * i * (1 + tag % 7) + 1.3 * tag)
    with torch.no_grad():
        t.copy_(v.reshape(t.shape).to(t.dtype))


def det_params(module: torch.nn.Module, a_mode='keep'):
    """Fill every parameter/buffer of a reference module deterministically.
    Conv weights ~ 1/sqrt(fan_in); BN gamma around 1, beta small; running
    mean small, running var around 1; `A` untouched unless a_mode == 'fill'."""
    for tag, (name, t) in enumerate(list(module.named_parameters()) + list(module.named_buffers())):
        if name.endswith('num_batches_tracked'):
            continue
        if name.endswith('spatialConv.A') or name == 'A':
            if a_mode == 'fill':
                det_fill(t, tag, 0.2)
            continue
        if 'Masks' in name:
            continue
        if name.endswith('running_var'):
            det_fill(t, tag, 0.3, 1.0)
        elif name.endswith('running_mean'):
            det_fill(t, tag, 0.2)
        elif 'batch_n' in name and name.endswith('weight'):
            det_fill(t, tag, 0.3, 1.0)
        elif 'batch_n' in name and name.endswith('bias'):
            det_fill(t, tag, 0.2)
        elif name.endswith('bias'):
            det_fill(t, tag, 0.1)
        else:
            fan_in = t[0].numel() if t.dim() > 1 else t.numel()
            det_fill(t, tag, 1.5 / np.sqrt(fan_in))
