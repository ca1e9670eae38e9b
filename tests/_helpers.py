"""Shared test helpers (golden loading, tolerance checks)."""
import glob
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + '.npz'), allow_pickle=False)
    return {k: z[k] for k in z.files}


def block_case_names():
    return sorted(os.path.basename(f)[:-4] for f in glob.glob(os.path.join(GOLDEN, 'b_*.npz')))


def rel_err(got, ref, floor=1e-6):
    """max|got-ref| / max(max|ref|, floor): the norm-scaled per-tensor measure of
    SURVEY.md section 7.3-5 (a per-element relative error is meaningless for
    structurally-zero gradients)."""
    got = torch.as_tensor(np.asarray(got)).double() if not torch.is_tensor(got) else got.detach().double().cpu()
    ref = torch.as_tensor(np.asarray(ref)).double() if not torch.is_tensor(ref) else ref.detach().double().cpu()
    assert got.shape == ref.shape, (got.shape, ref.shape)
    return float((got - ref).abs().max() / max(float(ref.abs().max()), floor))


def params_from_golden(g, dtype=torch.float32):
    return {k[2:]: torch.tensor(v).to(dtype) if v.dtype.kind == 'f' else torch.tensor(v)
            for k, v in g.items() if k.startswith('p.')}
