#!/usr/bin/env python
"""Generate the committed golden fixtures by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

The reference ships no golden vectors, so these are made by
importing its modules as they lie under /root/reference/src:
`network.st_graphconv` imports cleanly; `data.adjacency` and `lightning_model`
need matplotlib / seaborn / pytorch_lightning, which are absent here, so stub
modules are registered in sys.modules first (no reference file is edited or
copied).  Outputs: tests/golden/*.npz (small; committed).
"""
import argparse
import contextlib
import io
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
REF_SRC = '/root/reference/src'


def install_stubs():
    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m
    mod('matplotlib'); mod('matplotlib.pyplot')
    sys.modules['matplotlib'].pyplot = sys.modules['matplotlib.pyplot']
    mod('seaborn', set=lambda *a, **k: None)

    class _LM(torch.nn.Module):
        pass

    class _Trainer:
        @staticmethod
        def add_argparse_args(p):
            return p

    class _ES:
        def __init__(self, *a, **k):
            pass
    pl = mod('pytorch_lightning', Trainer=_Trainer)
    core = mod('pytorch_lightning.core')
    lightning = mod('pytorch_lightning.core.lightning', LightningModule=_LM)
    cb = mod('pytorch_lightning.callbacks', EarlyStopping=_ES)
    pl.core = core; core.lightning = lightning; pl.callbacks = cb
    if REF_SRC not in sys.path:
        sys.path.insert(0, REF_SRC)


from detfill import det_fill, det_params  # noqa: E402  (tests/golden/detfill.py)


BLOCK_CASES = [
    # name,        c_in, c_out, T,  V,  K, stride, residual, training, A kind
    ('b_def_train',   4,  8, 12, 25, 3, 1, False, True,  'dense'),
    ('b_def_eval',    4,  8, 12, 25, 3, 1, False, False, 'dense'),
    ('b_def_s2',      8, 16, 13, 25, 2, 2, False, True,  'dense'),
    ('b_def_k1_v18',  8,  8, 10, 18, 1, 1, False, True,  'dense'),
    ('b_res_id',      8,  8, 12, 25, 3, 1, True,  True,  'dense'),
    ('b_res_conv',    8, 16, 12, 18, 3, 2, True,  True,  'dense'),
    ('b_res_eval',    8, 16, 12, 25, 2, 2, True,  False, 'dense'),
    ('b_def_refA',    3,  8, 12, 25, 3, 1, False, True,  'spatial'),   # the reference's own huge-magnitude A
    ('b_def_refA_uni', 3, 8, 12, 25, 1, 1, False, True,  'uni'),
]


def make_blocks(out):
    from network.st_graphconv import SpatialTemporalConv
    from data import adjacency as adj
    N = 3
    for (name, ci, co, T, V, K, s, res, training, akind) in BLOCK_CASES:
        torch.manual_seed(hash(name) % 1000 if False else sum(map(ord, name)))
        if akind == 'dense':
            A = torch.randn(K, V, V) * 0.2
        elif akind == 'spatial':
            dist_file = os.path.join(HERE, '_dist.npy')
            np.save(dist_file, np.random.RandomState(0).rand(25))
            A = adj.get_normalized_adjacency_matrices(adj.Strategy.SPATIAL_CONFIGURATION, 1, distance_file=dist_file)
            os.remove(dist_file)
        else:
            A = adj.get_normalized_adjacency_matrices(adj.Strategy.UNI_LABELING, 1)
        with contextlib.redirect_stdout(io.StringIO()):
            m = SpatialTemporalConv(ci, co, A, 9, s, 4, dropout_rate=0, residual=res)
        det_params(m)
        m.train(training)
        x = torch.randn(N, ci, T, V, requires_grad=True)
        y = m(x)
        gy = torch.randn_like(y)
        (y * gy).sum().backward()
        rec = {'x': x.detach().numpy(), 'gy': gy.numpy(), 'y': y.detach().numpy(), 'dx': x.grad.numpy(),
               'A_init': A.numpy()}
        # parameters BEFORE (values are deterministic, stored anyway so the fixture is self-contained)
        with contextlib.redirect_stdout(io.StringIO()):
            m0 = SpatialTemporalConv(ci, co, A, 9, s, 4, dropout_rate=0, residual=res)
        det_params(m0)
        for k_, v_ in m0.state_dict().items():
            rec['p.' + k_] = v_.numpy()
        for k_, v_ in m.named_parameters():
            rec['g.' + k_] = v_.grad.numpy()
        for k_, v_ in m.named_buffers():
            rec['post.' + k_] = v_.numpy()
        rec['meta'] = np.array([ci, co, T, V, K, s, int(res), int(training)], dtype=np.int64)
        np.savez_compressed(os.path.join(out, f'{name}.npz'), **rec)
        print('wrote', name, 'y', tuple(y.shape), 'max|y|', float(y.abs().max()))


def make_adjacency(out):
    from data import adjacency as adj
    rec = {}
    dist = np.random.RandomState(0).rand(25)
    dist_file = os.path.join(HERE, '_dist.npy')
    np.save(dist_file, dist)
    rec['distances'] = dist
    for strat in range(4):
        for d in (1, 2):
            A = adj.get_normalized_adjacency_matrices(strat, d, distance_file=dist_file)
            rec[f'A_s{strat}_d{d}'] = A.numpy()
            raw = adj.create_adjacency_matrices(strat, d, distance_file=dist_file)
            rec[f'raw_s{strat}_d{d}'] = torch.stack(raw).numpy()
    os.remove(dist_file)
    np.savez_compressed(os.path.join(out, 'adjacency.npz'), **rec)
    print('wrote adjacency')


def summarize(t: torch.Tensor, nsample=64):
    """Compact fingerprint of a big tensor: sum, abs-sum, max-abs and `nsample`
    strided samples."""
    f = t.detach().double().flatten()
    idx = torch.linspace(0, f.numel() - 1, min(nsample, f.numel())).long()
    return np.concatenate([[f.sum().item(), f.abs().sum().item(), f.abs().max().item()], f[idx].numpy()])


def make_model(out):
    import lightning_model as lm
    dist_file = os.path.join(HERE, '_dist.npy')
    np.save(dist_file, np.random.RandomState(0).rand(25))
    for name, part, res in (('m_spatial', 2, False), ('m_uni_res', 0, True)):
        args = ['--C_in', '3', '--nr_classes', '60', '--partitioning', str(part), '--distance_file', dist_file]
        hp = lm.build_argument_parser().parse_args(args)
        hp.residual = res
        with contextlib.redirect_stdout(io.StringIO()):
            model = lm.L_STGCN(hp)
        # replace the reference's O(1e4..1e5) A by a sane dense one for one case, keep it for the other
        det_params(model, a_mode='fill' if res else 'keep')
        model.train()
        torch.manual_seed(7)
        N, T = 4, 20
        x = torch.randn(N, T, 25, 3)
        labels = torch.tensor([3, 17, 59, 0])
        rec = {'x': x.numpy(), 'labels': labels.numpy()}
        # three train-mode passes calibrate BN running stats, the third is recorded
        for _ in range(2):
            model(x)
        model.zero_grad()
        logits = model(x)
        loss = torch.nn.functional.cross_entropy(logits, labels)
        loss.backward()
        rec['logits_train'] = logits.detach().numpy()
        rec['loss_train'] = np.array(loss.item())
        for k_, v_ in model.named_parameters():
            if v_.grad is None:
                continue
            if v_.numel() <= 4096:
                rec['g.' + k_] = v_.grad.numpy()
            rec['gs.' + k_] = summarize(v_.grad)
        for k_, v_ in model.named_buffers():
            if 'running' in k_:
                rec['post.' + k_] = v_.numpy()
        model.eval()
        with torch.no_grad():
            rec['logits_eval'] = model(x).numpy()
        rec['A'] = model.conv[0].spatialConv.A.detach().numpy()
        rec['n_params'] = np.array(sum(p.numel() for p in model.parameters()))
        rec['state_keys'] = np.array(list(model.state_dict().keys()))
        rec['state_shapes'] = np.array([str(tuple(v.shape)) for v in model.state_dict().values()])
        np.savez_compressed(os.path.join(out, f'{name}.npz'), **rec)
        print('wrote', name, 'loss', loss.item(), 'max|logit|', float(logits.abs().max()))
    os.remove(dist_file)


if __name__ == '__main__':
    ap = argparse.ArgumentParser()
    ap.add_argument('--out', default=HERE)
    a = ap.parse_args()
    install_stubs()
    torch.set_num_threads(8)
    make_adjacency(a.out)
    make_blocks(a.out)
    make_model(a.out)
