"""CPU-side checks of the boundary: the C-ABI library builds/loads and exports every
symbol include/stgcn_b200.h declares; host logic (descriptor sizing, validation, module
state_dict layout, error behaviour without a GPU)."""
import ctypes as C
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, 'include', 'stgcn_b200.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(stgcn_[a-z0-9_]+)\s*\(', src)))


def test_library_exports_every_declared_symbol():
    from stgcn_b200 import _lib
    lib = _lib.lib()
    names = header_functions()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), n
    assert b'sm_100a' in lib.stgcn_version()


def test_sizing_and_validation_without_gpu():
    from stgcn_b200 import _lib
    lib = _lib.lib()
    d = _lib.BlockDesc(n=2, c_in=64, c_out=128, t=20, v=25, k=3, gamma=9, stride=2, pad=4, dtype=_lib.F32, training=1,
                       residual=_lib.RES_CONV, flags=0, dropout_p=0.0, bn_eps=1e-5, bn_momentum=0.1, seed=0, offset=0)
    assert lib.stgcn_block_t_out(C.byref(d)) == 10
    sb = lib.stgcn_block_saved_bytes(C.byref(d))
    assert sb >= 2 * 20 * 25 * 128 * 4
    assert lib.stgcn_block_workspace_bytes(C.byref(d), 0) > 0
    assert lib.stgcn_block_workspace_bytes(C.byref(d), 1) > lib.stgcn_block_workspace_bytes(C.byref(d), 0) // 4
    d.dtype = _lib.BF16
    assert lib.stgcn_block_saved_bytes(C.byref(d)) < sb
    # errors: bad V, bad residual combination, null pointers
    d.v = 40
    assert lib.stgcn_block_workspace_bytes(C.byref(d), 0) == 0
    assert b'V=40' in lib.stgcn_last_error()
    d.v = 25; d.residual = _lib.RES_IDENTITY
    assert lib.stgcn_block_saved_bytes(C.byref(d)) == 0
    d.residual = _lib.RES_NONE
    P = _lib.BlockParams()
    rc = lib.stgcn_block_forward(C.byref(d), None, C.byref(P), None, None, None, 0, None)
    assert rc == -1 and b'null' in lib.stgcn_last_error()


def test_modules_keep_reference_state_dict_layout():
    from stgcn_b200 import SpatialTemporalConv
    A = torch.randn(3, 25, 25)
    m = SpatialTemporalConv(64, 128, A, 9, 2, 4, dropout_rate=0, residual=True)
    keys = list(m.state_dict().keys())
    assert keys == ['apply_residual.weight', 'apply_residual.bias', 'batch_n.weight', 'batch_n.bias',
                    'batch_n.running_mean', 'batch_n.running_var', 'batch_n.num_batches_tracked', 'spatialConv.A',
                    'spatialConv.W.weight', 'spatialConv.W.bias', 'temporalConv.weight', 'temporalConv.bias',
                    'batch_n_2.weight', 'batch_n_2.bias', 'batch_n_2.running_mean', 'batch_n_2.running_var',
                    'batch_n_2.num_batches_tracked']
    assert m.spatialConv.W.weight.shape == (384, 64, 1, 1) and m.temporalConv.weight.shape == (128, 128, 9, 1)
    assert m.spatialConv.K == 3 and m.spatialConv.V == 25 and m.dropout is None and m.residual is True
    assert m.spatialConv.A.requires_grad and m.spatialConv.A.is_leaf


def test_no_cpu_fallback():
    from stgcn_b200 import SpatialConv, SpatialTemporalConv
    A = torch.randn(2, 25, 25)
    m = SpatialTemporalConv(4, 8, A, 9, 1, 4, dropout_rate=0)
    with pytest.raises(RuntimeError, match='CUDA'):
        m(torch.randn(1, 4, 12, 25))
    with pytest.raises(RuntimeError, match='CUDA'):
        SpatialConv(4, 8, A)(torch.randn(1, 4, 12, 25))


def test_graph_builder_matches_reference_fixture():
    import numpy as np
    from _helpers import load_golden, rel_err
    from stgcn_b200 import graph
    g = load_golden('adjacency')
    for strat in range(4):
        for d in (1, 2):
            raw = graph.create_adjacency_matrices(strat, d, distances=g['distances'])
            assert np.array_equal(torch.stack(raw).numpy(), g[f'raw_s{strat}_d{d}']), (strat, d)
            A = graph.get_normalized_adjacency_matrices(strat, d, distances=g['distances'])
            assert rel_err(A, g[f'A_s{strat}_d{d}']) < 1e-6
    A18 = graph.get_normalized_adjacency_matrices(graph.Strategy.DISTANCE, 2, skeleton='openpose18')
    assert A18.shape == (3, 18, 18)


def test_training_tail_and_input_entry_points_validate_arguments_without_gpu():
    """Argument errors are reported by return code + stgcn_last_error before anything is launched."""
    from stgcn_b200 import _lib
    lib = _lib.lib()
    # Adam: the step state is mandatory (the step number lives on the device)
    rc = lib.stgcn_adam_step(0, None, None, None, None, None, 1e-3, 0.9, 0.999, 1e-8, 0.0, 0, None, None)
    assert rc != 0 and b'adam_step' in lib.stgcn_last_error()
    # cross entropy: shapes and pointers
    assert lib.stgcn_cross_entropy_forward(0, 60, None, None, -100, None, None, None) != 0
    assert lib.stgcn_cross_entropy_forward(4, 60, None, None, -100, None, None, None) != 0
    assert b'cross_entropy' in lib.stgcn_last_error()
    assert lib.stgcn_cross_entropy_backward(4, 60, None, None, None, None) != 0
    # collate: bad joint / channel count, null buffers; an empty batch is a no-op
    assert lib.stgcn_collate_pad_augment(_lib.F32, 2, 10, 0, 3, None, None, None, None, None, None) != 0
    assert lib.stgcn_collate_pad_augment(_lib.F32, 2, 10, 25, 3, None, None, None, None, None, None) != 0
    assert b'collate_pad_augment' in lib.stgcn_last_error()
    assert lib.stgcn_collate_pad_augment(_lib.F32, 0, 10, 25, 3, None, None, None, None, None, None) == 0


def test_build_digest_does_not_depend_on_where_the_repository_lives(tmp_path):
    """The GPU box unpacks the repository under another path; a digest over absolute paths made every process there rebuild the
    shipped library (8 torchrun ranks at once).  Same file names + contents => same digest, wherever they are."""
    from stgcn_b200 import _build
    a, b = tmp_path / 'a', tmp_path / 'b' / 'deeper'
    a.mkdir(); b.mkdir(parents=True)
    for d in (a, b):
        (d / 'k.cu').write_text('__global__ void k() {}\n')
        (d / 'h.cuh').write_text('#pragma once\n')
    da = _build._digest([str(a / 'k.cu'), str(a / 'h.cuh')])
    db = _build._digest([str(b / 'h.cuh'), str(b / 'k.cu')])
    assert da == db
    (b / 'k.cu').write_text('__global__ void k() { }\n')
    assert _build._digest([str(b / 'h.cuh'), str(b / 'k.cu')]) != da
    assert _build.is_fresh()          # the library in the tree matches its sources (built by __graft_entry__.build())


def test_ctypes_structs_have_the_library_s_sizes():
    from stgcn_b200 import _lib
    lib = _lib.lib()
    assert lib.stgcn_abi_sizeof(0) == C.sizeof(_lib.BlockDesc) == 13 * 4 + 3 * 4 + 2 * 8
    assert lib.stgcn_abi_sizeof(1) == C.sizeof(_lib.BlockParams) == 23 * 8
    assert lib.stgcn_abi_sizeof(2) == C.sizeof(_lib.BlockGrads) == 11 * 8
    assert lib.stgcn_abi_sizeof(7) == 0
