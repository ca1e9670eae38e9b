"""stgcn_spatial_forward_static: the allocation-free variant of SpatialConv.forward (scratch in a static device array inside the
library; include/stgcn_b200.h).  No counterpart in the reference.

Called through the C ABI (ctypes), checked against the existing entry point stgcn_spatial_forward (caller-provided workspace)
on two calls in a row with different inputs (the second call finds the first call's contents in the shared array), in fp32
(CUDA-core / split path) and bf16 with C_in % 64 == 0 (tcgen05 path); shapes beyond the static limit must be rejected with a
clear error.  Collected last on purpose (file name).
"""
import ctypes as C

import pytest
import torch

from _helpers import rel_err

pytestmark = pytest.mark.gpu


def _desc(L, n, ci, co, t, v, k, dtype):
    return L.BlockDesc(n=n, c_in=ci, c_out=co, t=t, v=v, k=k, gamma=1, stride=1, pad=0, dtype=L.dtype_code(dtype), training=0,
                       residual=0, flags=0, dropout_p=0.0, bn_eps=1e-5, bn_momentum=0.1, seed=0, offset=0)


@pytest.mark.parametrize('dtype,ci,co', [(torch.float32, 16, 24), (torch.bfloat16, 64, 128)], ids=['float32', 'bfloat16-tcgen05'])
def test_static_scratch_spatial_forward_matches_workspace_entry_point(dtype, ci, co):
    from stgcn_b200 import _lib as L
    lib = L.lib()
    n, t, v, k = 3, 20, 25, 3
    dev = torch.device('cuda', 0)
    g = torch.Generator().manual_seed(17)
    A = (torch.randn(k, v, v, generator=g) * 0.3).to(dev)
    W = (torch.randn(k * co, ci, generator=g) * 0.1).to(dev)
    Wb = (torch.randn(k * co, generator=g) * 0.1).to(dev)
    d = _desc(L, n, ci, co, t, v, k, dtype)
    need = lib.stgcn_spatial_workspace_bytes(C.byref(d), 0)
    assert 0 < need <= lib.stgcn_static_scratch_bytes() == 16 << 20
    ws = torch.empty(need, dtype=torch.uint8, device=dev)
    st = L.stream_ptr(dev)
    tol = 1e-4 if dtype == torch.float32 else 2e-2          # the tolerances of the fp32 / bf16 paths everywhere in tests/
    for call in range(3):                                    # call 0 warms up (lazy module load); calls 1 and 2 are the two in a row
        x = (torch.randn(n, t, v, ci, generator=g) * (1.0 + call)).to(dev).to(dtype)
        ref = torch.empty((n, t, v, co), dtype=dtype, device=dev)
        got = torch.full((n, t, v, co), float('nan'), dtype=dtype, device=dev)
        L.check(lib.stgcn_spatial_forward(C.byref(d), L.ptr(x), L.ptr(A), L.ptr(W), L.ptr(Wb), L.ptr(ref), L.ptr(ws), ws.numel(), st),
                'spatial_forward')
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(dev)[0]        # nothing of torch allocates between here and the second reading
        L.check(lib.stgcn_spatial_forward_static(C.byref(d), L.ptr(x), L.ptr(A), L.ptr(W), L.ptr(Wb), L.ptr(got), st),
                'spatial_forward_static')
        torch.cuda.synchronize()
        if call > 0:                                         # (the first call may load kernels lazily)
            assert torch.cuda.mem_get_info(dev)[0] == free_before, 'the static entry point allocated device memory'
        assert torch.isfinite(got.float()).all()
        assert rel_err(got.float(), ref.float()) < tol, call
        assert float(ref.float().abs().max()) > 0.1          # a real signal was compared


def test_static_scratch_rejects_shapes_beyond_the_limit():
    from stgcn_b200 import _lib as L
    lib = L.lib()
    d = _desc(L, 64, 64, 64, 300, 25, 3, torch.bfloat16)     # XA alone is 64*300*25*192*2 B = 184 MB
    assert lib.stgcn_spatial_workspace_bytes(C.byref(d), 0) > lib.stgcn_static_scratch_bytes()
    x = torch.zeros(8, device='cuda')                         # never dereferenced: the size check comes first
    rc = lib.stgcn_spatial_forward_static(C.byref(d), L.ptr(x), L.ptr(x), L.ptr(x), L.ptr(x), L.ptr(x), L.stream_ptr(x.device))
    assert rc == -4                                           # STGCN_ERR_WORKSPACE
    msg = lib.stgcn_last_error().decode()
    assert 'static array holds 16777216' in msg and 'stgcn_spatial_forward' in msg
    torch.cuda.synchronize()
