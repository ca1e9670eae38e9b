"""Unit tests of the tcgen05 implicit-GEMM engine and descriptor probes (B200 only)."""
import ctypes as C
import json
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def ref_conv(x, w, bias, stride, pad, t_out):
    # x (nb,t,v,c) bf16 ; w (taps,n,c) bf16 -> (nb,t_out,v,n) fp32, computed in fp32 from the bf16 values
    nb, t, v, c = x.shape
    taps, n, _ = w.shape
    xf = torch.nn.functional.pad(x.float(), (0, 0, 0, 0, pad, pad + stride))
    out = torch.zeros(nb, t_out, v, n, device=x.device)
    for tap in range(taps):
        sl = xf[:, tap: tap + stride * (t_out - 1) + 1: stride]
        out += sl @ w[tap].float().t()
    if bias is not None:
        out += bias
    return out


CASES = [
    # nb, t, v, c, n_out, taps, stride, pad
    (2, 10, 25, 64, 64, 1, 1, 0),
    (2, 23, 25, 64, 64, 9, 1, 4),
    (3, 20, 25, 128, 128, 9, 2, 4),
    (2, 14, 18, 256, 256, 9, 1, 4),
    (2, 12, 25, 192, 64, 1, 1, 0),
    (2, 12, 25, 64, 192, 1, 1, 0),
    (2, 12, 18, 128, 384, 1, 1, 0),
    (1, 12, 25, 256, 768, 1, 1, 0),
    (2, 20, 25, 64, 128, 1, 2, 0),
    (40, 30, 25, 64, 64, 9, 1, 4),       # > 148 tiles: persistent loop, accumulator double-buffer
]


@pytest.mark.parametrize('case', CASES)
def test_umma_conv_gemm(case):
    from stgcn_b200 import _lib as L
    nb, t, v, c, n_out, taps, stride, pad = case
    torch.manual_seed(sum(case))
    t_out = (t + 2 * pad - taps) // stride + 1
    x = torch.randn(nb, t, v, c, device='cuda').bfloat16()
    w = (torch.randn(taps, n_out, c, device='cuda') / (c * taps) ** 0.5).bfloat16()
    bias = torch.randn(n_out, device='cuda')
    out = torch.full((nb, t_out, v, n_out), float('nan'), device='cuda', dtype=torch.bfloat16)
    L.check(L.lib().stgcn_umma_conv_gemm(nb, t, t_out, v, c, n_out, taps, stride, pad, L.ptr(x), L.ptr(w), L.ptr(bias),
                                         L.ptr(out), L.stream_ptr()), 'umma_conv_gemm')
    torch.cuda.synchronize()
    ref = ref_conv(x, w, bias, stride, pad, t_out)
    err = float((out.float() - ref).abs().max() / ref.abs().max())
    assert torch.isfinite(out.float()).all()
    assert err < 8e-3, err          # bf16 output rounding only (2^-9 relative)


def _probe(mode, row_off, base_off, a_rows):
    from stgcn_b200 import _lib as L
    torch.manual_seed(100 + row_off)
    cols = 64 if mode == 0 else 128
    a = torch.randn(a_rows, cols, device='cuda').bfloat16()
    b = torch.randn(64, 64, device='cuda').bfloat16()
    d = torch.zeros(128, 64, device='cuda')
    L.check(L.lib().stgcn_umma_probe(mode, row_off, base_off, a_rows, L.ptr(a), L.ptr(b), L.ptr(d), L.stream_ptr()), 'probe')
    torch.cuda.synchronize()
    if mode == 0:
        ref = a[row_off:row_off + 128].float() @ b.float().t()
    else:
        ref = a[row_off:row_off + 64].float().t() @ b.float()
    return float((d - ref).abs().max() / ref.abs().max())


def test_probe_baselines():
    assert _probe(0, 0, -1, 128) < 1e-5        # plain K-major, aligned
    assert _probe(0, 8, -1, 256) < 1e-5        # shifted by one whole swizzle atom
    assert _probe(1, 0, -1, 64) < 1e-5         # MN-major A and B


def test_probe_unaligned_descriptor_offsets_report():
    """Informational: records whether sub-atom row offsets work (decides the slab-reuse design)."""
    res = {}
    for mode, rows in ((0, 256), (1, 256)):
        for off in (1, 3, 18, 25, 50, 100):
            for bo in (-1, off % 8):
                try:
                    res[f'mode{mode}_off{off}_bo{bo}'] = _probe(mode, off, bo, rows)
                except RuntimeError as e:      # pragma: no cover
                    res[f'mode{mode}_off{off}_bo{bo}'] = str(e)
    os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
    with open(os.path.join(ROOT, 'gpurun_out', 'umma_probe.json'), 'w') as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))
