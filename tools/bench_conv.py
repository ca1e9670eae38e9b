#!/usr/bin/env python
"""Micro-benchmark of the tcgen05 implicit-GEMM engine at the block shapes of the benchmark model (B200 only)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from stgcn_b200 import _lib as L

def run(nb, t, v, c, n_out, taps, stride, pad, iters=20):
    t_out = (t + 2 * pad - taps) // stride + 1
    x = torch.randn(nb, t, v, c, device='cuda').bfloat16()
    w = (torch.randn(taps, n_out, c, device='cuda') / (c * taps) ** 0.5).bfloat16()
    bias = None if os.environ.get('NOBIAS') else torch.randn(n_out, device='cuda')
    out = torch.empty(nb, t_out, v, n_out, device='cuda', dtype=torch.bfloat16)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for i in range(iters + 3):
        flush.zero_()
        if i >= 3: ev[i - 3][0].record()
        L.check(L.lib().stgcn_umma_conv_gemm(nb, t, t_out, v, c, n_out, taps, stride, pad, L.ptr(x), L.ptr(w), L.ptr(bias) if bias is not None else None,
                                             L.ptr(out), L.stream_ptr()), 'umma_conv_gemm')
        if i >= 3: ev[i - 3][1].record()
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in ev)[iters // 2]
    fl = 2.0 * nb * t_out * v * c * n_out * taps
    print('nb=%d t=%d c=%d n=%d taps=%d s=%d : %.1f us  %.0f TFLOP/s' % (nb, t, c, n_out, taps, stride, ms * 1e3, fl / ms / 1e9))

for shape in [(64, 300, 25, 64, 64, 9, 1, 4), (64, 300, 25, 64, 128, 9, 2, 4), (64, 150, 25, 128, 128, 9, 1, 4),
              (64, 150, 25, 128, 256, 9, 2, 4), (64, 75, 25, 256, 256, 9, 1, 4), (64, 300, 25, 192, 64, 1, 1, 0),
              (64, 75, 25, 768, 256, 1, 1, 0),
              (64, 300, 25, 64, 192, 1, 1, 0), (64, 150, 25, 128, 384, 1, 1, 0), (64, 75, 25, 256, 768, 1, 1, 0)]:
    if len(sys.argv) > 1 and str(shape[3]) + 'x' + str(shape[4]) not in sys.argv[1:]: continue
    run(*shape, iters=int(os.environ.get('ITERS', '20')))
