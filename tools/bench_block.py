#!/usr/bin/env python
"""BASELINE cfg2: single ST-GCN block microbench sweep on one B200 -- forward + backward of one block, this library against the
reference's stock-module op sequence (oracle/torch_stock.py: nn.BatchNorm2d -> nn.Conv2d 1x1 -> einsum -> nn.Conv2d 9x1 -> BN ->
ReLU, st_graphconv.py:85-109) on the same GPU, over C_in/C_out 64 -> 256, T matched to depth (300 / 150 / 75), V in {18, 25},
K in {1, 3}, residual off / on, N_eff = 128, training mode.

    python tools/bench_block.py [--out profiles/r2_block_sweep.json] [--iters 10] [--n 128] [--quick]

Per case: ms (CUDA events, L2 flushed between iterations, flush time subtracted), algorithmic TFLOP/s (3 x forward FLOPs of
SURVEY.md section 8(d)), GB/s on the fused-minimum bytes (read x + write y forward; read x, dy, write dx backward; + the
batch-statistics re-read of y), and the speed-up over the stock modules in eager fp32 and under bf16 autocast.
"""
import argparse
import contextlib
import io
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

SHAPES = [(64, 64, 1, 300), (64, 128, 2, 300), (128, 128, 1, 150), (128, 256, 2, 150), (256, 256, 1, 75)]


def flops_fwd(n, ci, co, s, t, v, k, residual):
    to = (t - 1) // s + 1
    f = 2.0 * n * t * v * ci * k * co + 2.0 * n * k * co * t * v * v + 2.0 * n * to * v * 9 * co * co
    if residual and not (ci == co and s == 1):
        f += 2.0 * n * to * v * ci * co
    return f


def min_bytes(n, ci, co, s, t, v, esize=2):
    to = (t - 1) // s + 1
    xin, yout = n * t * v * ci * esize, n * to * v * co * esize
    return (xin + yout) + (yout + xin + yout + xin)


def timed(fn, iters, flush):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        flush.zero_(); fn()
    e1.record(); torch.cuda.synchronize()
    t_all = e0.elapsed_time(e1)
    e0.record()
    for _ in range(iters):
        flush.zero_()
    e1.record(); torch.cuda.synchronize()
    return (t_all - e0.elapsed_time(e1)) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--out', default=os.path.join(ROOT, 'profiles', 'r2_block_sweep.json'))
    ap.add_argument('--iters', type=int, default=10)
    ap.add_argument('--n', type=int, default=128)
    ap.add_argument('--quick', action='store_true', help='V=25 K=3 only')
    args = ap.parse_args()
    from oracle.torch_stock import StockBlock
    from stgcn_b200 import SpatialTemporalConv
    dev = torch.device('cuda', 0)
    flush = torch.empty(160 << 20, dtype=torch.uint8, device=dev)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:
        pass
    # throwaway case first: the first tcgen05 launches of a process pay module loading and clock ramp (the first row of an
    # earlier sweep read 2.0 ms where its twins read 0.7)
    with contextlib.redirect_stdout(io.StringIO()):
        w = SpatialTemporalConv(64, 64, torch.randn(3, 25, 25) * 0.2, 9, 1, 4, dropout_rate=0, compute_dtype=torch.bfloat16).to(dev).train()
    xw = torch.randn(args.n, 64, 300, 25, device=dev, dtype=torch.bfloat16).permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2).requires_grad_(True)
    for _ in range(5):
        w(xw).float().mean().backward()
    torch.cuda.synchronize()
    del w, xw
    rows = []
    combos = [(25, 3)] if args.quick else [(25, 3), (25, 1), (18, 3), (18, 1)]
    for v, k in combos:
        for ci, co, s, t in SHAPES:
            for residual in (False, True):
                n = args.n
                torch.manual_seed(0)
                A = torch.randn(k, v, v) * 0.2
                to = (t - 1) // s + 1
                x = torch.randn(n, t, v, ci, device=dev)
                gy = torch.randn(n, to, v, co, device=dev)
                case = {'c_in': ci, 'c_out': co, 'stride': s, 'T': t, 'V': v, 'K': k, 'residual': residual, 'n_eff': n}
                fl = 3 * flops_fwd(n, ci, co, s, t, v, k, residual)
                with contextlib.redirect_stdout(io.StringIO()):
                    ours = SpatialTemporalConv(ci, co, A, 9, s, 4, dropout_rate=0, residual=residual, compute_dtype=torch.bfloat16).to(dev).train()
                xb = x.to(torch.bfloat16).permute(0, 3, 1, 2).requires_grad_(True)       # logical (N,C,T,V), channels-last memory
                gyb = gy.to(torch.bfloat16).permute(0, 3, 1, 2)

                def ours_step():
                    xb.grad = None
                    ours.zero_grad(set_to_none=True)
                    ours(xb).backward(gyb)
                ms = timed(ours_step, args.iters, flush)
                case['ours_bf16'] = {'ms': ms, 'tflops': fl / (ms * 1e-3) / 1e12, 'gbs_on_min_bytes': min_bytes(n, ci, co, s, t, v) / (ms * 1e-3) / 1e9}
                del ours
                for tag, autocast in (('torch_eager_fp32', False), ('torch_autocast_bf16', True)):
                    try:
                        stock = StockBlock(ci, co, A, 9, s, 4, 0.0, residual).to(dev).train()
                        xs = x.permute(0, 3, 1, 2).contiguous().requires_grad_(True)
                        gys = gy.permute(0, 3, 1, 2).contiguous()

                        def stock_step():
                            xs.grad = None
                            stock.zero_grad(set_to_none=True)
                            with torch.autocast('cuda', dtype=torch.bfloat16, enabled=autocast):
                                y = stock(xs)
                            y.backward(gys.to(y.dtype))
                        ms_s = timed(stock_step, max(3, args.iters // 2), flush)
                        case[tag] = {'ms': ms_s, 'speedup_of_ours': ms_s / ms}
                        del stock, xs, gys
                    except Exception as e:
                        case[tag] = {'error': str(e)[:120]}
                    torch.cuda.empty_cache()
                rows.append(case)
                print(json.dumps(case), flush=True)
    out = {'what': 'BASELINE cfg2 block microbench: fwd+bwd of one SpatialTemporalConv, training mode, N_eff=%d' % args.n,
           'peaks': {k: peaks.get(k) for k in ('hbm_gbs', 'bf16_tflops', 'bf16_tflops_sustained')}, 'torch': torch.__version__,
           'rows': rows}
    with open(args.out, 'w') as f:
        json.dump(out, f, indent=1)


if __name__ == '__main__':
    main()
