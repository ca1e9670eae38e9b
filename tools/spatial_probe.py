"""Device time of the fused spatial kernel (umma_spatial.cu) in the eval-mode block forward, per shape, from the library's timing
hooks:  python tools/spatial_probe.py [n_eff] [iters]   (STGCN_SPATIAL_TS=0|1 selects the kernel variant)"""
import contextlib, io, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stgcn_b200 import SpatialTemporalConv, _lib
n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 10
v, k = 25, 3
flush = torch.empty(160 << 20, dtype=torch.uint8, device='cuda')
for ci, co, s, T in ((64, 64, 1, 300), (64, 128, 2, 300), (128, 128, 1, 150), (128, 256, 2, 150), (256, 256, 1, 75)):
    torch.manual_seed(0)
    with contextlib.redirect_stdout(io.StringIO()):
        m = SpatialTemporalConv(ci, co, torch.randn(k, v, v) * 0.2, 9, s, 4, dropout_rate=0, compute_dtype=torch.bfloat16).cuda().eval()
    x = torch.randn(n, T, v, ci, device='cuda', dtype=torch.bfloat16).permute(0, 3, 1, 2)
    with torch.no_grad():
        for _ in range(3):
            m(x)
        torch.cuda.synchronize()
        _lib.timing_reset(); _lib.timing_enable(True)
        for _ in range(iters):
            flush.zero_(); m(x)
        torch.cuda.synchronize(); _lib.timing_enable(False)
    prof = _lib.timing_read()
    rows = n * T * v
    line = '%3d->%3d s%d T%3d n=%d:' % (ci, co, s, T, n)
    for kname, (ms, work, cnt) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
        if cnt:
            line += '  %s %.1f us' % (kname, ms / cnt * 1e3)
    sp = prof.get('umma_spatial')
    if sp and sp[2]:
        us = sp[0] / sp[2] * 1e3
        line += '  | spatial: %.0f cycles/tile @1.9GHz, %.2f TB/s on x+S bytes' % (us * 1900 / (rows / 125 / 148), rows * (ci + co) * 2 / us / 1e6)
    print(line, flush=True)
