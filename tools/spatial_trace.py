"""Segment timers of CTA 0 of the tensor-memory spatial kernel (temporary instrumentation): python tools/spatial_trace.py"""
import contextlib, io, os, sys
os.environ['STGCN_SPATIAL_TRACE'] = '1'
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stgcn_b200 import SpatialTemporalConv
for ci, co, s, T in ((64, 64, 1, 300), (64, 128, 2, 300), (128, 128, 1, 150), (128, 256, 2, 150), (256, 256, 1, 75)):
    torch.manual_seed(0)
    with contextlib.redirect_stdout(io.StringIO()):
        m = SpatialTemporalConv(ci, co, torch.randn(3, 25, 25) * 0.2, 9, s, 4, dropout_rate=0, compute_dtype=torch.bfloat16).cuda().eval()
    x = torch.randn(128, T, 25, ci, device='cuda', dtype=torch.bfloat16).permute(0, 3, 1, 2)
    with torch.no_grad():
        for _ in range(2):
            m(x)
    torch.cuda.synchronize()
