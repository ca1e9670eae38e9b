"""Small bf16 cases for compute-sanitizer (tools/gpu_sanitize.sh): one default block (64,64,T=20), one strided block
(64,128,s2) and one residual block, forward + backward on the tcgen05 path with the side lane on, plus the static-scratch
entry point.  Prints SANITIZE_CASE_OK when every result is finite."""
import contextlib, io, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stgcn_b200 import SpatialTemporalConv

torch.manual_seed(0)
A = torch.randn(3, 25, 25) * 0.2
for ci, co, s, res in ((64, 64, 1, False), (64, 128, 2, False), (64, 64, 1, True)):
    with contextlib.redirect_stdout(io.StringIO()):
        m = SpatialTemporalConv(ci, co, A, 9, s, 4, dropout_rate=0, residual=res, compute_dtype=torch.bfloat16).cuda().train()
    x = torch.randn(2, ci, 20, 25).cuda().requires_grad_(True)
    y = m(x)
    y.float().square().mean().backward()
    torch.cuda.synchronize()
    assert torch.isfinite(y.float()).all() and torch.isfinite(x.grad).all()
    assert all(torch.isfinite(p.grad).all() for p in m.parameters() if p.grad is not None)
print('SANITIZE_CASE_OK')
