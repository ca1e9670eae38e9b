"""Per-kernel-class device time of one block's forward+backward (library timing hooks): python tools/block_profile.py ci co s T res"""
import contextlib, io, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stgcn_b200 import SpatialTemporalConv, _lib
ci, co, s, T, res = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
n, v, k = (int(sys.argv[6]) if len(sys.argv) > 6 else 128), 25, 3
torch.manual_seed(0)
with contextlib.redirect_stdout(io.StringIO()):
    m = SpatialTemporalConv(ci, co, torch.randn(k, v, v) * 0.2, 9, s, 4, dropout_rate=0, residual=bool(res), compute_dtype=torch.bfloat16).cuda().train()
x = torch.randn(n, T, v, ci, device='cuda', dtype=torch.bfloat16).permute(0, 3, 1, 2).requires_grad_(True)
to = (T - 1) // s + 1
gy = torch.randn(n, to, v, co, device='cuda', dtype=torch.bfloat16).permute(0, 3, 1, 2)
def step():
    x.grad = None; m.zero_grad(set_to_none=True); m(x).backward(gy)
for _ in range(3): step()
torch.cuda.synchronize()
_lib.timing_reset(); _lib.timing_enable(True)
l0 = _lib.launch_count()
for _ in range(5): step()
torch.cuda.synchronize(); _lib.timing_enable(False)
prof = _lib.timing_read()
print('block %d->%d s%d T%d res=%d n=%d: launches/step %d' % (ci, co, s, T, res, n, (_lib.launch_count() - l0) // 5))
for kname, (ms, work, cnt) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
    print('  %-16s %7.3f ms/step  %3d launches/step  %7.2f us per sample' % (kname, ms / 5, cnt // 5, ms / 5 * 1e3 / n))
