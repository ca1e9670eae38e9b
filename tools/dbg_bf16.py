import sys, os
sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.abspath(__file__)))]
sys.path.insert(0, os.path.join(sys.path[0], 'tests'))
import torch
from stgcn_b200 import SpatialTemporalConv, _lib
from _helpers import rel_err

def run(m, x, gy, training):
    m.train(training); m.zero_grad()
    xg = x.cuda().requires_grad_(True)
    y = m(xg)
    (y.float() * gy.cuda()).sum().backward()
    return y.detach().float().cpu(), xg.grad.float().cpu(), {k: v.grad.detach().cpu().clone() for k, v in m.named_parameters()}

for (ci, co, T, V, K, s, res) in [(64, 64, 20, 25, 3, 1, False), (64, 128, 20, 25, 3, 2, False), (64, 64, 14, 18, 1, 1, True)]:
    for training in (True, False):
        torch.manual_seed(1)
        A = torch.randn(K, V, V) * 0.2
        ref = SpatialTemporalConv(ci, co, A, 9, s, 4, dropout_rate=0, residual=res).cuda()
        x = torch.randn(3, ci, T, V); To = (T - 1) // s + 1; gy = torch.randn(3, co, To, V)
        sd = {k: v.clone() for k, v in ref.state_dict().items()}
        y0, dx0, g0 = run(ref, x, gy, training)
        for name, flags in (('bf16-simt', 1), ('bf16-umma', 0)):
            m = SpatialTemporalConv(ci, co, A, 9, s, 4, dropout_rate=0, residual=res, compute_dtype=torch.bfloat16).cuda()
            m.load_state_dict(sd); m.flags = flags; m.spatialConv.flags = flags
            y, dx, g = run(m, x, gy, training)
            print((ci, co, s, res, training), name, 'y %.4f dx %.4f' % (rel_err(y, y0), rel_err(dx, dx0)),
                  ' '.join('%s %.4f' % (k.split('.')[-2][:6] + '.' + k.split('.')[-1][:1], rel_err(g[k], g0[k])) for k in g0 if g0[k].abs().max() > 1e-6))
