import sys, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [R, os.path.join(R, 'tests')]
import torch
from stgcn_b200 import SpatialTemporalConv
from oracle import stgcn_oracle as O
from _helpers import rel_err
cases = [(64, 128, 20, 25, 3, 2, True)]
for (ci, co, T, V, K, s, res) in cases:
    torch.manual_seed(1234 + ci + co + T)
    A = torch.randn(K, V, V) * 0.2
    m = SpatialTemporalConv(ci, co, A, 9, s, 4, dropout_rate=0, residual=res).cuda().eval()
    with torch.no_grad():
        for bn in (m.batch_n, m.batch_n_2):
            bn.weight.uniform_(0.5, 1.5); bn.bias.uniform_(-0.3, 0.3); bn.running_mean.uniform_(-0.2, 0.2); bn.running_var.uniform_(0.6, 1.4)
    p = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    x = torch.randn(3, ci, T, V); To = (T - 1) // s + 1; gy = torch.randn(3, co, To, V)
    xg = x.cuda().requires_grad_(True); y = m(xg); (y * gy.cuda()).sum().backward()
    leaves = {k: v.requires_grad_(True) for k, v in p.items() if v.dtype.is_floating_point and 'running' not in k}
    xr = x.clone().requires_grad_(True); yr = O.block_forward(xr, p, O.BlockCfg(ci, co, 9, s, 4, 0.0, res), False); (yr * gy).sum().backward()
    d = (xg.grad.cpu() - xr.grad).abs()
    idx = (d > 1e-3 * xr.grad.abs().max()).nonzero()
    print((ci, co, T, s, res), 'y %.1e dx %.1e' % (rel_err(y, yr), rel_err(xg.grad, xr.grad)), 'bad', len(idx), 'frames', sorted(set(idx[:, 2].tolist()))[:30], 'n', sorted(set(idx[:,0].tolist())),
          ' '.join('%s %.1e' % (k[-12:], rel_err(dict(m.named_parameters())[k].grad.reshape(v.shape), v.grad)) for k, v in leaves.items()))
    torch.set_printoptions(precision=1, linewidth=250, sci_mode=True)
    print((d.amax(dim=(1, 3)) / xr.grad.abs().max()))
