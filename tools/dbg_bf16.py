import sys, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [R, os.path.join(R, 'tests')]
import torch
from stgcn_b200 import SpatialTemporalConv
from oracle import stgcn_oracle as O
from _helpers import rel_err
ci, co, T, V, K, s, res = 128, 256, 12, 18, 2, 2, False
for training in (False, True):
    torch.manual_seed(1234 + ci + co + T)
    A = torch.randn(K, V, V) * 0.2
    m = SpatialTemporalConv(ci, co, A, 9, s, 4, dropout_rate=0, residual=res, compute_dtype=torch.bfloat16).cuda()
    with torch.no_grad():
        for bn in (m.batch_n, m.batch_n_2):
            bn.weight.uniform_(0.5, 1.5); bn.bias.uniform_(-0.3, 0.3); bn.running_mean.uniform_(-0.2, 0.2); bn.running_var.uniform_(0.6, 1.4)
    p = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    x = torch.randn(3, ci, T, V); To = (T - 1) // s + 1; gy = torch.randn(3, co, To, V)
    m.train(training)
    xg = x.cuda().requires_grad_(True); y = m(xg); (y.float() * gy.cuda()).sum().backward()
    leaves = {k: v.requires_grad_(True) for k, v in p.items() if v.dtype.is_floating_point and 'running' not in k}
    xr = x.clone().requires_grad_(True); yr = O.block_forward_bf16(xr, p, O.BlockCfg(ci, co, 9, s, 4, 0.0, res), training); (yr * gy).sum().backward()
    yg = y.float().cpu()
    flips = ((yg > 0) != (yr > 0)).sum().item()
    print('training', training, 'y %.2e dx %.2e' % (rel_err(yg, yr), rel_err(xg.grad.float(), xr.grad)), 'mask flips', flips, 'of', yr.numel())
    for k, v in leaves.items():
        g = dict(m.named_parameters())[k].grad.reshape(v.shape).cpu()
        dd = (g - v.grad).abs()
        print('  %-28s %.2e  (n>2e-2: %d of %d)' % (k, rel_err(g, v.grad), int((dd > 2e-2 * v.grad.abs().max()).sum()), v.numel()))
