import sys, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [R, os.path.join(R, 'tests')]
import torch
from stgcn_b200 import SpatialTemporalConv
torch.manual_seed(5)
A = torch.rand(3, 25, 25) * 0.2
dt = torch.bfloat16
b1 = SpatialTemporalConv(64, 64, A, 9, 1, 4, dropout_rate=0, residual=False, compute_dtype=dt).cuda().train()
b2 = SpatialTemporalConv(64, 128, A, 9, 2, 4, dropout_rate=0, residual=False, compute_dtype=dt).cuda().train()
x = torch.randn(4, 64, 20, 25).cuda()
sd = {k: v.clone() for k, v in b2.state_dict().items()}
with torch.no_grad():
    for it in range(6):
        chain = it % 2 == 0
        if it in (2, 3):
            b1(x)          # shift the toggle parity by an odd number of participating launches? (4 per forward: even) -> use a lone block-0 style call
        b2.load_state_dict(sd)
        y1 = b1(x)
        if not chain:
            del y1._stgcn_colstats
        ref_mean = y1.float().mean((0, 2, 3)) * 0.1
        ref_var = y1.float().var((0, 2, 3), unbiased=True) * 0.1 + 0.9
        b2(y1)
        em = (b2.batch_n.running_mean - ref_mean).abs().max().item() / ref_mean.abs().max().item()
        ev = (b2.batch_n.running_var - ref_var).abs().max().item() / ref_var.abs().max().item()
        bad = ((b2.batch_n.running_mean - ref_mean).abs() > 1e-4 * ref_mean.abs().max()).nonzero().flatten().tolist()
        print('it', it, 'chain', chain, 'mean err %.2e var err %.2e' % (em, ev), 'bad channels', bad[:10])
