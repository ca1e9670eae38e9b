"""Run-to-run variation of the bf16 training gradients of the drop-in model (same weights, same batch, two runs)."""
import argparse, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stgcn_b200 import L_STGCN, cross_entropy

hp = argparse.Namespace(C_in=3, nr_classes=60, gamma=9, dropout_rate=0.0, residual=False, use_edge_importance=False,
                        max_mask_jitter=0.01, partitioning=2, d=1, distance_file=None, distances=None)
torch.manual_seed(0)
A = torch.randn(3, 25, 25, generator=torch.Generator().manual_seed(0)) * 0.2
for dt in (torch.bfloat16, torch.float32):
    model = L_STGCN(hp, compute_dtype=dt, A=A).cuda().train()
    state = {k: v.detach().clone() for k, v in model.state_dict().items()}
    g = torch.Generator().manual_seed(5)
    x = torch.randn(8, 60, 25, 3, generator=g).cuda(); y = torch.randint(0, 60, (8,), generator=g).cuda()
    runs = []
    for r in range(3):
        model.load_state_dict(state); model.zero_grad(set_to_none=True)
        lg = model(x); cross_entropy(lg, y).backward()
        runs.append((lg.detach().float().clone(), {k: p.grad.detach().clone() for k, p in model.named_parameters() if p.grad is not None}))
        del lg
    for a, b in ((0, 1), (0, 2)):
        la, ga = runs[a]; lb, gb = runs[b]
        worst = max(((float((ga[k] - gb[k]).abs().max() / gb[k].abs().max().clamp_min(1e-12)), k) for k in ga if not k.endswith('temporalConv.bias')))
        fa = torch.cat([v.flatten() for v in ga.values()]); fb = torch.cat([v.flatten() for v in gb.values()])
        print(dt, 'pingpong' if not os.environ.get('STGCN_NO_PINGPONG') else 'no-pingpong', 'runs', a, b,
              'logits maxdiff %.3g' % float((la - lb).abs().max() / lb.abs().max()),
              'grad worst tensor %.3g (%s)' % worst, 'global max-norm %.3g' % float((fa - fb).abs().max() / fb.abs().max()),
              'global l2 %.3g' % float((fa - fb).norm() / fb.norm()))
