import sys, os, struct
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from stgcn_b200.optim import Adam
from oracle import stgcn_oracle as O
g = torch.Generator().manual_seed(7)
base = torch.randn(5000, generator=g)
p = torch.nn.Parameter(base.clone().cuda())
opt = Adam([p], lr=0.01)
q, m, v = base.double(), torch.zeros(5000).double(), torch.zeros(5000).double()
for step in range(1, 5):
    gr = torch.randn(5000, generator=g)
    p.grad = gr.cuda()
    if len(sys.argv) > 1: torch.cuda.synchronize()
    opt.step()
    torch.cuda.synchronize()
    st = list(opt._dev_step.values())[0].cpu()
    raw = st.numpy().tobytes()
    print('step', step, 'dev', struct.unpack('qff', raw), 'expected', 0.01 / (1 - 0.9 ** step), 1 / (1 - 0.999 ** step) ** 0.5)
    q, m, v = O.adam_step(q, gr.double(), m, v, step, lr=0.01)
    print('   err p', float((p.detach().cpu().double() - q).abs().max()), 'err m', float((opt.state[p]['exp_avg'].cpu().double() - m).abs().max()),
          'err v', float((opt.state[p]['exp_avg_sq'].cpu().double() - v).abs().max()))
