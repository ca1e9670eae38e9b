"""CPU: the oracle's restatement of the training-step tail (cross entropy, Adam) pinned against torch itself --
the reference calls F.cross_entropy (lightning_model.py:202) and optim.Adam (lightning_model.py:196-197)."""
import torch

from oracle import stgcn_oracle as O


def test_cross_entropy_and_its_gradient_match_torch():
    g = torch.Generator().manual_seed(3)
    logits = (torch.randn(17, 60, generator=g) * 3).requires_grad_(True)
    labels = torch.randint(0, 60, (17,), generator=g)
    ref = torch.nn.functional.cross_entropy(logits, labels)
    ref.backward()
    assert abs(float(O.cross_entropy(logits.detach(), labels)) - float(ref)) < 1e-6
    assert (O.cross_entropy_grad(logits.detach(), labels) - logits.grad).abs().max() < 1e-7


def test_adam_restatement_matches_torch_optim_adam():
    g = torch.Generator().manual_seed(5)
    for wd in (0.0, 0.01):
        p0 = torch.randn(1000, generator=g)
        p = torch.nn.Parameter(p0.clone())
        opt = torch.optim.Adam([p], lr=0.01, weight_decay=wd)
        q, m, v = p0.clone().double(), torch.zeros(1000).double(), torch.zeros(1000).double()
        for step in range(1, 6):
            grad = torch.randn(1000, generator=g)
            p.grad = grad.clone()
            opt.step()
            q, m, v = O.adam_step(q, grad.double(), m, v, step, lr=0.01, weight_decay=wd)
            assert (p.detach().double() - q).abs().max() < 2e-6, (wd, step)
