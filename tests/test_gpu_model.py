"""Model-level GPU parity: the drop-in L_STGCN against the oracle and the reference's golden logits."""
import argparse

import numpy as np
import pytest
import torch

from _helpers import load_golden, rel_err
from detfill import det_params
from oracle import stgcn_oracle as O

pytestmark = pytest.mark.gpu


def hparams(residual, partitioning):
    return argparse.Namespace(C_in=3, nr_classes=60, gamma=9, dropout_rate=0.0, residual=residual,
                              use_edge_importance=False, max_mask_jitter=0.001, partitioning=partitioning, d=1,
                              distance_file=None, distances=np.random.RandomState(0).rand(25))


@pytest.mark.parametrize('name,residual', [('m_spatial', False), ('m_uni_res', True)])
def test_fp32_model_matches_reference_golden(name, residual):
    from stgcn_b200 import L_STGCN
    g = load_golden(name)
    model = L_STGCN(hparams(residual, 0 if residual else 2))
    det_params(model, a_mode='fill' if residual else 'keep')
    model = model.cuda().train()
    x = torch.tensor(g['x']).cuda(); labels = torch.tensor(g['labels']).cuda()
    for _ in range(2):
        model(x)
    model.zero_grad()
    logits = model(x)
    loss = torch.nn.functional.cross_entropy(logits, labels)
    loss.backward()
    assert rel_err(logits, g['logits_train']) < 2e-3
    assert abs(loss.item() - float(g['loss_train'])) < 1e-4
    assert np.array_equal(logits.argmax(1).cpu().numpy(), g['logits_train'].argmax(1))
    sd = model.state_dict()
    for k in g:
        if k.startswith('post.'):
            assert rel_err(sd[k[5:]], g[k]) < 2e-3, k
    params = dict(model.named_parameters())
    for k in g:
        if k.startswith('g.') and float(np.abs(g[k]).max()) > 1e-7:
            # the reference's own fp32 gradients of this deep tiny-batch net are only reproducible to
            # ~8e-2 against fp64 (tests/test_oracle_golden.py); block-level tests carry the tight bound
            assert rel_err(params[k[2:]].grad.reshape(g[k].shape), g[k]) < 0.12, k
    model.eval()
    with torch.no_grad():
        ev = model(x)
    assert rel_err(ev, g['logits_eval']) < 2e-3
    assert np.array_equal(ev.argmax(1).cpu().numpy(), g['logits_eval'].argmax(1))


@pytest.mark.parametrize('dtype,tol', [(torch.float32, 1e-4), (torch.bfloat16, 2e-2)])
def test_model_matches_oracle_sane_A(dtype, tol):
    """Random dense O(1) adjacency, default (non-residual) stack, N=4 T=40: logits / loss / argmax."""
    from stgcn_b200 import L_STGCN
    torch.manual_seed(0)
    A = torch.randn(3, 25, 25) * 0.15
    model = L_STGCN(hparams(False, 2), compute_dtype=dtype, A=A)
    state = {k: v.detach().clone() for k, v in model.state_dict().items()}
    model = model.cuda().train()
    x = torch.randn(4, 40, 25, 3); labels = torch.tensor([1, 5, 30, 59])
    cfgs = O.stack_cfgs(3, 9, 0.0, False)
    ref = O.model_forward(x, state, cfgs, True)
    got = model(x.cuda())
    assert rel_err(got, ref) < tol            # measured on B200: fp32 ~1e-6, bf16 4.8e-3
    if dtype == torch.float32:
        assert np.array_equal(got.argmax(1).cpu().numpy(), ref.argmax(1).numpy())
    # eval after the running stats moved once
    model.eval()
    with torch.no_grad():
        got_e = model(x.cuda())
        ref_e = O.model_forward(x, state, cfgs, False)
    assert rel_err(got_e, ref_e) < tol        # measured: bf16 6.4e-4


def test_state_dict_round_trip_and_adam_step():
    from stgcn_b200 import L_STGCN
    torch.manual_seed(0)
    m1 = L_STGCN(hparams(False, 0)).cuda()
    m2 = L_STGCN(hparams(False, 0)).cuda()
    m2.load_state_dict(m1.state_dict())
    x = torch.randn(2, 24, 25, 3).cuda()
    opt = torch.optim.Adam(m1.parameters(), lr=0.01)
    loss = torch.nn.functional.cross_entropy(m1(x), torch.tensor([1, 2]).cuda())
    loss.backward()
    assert all(p.grad is not None for n, p in m1.named_parameters())
    opt.step()
    assert torch.isfinite(torch.stack([p.abs().max() for p in m1.parameters()])).all()
