"""GPU parity of the training-step tail through the C ABI: fused cross entropy (loss + gradient) and the
multi-tensor Adam step, against the oracle restatements (pinned against torch in tests/test_oracle_train.py)."""
import pytest
import torch

from oracle import stgcn_oracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('n,classes', [(1, 60), (8, 60), (64, 60), (37, 400), (600, 7)])
def test_cross_entropy_forward_backward(n, classes):
    import stgcn_b200
    g = torch.Generator().manual_seed(n * 1000 + classes)
    logits = torch.randn(n, classes, generator=g) * 4
    labels = torch.randint(0, classes, (n,), generator=g)
    lg = logits.cuda().requires_grad_(True)
    loss = stgcn_b200.cross_entropy(lg, labels.cuda())
    (loss * 1.7).backward()
    ref = O.cross_entropy(logits.double(), labels)
    assert abs(float(loss) - float(ref)) < 1e-5 * max(1.0, abs(float(ref)))          # fp32 path: 1e-4 bound with margin
    gref = O.cross_entropy_grad(logits.double(), labels) * 1.7
    assert float((lg.grad.cpu().double() - gref).abs().max() / gref.abs().max()) < 1e-5


def test_cross_entropy_ignore_index_matches_torch():
    import stgcn_b200
    g = torch.Generator().manual_seed(11)
    logits = torch.randn(12, 60, generator=g)
    labels = torch.randint(0, 60, (12,), generator=g)
    labels[3] = -100; labels[7] = -100
    a = logits.clone().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(a, labels)
    ref.backward()
    b = logits.cuda().requires_grad_(True)
    out = stgcn_b200.cross_entropy(b, labels.cuda())
    out.backward()
    assert abs(float(out) - float(ref)) < 1e-5
    assert float((b.grad.cpu() - a.grad).abs().max()) < 1e-6


def test_cross_entropy_rejects_cpu_tensors():
    import stgcn_b200
    with pytest.raises(RuntimeError):
        stgcn_b200.cross_entropy(torch.randn(4, 6), torch.zeros(4, dtype=torch.int64))


@pytest.mark.parametrize('wd', [0.0, 0.01])
def test_adam_matches_oracle_over_steps(wd):
    """Ragged tensor sizes (1 element .. several chunks, an unaligned view) over 4 steps."""
    from stgcn_b200.optim import Adam
    g = torch.Generator().manual_seed(7)
    sizes = [1, 3, 64, 2047, 2048, 2049, 5000, 25 * 25 * 3, 384 * 64, 9 * 128 * 128]
    base = [torch.randn(s, generator=g) for s in sizes]
    params = [torch.nn.Parameter(b.clone().cuda()) for b in base]
    # one parameter living at a 4-byte-aligned (not 16-byte-aligned) address exercises the scalar path
    big = torch.randn(1001, generator=g)
    store = torch.zeros(1002).cuda(); store[1:] = big.cuda()
    odd = torch.nn.Parameter(store[1:]); base.append(big); params.append(odd)
    opt = Adam(params, lr=0.01, weight_decay=wd)
    ref = [(b.double(), torch.zeros_like(b).double(), torch.zeros_like(b).double()) for b in base]
    for step in range(1, 5):
        grads = [torch.randn(b.numel(), generator=g) for b in base]
        for p, gr in zip(params, grads):
            p.grad = gr.cuda()
        opt.step()
        ref = [O.adam_step(q, gr.double(), m, v, step, lr=0.01, weight_decay=wd) for (q, m, v), gr in zip(ref, grads)]
        for i, (p, (q, m, v)) in enumerate(zip(params, ref)):
            err = float((p.detach().cpu().double() - q).abs().max() / q.abs().max().clamp_min(1e-6))
            assert err < 1e-5, (wd, step, i, err)
            st = opt.state[p]
            assert float((st['exp_avg'].cpu().double() - m).abs().max()) < 1e-6
            assert float((st['exp_avg_sq'].cpu().double() - v).abs().max()) < 1e-6
    assert opt.steps_taken() == 4
    assert int(opt.state_dict()['state'][0]['step']) == 4


def test_adam_state_dict_loads_into_torch_adam():
    from stgcn_b200.optim import Adam
    p = torch.nn.Parameter(torch.randn(100).cuda())
    opt = Adam([p], lr=0.01)
    p.grad = torch.randn(100).cuda(); opt.step()
    q = torch.nn.Parameter(p.detach().clone())
    ref = torch.optim.Adam([q], lr=0.01)
    import copy
    ref.load_state_dict(copy.deepcopy(opt.state_dict()))      # torch aliases same-device tensors on load
    g2 = torch.randn(100).cuda()
    p.grad = g2.clone(); q.grad = g2.clone()
    opt.step(); ref.step()
    assert float((p - q).abs().max()) < 1e-6


def test_adam_rejects_cpu_parameters():
    from stgcn_b200.optim import Adam
    p = torch.nn.Parameter(torch.randn(10)); p.grad = torch.randn(10)
    with pytest.raises(RuntimeError):
        Adam([p]).step()


def _small_model(dtype):
    import argparse
    from stgcn_b200 import L_STGCN
    hp = argparse.Namespace(C_in=3, nr_classes=60, gamma=9, dropout_rate=0.0, residual=False, use_edge_importance=False,
                            max_mask_jitter=0.001, partitioning=2, d=1, distance_file=None, distances=None)
    torch.manual_seed(0)
    A = torch.randn(3, 25, 25, generator=torch.Generator().manual_seed(0)) * 0.15
    return L_STGCN(hp, compute_dtype=dtype, A=A).cuda().train()


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_captured_train_step_matches_eager_steps(dtype):
    """Three replays of the captured step == three eager steps (zero_grad, forward, cross entropy, backward, Adam) from the same
    start, and constructing the captured step leaves parameters, buffers and optimizer state untouched."""
    import stgcn_b200
    from stgcn_b200.optim import Adam
    g = torch.Generator().manual_seed(1)
    batches = [(torch.randn(4, 40, 25, 3, generator=g).cuda(), torch.randint(0, 60, (4,), generator=g).cuda()) for _ in range(3)]
    ma, mb = _small_model(dtype), _small_model(dtype)
    mb.load_state_dict(ma.state_dict())
    oa, ob = Adam(ma.parameters(), lr=1e-3), Adam(mb.parameters(), lr=1e-3)
    before = {k: v.clone() for k, v in mb.state_dict().items()}
    step = stgcn_b200.CapturedTrainStep(mb, ob, *batches[0])
    for k, v in mb.state_dict().items():
        assert torch.equal(v, before[k]), 'construction changed ' + k
    assert ob.steps_taken() == 0
    assert all(float(st['exp_avg'].abs().max()) == 0 for st in ob.state.values())
    la, lb = [], []
    for x, y in batches:
        oa.zero_grad(set_to_none=True)
        loss = stgcn_b200.cross_entropy(ma(x), y)
        loss.backward(); oa.step()
        la.append(float(loss))
        lb.append(float(step(x, y)))
    assert oa.steps_taken() == 3 and ob.steps_taken() == 3
    tol = 1e-4 if dtype == torch.float32 else 2e-2
    for a, b in zip(la, lb):
        assert abs(a - b) < tol * max(1.0, abs(a)), (la, lb)
    sa, sb = ma.state_dict(), mb.state_dict()
    assert int(sb['conv.0.batch_n.num_batches_tracked']) == int(sa['conv.0.batch_n.num_batches_tracked']) == 3
    for k in ('conv.0.batch_n.running_mean', 'conv.9.batch_n_2.running_var', 'fc_layer.bias'):
        d = float((sa[k].float() - sb[k].float()).abs().max() / sa[k].float().abs().max().clamp_min(1e-6))
        assert d < 50 * tol, (k, d)


def test_captured_forward_equals_eager_eval_forward():
    import stgcn_b200
    model = _small_model(torch.bfloat16)
    g = torch.Generator().manual_seed(2)
    xs = [torch.randn(4, 40, 25, 3, generator=g).cuda() for _ in range(3)]
    with torch.no_grad():
        model(xs[0])                                   # move the running statistics once
    model.eval()
    fwd = stgcn_b200.CapturedForward(model, xs[0])
    for x in xs:
        with torch.no_grad():
            ref = model(x).clone()
        got = fwd(x)
        assert float((got - ref).abs().max() / ref.abs().max()) < 1e-2        # same kernels; tiles may be walked the other way round
    with pytest.raises(RuntimeError):
        fwd(torch.randn(2, 40, 25, 3).cuda())


def test_captured_forward_serving_loop_matches_single_calls():
    """CapturedForward.stream: pinned host batches in, pinned host logits out, the copy of the next batch on a copy stream while the
    graph runs -- same numbers as one call per batch, in order, for an odd and an even number of batches and across two runs."""
    import stgcn_b200
    model = _small_model(torch.bfloat16)
    g = torch.Generator().manual_seed(4)
    hosts = [torch.randn(4, 40, 25, 3, generator=g).pin_memory() for _ in range(5)]
    with torch.no_grad():
        model(hosts[0].cuda())
    model.eval()
    fwd = stgcn_b200.CapturedForward(model, hosts[0].cuda())
    refs = [fwd(h.cuda()).float().cpu().clone() for h in hosts]
    for n in (5, 2, 1):
        got = [o.clone() for o in fwd.stream(hosts[:n])]
        assert len(got) == n
        for a, b in zip(got, refs):
            assert float((a - b).abs().max() / b.abs().max()) < 1e-2
    outs = [torch.empty(refs[0].shape).pin_memory() for _ in range(3)]
    res = list(fwd.stream(hosts[:3], out_host=outs))
    assert all(r is o for r, o in zip(res, outs))
    assert float((outs[2] - refs[2]).abs().max() / refs[2].abs().max()) < 1e-2
    with pytest.raises(RuntimeError):
        list(fwd.stream([torch.randn(4, 40, 25, 3)]))           # not pinned
