"""Shared test helpers (golden loading, tolerance checks)."""
import glob
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + '.npz'), allow_pickle=False)
    return {k: z[k] for k in z.files}


def block_case_names():
    return sorted(os.path.basename(f)[:-4] for f in glob.glob(os.path.join(GOLDEN, 'b_*.npz')))


def rel_err(got, ref, floor=1e-6):
    """max|got-ref| / max(max|ref|, floor): the norm-scaled per-tensor measure of
    SURVEY.md section 7.3-5 (a per-element relative error is meaningless for
    structurally-zero gradients)."""
    got = torch.as_tensor(np.asarray(got)).double() if not torch.is_tensor(got) else got.detach().double().cpu()
    ref = torch.as_tensor(np.asarray(ref)).double() if not torch.is_tensor(ref) else ref.detach().double().cpu()
    assert got.shape == ref.shape, (got.shape, ref.shape)
    val = float((got - ref).abs().max() / max(float(ref.abs().max()), floor))
    rec = os.environ.get('STGCN_RECORD_ERRS')            # tolerance audit: every measured error with the test that measured it
    if rec:
        import inspect
        fr = inspect.stack()[1]
        with open(rec, 'a') as f:
            f.write('%s\t%s:%d\t%.6g\n' % (os.environ.get('PYTEST_CURRENT_TEST', '?').split(' ')[0], os.path.basename(fr.filename), fr.lineno, val))
    return val


def params_from_golden(g, dtype=torch.float32):
    return {k[2:]: torch.tensor(v).to(dtype) if v.dtype.kind == 'f' else torch.tensor(v)
            for k, v in g.items() if k.startswith('p.')}


def check_dx_kink_tolerant(got, ref, tol, max_frames=10, kink_tol=0.5):
    """Compare an input gradient (N,C,T,V) in the norm-scaled max norm, tolerating ISOLATED ReLU-kink flips.

    A ReLU pre-activation that lies within floating-point noise of zero (|z| ~ 1e-7..1e-6 relative; with
    ~2e5 activations per test tensor that happens in roughly one case out of six) takes a different branch
    in two correct implementations that sum in a different order.  The gradient of that single element then
    differs by O(1), which shows up as a ~1e-2 error confined to the 1..9 frames its temporal taps reach
    (measured on B200: one frame of one sample off by 1.2e-2, every other frame at 3e-7; through an identity
    skip connection the flipped element's own upstream gradient lands in dx unattenuated, up to ~0.2 of max|dx|).  Such frames are
    counted, must be few and mildly off, and every other frame must meet `tol`.  Returns the number of
    tolerated frames (0 in the normal case) so callers can relax the parameter-gradient bound accordingly."""
    got = torch.as_tensor(np.asarray(got)).double() if not torch.is_tensor(got) else got.detach().double().cpu()
    ref = torch.as_tensor(np.asarray(ref)).double() if not torch.is_tensor(ref) else ref.detach().double().cpu()
    scale = max(float(ref.abs().max()), 1e-6)
    per_frame = (got - ref).abs().amax(dim=(1, 3)) / scale          # (N,T)
    bad = per_frame > tol
    n_bad = int(bad.sum())
    if n_bad:
        assert n_bad <= max_frames, 'dx: %d frames above %g (max %.3g)' % (n_bad, tol, float(per_frame.max()))
        assert float(per_frame.max()) < kink_tol, 'dx error %.3g too large for a kink flip' % float(per_frame.max())
    return n_bad


def assert_argmax_consistent(got, ref, exact=False):
    """north_star: "identical argmax predictions".  fp32 (exact=True): plain equality.  bf16: a row whose two best reference
    scores are closer than the measured error of the scores themselves has no well-defined winner at that precision; every
    row with a decisive margin (top-1 minus top-2 of the reference > 2 x max|got-ref|) must agree.  Returns (rows checked,
    rows total) so callers can require that the check had teeth."""
    got = got.detach().double().cpu(); ref = ref.detach().double().cpu()
    if exact:
        assert np.array_equal(got.argmax(1).numpy(), ref.argmax(1).numpy())
        return ref.shape[0], ref.shape[0]
    err = float((got - ref).abs().max())
    top2 = ref.topk(2, dim=1).values
    decisive = (top2[:, 0] - top2[:, 1]) > 2 * err
    assert torch.equal(got.argmax(1)[decisive], ref.argmax(1)[decisive])
    return int(decisive.sum()), ref.shape[0]


def bf16_grad_bound(tol, emul, fp32):
    """Bound for |CUDA bf16 gradient - bf16-emulating oracle gradient| / max|.| of ONE tensor: the stated tolerance, or -- when
    the emulation itself is further than that from the fp32 oracle for this tensor (ReLU-branch flips and cancellation in the
    BatchNorm backward sums: 5e-2 .. 1e-1 per block, measured and printed by the tests) -- that measured distance.  Two bf16
    realisations of the same arithmetic cannot be expected to agree better than either agrees with fp32."""
    return max(tol, rel_err(emul, fp32))
