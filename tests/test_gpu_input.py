"""GPU parity of the device input pipeline (stgcn_collate_pad_augment through DeviceCollate) against the oracle restatement of
the reference's collate / augmentation (pinned by tests/test_oracle_input.py) on the reference-made golden batch."""
import numpy as np
import pytest
import torch

from _helpers import load_golden
from oracle import stgcn_oracle as O

pytestmark = pytest.mark.gpu


def _batch(g):
    return [(g['collate_in%d' % i], np.array([i % 6])) for i in range(len(g['collate_lens']))]


def test_collate_fp32_is_bit_exact_with_reference_output():
    from stgcn_b200.pipeline import DeviceCollate
    g = load_golden('input_pipeline')
    x, labels = DeviceCollate('cuda:0')(_batch(g))
    assert x.shape == g['collate_out'].shape
    assert np.array_equal(x.cpu().numpy(), g['collate_out'])            # pure data movement: bit exact
    assert np.array_equal(labels.cpu().numpy(), g['collate_labels'])
    b2 = [(g['collate2_in%d' % i], np.array([1, 1])) for i in range(2)]
    x2, l2 = DeviceCollate('cuda:0')(b2)
    assert np.array_equal(x2.cpu().numpy(), g['collate2_out']) and np.array_equal(l2.cpu().numpy(), g['collate2_labels'])


@pytest.mark.parametrize('dtype,tol', [(torch.float32, 1e-6), (torch.bfloat16, 2e-2)])
def test_collate_with_augmentation_matches_oracle(dtype, tol):
    from stgcn_b200.pipeline import DeviceCollate, sample_augmentation
    g = load_golden('input_pipeline')
    batch = _batch(g)
    mats = [None if i == 2 else sample_augmentation(np.random.RandomState(i)) for i in range(len(batch))]
    x, _ = DeviceCollate('cuda:0', dtype=dtype)(batch, xforms=mats)
    ref_batch = [(x_ if m is None else O.augment_apply(x_, np.pad(m, ((0, 1), (0, 1)))), l) for (x_, l), m in zip(batch, mats)]
    ref, _ = O.loopy_pad_collate(ref_batch)
    err = np.abs(x.float().cpu().numpy().astype(np.float64) - ref).max() / np.abs(ref).max()
    assert err < tol, err


def test_collate_full_size_round_trip():
    """NTU-sized batch (64 sequences, up to 300 frames, 25 joints, 3 channels): every output frame equals the source frame it
    wraps to, and the confidence channel passes through the augmentation untouched."""
    from stgcn_b200.pipeline import DeviceCollate
    rng = np.random.RandomState(0)
    lens = rng.randint(40, 301, size=64); lens[5] = 300
    batch = [(rng.randn(1, int(t), 25, 3).astype(np.float32), np.array([0])) for t in lens]
    flip = np.array([[-1.0, 0.0], [0.0, 1.0]])
    x, _ = DeviceCollate('cuda:0')(batch, xforms=[flip] * 64)
    x = x.cpu().numpy()
    assert x.shape == (64, 300, 25, 3)
    for i in (0, 5, 17, 63):
        src = batch[i][0][0]
        idx = np.arange(300) % len(src)
        assert np.array_equal(x[i, :, :, 0], -src[idx, :, 0]) and np.array_equal(x[i, :, :, 1:], src[idx, :, 1:])


def test_collate_rejects_mixed_joint_counts():
    from stgcn_b200.pipeline import DeviceCollate
    with pytest.raises(RuntimeError):
        DeviceCollate('cuda:0')([(np.zeros((1, 4, 25, 2), np.float32), np.array([0])), (np.zeros((1, 4, 18, 2), np.float32), np.array([0]))])
