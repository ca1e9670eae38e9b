"""CPU: the oracle's restatement of the input pipeline (loop-padding collate, augmentation) against the golden fixture made by
the unmodified reference (tests/golden/make_golden_input.py), and the host-side sampler of the augmentation matrix."""
import numpy as np

from _helpers import load_golden
from oracle import stgcn_oracle as O


def _batches(g):
    lens = [int(t) for t in g['collate_lens']]
    b1 = [(g['collate_in%d' % i], np.array([i % 6])) for i in range(len(lens))]
    b2 = [(g['collate2_in%d' % i], np.array([1, 1])) for i in range(2)]
    return b1, b2


def test_loopy_pad_collate_matches_reference():
    g = load_golden('input_pipeline')
    b1, b2 = _batches(g)
    x, l = O.loopy_pad_collate(b1)
    assert np.array_equal(x, g['collate_out']) and np.array_equal(l, g['collate_labels'])
    x2, l2 = O.loopy_pad_collate(b2)
    assert np.array_equal(x2, g['collate2_out']) and np.array_equal(l2, g['collate2_labels'])


def test_augmentation_matrix_and_transform_match_reference():
    from stgcn_b200.pipeline import sample_augmentation
    g = load_golden('input_pipeline')
    kinds = set()
    for seed in range(12):
        T = O.augmentation_matrix(np.random.RandomState(seed))
        assert np.allclose(T[:2, :2], g['aug_T2_%d' % seed], atol=1e-12), seed
        assert np.allclose(O.augment_apply(g['aug_in'], T), g['aug_out_%d' % seed], atol=1e-12), seed
        # the product's host-side sampler draws the same 2x2 block from the same generator state
        assert np.allclose(sample_augmentation(np.random.RandomState(seed)), T[:2, :2], atol=1e-12), seed
        kinds.add(tuple(np.round(T[:2, :2], 6).reshape(-1)))
    assert len(kinds) >= 6          # the seeds cover rotations, scalings, flips and their combinations


def test_device_collate_and_captured_step_refuse_cpu():
    """No CPU fallback anywhere on the product path: the host-side wrappers raise before touching the library."""
    import pytest
    import torch
    from stgcn_b200.pipeline import DeviceCollate
    from stgcn_b200.training import CapturedTrainStep
    with pytest.raises(RuntimeError):
        DeviceCollate('cpu')([(np.zeros((1, 4, 25, 2), np.float32), np.array([0]))])
    lin = torch.nn.Linear(4, 2).train()
    with pytest.raises(RuntimeError):
        CapturedTrainStep(lin, torch.optim.SGD(lin.parameters(), lr=0.1), torch.zeros(2, 4), torch.zeros(2, dtype=torch.int64))
