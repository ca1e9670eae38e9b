#!/usr/bin/env python
"""Golden fixture for the input pipeline, produced by the UNMODIFIED reference (build container only):

    python tests/golden/make_golden_input.py        ->  tests/golden/input_pipeline.npz

Runs `data.util.loopy_pad_collate_fn` (src/data/util.py:33-47) on a ragged batch and `data.augmentation.augment_data`
(src/data/augmentation.py:8-69) under fixed numpy seeds.  The 3x3 matrix the reference draws is recovered from its
own output by transforming the two unit points (the function does not return it).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import install_stubs  # noqa: E402

install_stubs()
from data import util as ref_util  # noqa: E402
from data.augmentation import augment_data  # noqa: E402


def main():
    rng = np.random.RandomState(12)
    lens = [7, 19, 4, 19, 1, 11]
    batch = [(rng.randn(1, t, 25, 2).astype(np.float32), np.array([i % 6])) for i, t in enumerate(lens)]
    xx, labels = ref_util.loopy_pad_collate_fn(batch)
    out = {'collate_lens': np.array(lens), 'collate_out': xx.numpy(), 'collate_labels': labels.numpy()}
    for i, (x, _) in enumerate(batch):
        out['collate_in%d' % i] = x
    # two persons per element (N = 2 rows sharing one length)
    batch2 = [(rng.randn(2, t, 25, 3).astype(np.float32), np.array([1, 1])) for t in (5, 9)]
    xx2, labels2 = ref_util.loopy_pad_collate_fn(batch2)
    out['collate2_out'] = xx2.numpy(); out['collate2_labels'] = labels2.numpy()
    for i, (x, _) in enumerate(batch2):
        out['collate2_in%d' % i] = x
    # augmentation under fixed seeds: inputs, outputs, and the 2x2 block recovered from unit points
    seqs = rng.randn(3, 13, 25, 2)
    out['aug_in'] = seqs
    unit = np.zeros((1, 1, 25, 2)); unit[0, 0, 0] = (1.0, 0.0); unit[0, 0, 1] = (0.0, 1.0)
    for seed in range(12):
        np.random.seed(seed)
        out['aug_out_%d' % seed] = augment_data(seqs)
        np.random.seed(seed)
        u = augment_data(unit)
        out['aug_T2_%d' % seed] = np.stack([u[0, 0, 0], u[0, 0, 1]])      # rows: image of (1,0) and of (0,1) -> T[:2,:2]
    np.savez_compressed(os.path.join(HERE, 'input_pipeline.npz'), **out)
    print('wrote input_pipeline.npz', {k: v.shape for k, v in out.items() if k.startswith(('collate_out', 'aug_out_0'))})


if __name__ == '__main__':
    main()
