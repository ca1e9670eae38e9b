"""Skeleton graphs and partitioned adjacency (build-time input of the block).

Host-side, runs once per model construction.  Reproduces the reference's
`get_normalized_adjacency_matrices` (src/data/adjacency.py:34-200) including its
dense-alpha normalisation (alpha is added to every entry of diag(rowsum) before
the element-wise power, :180), generalised from the hard-wired 25-joint list
(src/data/util.py:156-180) to any adjacency list so a Kinetics-18 graph can be
built the same way.
"""
from enum import IntEnum

import torch


class Strategy(IntEnum):
    UNI_LABELING = 0
    DISTANCE = 1
    SPATIAL_CONFIGURATION = 2
    SYMMETRICAL = 3


def _edges_to_adj(n, edges):
    adj = {i: [] for i in range(n)}
    for a, b in edges:
        adj[a].append(b)
        adj[b].append(a)
    return adj


# OpenPose BODY_25 as wired in the reference (neighbour ORDER matters for nothing numerical, kept anyway)
BODY25 = {0: [1, 15, 16], 1: [0, 2, 5, 8], 2: [1, 3], 3: [2, 4], 4: [3], 5: [1, 6], 6: [5, 7], 7: [6],
          8: [1, 9, 12], 9: [8, 10], 10: [9, 11], 11: [10, 22, 24], 12: [8, 13], 13: [12, 14], 14: [13, 19, 21],
          15: [0, 17], 16: [0, 18], 17: [15], 18: [16], 19: [14, 20], 20: [19], 21: [14], 22: [11, 23], 23: [22],
          24: [11]}
BODY25_OPPOSITE = {2: 5, 3: 6, 4: 7, 5: 2, 6: 3, 7: 4, 9: 12, 10: 13, 11: 14, 12: 9, 13: 10, 14: 11, 15: 16, 16: 15,
                   17: 18, 18: 17, 19: 22, 20: 23, 21: 24, 22: 19, 23: 20, 24: 21}
# OpenPose COCO-18 (Kinetics-skeleton): neck=1, nose=0
OPENPOSE18 = _edges_to_adj(18, [(4, 3), (3, 2), (7, 6), (6, 5), (13, 12), (12, 11), (10, 9), (9, 8), (11, 5), (8, 2),
                                (5, 1), (2, 1), (0, 1), (15, 0), (14, 0), (17, 15), (16, 14)])
OPENPOSE18_OPPOSITE = {2: 5, 3: 6, 4: 7, 5: 2, 6: 3, 7: 4, 8: 11, 9: 12, 10: 13, 11: 8, 12: 9, 13: 10, 14: 15, 15: 14,
                       16: 17, 17: 16}
SKELETONS = {'body25': (BODY25, BODY25_OPPOSITE), 'openpose18': (OPENPOSE18, OPENPOSE18_OPPOSITE)}


def _hops(adj, root, d):
    """nodes within d hops (in discovery order) and the per-hop rings"""
    seen = [root]
    frontier = [root]
    rings = []
    for _ in range(d):
        nxt = []
        for cur in frontier:
            for j in adj[cur]:
                if j not in seen:
                    seen.append(j)
                    nxt.append(j)
        rings.append(nxt)
        frontier = nxt
    return seen, rings


def create_adjacency_matrices(strat=Strategy.UNI_LABELING, d=1, distances=None, skeleton='body25'):
    adj, opposite = SKELETONS[skeleton] if isinstance(skeleton, str) else skeleton
    V = len(adj)
    strat = int(strat)
    if strat == Strategy.UNI_LABELING:
        A = torch.zeros(V, V)
        for i in range(V):
            A[i, _hops(adj, i, d)[0]] = 1
        return [A]
    if strat in (Strategy.DISTANCE, Strategy.SYMMETRICAL):
        mats = [torch.eye(V)] + [torch.zeros(V, V) for _ in range(d)]
        for i in range(V):
            for hop, ring in enumerate(_hops(adj, i, d)[1]):
                for j in ring:
                    mats[hop + 1][i, j] = 1
                    if strat == Strategy.SYMMETRICAL and j in opposite:
                        mats[hop + 1][i, opposite[j]] = 1
            if strat == Strategy.SYMMETRICAL and i in opposite:
                mats[d][i, opposite[i]] = 1
        return mats
    if strat == Strategy.SPATIAL_CONFIGURATION:
        if distances is None:
            raise ValueError('Distance file not provided')
        mats = [torch.zeros(V, V) for _ in range(3)]
        for i in range(V):
            for j in _hops(adj, i, d)[0]:
                lab = 0 if distances[j] == distances[i] else (1 if distances[j] < distances[i] else 2)
                mats[lab][i, j] = 1
        return mats
    raise ValueError('unknown partitioning strategy %r' % (strat,))


def normalize(matrices, expo=-0.5, alpha=0.001, dense_alpha=True):
    """dense_alpha=True is the reference (adjacency.py:161-183): alpha is added to EVERY entry of diag(rowsum) before the
    element-wise power, which makes the result dense with entries ~1e4..1e5.  dense_alpha=False is the textbook
    D^-1/2 A D^-1/2 the reference's comment intends (alpha on the diagonal only; entries <= 0.5) -- an option beyond the
    reference, used for the Kinetics-18 benchmark graph."""
    out = torch.empty(len(matrices), matrices[0].shape[0], matrices[0].shape[1])
    for i, A in enumerate(matrices):
        if dense_alpha:
            lam = (torch.diag(A.sum(dim=1)) + alpha) ** expo      # dense: alpha lands off the diagonal too
        else:
            lam = torch.diag((A.sum(dim=1) + alpha) ** expo)
        out[i] = lam @ A @ lam
    return out


def normalize_on_device(matrices, device, expo=-0.5, alpha=0.001, dense_alpha=True):
    """`normalize` executed by the library on `device` (stgcn_adjacency_normalize): (K,V,V) fp32 CUDA tensor, V <= 32."""
    from . import _lib as L
    raw = torch.stack([m.float() for m in matrices]).contiguous().to(device)
    out = torch.empty_like(raw)
    L.check(L.lib().stgcn_adjacency_normalize(raw.shape[0], raw.shape[1], L.ptr(raw), float(alpha), float(expo), int(bool(dense_alpha)),
                                              L.ptr(out), L.stream_ptr(out.device)), 'adjacency_normalize')
    return out


def get_normalized_adjacency_matrices(strat=Strategy.UNI_LABELING, d=1, alpha=0.001, distance_file=None,
                                      distances=None, skeleton='body25', dense_alpha=True, device=None):
    """device=None: host arithmetic as in the reference (bit-for-bit, tests/test_oracle_golden.py); a CUDA device: the
    normalisation runs in the library on that device (same formula, fp32, within 1e-6 of the host result)."""
    if distances is None and distance_file:
        import numpy as np
        distances = np.load(distance_file)
    mats = create_adjacency_matrices(strat, d, distances, skeleton)
    if device is not None:
        return normalize_on_device(mats, device, alpha=alpha, dense_alpha=dense_alpha)
    return normalize(mats, alpha=alpha, dense_alpha=dense_alpha)
