"""Oracle: nearest-neighbour matcher (similarity, one-way nearest neighbours with ratio / distance tests,
mutual check, double-softmax log-assignment), restated functionally.

Test infrastructure only (see oracle/__init__.py).  Written from the behaviour of
/root/reference/gluefactory/models/matchers/nearest_neighbor_matcher.py (citations below); no text of that
file is used.  Pinned against reference-generated golden vectors (tests/golden/nn_*.npz, made by
tests/golden/make_golden_nn.py).
"""
from __future__ import annotations

import torch


def similarity(desc0, desc1):
    """sim[b, i, j] = <desc0[b, i], desc1[b, j]> (nearest_neighbor_matcher.py:52)."""
    return desc0.float() @ desc1.float().transpose(1, 2)


def one_way_nn(sim, ratio_thresh=None, distance_thresh=None):
    """For every row the column of the largest similarity, or -1 when a test fails (:14-25).

    Squared descriptor distance of unit vectors: d = 2 (1 - sim).  Ratio test: d_best <= ratio^2 d_second;
    distance test: d_best <= distance^2.  A threshold of None or 0 disables its test.  Ties go to the lowest
    index (torch.topk leaves the order of equal values unspecified)."""
    order = torch.sort(sim, dim=-1, descending=True, stable=True)
    best, idx = order.values[..., 0], order.indices[..., 0]
    d_best = 2 * (1 - best)
    keep = torch.ones_like(best, dtype=torch.bool)
    if ratio_thresh:
        d_second = 2 * (1 - order.values[..., 1])
        keep &= d_best <= (ratio_thresh**2) * d_second
    if distance_thresh:
        keep &= d_best <= distance_thresh**2
    return torch.where(keep, idx, torch.full_like(idx, -1))


def mutual(m0, m1):
    """Keep i -> j only when j -> i as well (:28-36)."""
    i0 = torch.arange(m0.shape[-1], device=m0.device).expand_as(m0)
    i1 = torch.arange(m1.shape[-1], device=m1.device).expand_as(m1)
    back0 = torch.gather(m1, -1, m0.clamp(min=0))
    back1 = torch.gather(m0, -1, m1.clamp(min=0))
    return (torch.where((m0 >= 0) & (back0 == i0), m0, torch.full_like(m0, -1)),
            torch.where((m1 >= 0) & (back1 == i1), m1, torch.full_like(m1, -1)))


def log_assignment(sim):
    """[B, M+1, N+1]: log_softmax over columns + log_softmax over rows inside, zeros in the last row / column (:59-61)."""
    b, m, n = sim.shape
    la = sim.new_zeros(b, m + 1, n + 1)
    la[:, :m, :n] = 2 * sim - torch.logsumexp(sim, -1, keepdim=True) - torch.logsumexp(sim, -2, keepdim=True)
    return la


def forward(desc0, desc1, ratio_thresh=None, distance_thresh=None, mutual_check=True, sim=None):
    """NearestNeighborMatcher._forward (:51-72).  `sim` may be supplied to teacher-force the later stages."""
    if sim is None:
        sim = similarity(desc0, desc1)
    m0 = one_way_nn(sim, ratio_thresh, distance_thresh)
    m1 = one_way_nn(sim.transpose(1, 2), ratio_thresh, distance_thresh)
    if mutual_check:
        m0, m1 = mutual(m0, m1)
    return {
        "matches0": m0,
        "matches1": m1,
        "matching_scores0": (m0 >= 0).float(),
        "matching_scores1": (m1 >= 0).float(),
        "similarity": sim,
        "log_assignment": log_assignment(sim),
    }
