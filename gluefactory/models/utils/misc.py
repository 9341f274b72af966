"""Sequence padding helpers (reference: gluefactory/models/utils/misc.py:19-70).

Only used on the extraction path when `force_num_keypoints` is set: keypoints are padded
with uniform-random coordinates drawn from the torch RNG, so results past the real
keypoints are not reproducible across implementations (documented in DESIGN.md).
"""
from typing import List, Optional, Tuple

import torch


def pad_to_length(x, length: int, pad_dim: int = -2, mode: str = "zeros", bounds: Tuple = (None, None)):
    have = x.shape[pad_dim]
    assert have <= length
    if have == length:
        return x
    shape = list(x.shape)
    shape[pad_dim] = length - have
    low, high = bounds
    if mode == "zeros":
        extra = torch.zeros(*shape, device=x.device, dtype=x.dtype)
    elif mode == "ones":
        extra = torch.ones(*shape, device=x.device, dtype=x.dtype)
    elif mode == "random":
        lo = x.min() if low is None else low
        hi = x.max() if high is None else high
        extra = torch.empty(*shape, device=x.device).uniform_(lo, hi)
    elif mode == "random_c":
        cols = []
        for c in range(shape[-1]):
            lo = x[..., c].min() if have > 0 else low
            hi = x[..., c].max() if have > 0 else high
            cols.append(torch.empty(*shape[:-1], 1, device=x.device).uniform_(lo, hi))
        extra = torch.cat(cols, dim=-1)
    else:
        raise ValueError(mode)
    return torch.cat([x, extra], dim=pad_dim)


def pad_and_stack(sequences: List[torch.Tensor], length: Optional[int] = None, pad_dim: int = -2, **kwargs):
    if length is None:
        length = max(x.shape[pad_dim] for x in sequences)
    return torch.stack([pad_to_length(x, length, pad_dim, **kwargs) for x in sequences], 0)
