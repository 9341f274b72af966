"""Nested-container tensor helpers (reference: gluefactory/utils/tensor.py:13-52)."""
from collections.abc import Mapping, Sequence

import numpy as np
import torch


def map_tensor(obj, func):
    """Apply `func` to every leaf of a nested dict / list; strings and None pass through."""
    if obj is None or isinstance(obj, (str, bytes)):
        return obj
    if isinstance(obj, Mapping):
        return {key: map_tensor(value, func) for key, value in obj.items()}
    if isinstance(obj, Sequence):
        return [map_tensor(value, func) for value in obj]
    return func(obj)


def batch_to_numpy(batch):
    return map_tensor(batch, lambda t: t.cpu().numpy())


def batch_to_device(batch, device, non_blocking=True):
    return map_tensor(batch, lambda t: t.to(device=device, non_blocking=non_blocking))


def rbd(data: dict) -> dict:
    """Drop the leading batch dimension of every indexable entry."""
    out = {}
    for key, value in data.items():
        out[key] = value[0] if isinstance(value, (torch.Tensor, np.ndarray, list)) else value
    return out


def index_batch(tensor_dict):
    n = len(next(iter(tensor_dict.values())))
    for i in range(n):
        yield map_tensor(tensor_dict, lambda t: t[i])
