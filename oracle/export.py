"""Oracle: the per-pair export step, restated in numpy.

Test infrastructure only (see oracle/__init__.py).  Restates the renormalisation and conversion block of
/root/reference/gluefactory/utils/export_predictions.py:40-77: keypoints multiplied by 1 / scales of their view,
batch element 0 taken, optional fp16 conversion of the float32 arrays, one group per pair name.  Pinned:
tests/golden/make_golden_io.py runs the UNMODIFIED `export_predictions` on recorded predictions (a replay model and
an in-memory stand-in for the absent h5py, oracle/_shim/h5py.py) and stores what it wrote (export_*.npz).
"""
from __future__ import annotations

import numpy as np
import torch


def export_records(pred: dict, data: dict, keys="*", optional_keys=(), as_half=False) -> dict:
    """One batch element (index 0), as the reference's loop body produces it: {key: numpy array}."""
    if keys != "*":
        missing = set(keys) - set(pred.keys())
        if missing:
            raise ValueError(f"Missing key {missing}")
        pred = {k: v for k, v in pred.items() if k in list(keys) + list(optional_keys)}
    assert len(pred) > 0
    out = {}
    for k, v in pred.items():
        for prefix in ("keypoints", "lines", "orig_lines"):
            if k.startswith(prefix):
                idx = k.replace(prefix, "")
                scales = 1.0 / (data["scales"] if len(idx) == 0 else data[f"view{idx}"]["scales"])
                v = v * scales[None]
                break
        out[k] = v[0].cpu().numpy()
    if as_half:
        for k in out:
            if out[k].dtype == np.float32:
                out[k] = out[k].astype(np.float16)
    return out


def export_batch(pred: dict, data: dict, **kw):
    """Per-pair records of a batch of B pairs: what B iterations of the reference loop at batch size 1 write."""
    B = next(iter(pred.values())).shape[0]
    recs = []
    for b in range(B):
        p = {k: v[b : b + 1] for k, v in pred.items()}
        d = {}
        for k, v in data.items():
            if isinstance(v, dict):
                d[k] = {kk: (vv[b : b + 1] if torch.is_tensor(vv) else vv) for kk, vv in v.items()}
            elif torch.is_tensor(v):
                d[k] = v[b : b + 1]
        recs.append(export_records(p, d, **kw))
    return recs
