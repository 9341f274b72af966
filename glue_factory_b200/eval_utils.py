"""Match-quality metrics on the B200: drop-ins for `eval_matches_homography`,
`eval_matches_epipolar` and `eval_homography_dlt` of `gluefactory.eval.utils`
(reference: gluefactory/eval/utils.py:40-96,176-196).

The predictions stay on the device between the matcher and the metric: one kernel launch per metric for the whole
batch (one CTA per pair) instead of a Python loop over batch items with boolean indexing and `.item()` per figure.
Batched inputs return a dict of lists exactly like `eval_per_batch_item` (:30-37); un-batched inputs return
scalars like the reference.  RANSAC estimators (`eval_homography_robust`, `eval_relative_pose_robust`) stay
third-party CPU code and are out of scope.
"""
import torch

from . import ops


def _batched(x, nd):
    return x if x.dim() == nd + 1 else x[None]


def _pred(pred):
    for k in ("keypoints0", "keypoints1", "matches0", "matching_scores0"):
        assert k in pred.keys()
    kp0, kp1 = _batched(pred["keypoints0"], 2), _batched(pred["keypoints1"], 2)
    m0, s0 = _batched(pred["matches0"], 1), _batched(pred["matching_scores0"], 1)
    if not kp0.is_cuda:
        raise RuntimeError("glue_factory_b200 metrics run on CUDA tensors only (no CPU fallback)")
    return kp0, kp1, m0.to(torch.int64), s0


def _result(table, names, batched, int_keys=("num_matches",)):
    rows = table.cpu().tolist()  # the single read-back of the metric
    out = {n: [int(r[i]) if n in int_keys else r[i] for r in rows] for i, n in enumerate(names)}
    return out if batched else {k: v[0] for k, v in out.items()}


def eval_matches_homography(data: dict, pred: dict) -> dict:
    """prec@1px / prec@3px of the symmetric transfer error against H_0to1, num_matches, num_keypoints (:76-96)."""
    assert "H_0to1" in data.keys()
    H = data["H_0to1"]
    batched = H.dim() > 2
    kp0, kp1, m0, _ = _pred(pred)
    with torch.cuda.device(kp0.device):
        table = ops.eval_matches_homography(kp0, kp1, m0, _batched(H, 2).to(kp0.device))
    return _result(table, ["prec@1px", "prec@3px", "num_matches", "num_keypoints"], batched)


def _camera_params(cam, dev):
    """(fx, fy, cx, cy) rows from a reference Camera wrapper (`_data` = w, h, fx, fy, cx, cy[, dist]) or a tensor."""
    d = cam._data if hasattr(cam, "_data") else cam
    d = d if d.dim() == 2 else d[None]
    if d.shape[-1] != 6:
        raise NotImplementedError("the B200 epipolar metric supports pinhole cameras (no distortion parameters)")
    return d[:, 2:6].to(device=dev, dtype=torch.float32).contiguous()


def _pose_params(T, dev):
    """R (row-major) | t rows from a reference Pose wrapper (`_data` = R flattened, t) or a (…,12) tensor."""
    d = T._data if hasattr(T, "_data") else T
    d = d if d.dim() == 2 else d[None]
    return d.to(device=dev, dtype=torch.float32).contiguous()


def eval_matches_epipolar(data: dict, pred: dict) -> dict:
    """epi_prec@{1e-4,5e-4,1e-3} of the symmetric epipolar distance in normalised coordinates (:40-73)."""
    for k in ("view0", "view1", "T_0to1"):
        assert k in data.keys()
    kp0, kp1, m0, _ = _pred(pred)
    batched = pred["keypoints0"].dim() == 3
    dev = kp0.device
    with torch.cuda.device(dev):
        table = ops.eval_matches_epipolar(kp0, kp1, m0, _camera_params(data["view0"]["camera"], dev),
                                          _camera_params(data["view1"]["camera"], dev), _pose_params(data["T_0to1"], dev))
    return _result(table, ["epi_prec@1e-4", "epi_prec@5e-4", "epi_prec@1e-3", "num_matches", "num_keypoints"], batched)


def eval_homography_dlt(data: dict, pred: dict) -> dict:
    """Corner error of the score-weighted DLT homography of the matched points (:176-196)."""
    H_gt = data["H_0to1"]
    batched = H_gt.dim() > 2
    kp0, kp1, m0, s0 = _pred(pred)
    dev = kp0.device
    size = data["view0"]["image_size"]
    with torch.cuda.device(dev):
        _, err = ops.eval_homography_dlt(kp0, kp1, m0, s0, _batched(H_gt, 2).to(dev), _batched(size, 1).to(dev))
    return _result(err[:, None], ["H_error_dlt"], batched)
