"""Oracle: match-quality metrics, restated with plain torch ops (CPU, fp32).

Test infrastructure only (see oracle/__init__.py).  Citations are to /root/reference/gluefactory/.

* `get_matches_scores`, `eval_matches_homography`, `eval_matches_epipolar`, `eval_homography_dlt` restate
  eval/utils.py:21-27,40-96,176-196; `sym_homography_error`, `homography_corner_error` geometry/homography.py:319-347;
  `T_to_E`, `sym_epipolar_distance` geometry/epipolar.py:7-9,32-56; pinhole `image2cam` geometry/wrappers.py:371-397.
  Pinned: tests/golden/make_golden_io.py runs the UNMODIFIED reference functions (eval_*.npz).
* `find_homography_dlt` / `normalize_points` restate the third-party `kornia.geometry.homography` /
  `kornia.geometry.epipolar` functions the reference imports at eval/utils.py:3 (kornia: un-pinned dependency,
  pyproject.toml:31, not vendored, not installed here), from kornia's published algorithm: Hartley normalisation
  (mean, sqrt(2) / mean distance), the 2N x 9 DLT system, A^T diag(w, w) A, SVD, last right-singular vector,
  de-normalisation, division by H[2,2] + eps.  PARITY UNPINNED for this function.
"""
from __future__ import annotations

import torch


def to_homogeneous(p):
    return torch.cat([p, p.new_ones(p.shape[:-1] + (1,))], -1)


def from_homogeneous(p, eps=0.0):
    return p[..., :-1] / (p[..., -1:] + eps)


def get_matches_scores(kpts0, kpts1, matches0, mscores0):
    """(eval/utils.py:21-27)"""
    m0 = matches0 > -1
    m1 = matches0[m0]
    return kpts0[m0], kpts1[m1], mscores0[m0]


def sym_homography_error(kpts0, kpts1, T_0to1):
    """(geometry/homography.py:319-328)"""
    kpts0_1 = from_homogeneous(to_homogeneous(kpts0) @ T_0to1.transpose(-1, -2))
    dist0_1 = ((kpts0_1 - kpts1) ** 2).sum(-1).sqrt()
    kpts1_0 = from_homogeneous(to_homogeneous(kpts1) @ torch.pinverse(T_0to1.transpose(-1, -2)))
    dist1_0 = ((kpts1_0 - kpts0) ** 2).sum(-1).sqrt()
    return (dist0_1 + dist1_0) / 2.0


def homography_corner_error(T, T_gt, image_size):
    """(geometry/homography.py:341-347)"""
    W, H = image_size[..., 0], image_size[..., 1]
    corners0 = torch.Tensor([[0, 0], [W, 0], [W, H], [0, H]]).float().to(T)
    corners1_gt = from_homogeneous(to_homogeneous(corners0) @ T_gt.transpose(-1, -2))
    corners1 = from_homogeneous(to_homogeneous(corners0) @ T.transpose(-1, -2))
    return torch.sqrt(((corners1 - corners1_gt) ** 2).sum(-1)).mean(-1)


def eval_matches_homography(H_gt, kp0, kp1, m0, scores0):
    """(eval/utils.py:76-96) for one pair"""
    pts0, pts1, _ = get_matches_scores(kp0, kp1, m0, scores0)
    err = sym_homography_error(pts0, pts1, H_gt)
    return {"prec@1px": (err < 1).float().mean().nan_to_num().item(),
            "prec@3px": (err < 3).float().mean().nan_to_num().item(),
            "num_matches": pts0.shape[0], "num_keypoints": (kp0.shape[0] + kp1.shape[0]) / 2.0}


def skew_symmetric(v):
    z = torch.zeros_like(v[..., 0])
    return torch.stack([z, -v[..., 2], v[..., 1], v[..., 2], z, -v[..., 0], -v[..., 1], v[..., 0], z], -1).reshape(
        v.shape[:-1] + (3, 3))


def T_to_E(R, t):
    """(geometry/epipolar.py:7-9)"""
    return skew_symmetric(t) @ R


def sym_epipolar_distance(p0, p1, E, squared=True):
    """(geometry/epipolar.py:32-56)"""
    if p0.shape[-2] == 0:
        return torch.zeros(p0.shape[:-1]).to(p0)
    p0, p1 = to_homogeneous(p0), to_homogeneous(p1)
    p1_E_p0 = torch.einsum("...ni,...ij,...nj->...n", p1, E, p0)
    E_p0 = torch.einsum("...ij,...nj->...ni", E, p0)
    Et_p1 = torch.einsum("...ij,...ni->...nj", E, p1)
    d0 = (E_p0[..., 0] ** 2 + E_p0[..., 1] ** 2).clamp(min=1e-6)
    d1 = (Et_p1[..., 0] ** 2 + Et_p1[..., 1] ** 2).clamp(min=1e-6)
    if squared:
        return p1_E_p0**2 * (1 / d0 + 1 / d1)
    return p1_E_p0.abs() * (1 / d0.sqrt() + 1 / d1.sqrt()) / 2


def eval_matches_epipolar(cam0, cam1, R, t, kp0, kp1, m0, scores0):
    """(eval/utils.py:40-73) for one pair and pinhole cameras cam = (fx, fy, cx, cy)
    (Camera.image2cam = (p - c) / f without distortion, geometry/wrappers.py:371-397)."""
    pts0, pts1, _ = get_matches_scores(kp0, kp1, m0, scores0)
    p0 = (pts0 - cam0[2:4]) / cam0[0:2]
    p1 = (pts1 - cam1[2:4]) / cam1[0:2]
    err = sym_epipolar_distance(p0[None], p1[None], T_to_E(R, t), squared=False)[0]
    return {"epi_prec@1e-4": (err < 1e-4).float().mean().item(), "epi_prec@5e-4": (err < 5e-4).float().mean().item(),
            "epi_prec@1e-3": (err < 1e-3).float().mean().item(), "num_matches": pts0.shape[0],
            "num_keypoints": (kp0.shape[0] + kp1.shape[0]) / 2.0}


def normalize_points(points, eps=1e-8):
    """kornia.geometry.epipolar.normalize_points on (B,N,2)"""
    x_mean = points.mean(1, keepdim=True)
    scale = (points - x_mean).norm(dim=-1, p=2).mean(dim=-1)
    scale = torch.sqrt(torch.tensor(2.0)) / (scale + eps)
    ones, zeros = torch.ones_like(scale), torch.zeros_like(scale)
    transform = torch.stack([scale, zeros, -scale * x_mean[..., 0, 0], zeros, scale, -scale * x_mean[..., 0, 1],
                             zeros, zeros, ones], -1).view(-1, 3, 3)
    return (to_homogeneous(points) @ transform.transpose(-1, -2))[..., :2], transform


def find_homography_dlt(points1, points2, weights=None, solver="svd"):
    """kornia.geometry.homography.find_homography_dlt on (B,N,2) points and (B,N) weights."""
    if points1.shape[1] < 4:
        raise AssertionError(points1.shape)
    eps = 1e-8
    p1, t1 = normalize_points(points1)
    p2, t2 = normalize_points(points2)
    x1, y1 = torch.chunk(p1, dim=-1, chunks=2)
    x2, y2 = torch.chunk(p2, dim=-1, chunks=2)
    ones, zeros = torch.ones_like(x1), torch.zeros_like(x1)
    ax = torch.cat([zeros, zeros, zeros, -x1, -y1, -ones, y2 * x1, y2 * y1, y2], dim=-1)
    ay = torch.cat([x1, y1, ones, zeros, zeros, zeros, -x2 * x1, -x2 * y1, -x2], dim=-1)
    A = torch.cat((ax, ay), dim=-1).reshape(ax.shape[0], -1, ax.shape[-1])
    if weights is None:
        A = A.transpose(-2, -1) @ A
    else:
        w_diag = torch.diag_embed(weights.unsqueeze(dim=-1).repeat(1, 1, 2).reshape(weights.shape[0], -1))
        A = A.transpose(-2, -1) @ w_diag @ A
    _, _, Vh = torch.linalg.svd(A)
    H = Vh.mH[..., -1].view(-1, 3, 3)
    H = t2.inverse() @ (H @ t1)
    return H / (H[..., -1:, -1:] + eps)


def eval_homography_dlt(H_gt, image_size, kp0, kp1, m0, scores0):
    """(eval/utils.py:176-196) for one pair; returns (H_error_dlt, H)"""
    pts0, pts1, scores = get_matches_scores(kp0, kp1, m0, scores0)
    try:
        h = find_homography_dlt(pts0[None], pts1[None], scores.to(pts0)[None])[0]
    except AssertionError:
        h = torch.ones_like(H_gt) * float("inf")
    return homography_corner_error(h, H_gt, image_size).item(), h
