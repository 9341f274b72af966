"""Oracle: image pre-processing, restated with plain torch ops (CPU, fp32).

Test infrastructure only (see oracle/__init__.py).  Two layers:

* `ImagePreprocessor.__call__` / `get_new_image_size` / `numpy_image_to_torch` restate
  /root/reference/gluefactory/utils/image.py:12-101,112-121.  Pinned: tests/golden/make_golden_io.py runs the
  UNMODIFIED reference class and stores its outputs (prep_*.npz).
* `kornia_resize` / `gaussian_blur2d` restate the third-party dependency the reference calls at image.py:37-44:
  `kornia.geometry.transform.resize` (kornia is listed un-pinned in the reference's pyproject.toml:31, is not
  vendored under /root/reference and is not installed here).  Restated from kornia's published algorithm
  (kornia >= 0.6.5, `geometry/transform/affwarp.py::resize`, `filters/gaussian.py`, `filters/kernels.py`):
  blur only when down-scaling, sigma = max((factor - 1) / 2, 0.001) per axis, kernel size int(max(4 sigma, 3)) made
  odd, separable Gaussian with `reflect` border (horizontal pass first), then
  `F.interpolate(size, mode, align_corners)`.  PARITY UNPINNED for this layer: the golden generator routes the
  reference's `kornia.geometry.transform.resize` call to this function through oracle/_shim/kornia.
"""
from __future__ import annotations

import collections.abc as collections

import numpy as np
import torch
import torch.nn.functional as F


def gaussian_taps(ksize: int, sigma: float) -> torch.Tensor:
    """kornia.filters.kernels.gaussian: exp(-x^2 / (2 sigma^2)) on x = arange(k) - k // 2, normalised (fp32)."""
    x = torch.arange(ksize, dtype=torch.float32) - ksize // 2
    if ksize % 2 == 0:
        x = x + 0.5
    g = torch.exp(-x.pow(2.0) / (2 * torch.tensor(float(sigma), dtype=torch.float32).pow(2.0)))
    return g / g.sum()


def blur_plan(in_hw, out_hw, antialias: bool):
    """(ky, sigma_y, kx, sigma_x) of kornia's anti-alias blur, or None when resize() does not blur."""
    fy, fx = in_hw[0] / out_hw[0], in_hw[1] / out_hw[1]
    if not (antialias and max(fy, fx) > 1):
        return None
    sy, sx = max((fy - 1.0) / 2.0, 0.001), max((fx - 1.0) / 2.0, 0.001)
    ky, kx = int(max(2.0 * 2 * sy, 3)), int(max(2.0 * 2 * sx, 3))
    ky += 1 - ky % 2
    kx += 1 - kx % 2
    return ky, sy, kx, sx


def gaussian_blur2d(x, ky, sy, kx, sx):
    """kornia.filters.gaussian_blur2d(border_type='reflect', separable=True) on (B,C,H,W)."""
    b, c, h, w = x.shape
    tx, ty = gaussian_taps(kx, sx), gaussian_taps(ky, sy)
    x = x.reshape(b * c, 1, h, w)
    x = F.conv2d(F.pad(x, (kx // 2, kx // 2, 0, 0), mode="reflect"), tx.view(1, 1, 1, kx))
    x = F.conv2d(F.pad(x, (0, 0, ky // 2, ky // 2), mode="reflect"), ty.view(1, 1, ky, 1))
    return x.reshape(b, c, h, w)


def side_to_image_size(side_size: int, aspect_ratio: float, side: str):
    """kornia.geometry.transform.affwarp._side_to_image_size"""
    if side not in ("short", "long", "vert", "horz"):
        raise ValueError(f"side can be one of 'short', 'long', 'vert', and 'horz'. Got '{side}'")
    if side == "vert":
        return side_size, int(side_size * aspect_ratio)
    if side == "horz":
        return int(side_size / aspect_ratio), side_size
    if (side == "short") ^ (aspect_ratio < 1.0):
        return side_size, int(side_size * aspect_ratio)
    return int(side_size / aspect_ratio), side_size


def kornia_resize(input, size, interpolation="bilinear", align_corners=None, side="short", antialias=False):
    """kornia.geometry.transform.resize on (C,H,W) or (B,C,H,W) float tensors."""
    squeeze = input.dim() == 3
    x = input[None] if squeeze else input
    h, w = x.shape[-2:]
    if isinstance(size, int):
        size = side_to_image_size(size, w / h, side)
    size = tuple(int(s) for s in size)
    if size != (h, w):
        plan = blur_plan((h, w), size, antialias)
        if plan is not None:
            x = gaussian_blur2d(x, *plan)
        x = F.interpolate(x, size=size, mode=interpolation, align_corners=align_corners)
    return x[0] if squeeze else x


def numpy_image_to_torch(image: np.ndarray) -> torch.Tensor:
    """(image.py:112-121) HxWxC or HxW uint8 -> CxHxW float32 in [0, 1] (division in float64)."""
    if image.ndim == 3:
        image = image.transpose((2, 0, 1))
    elif image.ndim == 2:
        image = image[None]
    else:
        raise ValueError(f"Not an image: {image.shape}")
    return torch.tensor(image / 255.0, dtype=torch.float)


DEFAULT_CONF = dict(resize=None, edge_divisible_by=None, side="long", interpolation="bilinear", align_corners=None,
                    antialias=True, square_pad=False, add_padding_mask=False)


def get_new_image_size(conf, h: int, w: int):
    """(image.py:72-99)"""
    if isinstance(conf["resize"], collections.Iterable):
        assert len(conf["resize"]) == 2
        return tuple(conf["resize"])
    size = side_to_image_size(conf["resize"], w / h, conf["side"])
    if conf["edge_divisible_by"] is not None:
        df = conf["edge_divisible_by"]
        size = [int(x // df * df) for x in size]
    return size


def preprocess(img: torch.Tensor, conf=None) -> dict:
    """ImagePreprocessor.__call__ (image.py:30-67) on a CPU float tensor (C,H,W)."""
    c = dict(DEFAULT_CONF)
    c.update(conf or {})
    h, w = img.shape[-2:]
    size = h, w
    if c["resize"] is not None:
        size = get_new_image_size(c, h, w)
        img = kornia_resize(img, size, side=c["side"], antialias=c["antialias"], align_corners=c["align_corners"],
                            interpolation=c["interpolation"])
    scale = torch.Tensor([img.shape[-1] / w, img.shape[-2] / h]).to(img)
    data = {"scales": scale, "image_size": np.array(size[::-1]), "transform": np.diag([scale[0], scale[1], 1]),
            "original_image_size": np.array([w, h])}
    if c["square_pad"]:
        sl = max(img.shape[-2:])
        data["image"] = torch.zeros(*img.shape[:-2], sl, sl, dtype=img.dtype)
        data["image"][:, : img.shape[-2], : img.shape[-1]] = img
        if c["add_padding_mask"]:
            data["padding_mask"] = torch.zeros(*img.shape[:-3], 1, sl, sl, dtype=torch.bool)
            data["padding_mask"][:, : img.shape[-2], : img.shape[-1]] = True
    else:
        data["image"] = img
    return data
