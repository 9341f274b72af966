"""Image pre-processing on the B200: drop-in for `gluefactory.utils.image.ImagePreprocessor`
(reference: gluefactory/utils/image.py:12-101; used by datasets/image_pairs.py:59-90 in CPU workers).

Same `default_conf`, same `get_new_image_size`, same output dict (`image`, `scales`, `image_size`, `transform`,
`original_image_size`, optional `padding_mask`).  The resize (kornia.geometry.transform.resize in the reference:
Gaussian anti-alias blur when down-scaling, then bilinear interpolation), the square padding and the uint8 -> float
conversion run in the `prep_*` kernels of libb200glue.so.  In addition to the reference's float (C,H,W) tensor the
call accepts a uint8 (H,W,C) / (H,W) tensor as cv2 reads it (the conversion of `numpy_image_to_torch` is then fused)
and batches of equally sized images.  No CPU fallback: CPU tensors raise.
"""
import collections.abc as collections
from pathlib import Path
from typing import Optional

import numpy as np
import torch

from . import ops
from ._conf import get_conf_backend

_, OmegaConf = get_conf_backend()


def gaussian_taps(ksize: int, sigma: float) -> torch.Tensor:
    """Normalised Gaussian window as kornia builds it (filters/kernels.py `gaussian`), fp32 on the host."""
    x = torch.arange(ksize, dtype=torch.float32) - ksize // 2
    if ksize % 2 == 0:
        x = x + 0.5
    g = torch.exp(-x.pow(2.0) / (2 * torch.tensor(float(sigma), dtype=torch.float32).pow(2.0)))
    return g / g.sum()


def antialias_taps(in_hw, out_hw, antialias: bool):
    """(taps_y, taps_x) of kornia's resize anti-alias blur, or (None, None): only when down-scaling,
    sigma = max((factor - 1) / 2, 0.001) per axis, kernel size int(max(4 sigma, 3)) made odd."""
    fy, fx = in_hw[0] / out_hw[0], in_hw[1] / out_hw[1]
    if not (antialias and max(fy, fx) > 1):
        return None, None
    sy, sx = max((fy - 1.0) / 2.0, 0.001), max((fx - 1.0) / 2.0, 0.001)
    ky, kx = int(max(2.0 * 2 * sy, 3)), int(max(2.0 * 2 * sx, 3))
    ky += 1 - ky % 2
    kx += 1 - kx % 2
    return gaussian_taps(ky, sy), gaussian_taps(kx, sx)


class ImagePreprocessor:
    default_conf = {
        "resize": None,  # target edge length, None for no resizing
        "edge_divisible_by": None,
        "side": "long",
        "interpolation": "bilinear",
        "align_corners": None,
        "antialias": True,
        "square_pad": False,
        "add_padding_mask": False,
    }

    def __init__(self, conf) -> None:
        default_conf = OmegaConf.create(self.default_conf)
        OmegaConf.set_struct(default_conf, True)
        self.conf = OmegaConf.merge(default_conf, conf)

    def __call__(self, img: torch.Tensor, interpolation: Optional[str] = None) -> dict:
        """Resize and preprocess an image, return image and resize scale (image.py:30-67)."""
        if not img.is_cuda:
            raise RuntimeError("glue_factory_b200 ImagePreprocessor runs on CUDA tensors only (no CPU fallback)")
        c = self.conf
        u8 = img.dtype == torch.uint8
        if u8:  # (H,W), (H,W,C) or (B,H,W,C) as cv2 reads it
            src = img[..., None] if img.dim() == 2 else img
            batched = src.dim() == 4
            src = src if batched else src[None]
            h, w = src.shape[1:3]
        else:   # (C,H,W) or (B,C,H,W) float in [0, 1]
            batched = img.dim() == 4
            src = (img if batched else img[None]).float()
            h, w = src.shape[-2:]
        size = h, w
        ty = tx = None
        if c.resize is not None:
            mode = c.interpolation if interpolation is None else interpolation
            if mode != "bilinear" or c.align_corners:
                raise NotImplementedError("the B200 resize kernel implements bilinear interpolation, align_corners=False")
            size = tuple(int(s) for s in self.get_new_image_size(h, w))
            if size != (h, w):
                ty, tx = antialias_taps((h, w), size, bool(c.antialias))
        pad = (max(size), max(size)) if c.square_pad else size
        with torch.cuda.device(img.device):
            image, mask = ops.prep_resize(src, size, tx, ty, pad_hw=pad, mask=bool(c.square_pad and c.add_padding_mask))
        scale = torch.Tensor([size[1] / w, size[0] / h]).to(image)
        sc = scale.cpu()
        data = {
            "scales": scale,
            "image_size": np.array(size[::-1]),
            "transform": np.diag([sc[0], sc[1], 1]),
            "original_image_size": np.array([w, h]),
            "image": image if batched else image[0],
        }
        if mask is not None:
            m = mask.bool()[:, None]
            data["padding_mask"] = m if batched else m[0]
        return data

    def load_image(self, image_path: Path, device="cuda") -> dict:
        """Reads with cv2 on the host, ships the uint8 pixels (a quarter of the float bytes) and converts on the device."""
        return self(torch.from_numpy(np.ascontiguousarray(read_image(image_path))).to(device))

    def get_new_image_size(self, h: int, w: int):
        """Target (height, width) for an h x w image (image.py:72-99): an explicit pair wins; otherwise `resize` is the
        length of the side named by `side`, the other side follows the aspect ratio (truncated), and both are rounded
        down to a multiple of `edge_divisible_by`."""
        target = self.conf.resize
        if isinstance(target, collections.Iterable):
            assert len(target) == 2
            return tuple(target)
        out = _fit_side(int(target), h, w, self.conf.side)
        step = self.conf.edge_divisible_by
        return out if step is None else [int(v // step * step) for v in out]


def _fit_side(length: int, h: int, w: int, side: str):
    """(height, width) with the chosen side equal to `length` (kornia's `side` convention: short / long / vert / horz)."""
    if side not in ("short", "long", "vert", "horz"):
        raise ValueError(f"side can be one of 'short', 'long', 'vert', and 'horz'. Got '{side}'")
    ratio = w / h
    if side in ("short", "long"):  # which axis is the short / long one decides; a square image counts as landscape
        side = "vert" if (side == "short") != (ratio < 1.0) else "horz"
    return (length, int(length * ratio)) if side == "vert" else (int(length / ratio), length)


def read_image(path: Path, grayscale: bool = False) -> np.ndarray:
    """Read an image from path as RGB or grayscale (image.py:102-113)."""
    import cv2

    path = Path(path)
    if not path.exists():
        raise FileNotFoundError(f"No image at path {path}.")
    pixels = cv2.imread(str(path), cv2.IMREAD_GRAYSCALE if grayscale else cv2.IMREAD_COLOR)
    if pixels is None:
        raise IOError(f"Could not read image at {path}.")
    return pixels if grayscale else pixels[..., ::-1]  # BGR -> RGB


def numpy_image_to_torch(image: np.ndarray) -> torch.Tensor:
    """(image.py:112-121)"""
    if image.ndim == 3:
        image = image.transpose((2, 0, 1))
    elif image.ndim == 2:
        image = image[None]
    else:
        raise ValueError(f"Not an image: {image.shape}")
    return torch.tensor(image / 255.0, dtype=torch.float)


def load_image(path: Path, grayscale=False) -> torch.Tensor:
    return numpy_image_to_torch(read_image(path, grayscale=grayscale))
