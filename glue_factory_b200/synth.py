"""Seeded synthetic inputs and conditioned random-init weights.

PyTorch's default init is a degenerate test input for this path (flat heatmap, parallel
descriptors, uniform assignment), so both the reference/oracle and the CUDA path are fed
the *same* conditioned weights built here.  Everything is derived from integer PCG64
streams and float64 add/mul arithmetic only, then rounded to fp32, so the tensors are
bit-identical on any host (no libm, no vectorised transcendental in the recipe).

State-dict schemas follow the reference modules
(reference: gluefactory/models/extractors/superpoint_open.py:59-115,
 gluefactory/models/matchers/lightglue.py:333-357).
"""
from __future__ import annotations

import hashlib
import math
from collections import OrderedDict

import numpy as np
import torch
import torch.nn.functional as F

SP_CHANNELS = [64, 64, 128, 128, 256]


def _normal(rng: np.random.Generator, shape, std=1.0) -> torch.Tensor:
    """Approximately normal samples: centred sum of 4 uniforms, unit variance, |z| <= 3.47."""
    n = int(np.prod(shape))
    u = (rng.integers(0, 1 << 32, size=(4, n), dtype=np.uint64).astype(np.float64) + 0.5) / float(1 << 32)
    z = (u.sum(0) - 2.0) * math.sqrt(3.0) * std
    return torch.from_numpy(z.reshape(shape)).to(torch.float32)


def _uniform(rng: np.random.Generator, shape, lo=0.0, hi=1.0) -> torch.Tensor:
    u = (rng.integers(0, 1 << 32, size=int(np.prod(shape)), dtype=np.uint64).astype(np.float64) + 0.5) / float(1 << 32)
    return torch.from_numpy((lo + (hi - lo) * u).reshape(shape))  # float64


# ----------------------------------------------------------------------------------------
# synthetic images
# ----------------------------------------------------------------------------------------
def synthetic_image(seed: int, height: int, width: int) -> torch.Tensor:
    """Band-limited noise: octaves at 1/32, 1/16, 1/8, 1/4 resolution (weights 8:4:2:1), bilinearly
    upsampled, min-max normalised, (1,H,W) fp32."""
    rng = np.random.default_rng(seed)
    acc = torch.zeros(1, 1, height, width, dtype=torch.float64)
    for weight, div in ((8.0, 32), (4.0, 16), (2.0, 8), (1.0, 4)):
        h, w = max(2, height // div), max(2, width // div)
        octave = (_uniform(rng, (1, 1, h, w)) - 0.5) * math.sqrt(12.0)
        if (h, w) != (height, width):
            octave = F.interpolate(octave, size=(height, width), mode="bilinear", align_corners=False)
        acc += weight * octave
    acc = (acc - acc.min()) / (acc.max() - acc.min())
    return acc[0].to(torch.float32)


def _homography(rng: np.random.Generator, height: int, width: int, max_rot_deg=3.0, max_jitter=0.02) -> torch.Tensor:
    """Seeded homography: rotation about the centre, 0.9 zoom and per-corner jitter (fraction of the side)."""
    rot_lim = math.radians(max_rot_deg)
    ang = float(_uniform(rng, (1,), -rot_lim, rot_lim)[0])
    jitter = _uniform(rng, (4, 2), -max_jitter, max_jitter).numpy() * np.array([width, height])
    cx, cy = width / 2.0, height / 2.0
    src = np.array([[0, 0], [width, 0], [width, height], [0, height]], dtype=np.float64)
    c, s = math.cos(ang), math.sin(ang)
    rot = np.array([[c, -s], [s, c]])
    dst = (src - [cx, cy]) @ rot.T * 0.95 + [cx, cy] + jitter
    rows = []
    for (x, y), (u, v) in zip(src, dst):
        rows.append([x, y, 1, 0, 0, 0, -u * x, -u * y, -u])
        rows.append([0, 0, 0, x, y, 1, -v * x, -v * y, -v])
    _, _, vt = np.linalg.svd(np.asarray(rows))
    hmat = vt[-1].reshape(3, 3)
    return torch.from_numpy(hmat / hmat[2, 2])


def synthetic_pair(seed: int, height: int, width: int):
    """(image0, image1, H_0to1): image1 is image0 warped by a seeded homography + N(0,0.005) noise."""
    img0 = synthetic_image(seed, height, width)
    rng = np.random.default_rng(seed + 7919)
    h01 = _homography(rng, height, width)
    hinv = torch.linalg.inv(h01)
    ys, xs = torch.meshgrid(
        torch.arange(height, dtype=torch.float64) + 0.5,
        torch.arange(width, dtype=torch.float64) + 0.5,
        indexing="ij",
    )
    pts = torch.stack([xs, ys, torch.ones_like(xs)], -1) @ hinv.T
    px, py = pts[..., 0] / pts[..., 2], pts[..., 1] / pts[..., 2]
    grid = torch.stack([px / width * 2 - 1, py / height * 2 - 1], -1)[None]
    warped = F.grid_sample(
        img0[None].to(torch.float64), grid, mode="bilinear", padding_mode="border", align_corners=False
    )[0]
    noise = _normal(rng, (1, height, width), 0.005).to(torch.float64)
    img1 = (warped + noise).clamp(0.0, 1.0).to(torch.float32)
    return img0, img1, h01.to(torch.float32)


def synthetic_batch(first_pair: int, num_pairs: int, height: int, width: int, seed: int = 1234):
    """Stacked views: image0 (P,1,H,W), image1 (P,1,H,W), H (P,3,3)."""
    im0, im1, hs = zip(*(synthetic_pair(seed + first_pair + i, height, width) for i in range(num_pairs)))
    return torch.stack(im0), torch.stack(im1), torch.stack(hs)


# ----------------------------------------------------------------------------------------
# conditioned weights
# ----------------------------------------------------------------------------------------
def _vgg_block_entries(sd, prefix, rng, c_in, c_out, k, gain):
    fan_in = c_in * k * k
    sd[f"{prefix}.conv.weight"] = _normal(rng, (c_out, c_in, k, k), gain / math.sqrt(fan_in))
    sd[f"{prefix}.conv.bias"] = torch.zeros(c_out)
    sd[f"{prefix}.bn.weight"] = torch.ones(c_out)
    sd[f"{prefix}.bn.bias"] = torch.zeros(c_out)
    sd[f"{prefix}.bn.running_mean"] = torch.zeros(c_out)
    sd[f"{prefix}.bn.running_var"] = torch.ones(c_out)
    sd[f"{prefix}.bn.num_batches_tracked"] = torch.tensor(0, dtype=torch.long)


def _calibrate_superpoint_bn(sd, images: torch.Tensor):
    """Set every BN's running statistics to the activation statistics it sees on `images`.

    Runs in float64 so the fp32-rounded result does not depend on host thread count or ISA.
    Layer semantics: conv -> ReLU -> BN(eval, eps=1e-3) (reference: superpoint_open.py:59-75).
    """
    x = images.to(torch.float64)

    def block(x, prefix, relu=True):
        w = sd[f"{prefix}.conv.weight"].to(torch.float64)
        b = sd[f"{prefix}.conv.bias"].to(torch.float64)
        x = F.conv2d(x, w, b, padding=(w.shape[-1] - 1) // 2)
        if relu:
            x = F.relu(x)
        mean = x.mean((0, 2, 3))
        var = x.var((0, 2, 3), unbiased=False)
        # snap to a 2^-20 grid: sums that cancel to ~1e-8 would otherwise carry reduction-order noise
        grid = float(1 << 20)
        sd[f"{prefix}.bn.running_mean"] = (torch.round(mean * grid) / grid).to(torch.float32)
        sd[f"{prefix}.bn.running_var"] = (torch.round(var * grid) / grid).clamp_min(1e-4).to(torch.float32)
        m32 = sd[f"{prefix}.bn.running_mean"].to(torch.float64)
        v32 = sd[f"{prefix}.bn.running_var"].to(torch.float64)
        g = sd[f"{prefix}.bn.weight"].to(torch.float64)
        beta = sd[f"{prefix}.bn.bias"].to(torch.float64)
        return (x - m32[None, :, None, None]) / torch.sqrt(v32[None, :, None, None] + 1e-3) * g[
            None, :, None, None
        ] + beta[None, :, None, None]

    for stage in range(4):
        x = block(x, f"backbone.{stage}.0")
        x = block(x, f"backbone.{stage}.1")
        if stage < 3:
            x = F.max_pool2d(x, 2, 2)
    d = block(x, "descriptor.0")
    block(d, "descriptor.1", relu=False)
    s = block(x, "detector.0")
    block(s, "detector.1", relu=False)


def superpoint_open_state_dict(seed: int = 0, detector_gain: float = 1.5) -> "OrderedDict[str, torch.Tensor]":
    """Conditioned SuperPoint-open weights, 84 tensors in the reference's key order."""
    rng = np.random.default_rng(1000 + seed)
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    chans = [1] + SP_CHANNELS[:-1]
    for stage in range(4):
        _vgg_block_entries(sd, f"backbone.{stage}.0", rng, chans[stage], chans[stage + 1], 3, math.sqrt(2.0))
        _vgg_block_entries(sd, f"backbone.{stage}.1", rng, chans[stage + 1], chans[stage + 1], 3, math.sqrt(2.0))
    c = SP_CHANNELS[-1]
    _vgg_block_entries(sd, "detector.0", rng, chans[-1], c, 3, math.sqrt(2.0))
    _vgg_block_entries(sd, "detector.1", rng, c, 65, 1, 1.0)
    _vgg_block_entries(sd, "descriptor.0", rng, chans[-1], c, 3, math.sqrt(2.0))
    _vgg_block_entries(sd, "descriptor.1", rng, c, 256, 1, 1.0)
    # calibration set: two smooth images and two hard-edged ones (thresholded noise), so that the
    # statistics also cover natural images with step edges (the boat pair of configuration 1)
    smooth = [synthetic_image(900 + i, 96, 96) for i in range(4)]
    edged = [(im > 0.5).to(torch.float32) * 0.8 + 0.1 * smooth[i - 2] for i, im in enumerate(smooth[2:], 2)]
    calib = torch.stack(smooth[:2] + edged)
    _calibrate_superpoint_bn(sd, calib)
    # peaked 65-way softmax: scale the (BN-normalised) detector logits
    sd["detector.1.bn.weight"] = torch.full((65,), float(detector_gain))
    return sd


def _linear_entries(sd, prefix, rng, out_f, in_f, gain=1.0, bias_std=0.0):
    sd[f"{prefix}.weight"] = _normal(rng, (out_f, in_f), gain / math.sqrt(in_f))
    sd[f"{prefix}.bias"] = _normal(rng, (out_f,), bias_std) if bias_std > 0 else torch.zeros(out_f)


def lightglue_state_dict(
    seed: int = 0,
    n_layers: int = 9,
    dim: int = 256,
    qk_gain: float = 16.0,
    final_gain: float = 14.0,
    ffn_out_gain: float = 0.04,
) -> "OrderedDict[str, torch.Tensor]":
    """Conditioned LightGlue weights in the reference's key order (lightglue.py:333-357)."""
    rng = np.random.default_rng(2000 + seed)
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    head_dim = dim // 4
    sd["posenc.Wr.weight"] = _normal(rng, (head_dim // 2, 2), 1.0)
    for i in range(n_layers):
        p = f"transformers.{i}.self_attn"
        _linear_entries(sd, f"{p}.Wqkv", rng, 3 * dim, dim, qk_gain, 0.02)
        _linear_entries(sd, f"{p}.out_proj", rng, dim, dim, 1.0, 0.02)
        _linear_entries(sd, f"{p}.ffn.0", rng, 2 * dim, 2 * dim, 1.0, 0.02)
        sd[f"{p}.ffn.1.weight"] = torch.ones(2 * dim) + _normal(rng, (2 * dim,), 0.05)
        sd[f"{p}.ffn.1.bias"] = _normal(rng, (2 * dim,), 0.05)
        _linear_entries(sd, f"{p}.ffn.3", rng, dim, 2 * dim, ffn_out_gain, 0.002)
        p = f"transformers.{i}.cross_attn"
        _linear_entries(sd, f"{p}.to_qk", rng, dim, dim, qk_gain, 0.02)
        _linear_entries(sd, f"{p}.to_v", rng, dim, dim, 1.0, 0.02)
        _linear_entries(sd, f"{p}.to_out", rng, dim, dim, 1.0, 0.02)
        _linear_entries(sd, f"{p}.ffn.0", rng, 2 * dim, 2 * dim, 1.0, 0.02)
        sd[f"{p}.ffn.1.weight"] = torch.ones(2 * dim) + _normal(rng, (2 * dim,), 0.05)
        sd[f"{p}.ffn.1.bias"] = _normal(rng, (2 * dim,), 0.05)
        _linear_entries(sd, f"{p}.ffn.3", rng, dim, 2 * dim, ffn_out_gain, 0.002)
    for i in range(n_layers):
        p = f"log_assignment.{i}"
        _linear_entries(sd, f"{p}.matchability", rng, 1, dim, 24.0, 0.0)
        sd[f"{p}.matchability.bias"] = torch.full((1,), 2.0)
        _linear_entries(sd, f"{p}.final_proj", rng, dim, dim, final_gain, 0.02)
    for i in range(n_layers - 1):
        p = f"token_confidence.{i}.token.0"
        _linear_entries(sd, p, rng, 1, dim, 24.0, 0.0)
        sd[f"{p}.bias"] = torch.full((1,), -2.0 + 5.0 * i / max(1, n_layers - 2))
    return sd


def superpoint_nonfree_state_dict(seed: int = 0, detector_gain: float = 2.0) -> "OrderedDict[str, torch.Tensor]":
    """Conditioned MagicLeap-SuperPoint weights (conv -> ReLU, no batch norm), 24 tensors with the
    reference's key names (reference: gluefactory_nonfree/superpoint.py:175-196)."""
    rng = np.random.default_rng(3000 + seed)
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()

    def conv(name, c_in, c_out, k, gain):
        sd[f"{name}.weight"] = _normal(rng, (c_out, c_in, k, k), gain / math.sqrt(c_in * k * k))
        sd[f"{name}.bias"] = _normal(rng, (c_out,), 0.02)

    g = math.sqrt(2.0)
    conv("conv1a", 1, 64, 3, 2.0 * g)  # images live in [0, 1]: lift the first layer
    conv("conv1b", 64, 64, 3, g)
    conv("conv2a", 64, 64, 3, g)
    conv("conv2b", 64, 64, 3, g)
    conv("conv3a", 64, 128, 3, g)
    conv("conv3b", 128, 128, 3, g)
    conv("conv4a", 128, 128, 3, g)
    conv("conv4b", 128, 128, 3, g)
    conv("convPa", 128, 256, 3, g)
    conv("convPb", 256, 65, 1, 1.0)  # rescaled below to a logit std of `detector_gain`
    conv("convDa", 128, 256, 3, g)
    conv("convDb", 256, 256, 1, 1.0)
    # Without batch norm the all-positive ReLU features share a large common component, which makes every
    # descriptor nearly parallel.  Centre the descriptor head on a small calibration set (float64, snapped
    # to a 2^-20 grid so the fp32 result does not depend on the host), and scale the detector logits.
    calib = torch.stack([synthetic_image(900 + i, 96, 96) for i in range(4)]).to(torch.float64)
    x = calib
    for stage, names in enumerate((("conv1a", "conv1b"), ("conv2a", "conv2b"), ("conv3a", "conv3b"), ("conv4a", "conv4b"))):
        for n in names:
            x = F.relu(F.conv2d(x, sd[f"{n}.weight"].double(), sd[f"{n}.bias"].double(), padding=1))
        if stage < 3:
            x = F.max_pool2d(x, 2, 2)
    grid = float(1 << 20)
    d = F.conv2d(F.relu(F.conv2d(x, sd["convDa.weight"].double(), sd["convDa.bias"].double(), padding=1)),
                 sd["convDb.weight"].double())
    sd["convDb.bias"] = (-torch.round(d.mean((0, 2, 3)) * grid) / grid).to(torch.float32)
    pa = F.relu(F.conv2d(x, sd["convPa.weight"].double(), sd["convPa.bias"].double(), padding=1))
    logits = F.conv2d(pa, sd["convPb.weight"].double())
    std = float(torch.round(logits.std() * 1024) / 1024)
    sd["convPb.weight"] = (sd["convPb.weight"].double() * (detector_gain / std)).to(torch.float32)
    sd["convPb.bias"] = (-torch.round(logits.mean((0, 2, 3)) * (detector_gain / std) * grid) / grid).to(torch.float32)
    return sd


def superglue_state_dict(
    seed: int = 0,
    n_layers: int = 18,
    dim: int = 256,
    qk_gain: float = 12.0,
    final_gain: float = 12.0,
    mlp_out_gain: float = 0.02,
    bin_score: float = 1.0,
) -> "OrderedDict[str, torch.Tensor]":
    """Conditioned SuperGlue weights with the reference's key names and Conv1d shapes
    (reference: gluefactory_nonfree/superglue.py:63-72,89-140,230-245)."""
    rng = np.random.default_rng(4000 + seed)
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()

    def conv1d(name, c_out, c_in, gain=1.0, bias_std=0.02):
        sd[f"{name}.weight"] = _normal(rng, (c_out, c_in, 1), gain / math.sqrt(c_in))
        sd[f"{name}.bias"] = _normal(rng, (c_out,), bias_std) if bias_std > 0 else torch.zeros(c_out)

    def bn1d(name, c):
        sd[f"{name}.weight"] = torch.ones(c) + _normal(rng, (c,), 0.05)
        sd[f"{name}.bias"] = _normal(rng, (c,), 0.05)
        sd[f"{name}.running_mean"] = _normal(rng, (c,), 0.05)
        sd[f"{name}.running_var"] = torch.ones(c) + _uniform(rng, (c,), -0.2, 0.2).to(torch.float32)
        sd[f"{name}.num_batches_tracked"] = torch.tensor(0, dtype=torch.long)

    chans = [3, 32, 64, 128, 256, dim]
    for i in range(1, len(chans)):
        last = i == len(chans) - 1
        conv1d(f"kenc.encoder.{3 * (i - 1)}", chans[i], chans[i - 1], 0.25 if last else math.sqrt(2.0), 0.0 if last else 0.02)
        if not last:
            bn1d(f"kenc.encoder.{3 * (i - 1) + 1}", chans[i])
    for i in range(n_layers):
        p = f"gnn.layers.{i}"
        conv1d(f"{p}.attn.merge", dim, dim)
        conv1d(f"{p}.attn.proj.0", dim, dim, qk_gain)
        conv1d(f"{p}.attn.proj.1", dim, dim, qk_gain)
        conv1d(f"{p}.attn.proj.2", dim, dim)
        conv1d(f"{p}.mlp.0", 2 * dim, 2 * dim, math.sqrt(2.0))
        bn1d(f"{p}.mlp.1", 2 * dim)
        conv1d(f"{p}.mlp.3", dim, 2 * dim, mlp_out_gain, 0.0)
    conv1d("final_proj", dim, dim, final_gain)
    sd["bin_score"] = torch.tensor(float(bin_score))
    return sd


def state_dict_sha256(sd) -> str:
    h = hashlib.sha256()
    for key, value in sd.items():
        h.update(key.encode())
        h.update(value.detach().cpu().contiguous().numpy().tobytes())
    return h.hexdigest()
