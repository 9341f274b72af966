"""Oracle: MagicLeap SuperPoint (gluefactory_nonfree) forward, restated functionally on a
reference-schema state dict (conv1a ... convDb, no batch norm).

Test infrastructure only (see oracle/__init__.py).  Written from the behaviour of
/root/reference/gluefactory_nonfree/superpoint.py (citations below); no text of that file is used.
NMS, border handling in the full-image case, thresholding and top-k share the SuperPoint-open oracle.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import sp_open


def _conv(sd, name, x, relu=True):
    w, b = sd[f"{name}.weight"], sd[f"{name}.bias"]
    x = F.conv2d(x, w, b, padding=(w.shape[-1] - 1) // 2)
    return F.relu(x) if relu else x


def backbone(sd, image):
    """conv -> ReLU pairs with 2x2 max-pools after the first three pairs (superpoint.py:209-219)."""
    x = image
    for stage, names in enumerate((("conv1a", "conv1b"), ("conv2a", "conv2b"), ("conv3a", "conv3b"), ("conv4a", "conv4b"))):
        for n in names:
            x = _conv(sd, n, x)
        if stage < 3:
            x = F.max_pool2d(x, 2, 2)
    return x


def dense_heads(sd, feats):
    """Score map (softmax over 65, dustbin dropped, depth-to-space) and unit-norm dense descriptors
    (superpoint.py:222-238)."""
    logits = _conv(sd, "convPb", _conv(sd, "convPa", feats), relu=False)
    heat = sp_open.decode_scores(logits, 8)
    dense = F.normalize(_conv(sd, "convDb", _conv(sd, "convDa", feats), relu=False), p=2, dim=1)
    return heat, dense


def remove_borders(scores, pad, image_size=None):
    """Top / left frame always; bottom / right frame from the per-image (w, h) when given, else from
    the map size (superpoint.py:244-254)."""
    if not pad:
        return scores
    scores = scores.clone()
    scores[:, :pad] = -1
    scores[:, :, :pad] = -1
    if image_size is not None:
        for i in range(scores.shape[0]):
            w, h = image_size[i]
            scores[i, int(h.item()) - pad:] = -1
            scores[i, :, int(w.item()) - pad:] = -1
    else:
        scores[:, -pad:] = -1
        scores[:, :, -pad:] = -1
    return scores


def sample_descriptors(keypoints, dense, stride=8, legacy=True):
    """keypoints: integer pixel positions (x, y) as floats.  legacy: the original MagicLeap grid with
    align_corners=True (superpoint.py:117-133); else the corrected one (:138-149)."""
    b, c, h, w = dense.shape
    if legacy:
        k = keypoints - stride / 2 + 0.5
        k = k / keypoints.new_tensor([w * stride - stride / 2 - 0.5, h * stride - stride / 2 - 0.5])
        g = k * 2 - 1
        d = F.grid_sample(dense, g.view(b, 1, -1, 2), mode="bilinear", align_corners=True)
    else:
        g = keypoints / (keypoints.new_tensor([w, h]) * stride) * 2 - 1
        d = F.grid_sample(dense, g.view(b, 1, -1, 2), mode="bilinear", align_corners=False)
    return F.normalize(d.reshape(b, c, -1), p=2, dim=1)


DEFAULT_CONF = dict(nms_radius=4, detection_threshold=0.005, max_num_keypoints=-1, remove_borders=4,
                    legacy_sampling=True)


def forward(sd, image, conf=None, image_size=None):
    """Sparse outputs (superpoint.py:202-350) for batches whose images yield equal keypoint counts."""
    c = dict(DEFAULT_CONF)
    c.update(conf or {})
    image = sp_open.to_gray(image)
    heat, dense = dense_heads(sd, backbone(sd, image))
    scores = remove_borders(sp_open.nms(heat, c["nms_radius"]), c["remove_borders"], image_size)
    k = c["max_num_keypoints"] if c["max_num_keypoints"] and c["max_num_keypoints"] > 0 else None
    kps, scs = [], []
    for i in range(image.shape[0]):
        xy, sc, _ = sp_open.select_keypoints(scores[i], c["detection_threshold"], k)
        kps.append(xy)
        scs.append(sc)
    keypoints, kscores = torch.stack(kps, 0), torch.stack(scs, 0)
    desc = sample_descriptors(keypoints, dense, 8, c["legacy_sampling"])
    return {
        "keypoints": keypoints + 0.5,
        "keypoint_scores": kscores,
        "descriptors": desc.transpose(-1, -2),
        "dense_scores": heat,
        "dense_descriptors": dense,
    }
