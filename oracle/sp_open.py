"""Oracle: SuperPoint-open forward, restated functionally on a reference-schema state dict.

Test infrastructure only (see oracle/__init__.py).  Citations are to
/root/reference/gluefactory/models/extractors/superpoint_open.py.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

BN_EPS = 1e-3  # superpoint_open.py:66


def vgg_block(sd, prefix, x, relu=True):
    """conv -> (ReLU) -> BatchNorm in eval mode; note BN comes AFTER the ReLU (:59-75)."""
    w, b = sd[f"{prefix}.conv.weight"], sd[f"{prefix}.conv.bias"]
    x = F.conv2d(x, w, b, stride=1, padding=(w.shape[-1] - 1) // 2)
    if relu:
        x = F.relu(x)
    return F.batch_norm(
        x,
        sd[f"{prefix}.bn.running_mean"],
        sd[f"{prefix}.bn.running_var"],
        sd[f"{prefix}.bn.weight"],
        sd[f"{prefix}.bn.bias"],
        training=False,
        eps=BN_EPS,
    )


def to_gray(image):
    """RGB -> gray with the fixed luma weights (:125-127); single-channel passes through."""
    if image.shape[1] == 3:
        w = image.new_tensor([0.299, 0.587, 0.114]).view(1, 3, 1, 1)
        image = (image * w).sum(1, keepdim=True)
    return image


def backbone(sd, image, n_stages=4):
    """Four stages of two VGG blocks; 2x2 max-pool after the first three (:99-105,128)."""
    x = image
    for s in range(n_stages):
        x = vgg_block(sd, f"backbone.{s}.0", x)
        x = vgg_block(sd, f"backbone.{s}.1", x)
        if s < n_stages - 1:
            x = F.max_pool2d(x, kernel_size=2, stride=2)
    return x


def descriptor_head(sd, features):
    """3x3 block, 1x1 block without ReLU, then channel-wise L2 normalisation (:112-115,129-131)."""
    d = vgg_block(sd, "descriptor.1", vgg_block(sd, "descriptor.0", features), relu=False)
    return F.normalize(d, p=2, dim=1)


def detector_logits(sd, features):
    """(:108-111,134)"""
    return vgg_block(sd, "detector.1", vgg_block(sd, "detector.0", features), relu=False)


def decode_scores(logits, stride=8):
    """65-way softmax, drop the dustbin, depth-to-space (:135-140)."""
    prob = F.softmax(logits, 1)[:, :-1]
    b, _, h, w = prob.shape
    prob = prob.permute(0, 2, 3, 1).reshape(b, h, w, stride, stride)
    return prob.permute(0, 1, 3, 2, 4).reshape(b, h * stride, w * stride)


def nms(scores, radius):
    """Plateau-preserving iterative max-pool NMS with two refinement rounds (:34-49)."""
    assert radius >= 0
    k = 2 * radius + 1

    def pool(x):
        return F.max_pool2d(x, kernel_size=k, stride=1, padding=radius)

    keep = scores == pool(scores)
    for _ in range(2):
        near_kept = pool(keep.float()) > 0
        rest = scores.masked_fill(near_kept, 0.0)
        keep = keep | ((rest == pool(rest)) & ~near_kept)
    return scores * keep  # where(keep, scores, 0)


def remove_borders(scores, pad):
    """Frame of `pad` pixels set to -1 (:147-152)."""
    if pad:
        scores = scores.clone()
        scores[:, :pad] = -1
        scores[:, :, :pad] = -1
        scores[:, -pad:] = -1
        scores[:, :, -pad:] = -1
    return scores


def select_keypoints(score_map, threshold, max_kpts):
    """Candidates above threshold in row-major order; top-k by score when more than k (:52-56,155-179).

    Tie order is made total (score descending, then linear index ascending); `torch.topk`
    leaves it unspecified.  Returns (xy float (n,2) WITHOUT the +0.5, scores (n,), linear idx).
    """
    h, w = score_map.shape
    flat = score_map.reshape(-1)
    idx = torch.nonzero(flat > threshold).squeeze(1)
    sc = flat[idx]
    if max_kpts is not None and max_kpts < idx.numel():
        order = torch.sort(sc, descending=True, stable=True).indices[:max_kpts]
        idx, sc = idx[order], sc[order]
    xy = torch.stack([idx % w, idx // w], -1).float()
    return xy, sc, idx


def sample_descriptors(keypoints, dense, stride=8):
    """Bilinear sampling at keypoints (pixel units, before the +0.5) then L2-normalise (:20-31)."""
    b, c, h, w = dense.shape
    g = (keypoints + 0.5) / (keypoints.new_tensor([w, h]) * stride)
    g = g * 2 - 1
    d = F.grid_sample(dense, g.view(b, 1, -1, 2), mode="bilinear", align_corners=False)
    return F.normalize(d.reshape(b, c, -1), p=2, dim=1)


DEFAULT_CONF = dict(
    nms_radius=4, max_num_keypoints=None, detection_threshold=0.005, remove_borders=4, stride=8
)


def forward(sd, image, conf=None, return_stages=False):
    """End-to-end (:123-217) for batches whose images yield equal keypoint counts.

    Returns the reference's prediction dict; with `return_stages` also every stage tensor
    used by the teacher-forced kernel tests.
    """
    c = dict(DEFAULT_CONF)
    c.update(conf or {})
    image = to_gray(image)
    feats = backbone(sd, image)
    dense = descriptor_head(sd, feats)
    logits = detector_logits(sd, feats)
    heat = decode_scores(logits, c["stride"])
    suppressed = nms(heat, c["nms_radius"])
    bordered = remove_borders(suppressed, c["remove_borders"])
    kps, scs, idxs = [], [], []
    for i in range(image.shape[0]):
        xy, sc, idx = select_keypoints(bordered[i], c["detection_threshold"], c["max_num_keypoints"])
        kps.append(xy)
        scs.append(sc)
        idxs.append(idx)
    keypoints = torch.stack(kps, 0)
    scores = torch.stack(scs, 0)
    desc = sample_descriptors(keypoints, dense, c["stride"])
    pred = {
        "keypoints": keypoints + 0.5,
        "keypoint_scores": scores,
        "descriptors": desc.transpose(-1, -2),
        "heatmap": heat,
    }
    if return_stages:
        pred["_stages"] = {
            "features": feats,
            "dense_descriptors": dense,
            "logits": logits,
            "nms": suppressed,
            "bordered": bordered,
            "linear_index": torch.stack(idxs, 0),
        }
    return pred
