"""Oracle: SuperGlue forward (keypoint encoder, attentional GNN, log-domain Sinkhorn, mutual-NN
filter), restated functionally on a reference-schema state dict.

Test infrastructure only (see oracle/__init__.py).  Written from the behaviour of
/root/reference/gluefactory_nonfree/superglue.py (citations below); no text of that file is used.
Tensors are channel-first (B, C, N) as in the reference.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

BN_EPS = 1e-5  # nn.BatchNorm1d default


def _conv1d(sd, name, x):
    return F.conv1d(x, sd[f"{name}.weight"], sd[f"{name}.bias"])


def _bn(sd, name, x):
    return F.batch_norm(x, sd[f"{name}.running_mean"], sd[f"{name}.running_var"], sd[f"{name}.weight"],
                        sd[f"{name}.bias"], training=False, eps=BN_EPS)


def mlp(sd, prefix, x, n_layers):
    """Conv1d(k=1) layers with BatchNorm1d + ReLU between them, none after the last (superglue.py:63-72)."""
    for i in range(n_layers):
        x = _conv1d(sd, f"{prefix}.{3 * i}", x)
        if i < n_layers - 1:
            x = F.relu(_bn(sd, f"{prefix}.{3 * i + 1}", x))
    return x


def normalize_keypoints(kpts, size):
    """(kpts - size / 2) / (0.7 * max(size)), size = (w, h) per batch element (superglue.py:75-86)."""
    shift = size.float() / 2
    scale = size.max(1).values.float() * 0.7
    return (kpts - shift[:, None]) / scale[:, None, None]


def keypoint_encoder(sd, kpts, scores):
    """MLP on (x, y, score) (superglue.py:89-102)."""
    return mlp(sd, "kenc.encoder", torch.cat([kpts.transpose(1, 2), scores.unsqueeze(1)], 1), 5)


def multi_head_attention(sd, prefix, x, source, heads=4):
    """q/k/v projections, channels interleaved as (dim, head); softmax(q.k / sqrt(dim)) v; merge
    (superglue.py:105-128)."""
    b, d_model, _ = x.shape
    dim = d_model // heads
    q = _conv1d(sd, f"{prefix}.proj.0", x).view(b, dim, heads, -1)
    k = _conv1d(sd, f"{prefix}.proj.1", source).view(b, dim, heads, -1)
    v = _conv1d(sd, f"{prefix}.proj.2", source).view(b, dim, heads, -1)
    prob = F.softmax(torch.einsum("bdhn,bdhm->bhnm", q, k) / dim**0.5, dim=-1)
    out = torch.einsum("bhnm,bdhm->bdhn", prob, v)
    return _conv1d(sd, f"{prefix}.merge", out.contiguous().view(b, d_model, -1))


def propagation(sd, prefix, x, source):
    """message = attention(x, source, source); mlp(cat[x, message]) (superglue.py:131-140)."""
    message = multi_head_attention(sd, f"{prefix}.attn", x, source)
    return mlp(sd, f"{prefix}.mlp", torch.cat([x, message], 1), 2)


def gnn(sd, desc0, desc1, layer_names):
    """Alternating self / cross layers with residual updates (superglue.py:143-170)."""
    for i, name in enumerate(layer_names):
        p = f"gnn.layers.{i}"
        if name == "self":
            d0, d1 = propagation(sd, p, desc0, desc0), propagation(sd, p, desc1, desc1)
        else:
            d0, d1 = propagation(sd, p, desc0, desc1), propagation(sd, p, desc1, desc0)
        desc0, desc1 = desc0 + d0, desc1 + d1
    return desc0, desc1


def log_optimal_transport(cost, alpha, iters):
    """Dustbin-augmented couplings, `iters` log-domain Sinkhorn iterations, result times (M + N)
    (superglue.py:173-201)."""
    b, m, n = cost.shape
    z = torch.cat([torch.cat([cost, alpha.expand(b, m, 1)], -1), alpha.expand(b, 1, n + 1)], 1)
    norm = -math.log(m + n)
    log_mu = torch.cat([torch.full((m,), norm), torch.tensor([math.log(n) + norm])]).to(cost)[None].expand(b, -1)
    log_nu = torch.cat([torch.full((n,), norm), torch.tensor([math.log(m) + norm])]).to(cost)[None].expand(b, -1)
    u, v = torch.zeros_like(log_mu), torch.zeros_like(log_nu)
    for _ in range(iters):
        u = log_mu - torch.logsumexp(z + v.unsqueeze(1), dim=2)
        v = log_nu - torch.logsumexp(z + u.unsqueeze(2), dim=1)
    return z + u.unsqueeze(2) + v.unsqueeze(1) - norm


def filter_matches(scores, threshold):
    """Mutual nearest neighbours on the inner block, score = exp(max), threshold (superglue.py:288-298).
    arg-max ties resolve to the lowest index."""
    inner = scores[:, :-1, :-1]
    max0, max1 = inner.max(2), inner.max(1)
    m0, m1 = max0.indices, max1.indices
    ar0 = torch.arange(m0.shape[1])[None]
    ar1 = torch.arange(m1.shape[1])[None]
    mutual0 = ar0 == m1.gather(1, m0)
    mutual1 = ar1 == m0.gather(1, m1)
    ms0 = torch.where(mutual0, max0.values.exp(), torch.zeros(()))
    ms1 = torch.where(mutual1, ms0.gather(1, m1), torch.zeros(()))
    valid0 = mutual0 & (ms0 > threshold)
    valid1 = mutual1 & valid0.gather(1, m1)
    return torch.where(valid0, m0, -1), torch.where(valid1, m1, -1), ms0, ms1


DEFAULT_CONF = dict(GNN_layers=["self", "cross"] * 9, num_sinkhorn_iterations=50, filter_threshold=0.2,
                    descriptor_dim=256)


def forward(sd, kpts0, kpts1, desc0, desc1, scores0, scores1, size0, size1, conf=None):
    """SuperGlue._forward (superglue.py:253-307).  desc: (B, N, 256); size: (B, 2) = (w, h)."""
    c = dict(DEFAULT_CONF)
    c.update(conf or {})
    d0, d1 = desc0.transpose(-1, -2), desc1.transpose(-1, -2)
    d0 = d0 + keypoint_encoder(sd, normalize_keypoints(kpts0, size0), scores0)
    d1 = d1 + keypoint_encoder(sd, normalize_keypoints(kpts1, size1), scores1)
    d0, d1 = gnn(sd, d0, d1, c["GNN_layers"])
    m0, m1 = _conv1d(sd, "final_proj", d0), _conv1d(sd, "final_proj", d1)
    cost = torch.einsum("bdn,bdm->bnm", m0, m1) / c["descriptor_dim"] ** 0.5
    scores = log_optimal_transport(cost, sd["bin_score"], c["num_sinkhorn_iterations"])
    mt0, mt1, ms0, ms1 = filter_matches(scores, c["filter_threshold"])
    return {
        "sinkhorn_cost": cost,
        "log_assignment": scores,
        "matches0": mt0,
        "matches1": mt1,
        "matching_scores0": ms0,
        "matching_scores1": ms1,
        "_desc0": d0,
        "_desc1": d1,
    }
