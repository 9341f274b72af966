"""Oracle: LightGlue forward, restated functionally on a reference-schema state dict.

Test infrastructure only (see oracle/__init__.py).  Citations are to
/root/reference/gluefactory/models/matchers/lightglue.py.

Adaptive depth/width: the reference's branch raises as shipped (`confidence_thresholds`
is never assigned, :538,:550; empty `ref_descriptors` stack on early exit, :518).  The
oracle uses the formula of `confidence_threshold` (:527-530) for the thresholds and returns
the exit-layer descriptors as `ref_descriptors`; the golden generator patches the reference
instance the same way.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

N_HEADS = 4


def normalize_keypoints(kpts, size=None):
    """(:21-33) centre on size/2, divide by half the longer side; extent-based when size is None."""
    if size is None:
        size = 1 + kpts.max(-2).values - kpts.min(-2).values
    elif not isinstance(size, torch.Tensor):
        size = torch.tensor(size, device=kpts.device, dtype=kpts.dtype)
    size = size.to(kpts)
    shift = size / 2
    scale = size.max(-1).values / 2
    return (kpts - shift[..., None, :]) / scale[..., None, None]


def positional_encoding(sd, kpts):
    """(:46-59) returns (cos, sin), each (B,1,N,64) with every frequency duplicated x2."""
    proj = kpts @ sd["posenc.Wr.weight"].t()
    cos, sin = torch.cos(proj), torch.sin(proj)
    return cos.repeat_interleave(2, -1).unsqueeze(1), sin.repeat_interleave(2, -1).unsqueeze(1)


def _rotary(t, cos, sin):
    """(:36-43) pairs (2i,2i+1) -> (-x_{2i+1}, x_{2i})."""
    even, odd = t[..., 0::2], t[..., 1::2]
    rot = torch.stack((-odd, even), -1).flatten(-2)
    return t * cos + rot * sin


def _linear(sd, name, x):
    return F.linear(x, sd[f"{name}.weight"], sd[f"{name}.bias"])


def _ffn(sd, prefix, x, msg):
    """Linear(512,512) -> LayerNorm(512) -> GELU(erf) -> Linear(512,256) on cat[x,msg] (:138-143)."""
    h = _linear(sd, f"{prefix}.ffn.0", torch.cat([x, msg], -1))
    h = F.layer_norm(h, (h.shape[-1],), sd[f"{prefix}.ffn.1.weight"], sd[f"{prefix}.ffn.1.bias"])
    return _linear(sd, f"{prefix}.ffn.3", F.gelu(h))


def self_block(sd, prefix, x, enc):
    """(:144-157) Wqkv packs (head, dim, {q,k,v}) innermost-last."""
    b, n, _ = x.shape
    qkv = _linear(sd, f"{prefix}.Wqkv", x).view(b, n, N_HEADS, -1, 3).permute(0, 2, 1, 3, 4)
    q, k, v = qkv[..., 0], qkv[..., 1], qkv[..., 2]
    q, k = _rotary(q, *enc), _rotary(k, *enc)
    att = torch.softmax(q @ k.transpose(-1, -2) * q.shape[-1] ** -0.5, -1)
    ctx = (att @ v).transpose(1, 2).reshape(b, n, -1)
    return x + _ffn(sd, prefix, x, _linear(sd, f"{prefix}.out_proj", ctx))


def cross_block(sd, prefix, x0, x1):
    """(:186-215) one similarity, softmax along both axes, shared projections and FFN."""

    def heads(t):
        return t.view(t.shape[0], t.shape[1], N_HEADS, -1).transpose(1, 2)

    qk0, qk1 = heads(_linear(sd, f"{prefix}.to_qk", x0)), heads(_linear(sd, f"{prefix}.to_qk", x1))
    v0, v1 = heads(_linear(sd, f"{prefix}.to_v", x0)), heads(_linear(sd, f"{prefix}.to_v", x1))
    s = (qk0.shape[-1] ** -0.5) ** 0.5
    sim = (qk0 * s) @ (qk1 * s).transpose(-1, -2)
    m0 = torch.softmax(sim, -1) @ v1
    m1 = torch.softmax(sim.transpose(-1, -2), -1) @ v0

    def merge(t):
        return t.transpose(1, 2).reshape(t.shape[0], t.shape[2], -1)

    m0, m1 = _linear(sd, f"{prefix}.to_out", merge(m0)), _linear(sd, f"{prefix}.to_out", merge(m1))
    return x0 + _ffn(sd, prefix, x0, m0), x1 + _ffn(sd, prefix, x1, m1)


def transformer_layer(sd, i, d0, d1, enc0, enc1):
    """(:224-238)"""
    d0 = self_block(sd, f"transformers.{i}.self_attn", d0, enc0)
    d1 = self_block(sd, f"transformers.{i}.self_attn", d1, enc1)
    return cross_block(sd, f"transformers.{i}.cross_attn", d0, d1)


def log_assignment(sd, i, d0, d1):
    """(:250-281) similarity of projected descriptors, dual log-softmax + matchability terms."""
    p = f"log_assignment.{i}"
    m0, m1 = _linear(sd, f"{p}.final_proj", d0), _linear(sd, f"{p}.final_proj", d1)
    dim = m0.shape[-1]
    sim = (m0 / dim**0.25) @ (m1 / dim**0.25).transpose(-1, -2)
    z0, z1 = _linear(sd, f"{p}.matchability", d0), _linear(sd, f"{p}.matchability", d1)
    b, m, n = sim.shape
    out = sim.new_zeros(b, m + 1, n + 1)
    out[:, :m, :n] = (
        F.log_softmax(sim, 2) + F.log_softmax(sim, 1) + F.logsigmoid(z0) + F.logsigmoid(z1).transpose(1, 2)
    )
    out[:, :m, n] = F.logsigmoid(-z0.squeeze(-1))
    out[:, m, :n] = F.logsigmoid(-z1.squeeze(-1))
    return out, sim


def filter_matches(scores, th):
    """(:287-303) mutual nearest neighbours on the inner block, exp of the row maximum as score."""
    inner = scores[:, :-1, :-1]
    max0, max1 = inner.max(2), inner.max(1)
    m0, m1 = max0.indices, max1.indices
    i0 = torch.arange(m0.shape[1], device=m0.device)[None]
    i1 = torch.arange(m1.shape[1], device=m1.device)[None]
    mutual0 = i0 == m1.gather(1, m0)
    mutual1 = i1 == m0.gather(1, m1)
    e0 = max0.values.exp()
    ms0 = torch.where(mutual0, e0, e0.new_zeros(()))
    ms1 = torch.where(mutual1, ms0.gather(1, m1), e0.new_zeros(()))
    valid0 = mutual0 & (ms0 > th)
    valid1 = mutual1 & valid0.gather(1, m1)
    return torch.where(valid0, m0, -1), torch.where(valid1, m1, -1), ms0, ms1


def confidence_threshold(i, n_layers):
    """(:527-530)"""
    return float(np.clip(0.8 + 0.1 * np.exp(-4.0 * i / n_layers), 0, 1))


DEFAULT_CONF = dict(n_layers=9, depth_confidence=-1, width_confidence=-1, filter_threshold=0.0)


def forward(sd, kpts0, kpts1, desc0, desc1, size0=None, size1=None, conf=None, return_stages=False):
    """(:399-525) inference forward; adaptive depth/width require batch size 1 as in the reference."""
    c = dict(DEFAULT_CONF)
    c.update(conf or {})
    L = c["n_layers"]
    b, m, _ = kpts0.shape
    n = kpts1.shape[1]
    k0 = normalize_keypoints(kpts0.float(), size0)
    k1 = normalize_keypoints(kpts1.float(), size1)
    d0, d1 = desc0.contiguous(), desc1.contiguous()
    enc0, enc1 = positional_encoding(sd, k0), positional_encoding(sd, k1)
    early, prune = c["depth_confidence"] > 0, c["width_confidence"] > 0
    if prune:
        ind0, ind1 = torch.arange(m, device=kpts0.device)[None], torch.arange(n, device=kpts0.device)[None]
        prune0, prune1 = torch.ones_like(ind0), torch.ones_like(ind1)
    stages = []
    tok0 = tok1 = None
    for i in range(L):
        d0, d1 = transformer_layer(sd, i, d0, d1, enc0, enc1)
        if return_stages:
            stages.append((d0, d1))
        if i == L - 1:
            break
        if early:
            assert b == 1
            p = f"token_confidence.{i}.token.0"
            tok0 = torch.sigmoid(_linear(sd, p, d0)).squeeze(-1)
            tok1 = torch.sigmoid(_linear(sd, p, d1)).squeeze(-1)
            th = confidence_threshold(i, L)
            conf_all = torch.cat([tok0, tok1], -1)
            ratio = 1.0 - (conf_all < th).float().sum() / (m + n)
            if ratio > c["depth_confidence"]:
                break
        if prune:
            assert b == 1
            th = confidence_threshold(i, L)
            for side in (0, 1):
                d = d0 if side == 0 else d1
                tok = tok0 if side == 0 else tok1
                match = torch.sigmoid(_linear(sd, f"log_assignment.{i}.matchability", d)).squeeze(-1)
                keep = match > (1 - c["width_confidence"])
                if tok is not None:
                    keep = keep | (tok <= th)
                sel = torch.where(keep)[1]
                if side == 0:
                    ind0, d0 = ind0.index_select(1, sel), d0.index_select(1, sel)
                    enc0 = tuple(e.index_select(-2, sel) for e in enc0)
                    prune0[:, ind0] += 1
                else:
                    ind1, d1 = ind1.index_select(1, sel), d1.index_select(1, sel)
                    enc1 = tuple(e.index_select(-2, sel) for e in enc1)
                    prune1[:, ind1] += 1
    scores, sim = log_assignment(sd, i, d0, d1)
    m0, m1, ms0, ms1 = filter_matches(scores, c["filter_threshold"])
    if prune:
        m0_ = torch.full((b, m), -1, dtype=m0.dtype, device=m0.device)
        m1_ = torch.full((b, n), -1, dtype=m1.dtype, device=m0.device)
        m0_[:, ind0] = torch.where(m0 == -1, -1, ind1.gather(1, m0.clamp(min=0)))
        m1_[:, ind1] = torch.where(m1 == -1, -1, ind0.gather(1, m1.clamp(min=0)))
        ms0_, ms1_ = torch.zeros(b, m, device=m0.device), torch.zeros(b, n, device=m0.device)
        ms0_[:, ind0] = ms0
        ms1_[:, ind1] = ms1
        m0, m1, ms0, ms1 = m0_, m1_, ms0_, ms1_
    else:
        prune0 = torch.ones_like(ms0) * L
        prune1 = torch.ones_like(ms1) * L
    pred = {
        "matches0": m0,
        "matches1": m1,
        "matching_scores0": ms0,
        "matching_scores1": ms1,
        "ref_descriptors0": d0[:, None],
        "ref_descriptors1": d1[:, None],
        "log_assignment": scores,
        "prune0": prune0,
        "prune1": prune1,
    }
    if return_stages:
        pred["_stages"] = {"layers": stages, "enc0": enc0, "enc1": enc1, "sim": sim, "exit_layer": i}
    return pred
