"""Adaptive depth (early exit) and width (point pruning) for the B200 LightGlue, batch size 1.

Semantics follow the reference's eval branch (gluefactory/models/matchers/lightglue.py:470-511) with
`confidence_thresholds[i] = confidence_threshold(i)` (:527-530), which the reference forgets to
assign, and the exit-layer descriptors returned as `ref_descriptors` (the reference's empty
`torch.stack` would raise, :518).  Like the reference, the stop decision and the new keypoint
counts are read on the host (12 bytes per layer); all arithmetic is in kernels: token-confidence /
matchability logits (`rowdot`), `count_below`, ordered compaction (`prune`), `scatter_matches`.
"""
import numpy as np
import torch

from .. import ops


def _f32(x: float) -> float:
    return float(np.float32(x))


def adaptive_forward(model, data, kpts0, kpts1, desc0, desc1, size0, size1):
    from .lightglue import _size

    conf = model.conf
    B, M, _ = kpts0.shape
    N = kpts1.shape[1]
    assert B == 1  # lightglue.py:474,479
    dev = kpts0.device
    L = conf.n_layers
    early, prune = conf.depth_confidence > 0, conf.width_confidence > 0
    P = model._pack()
    T = M + N
    ws = model._workspace(T, dev)
    alt = {k: torch.empty_like(ws[k]) for k in ("x32", "xm", "cs")}  # compaction target (ping-pong)
    ops.lg_ingest(desc0, ws["x32"], ws["xm"], 512, row0=0)
    ops.lg_ingest(desc1, ws["x32"], ws["xm"], 512, row0=M)
    ops.lg_posenc(kpts0.float().contiguous(), _size(size0, dev, 1), P["wr"], ws["cs"], row0=0)
    ops.lg_posenc(kpts1.float().contiguous(), _size(size1, dev, 1), P["wr"], ws["cs"], row0=M)

    ind = torch.cat([torch.arange(M, dtype=torch.int32), torch.arange(N, dtype=torch.int32)]).to(dev)
    ind_alt = torch.empty_like(ind)
    prune0 = torch.ones(1, M, dtype=torch.int64, device=dev)
    prune1 = torch.ones(1, N, dtype=torch.int64, device=dev)
    pos = torch.empty(T, dtype=torch.int32, device=dev)
    counts = torch.zeros(4, dtype=torch.int32, device=dev)
    conf_buf = torch.empty(T, dtype=torch.float32, device=dev)
    match_buf = torch.empty(T, dtype=torch.float32, device=dev)

    m, n = M, N  # live keypoint counts
    exit_layer = L - 1
    for i in range(L):
        t = m + n
        seqs_self = ops.make_attn_seqs([(0, m, 0, m), (m, n, m, n)], dev)
        seqs_cross = ops.make_attn_seqs([(0, m, m, n), (m, n, 0, m)], dev)
        model._layer(P["layers"][i], ws, t, seqs_self, seqs_cross, 2, max(m, n),
                     work_self=1024.0 * (m * m + n * n), work_cross=1536.0 * m * n)
        if i == L - 1:
            break
        thr = _f32(model.confidence_thresholds[i])
        have_conf = False
        if early:
            tok = P["token"][i]
            ops.lg_rowdot(ws["x32"], t, tok["w"], tok["b"], conf_buf, sigmoid=True)
            have_conf = True
            ops.lg_count_below(conf_buf, t, thr, counts[2:3])
            below = int(counts[2].item())  # host decision, as in check_if_stop (:541-552)
            ratio = torch.tensor(1.0) - torch.tensor(float(below)) / (M + N)
            if bool(ratio > conf.depth_confidence):
                exit_layer = i
                break
        if prune:
            a = P["assign"][i]
            ops.lg_rowdot(ws["x32"], t, a["wm"], a["bm"], match_buf, sigmoid=True)
            ops.lg_prune(match_buf, conf_buf if have_conf else None, m, n, _f32(1 - conf.width_confidence), thr, pos,
                         counts, ws, alt, ind, ind_alt, prune0, prune1)
            m, n = (int(v) for v in counts[:2].tolist())
            for k in ("x32", "xm", "cs"):
                ws[k], alt[k] = alt[k], ws[k]
            ind, ind_alt = ind_alt, ind
            if m == 0 or n == 0:
                raise RuntimeError("adaptive width pruned every keypoint of an image")

    t = m + n
    scores, m0, m1, ms0, ms1, _ = model._assign(P["assign"][exit_layer], ws, 1, m, n, t)
    if prune:
        out_m0 = torch.full((1, M), -1, dtype=torch.int64, device=dev)
        out_m1 = torch.full((1, N), -1, dtype=torch.int64, device=dev)
        out_ms0 = torch.zeros(1, M, dtype=torch.float32, device=dev)
        out_ms1 = torch.zeros(1, N, dtype=torch.float32, device=dev)
        ops.lg_scatter_matches(m0, m1, ms0, ms1, ind, m, n, out_m0, out_m1, out_ms0, out_ms1)
        m0, m1, ms0, ms1 = out_m0, out_m1, out_ms0, out_ms1
    else:
        prune0 = torch.ones_like(ms0) * L
        prune1 = torch.ones_like(ms1) * L
    x = ws["x32"]
    return {
        "matches0": m0,
        "matches1": m1,
        "matching_scores0": ms0,
        "matching_scores1": ms1,
        "ref_descriptors0": x[:m].view(1, 1, m, 256),
        "ref_descriptors1": x[m:t].view(1, 1, n, 256),
        "log_assignment": scores,
        "prune0": prune0,
        "prune1": prune1,
    }
