"""Adaptive depth (early exit) and width (point pruning) for the B200 LightGlue, any batch size, with every
decision taken on the device.

Semantics follow the reference's eval branch (gluefactory/models/matchers/lightglue.py:470-511) with
`confidence_thresholds[i] = confidence_threshold(i)` (:527-530), which the reference forgets to assign, and the
exit-layer descriptors returned as `ref_descriptors` (the reference's empty `torch.stack` would raise, :518).  The
reference asserts batch size 1 and reads the stop decision and the kept indices on the host after every layer.
Here each pair of the batch carries (live lengths, active flag, exit layer) in device memory:

  * token rows keep a fixed layout (image 0 of pair b: rows [b M, (b+1) M), image 1: [B M + b N, ...)); pruning
    compacts inside a segment, the live tokens are its first m_b / n_b rows;
  * the attention kernels take the live lengths through `len_override`, the GEMMs skip 128-row tiles without a live
    token through `tile_live`, so executed work follows the surviving tokens without the host knowing them;
  * a pair that satisfies `check_if_stop` after layer i is projected with head i and snapshot (projected
    descriptors, matchability logits, residual stream, index map, lengths); the other pairs go on;
  * one variable-size assignment (`b200_lg_assignment_live`) and one scatter close the pass.

There is no host read-back inside the layer loop.  The single synchronisation is at the very end, to slice the
outputs of a B = 1 call to the reference's (1, m'+1, n'+1) shapes and to report a fully pruned image.
For B > 1 (no reference counterpart: it asserts B == 1) `log_assignment` / `ref_descriptors` stay padded to
(B, M+1, N+1) / (B, 1, M, 256) with the live block first, and `kept0` / `kept1` give the live sizes; a B = 1 call
over each pair returns the identical numbers (tests/test_gpu_adaptive.py).
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
    dev = kpts0.device
    L = conf.n_layers
    early, prune = conf.depth_confidence > 0, conf.width_confidence > 0
    P = model._pack()
    T = B * (M + N)
    off1 = B * M
    i32 = dict(dtype=torch.int32, device=dev)
    f32 = dict(dtype=torch.float32, device=dev)

    ws = model._workspace(T, dev, zero=True)
    alt = {k: torch.zeros_like(ws[k]) for k in ("x32", "xm", "cs")} if prune else None  # compaction target (ping-pong)
    ops.lg_ingest(desc0, ws["x32"], ws["xm"], 512, row0=0)
    ops.lg_ingest(desc1, ws["x32"], ws["xm"], 512, row0=off1)
    ops.lg_posenc(kpts0.float().contiguous(), _size(size0, dev, B), P["wr"], ws["cs"], row0=0)
    ops.lg_posenc(kpts1.float().contiguous(), _size(size1, dev, B), P["wr"], ws["cs"], row0=off1)

    ind = torch.cat([torch.arange(M, **i32).repeat(B), torch.arange(N, **i32).repeat(B)])
    ind_alt = torch.zeros_like(ind)
    prune0 = torch.ones(B, M, dtype=torch.int64, device=dev)
    prune1 = torch.ones(B, N, dtype=torch.int64, device=dev)
    state = torch.cat([torch.full((B,), M, **i32), torch.full((B,), N, **i32), torch.ones(B, **i32),
                       torch.full((B,), L - 1, **i32), torch.zeros(B, **i32)])
    lens = torch.empty(2 * B, **i32)
    tile_live = torch.empty((T + 127) // 128, **i32)
    hist = torch.zeros(L, 2 * B, **i32)      # live lengths entering each layer (executed-FLOPs report)
    newc = torch.empty(2 * B, **i32)
    pos = torch.empty(T, **i32)
    conf_buf = torch.empty(T, **f32)
    match_buf = torch.empty(T, **f32)
    md_final = torch.zeros(T, 256, **f32)
    z_final = torch.zeros(T, **f32)
    x_final = torch.zeros(T, 256, **f32)
    ind_final = torch.zeros(T, **i32)
    final_len = torch.zeros(2 * B, **i32)

    seqs_self, seqs_cross = model._seqs(B, M, N, dev)
    ops.lg_adaptive_commit(state, None, B, M, N, lens, tile_live, hist[0])

    for i in range(L):
        # work figures are the padded upper bound; the executed share follows `hist` (LightGlue.adaptive_stats)
        model._layer(P["layers"][i], ws, T, seqs_self, seqs_cross, 2 * B, max(M, N),
                     work_self=1024.0 * B * (M * M + N * N), work_cross=1536.0 * B * M * N, lens=lens,
                     tile_live=tile_live)
        if i == L - 1:
            break
        thr = _f32(model.confidence_thresholds[i])
        if early:
            tok = P["token"][i]
            ops.lg_rowdot(ws["x32"], T, tok["w"], tok["b"], conf_buf, sigmoid=True)
            ops.lg_adaptive_decide(conf_buf, B, M, N, thr, _f32(conf.depth_confidence), i, state)
            # the pairs that stop here are matched with this layer's assignment head (:496)
            md32, z = model._project(P["assign"][i], ws, T, tile_live)
            ops.lg_adaptive_snapshot(state, B, M, N, 0, md32, md_final, z, z_final, ws["x32"], x_final, ind, ind_final,
                                     final_len)
        if prune:
            a = P["assign"][i]
            ops.lg_rowdot(ws["x32"], T, a["wm"], a["bm"], match_buf, sigmoid=True)
            ops.lg_adaptive_prune(match_buf, conf_buf if early else None, B, M, N, _f32(1 - conf.width_confidence), thr,
                                  state, pos, newc, ws, alt, ind, ind_alt, prune0, prune1)
            ops.lg_adaptive_commit(state, newc, B, M, N, lens, tile_live, hist[i + 1])
            for k in ("x32", "xm", "cs"):
                ws[k], alt[k] = alt[k], ws[k]
            ind, ind_alt = ind_alt, ind
        else:
            ops.lg_adaptive_commit(state, None, B, M, N, lens, tile_live, hist[i + 1])

    # pairs still active after the last layer use the last head
    md32, z = model._project(P["assign"][L - 1], ws, T, tile_live)
    ops.lg_adaptive_snapshot(state, B, M, N, 1, md32, md_final, z, z_final, ws["x32"], x_final, ind, ind_final, final_len)
    sim = model._similarity(md_final, B, M, N, T)
    scores, m0, m1, ms0, ms1 = ops.lg_assignment_live(sim, z_final[:off1], z_final[off1:], final_len,
                                                      conf.filter_threshold)
    if prune:
        out_m0 = torch.full((B, M), -1, dtype=torch.int64, device=dev)
        out_m1 = torch.full((B, N), -1, dtype=torch.int64, device=dev)
        out_ms0 = torch.zeros(B, M, **f32)
        out_ms1 = torch.zeros(B, N, **f32)
        ops.lg_adaptive_scatter(m0, m1, ms0, ms1, ind_final, final_len, B, M, N, out_m0, out_m1, out_ms0, out_ms1)
        m0, m1, ms0, ms1 = out_m0, out_m1, out_ms0, out_ms1
    else:
        prune0 = torch.full_like(ms0, L)
        prune1 = torch.full_like(ms1, L)
    model.adaptive_stats = {"hist": hist, "state": state, "final_len": final_len, "M": M, "N": N}  # device tensors

    # the one read-back of the pass: final live sizes (output shapes of a B = 1 call, fully pruned images)
    kept = final_len.tolist()
    if min(kept) == 0:
        raise RuntimeError("adaptive width pruned every keypoint of an image")
    ref0 = x_final[:off1].view(B, 1, M, 256)
    ref1 = x_final[off1:].view(B, 1, N, 256)
    if B == 1:
        m, n = kept
        scores = scores[:, : m + 1, : n + 1].contiguous()
        ref0, ref1 = ref0[:, :, :m], ref1[:, :, :n]
    pred = {
        "matches0": m0,
        "matches1": m1,
        "matching_scores0": ms0,
        "matching_scores1": ms1,
        "ref_descriptors0": ref0,
        "ref_descriptors1": ref1,
        "log_assignment": scores,
        "prune0": prune0,
        "prune1": prune1,
    }
    if B > 1:
        pred["kept0"], pred["kept1"] = final_len[:B], final_len[B:]
    return pred


def executed_flops(stats, n_layers=9):
    """Algorithmic FLOPs a pass actually executed, from the per-layer live lengths (per layer
    (m+n) * 2 490 368 of projections / FFN + 1024 (m^2 + n^2) self + 1536 m n cross; assignment (m+n) 131 072 + 512 m n)
    against the same sums at the full sizes.  One host read-back."""
    hist = stats["hist"].cpu().numpy().astype(np.float64)
    fin = stats["final_len"].cpu().numpy().astype(np.float64)
    B = hist.shape[1] // 2
    M, N = float(stats["M"]), float(stats["N"])

    def layer(m, n):
        return (m + n) * 2490368.0 + 1024.0 * (m * m + n * n) + 1536.0 * m * n

    def assign(m, n):
        return (m + n) * 131072.0 + 512.0 * m * n

    done = sum(layer(hist[i, :B], hist[i, B:]).sum() for i in range(hist.shape[0])) + assign(fin[:B], fin[B:]).sum()
    full = B * (n_layers * layer(M, N) + assign(M, N))
    layers_run = (hist[:, :B] > 0).sum(0)
    return {"executed_gflop": done / 1e9, "full_gflop": full / 1e9, "share": done / full,
            "layers_run_mean": float(layers_run.mean()), "kept_mean": float((fin[:B] + fin[B:]).mean() / (M + N))}
