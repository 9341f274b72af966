"""Launch wrappers: torch tensors in, raw pointers to the C ABI (include/b200glue.h).

torch is used for device memory and streams only; every computation here is a kernel of
libb200glue.so.  All wrappers enqueue on torch's current CUDA stream and never synchronise.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from . import _lib
from ._lib import (  # noqa: F401
    EPI_BIAS_BF16, EPI_CONV_BN_L2, EPI_CONV_BN_SOFTMAX_D2S, EPI_CONV_RELU_BN, EPI_F32, EPI_LN_GELU,
    EPI_QKV_ROTARY, EPI_RESID, AttnDesc, AttnSeq, GemmDesc,
)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def on_device(key):
    """Decorator for a model's `_forward(self, data)`: the kernels launch on the CURRENT device's stream and the
    library keeps per-device launch state, so the device of `data[key]` is made current for the whole pass
    (a model moved with `.to('cuda:1')` then works whatever the caller's current device is)."""

    def wrap(fn):
        def inner(self, data):
            t = data[key]
            if not t.is_cuda:
                return fn(self, data)  # the body reports CPU tensors itself (there is no CPU fallback)
            with torch.cuda.device(t.device):
                return fn(self, data)

        inner.__name__, inner.__doc__ = fn.__name__, fn.__doc__
        return inner

    return wrap


# ---- launch accounting (bench.py): number of kernels launched, optional CUDA-event timing per tag --
class _Stats:
    launches = 0
    profile = None  # None, or a list collecting (tag, work, start_event, end_event)
    only = None     # None = every launch is bracketed when profiling, else the set of tags to bracket


def reset_stats(profile=False, only=None):
    """`only`: bracket just these tags (two event records per launch cost ~1 us of stream time each, which
    shows in a step of 130 launches; the timed region of bench.py brackets the dominant kernel alone)."""
    _Stats.launches = 0
    _Stats.profile = [] if profile else None
    _Stats.only = set(only) if only else None


def launches():
    return _Stats.launches


def profile_records():
    return _Stats.profile or []


class _timed:
    """Counts `n` kernel launches; when profiling, brackets them with CUDA events on the current stream."""

    def __init__(self, tag, n=1, work=0.0):
        self.tag, self.n, self.work = tag, n, work

    def __enter__(self):
        _Stats.launches += self.n
        self.on = _Stats.profile is not None and (_Stats.only is None or self.tag in _Stats.only)
        if self.on:
            self.t0 = torch.cuda.Event(enable_timing=True)
            self.t1 = torch.cuda.Event(enable_timing=True)
            self.t0.record()

    def __exit__(self, *exc):
        if self.on:
            self.t1.record()
            _Stats.profile.append((self.tag, self.work, self.t0, self.t1))
        return False


def _ptr(t, offset_elems: int = 0):
    if t is None:
        return None
    return t.data_ptr() + offset_elems * t.element_size()


def _require_cuda(*tensors):
    cur = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise _lib.B200Error("b200glue kernels need CUDA tensors (there is no CPU fallback)")
        if cur is None:
            cur = torch.cuda.current_device()
        if t.device.index != cur:
            raise _lib.B200Error(f"tensor on cuda:{t.device.index} but the current device is cuda:{cur}: wrap the call "
                                 "in torch.cuda.device(tensor.device) (the models' forward does)")


def gemm(a, w, *, M, N, K, lda, block_n, epilogue, a_rows=None, a_col0=0, ldb=None, batch=1,
         a_batch_rows=0, b_batch_rows=0, b_row0=0, b_rows=None, conv=None, bias=None, scale=None, shift=None,
         relu=0, pool=0, alpha=1.0, resid=None, cossin=None, rot_cols=0, ln_gamma=None, ln_beta=None,
         ln_eps=1e-5, out16=None, out16_col0=0, ld16=0, out32=None, ld32=0, out32_batch_stride=0,
         cells=(0, 0), tile_live=None):
    """C = epilogue(A . W^T) on the tcgen05 path.  `a`/`w` are bf16 tensors; views are expressed with
    (a_col0, lda) / (b_row0, ldb) so that a column slice of a wider matrix can be an operand."""
    _require_cuda(a, w, out16, out32)
    d = GemmDesc()
    d.a = _ptr(a, a_col0)
    d.lda = lda
    d.a_rows = a_rows if a_rows is not None else a.numel() // lda
    ldb = ldb if ldb is not None else K
    d.b = _ptr(w, b_row0 * ldb)
    d.ldb = ldb
    d.b_rows = b_rows if b_rows is not None else (w.numel() // ldb - b_row0)
    d.M, d.N, d.K, d.batch = M, N, K, batch
    d.a_batch_rows, d.b_batch_rows = a_batch_rows, b_batch_rows
    d.block_n, d.epilogue = block_n, epilogue
    if conv is not None:
        d.conv, d.H, d.W, d.Cin = 1, conv[0], conv[1], conv[2]
    d.bias, d.scale, d.shift = _ptr(bias), _ptr(scale), _ptr(shift)
    d.relu, d.pool, d.alpha = relu, pool, alpha
    d.resid, d.cossin, d.rot_cols = _ptr(resid), _ptr(cossin), rot_cols
    d.ln_gamma, d.ln_beta, d.ln_eps = _ptr(ln_gamma), _ptr(ln_beta), ln_eps
    d.out16, d.ld16 = _ptr(out16, out16_col0), ld16
    d.out32, d.ld32, d.out32_batch_stride = _ptr(out32), ld32, out32_batch_stride
    d.cells_h, d.cells_w = cells
    d.tile_live = _ptr(tile_live)
    if conv is not None:
        tag, work = f"conv3x3_{conv[2]}to{N}", 2.0 * batch * conv[0] * conv[1] * N * K
    else:
        tag, work = f"gemm_epi{epilogue}_n{N}_k{K}", 2.0 * batch * M * N * K
    # kernels behind the call: C_in = 64 -> 128 runs as two halo-patch passes (csrc/api.cu)
    launches = 2 if (conv is not None and conv[2] == 64 and N == 128 and epilogue == EPI_CONV_RELU_BN and
                     os.environ.get("B200_CONV_PATCH", "1") != "0" and
                     os.environ.get("B200_CONV_PATCH_SPLIT", "1") != "0") else 1
    with _timed(tag, launches, work):
        _lib.check(_lib.lib().b200_gemm(C.byref(d), _stream()))


def make_attn_seqs(entries, device):
    """entries: iterable of (q_row0, q_len, kv_row0, kv_len[, q_slot, kv_slot]) -> device int32 [n, 6]."""
    rows = [list(e) + [0] * (6 - len(e)) for e in entries]
    return torch.tensor(rows, dtype=torch.int32).to(device, non_blocking=True)


def attention(q, k, v, seqs, *, num_seqs, max_q_len, heads, ldq, ldk, ldv, q_col0, k_col0, v_col0, scale,
              out, ld_out, out_col0=0, len_override=None, attn_work=0.0):
    _require_cuda(q, k, v, seqs, out)
    d = AttnDesc()
    d.q, d.q_rows, d.ldq, d.q_col0 = _ptr(q), q.numel() // ldq, ldq, q_col0
    d.k, d.k_rows, d.ldk, d.k_col0 = _ptr(k), k.numel() // ldk, ldk, k_col0
    d.v, d.v_rows, d.ldv, d.v_col0 = _ptr(v), v.numel() // ldv, ldv, v_col0
    d.seqs, d.num_seqs, d.max_q_len, d.heads = _ptr(seqs), num_seqs, max_q_len, heads
    d.len_override = _ptr(len_override)
    d.scale = scale
    d.out, d.ld_out, d.out_col0 = _ptr(out), ld_out, out_col0
    with _timed("attention", 1, attn_work):
        _lib.check(_lib.lib().b200_attention(C.byref(d), _stream()))


# ---- SuperPoint-open ----------------------------------------------------------------------------
def sp_conv1a(image, w, bias, scale, shift, out):
    _require_cuda(image, out)
    b, c, h, wd = image.shape
    fn = _lib.lib().b200_sp_conv1a_u8 if image.dtype == torch.uint8 else _lib.lib().b200_sp_conv1a
    with _timed("sp_conv1a", 1, 2.0 * b * h * wd * 64 * 9):
        _lib.check(fn(_ptr(image), b, c, h, wd, _ptr(w), _ptr(bias), _ptr(scale), _ptr(shift),
                                             _ptr(out), _stream()))


def sp_conv_stem(image, w1, host_params, w2, pool, out):
    """First VGG block in one kernel: conv 1->64 (+ReLU+BN) -> conv 64->64 (+ReLU+BN) -> optional 2x2 max pool.
    host_params: CPU float32 tensor [6, 64] = bias1 | scale1 | shift1 | bias2 | scale2 | shift2."""
    _require_cuda(image, out, w2, w1)
    if host_params.is_cuda or host_params.dtype != torch.float32 or tuple(host_params.shape) != (6, 64):
        raise ValueError("sp_conv_stem: host_params must be a CPU float32 tensor of shape (6, 64)")
    host_params = host_params.contiguous()
    b, c, h, wd = image.shape
    fn = _lib.lib().b200_sp_conv_stem_u8 if image.dtype == torch.uint8 else _lib.lib().b200_sp_conv_stem
    with _timed("sp_conv_stem", 1, 2.0 * b * h * wd * 64 * (9 + 576)):
        _lib.check(fn(_ptr(image), b, c, h, wd, _ptr(w1), host_params.data_ptr(), _ptr(w2),
                                                int(pool), _ptr(out), _stream()))


def sp_nms_border(scores, radius, pad):
    _require_cuda(scores)
    b, h, w = scores.shape
    out = torch.empty_like(scores)
    with _timed("sp_nms_border", 1, 8.0 * b * h * w):
        _lib.check(_lib.lib().b200_sp_nms_border(_ptr(scores), b, h, w, int(radius), int(pad or 0), _ptr(out), _stream()))
    return out


CANDIDATE_CAPACITY = 1 << 16  # surviving maxima kept per image by the fast detection path


def sp_detect_fast(scores, radius, pad, threshold, k, image_hw=None):
    """NMS + border removal + top-k through the candidate list (nms_radius 1..4).
    image_hw: optional int32 device tensor [B,2] = (h, w) of the valid region (per-image bottom / right border).
    Returns (nms_map, keypoints, scores, indices, counts)."""
    _require_cuda(scores)
    b, h, w = scores.shape
    dev = scores.device
    nms = torch.empty_like(scores)
    cand = torch.empty(b, CANDIDATE_CAPACITY, dtype=torch.int64, device=dev)
    state = torch.empty(2 * b, dtype=torch.int32, device=dev)
    with _timed("sp_nms_border", 2, 8.0 * b * h * w):
        _lib.check(_lib.lib().b200_sp_nms_border_list(_ptr(scores), b, h, w, int(radius), int(pad or 0), _ptr(image_hw),
                                                      _ptr(nms), _ptr(cand), _ptr(state), CANDIDATE_CAPACITY, _stream()))
    kpts = torch.empty(b, k, 2, dtype=torch.float32, device=dev)
    ksc = torch.empty(b, k, dtype=torch.float32, device=dev)
    idx = torch.empty(b, k, dtype=torch.int32, device=dev)
    counts = torch.empty(b, dtype=torch.int32, device=dev)
    with _timed("sp_select_topk", 2, 4.0 * b * h * w):
        _lib.check(_lib.lib().b200_sp_select_topk_list(_ptr(nms), b, h, w, float(threshold), int(k), _ptr(cand),
                                                       _ptr(state), CANDIDATE_CAPACITY, _ptr(kpts), _ptr(ksc),
                                                       _ptr(idx), _ptr(counts), _stream()))
    return nms, kpts, ksc, idx, counts


def sp_count_candidates(nms_map, threshold):
    b, h, w = nms_map.shape
    counts = torch.empty(b, dtype=torch.int32, device=nms_map.device)
    with _timed("sp_count", 1, 0.0):
        _lib.check(_lib.lib().b200_sp_count_candidates(_ptr(nms_map), b, h, w, float(threshold), _ptr(counts), _stream()))
    return counts


def sp_select_topk(nms_map, threshold, k):
    _require_cuda(nms_map)
    b, h, w = nms_map.shape
    dev = nms_map.device
    kpts = torch.empty(b, k, 2, dtype=torch.float32, device=dev)
    scores = torch.empty(b, k, dtype=torch.float32, device=dev)
    idx = torch.empty(b, k, dtype=torch.int32, device=dev)
    counts = torch.empty(b, dtype=torch.int32, device=dev)
    with _timed("sp_select_topk", 1, 4.0 * b * h * w):
        _lib.check(_lib.lib().b200_sp_select_topk(_ptr(nms_map), b, h, w, float(threshold), int(k), _ptr(kpts),
                                                  _ptr(scores), _ptr(idx), _ptr(counts), _stream()))
    return kpts, scores, idx, counts


def sp_sample_descriptors(dense_nhwc, keypoints, stride=8, mode=0):
    """mode 0: SuperPoint-open grid, 1: nonfree fix_sampling, 2: nonfree legacy (align_corners=True)."""
    _require_cuda(dense_nhwc, keypoints)
    b, hc, wc, c = dense_nhwc.shape
    k = keypoints.shape[1]
    out = torch.empty(b, k, c, dtype=torch.float32, device=dense_nhwc.device)
    with _timed("sp_sample_desc", 1, b * k * c * 4 * 5.0):
        _lib.check(_lib.lib().b200_sp_sample_descriptors_mode(_ptr(dense_nhwc), b, hc, wc, c, _ptr(keypoints), k,
                                                              int(stride), int(mode), _ptr(out), _stream()))
    return out


# ---- LightGlue ----------------------------------------------------------------------------------
def lg_ingest(desc, x32, xm, ld_xm, row0=0):
    rows = desc.numel() // 256
    with _timed("lg_ingest", 1, rows * 256 * 10.0):
        _lib.check(_lib.lib().b200_lg_ingest(_ptr(desc), rows, _ptr(x32, row0 * 256), _ptr(xm, row0 * ld_xm), ld_xm,
                                             _stream()))


def lg_posenc(kpts, image_size, wr, cossin, row0=0):
    b, n, _ = kpts.shape
    scratch = torch.empty(b, 2, dtype=torch.float32, device=kpts.device) if image_size is None else None
    with _timed("lg_posenc", 1 if image_size is not None else 2, b * n * 264.0):
        _lib.check(_lib.lib().b200_lg_posenc(_ptr(kpts), _ptr(image_size), b, n, _ptr(wr), _ptr(scratch),
                                             _ptr(cossin, row0 * 64), _stream()))


def lg_rowdot(x32, rows, w, bias, out, sigmoid=False, row0=0):
    with _timed("lg_rowdot", 1, rows * 1028.0):
        _lib.check(_lib.lib().b200_lg_rowdot(_ptr(x32, row0 * 256), rows, _ptr(w), _ptr(bias), int(sigmoid),
                                             _ptr(out, row0), _stream()))


def lg_assignment(sim, z0, z1, threshold):
    b, m, n = sim.shape
    dev = sim.device
    ws = torch.empty(_lib.lib().b200_lg_assignment_workspace(b, m, n), dtype=torch.uint8, device=dev)
    scores = torch.empty(b, m + 1, n + 1, dtype=torch.float32, device=dev)
    m0 = torch.empty(b, m, dtype=torch.int64, device=dev)
    m1 = torch.empty(b, n, dtype=torch.int64, device=dev)
    ms0 = torch.empty(b, m, dtype=torch.float32, device=dev)
    ms1 = torch.empty(b, n, dtype=torch.float32, device=dev)
    with _timed("lg_assignment", 5, 12.0 * b * m * n):
        _lib.check(_lib.lib().b200_lg_assignment(_ptr(sim), _ptr(z0), _ptr(z1), b, m, n, float(threshold), _ptr(ws),
                                                 _ptr(scores), _ptr(m0), _ptr(m1), _ptr(ms0), _ptr(ms1), _stream()))
    return scores, m0, m1, ms0, ms1


# ---- adaptive depth / width ---------------------------------------------------------------------
def lg_count_below(conf, n, threshold, out):
    with _timed("lg_count_below", 1, n * 4.0):
        _lib.check(_lib.lib().b200_lg_count_below(_ptr(conf), int(n), float(threshold), _ptr(out), _stream()))


def lg_prune(match, conf, n0, n1, match_thr, conf_thr, pos, counts, src, dst, ind_src, ind_dst, prune0, prune1):
    """src / dst: dicts with x32, xm, cs (see LightGlue._workspace)."""
    with _timed("lg_prune", 2, (n0 + n1) * 3600.0):
        _lib.check(_lib.lib().b200_lg_prune(_ptr(match), _ptr(conf), int(n0), int(n1), float(match_thr), float(conf_thr),
                                            _ptr(pos), _ptr(counts), _ptr(src["x32"]), _ptr(dst["x32"]), _ptr(src["xm"]),
                                            _ptr(dst["xm"]), 512, _ptr(src["cs"]), _ptr(dst["cs"]), _ptr(ind_src),
                                            _ptr(ind_dst), _ptr(prune0), _ptr(prune1), _stream()))


def lg_scatter_matches(m0, m1, ms0, ms1, ind, n0, n1, out_m0, out_m1, out_ms0, out_ms1):
    with _timed("lg_scatter_matches", 1, (n0 + n1) * 40.0):
        _lib.check(_lib.lib().b200_lg_scatter_matches(_ptr(m0), _ptr(m1), _ptr(ms0), _ptr(ms1), _ptr(ind), int(n0), int(n1),
                                                      _ptr(out_m0), _ptr(out_m1), _ptr(out_ms0), _ptr(out_ms1), _stream()))


# ---- SuperGlue configuration ----------------------------------------------------------------------
def sg_kenc_head(kpts, scores, size, w1, b1, w2, b2, out, row0=0):
    """normalize_keypoints + first two keypoint-encoder layers -> bf16 rows [row0, row0 + B*n) of out [T,64]."""
    _require_cuda(kpts, scores, size, out)
    b, n, _ = kpts.shape
    with _timed("sg_kenc_head", 1, b * n * 140.0):
        _lib.check(_lib.lib().b200_sg_keypoint_encoder_head(_ptr(kpts), _ptr(scores), _ptr(size), b, n, _ptr(w1), _ptr(b1),
                                                            _ptr(w2), _ptr(b2), _ptr(out, row0 * 64), _stream()))


def sg_split_hilo(x32, rows, split_row, out):
    with _timed("sg_split_hilo", 1, rows * 256 * 10.0):
        _lib.check(_lib.lib().b200_sg_split_hilo(_ptr(x32), int(rows), int(split_row), _ptr(out), _stream()))


def split_hilo(x32, rows, cols, split_row, out):
    _require_cuda(x32, out)
    with _timed("split_hilo", 1, rows * cols * 10.0):
        _lib.check(_lib.lib().b200_split_hilo(_ptr(x32), int(rows), int(cols), int(split_row), _ptr(out), _stream()))


def nn_match(sim, ratio_thresh, distance_thresh, mutual_check):
    """find_nn both ways + mutual check + log_assignment: (log_assignment [B,M+1,N+1], m0, m1, ms0, ms1)."""
    _require_cuda(sim)
    b, m, n = sim.shape
    dev = sim.device
    ws = torch.empty(_lib.lib().b200_nn_match_workspace(b, m, n), dtype=torch.uint8, device=dev)
    la = torch.empty(b, m + 1, n + 1, dtype=torch.float32, device=dev)
    m0 = torch.empty(b, m, dtype=torch.int64, device=dev)
    m1 = torch.empty(b, n, dtype=torch.int64, device=dev)
    ms0 = torch.empty(b, m, dtype=torch.float32, device=dev)
    ms1 = torch.empty(b, n, dtype=torch.float32, device=dev)
    ratio_sq = float(ratio_thresh) ** 2 if ratio_thresh else 0.0
    dist_sq = float(distance_thresh) ** 2 if distance_thresh else 0.0
    with _timed("nn_match", 4, 12.0 * b * m * n):
        _lib.check(_lib.lib().b200_nn_match(_ptr(sim), b, m, n, ratio_sq, dist_sq, int(bool(mutual_check)), _ptr(ws),
                                            _ptr(la), _ptr(m0), _ptr(m1), _ptr(ms0), _ptr(ms1), _stream()))
    return la, m0, m1, ms0, ms1


# Pairs per Sinkhorn launch, 0 = one launch for the whole batch (default).  Measured (32 pairs x 2049^2 x 100 iterations,
# scripts/bench_superglue.py): waves of 8 / 6 pairs, whose couplings (134 / 100 MB) stay in the 126 MB L2 across the
# iterations, take 24.8 / 24.7 ms against 24.9 ms for one launch, waves of 4 / 3 / 2 pairs 34.5 / 50.6 / 58.5 ms (too few
# row blocks per SM, one grid barrier per iteration and wave): the kernel is bound by its exponentials and issue slots
# (ncu: alu 40 %, issue 56 %, dram 18 %), not by HBM, so L2 residency buys nothing.
SINKHORN_WAVE = int(os.environ.get("B200GLUE_SINKHORN_WAVE", "0"))


def sg_sinkhorn_assignment(cost, bin_score, iters, threshold, wave=None):
    """log_optimal_transport + mutual-NN filter: (scores [B,M+1,N+1], m0, m1, ms0, ms1)."""
    _require_cuda(cost, bin_score)
    b, m, n = cost.shape
    dev = cost.device
    wave = SINKHORN_WAVE if wave is None else wave
    if wave <= 0 or (m + 1) * (n + 1) * 4 * wave > 96 << 20:  # problems too large for L2 anyway: one launch
        wave = b
    wave = max(1, min(wave, b))
    ws = torch.empty(_lib.lib().b200_sg_sinkhorn_workspace(wave, m, n), dtype=torch.uint8, device=dev)
    scores = torch.empty(b, m + 1, n + 1, dtype=torch.float32, device=dev)
    m0 = torch.empty(b, m, dtype=torch.int64, device=dev)
    m1 = torch.empty(b, n, dtype=torch.int64, device=dev)
    ms0 = torch.empty(b, m, dtype=torch.float32, device=dev)
    ms1 = torch.empty(b, n, dtype=torch.float32, device=dev)
    # algorithmic bytes: one pass over the couplings per iteration (+ the final write)
    with _timed("sg_sinkhorn", 4 * ((b + wave - 1) // wave), 4.0 * b * (m + 1) * (n + 1) * (iters + 2)):
        for b0 in range(0, b, wave):
            nb = min(wave, b - b0)
            _lib.check(_lib.lib().b200_sg_sinkhorn_assignment(
                _ptr(cost, b0 * m * n), _ptr(bin_score), nb, m, n, int(iters), float(threshold), _ptr(ws),
                _ptr(scores, b0 * (m + 1) * (n + 1)), _ptr(m0, b0 * m), _ptr(m1, b0 * n), _ptr(ms0, b0 * m),
                _ptr(ms1, b0 * n), _stream()))
    return scores, m0, m1, ms0, ms1


# ---- rows either side of the path: pre-processing, export packing, match metrics ----------------------
def prep_resize(src, out_hw, taps_x=None, taps_y=None, pad_hw=None, mask=False):
    """ImagePreprocessor resize on the device.  src: float32 (B,C,H,W) in [0,1] or uint8 (B,H,W,C).
    taps_x / taps_y: CPU float32 Gaussian taps (odd lengths) or None for no anti-alias blur.
    Returns (image (B,C,pad_h,pad_w) float32, mask (B,pad_h,pad_w) uint8 or None)."""
    _require_cuda(src)
    u8 = src.dtype == torch.uint8
    if u8:
        b, h, w, c = src.shape
    else:
        if src.dtype != torch.float32:
            raise ValueError("prep_resize: the image must be float32 (B,C,H,W) or uint8 (B,H,W,C)")
        b, c, h, w = src.shape
    src = src.contiguous()
    oh, ow = int(out_hw[0]), int(out_hw[1])
    ph, pw = (oh, ow) if pad_hw is None else (int(pad_hw[0]), int(pad_hw[1]))
    dev = src.device
    out = torch.empty(b, c, ph, pw, dtype=torch.float32, device=dev)
    msk = torch.empty(b, ph, pw, dtype=torch.uint8, device=dev) if mask else None
    blur = taps_x is not None and taps_y is not None and (len(taps_x) > 1 or len(taps_y) > 1)
    tmp = torch.empty(b, c, h, w, dtype=torch.float32, device=dev) if blur else None
    tx = taps_x.detach().to("cpu", torch.float32).contiguous() if blur else None
    ty = taps_y.detach().to("cpu", torch.float32).contiguous() if blur else None
    # algorithmic bytes: read the source once, write the result once (the blur scratch is extra traffic)
    work = float(src.numel() * src.element_size() + out.numel() * 4)
    with _timed("prep_resize", 2 if blur else 1, work):
        _lib.check(_lib.lib().b200_prep_resize(
            _ptr(src), 1 if u8 else 0, b, c, h, w, oh, ow, tx.data_ptr() if blur else None, len(tx) if blur else 1,
            ty.data_ptr() if blur else None, len(ty) if blur else 1, _ptr(tmp), _ptr(out), ph, pw, _ptr(msk), _stream()))
    return out, msk


def export_pack(fields, B, record_bytes, as_half, records):
    """fields: list of (tensor, kind, elems_per_pair, dst_offset, scales or None); records: uint8 (B, record_bytes)."""
    arr = (_lib.ExportField * len(fields))()
    keep = []
    for i, (t, kind, n, off, scales) in enumerate(fields):
        _require_cuda(t, scales)
        t = t.contiguous()
        keep.append(t)
        arr[i].src, arr[i].kind, arr[i].elems_per_pair, arr[i].dst_offset = _ptr(t), kind, n, off
        arr[i].scales = _ptr(scales)
    with _timed("export_pack", 1, float(B * record_bytes * 2)):
        _lib.check(_lib.lib().b200_export_pack(arr, len(fields), B, int(bool(as_half)), _ptr(records), record_bytes,
                                               _stream()))


def eval_matches_homography(kp0, kp1, m0, H):
    """(B,M,2), (B,N,2), (B,M) int64, (B,3,3) -> (B,4) = prec@1px, prec@3px, num_matches, num_keypoints."""
    _require_cuda(kp0, kp1, m0, H)
    B, M, _ = kp0.shape
    out = torch.empty(B, 4, dtype=torch.float32, device=kp0.device)
    with _timed("eval_homography", 1, B * (M * 24.0)):
        _lib.check(_lib.lib().b200_eval_matches_homography(_ptr(kp0.float().contiguous()), _ptr(kp1.float().contiguous()),
                                                           _ptr(m0.contiguous()), B, M, kp1.shape[1],
                                                           _ptr(H.float().reshape(B, 9).contiguous()), _ptr(out), _stream()))
    return out


def eval_matches_epipolar(kp0, kp1, m0, cam0, cam1, T01):
    """cam: (B,4) = fx, fy, cx, cy; T01: (B,12) = R row-major | t -> (B,5)."""
    _require_cuda(kp0, kp1, m0, cam0, cam1, T01)
    B, M, _ = kp0.shape
    out = torch.empty(B, 5, dtype=torch.float32, device=kp0.device)
    with _timed("eval_epipolar", 1, B * (M * 24.0)):
        _lib.check(_lib.lib().b200_eval_matches_epipolar(_ptr(kp0.float().contiguous()), _ptr(kp1.float().contiguous()),
                                                         _ptr(m0.contiguous()), B, M, kp1.shape[1],
                                                         _ptr(cam0.float().contiguous()), _ptr(cam1.float().contiguous()),
                                                         _ptr(T01.float().contiguous()), _ptr(out), _stream()))
    return out


def eval_homography_dlt(kp0, kp1, m0, ms0, H_gt, image_size):
    """Weighted DLT on the matched points + corner error: (H (B,3,3), error (B,))."""
    _require_cuda(kp0, kp1, m0, ms0, H_gt, image_size)
    B, M, _ = kp0.shape
    dev = kp0.device
    H = torch.empty(B, 3, 3, dtype=torch.float32, device=dev)
    err = torch.empty(B, dtype=torch.float32, device=dev)
    with _timed("eval_dlt", 1, B * (M * 28.0)):
        _lib.check(_lib.lib().b200_eval_homography_dlt(_ptr(kp0.float().contiguous()), _ptr(kp1.float().contiguous()),
                                                       _ptr(m0.contiguous()), _ptr(ms0.float().contiguous()), B, M,
                                                       kp1.shape[1], _ptr(H_gt.float().reshape(B, 9).contiguous()),
                                                       _ptr(image_size.float().reshape(B, 2).contiguous()), _ptr(H),
                                                       _ptr(err), _stream()))
    return H, err


# ---- adaptive depth / width for a batch of pairs (state on the device) ------------------------------------
def lg_adaptive_decide(conf, B, M, N, threshold, depth_conf, layer, state):
    with _timed("lg_adaptive_decide", 1, B * (M + N) * 4.0):
        _lib.check(_lib.lib().b200_lg_adaptive_decide(_ptr(conf), B, M, N, float(threshold), float(depth_conf), int(layer),
                                                      _ptr(state), _stream()))


def lg_adaptive_prune(match, conf, B, M, N, match_thr, conf_thr, state, pos, newc, src, dst, ind_src, ind_dst, prune0,
                      prune1):
    with _timed("lg_adaptive_prune", 2, B * (M + N) * 3600.0):
        _lib.check(_lib.lib().b200_lg_adaptive_prune(
            _ptr(match), _ptr(conf), B, M, N, float(match_thr), float(conf_thr), _ptr(state), _ptr(pos), _ptr(newc),
            _ptr(src["x32"]), _ptr(dst["x32"]), _ptr(src["xm"]), _ptr(dst["xm"]), 512, _ptr(src["cs"]), _ptr(dst["cs"]),
            _ptr(ind_src), _ptr(ind_dst), _ptr(prune0), _ptr(prune1), _stream()))


def lg_adaptive_commit(state, newc, B, M, N, lens, tile_live, hist_next=None):
    with _timed("lg_adaptive_commit", 2 if newc is not None else 1, 0.0):
        _lib.check(_lib.lib().b200_lg_adaptive_commit(_ptr(state), _ptr(newc), B, M, N, _ptr(lens), _ptr(tile_live),
                                                      _ptr(hist_next), _stream()))


def lg_adaptive_snapshot(state, B, M, N, final_pass, md, md_out, z, z_out, x, x_out, ind, ind_out, final_len):
    with _timed("lg_adaptive_snapshot", 1 if final_pass else 2, B * (M + N) * 4100.0):
        _lib.check(_lib.lib().b200_lg_adaptive_snapshot(_ptr(state), B, M, N, int(final_pass), _ptr(md), _ptr(md_out),
                                                        _ptr(z), _ptr(z_out), _ptr(x), _ptr(x_out), _ptr(ind),
                                                        _ptr(ind_out), _ptr(final_len), _stream()))


def lg_assignment_live(sim, z0, z1, live, threshold):
    """lg_assignment on padded problems with device-side live sizes; scores is zero-filled outside the live blocks."""
    b, m, n = sim.shape
    dev = sim.device
    ws = torch.empty(_lib.lib().b200_lg_assignment_workspace(b, m, n), dtype=torch.uint8, device=dev)
    scores = torch.zeros(b, m + 1, n + 1, dtype=torch.float32, device=dev)
    m0 = torch.empty(b, m, dtype=torch.int64, device=dev)
    m1 = torch.empty(b, n, dtype=torch.int64, device=dev)
    ms0 = torch.empty(b, m, dtype=torch.float32, device=dev)
    ms1 = torch.empty(b, n, dtype=torch.float32, device=dev)
    with _timed("lg_assignment", 5, 12.0 * b * m * n):
        _lib.check(_lib.lib().b200_lg_assignment_live(_ptr(sim), _ptr(z0), _ptr(z1), b, m, n, _ptr(live), float(threshold),
                                                      _ptr(ws), _ptr(scores), _ptr(m0), _ptr(m1), _ptr(ms0), _ptr(ms1),
                                                      _stream()))
    return scores, m0, m1, ms0, ms1


def lg_adaptive_scatter(m0, m1, ms0, ms1, ind, final_len, B, M, N, out_m0, out_m1, out_ms0, out_ms1):
    with _timed("lg_scatter_matches", 1, B * (M + N) * 40.0):
        _lib.check(_lib.lib().b200_lg_adaptive_scatter(_ptr(m0), _ptr(m1), _ptr(ms0), _ptr(ms1), _ptr(ind), _ptr(final_len),
                                                       B, M, N, _ptr(out_m0), _ptr(out_m1), _ptr(out_ms0), _ptr(out_ms1),
                                                       _stream()))
