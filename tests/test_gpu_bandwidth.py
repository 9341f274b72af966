"""GPU parity tests of the bandwidth-bound kernels against the CPU oracle (teacher-forced: every
kernel is fed the oracle's input for its stage).  Integer / comparison work must be bit-exact."""
import numpy as np
import pytest
import torch

from oracle import lightglue as olg
from oracle import sp_open as osp

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda:0")


def _scores(seed, b, h, w, plateaus=False):
    g = torch.Generator().manual_seed(seed)
    s = torch.rand(b, h, w, generator=g)
    if plateaus:  # quantise so that equal-score plateaus and ties occur
        s = torch.round(s * 12) / 12
    return s


@pytest.mark.parametrize("r", [0, 1, 3, 4])
@pytest.mark.parametrize("shape", [(1, 64, 64), (2, 97, 131), (1, 200, 70)])
@pytest.mark.parametrize("plateaus", [False, True])
def test_nms_border_bit_exact(dev, r, shape, plateaus):
    from glue_factory_b200 import ops

    s = _scores(r * 7 + shape[1], *shape, plateaus=plateaus)
    want = osp.remove_borders(osp.nms(s, r), 4)
    got = ops.sp_nms_border(s.to(dev), r, 4).cpu()
    assert torch.equal(got, want), f"{(got != want).sum().item()} pixels differ"


def test_nms_no_border(dev):
    from glue_factory_b200 import ops

    s = _scores(5, 1, 80, 96)
    assert torch.equal(ops.sp_nms_border(s.to(dev), 2, 0).cpu(), osp.nms(s, 2))


@pytest.mark.parametrize("thr,k", [(0.005, 50), (0.0, 300), (-1.0, 128), (0.5, 4096), (2.0, 16)])
@pytest.mark.parametrize("plateaus", [False, True])
def test_select_topk_bit_exact(dev, thr, k, plateaus):
    from glue_factory_b200 import ops

    s = _scores(11, 3, 72, 88, plateaus=plateaus)
    m = osp.remove_borders(osp.nms(s, 2), 4)
    kpts, scores, idx, counts = ops.sp_select_topk(m.to(dev), thr, k)
    torch.cuda.synchronize()
    for b in range(3):
        xy, sc, li = osp.select_keypoints(m[b], thr, k)
        n = li.numel()
        assert counts[b].item() == n
        assert torch.equal(idx[b, :n].cpu().long(), li)
        assert torch.equal(scores[b, :n].cpu(), sc)
        assert torch.equal(kpts[b, :n].cpu(), xy + 0.5)
    cnt = ops.sp_count_candidates(m.to(dev), thr).cpu()
    assert cnt.tolist() == [(m[b] > thr).sum().item() for b in range(3)]


def test_select_topk_large_map_with_massive_ties(dev):
    """detection_threshold = -1 on a map that is mostly zeros: the k-th score is 0 with ~1e5 ties."""
    from glue_factory_b200 import ops

    s = _scores(3, 2, 320, 384)
    m = osp.remove_borders(osp.nms(s, 4), 4)
    k = 2048
    kpts, scores, idx, counts = ops.sp_select_topk(m.to(dev), -1.0, k)
    for b in range(2):
        xy, sc, li = osp.select_keypoints(m[b], -1.0, k)
        assert counts[b].item() == k
        assert torch.equal(idx[b].cpu().long(), li)
        assert torch.equal(scores[b].cpu(), sc)


def _check_detect_fast(dev, s, r, pad, thr, k):
    from glue_factory_b200 import ops

    want_map = osp.remove_borders(osp.nms(s, r), pad) if pad else osp.nms(s, r)
    nms, kpts, scores, idx, counts = ops.sp_detect_fast(s.to(dev), r, pad, thr, k)
    torch.cuda.synchronize()
    assert torch.equal(nms.cpu(), want_map), f"{(nms.cpu() != want_map).sum().item()} NMS pixels differ"
    for b in range(s.shape[0]):
        xy, sc, li = osp.select_keypoints(want_map[b], thr, k)
        n = li.numel()
        assert counts[b].item() == n
        assert torch.equal(idx[b, :n].cpu().long(), li)
        assert torch.equal(scores[b, :n].cpu(), sc)
        assert torch.equal(kpts[b, :n].cpu(), xy + 0.5)


@pytest.mark.parametrize("r", [1, 2, 3, 4])
@pytest.mark.parametrize("shape", [(1, 64, 64), (2, 97, 131), (1, 200, 70), (2, 300, 260)])
@pytest.mark.parametrize("plateaus", [False, True])
def test_detect_fast_bit_exact(dev, r, shape, plateaus):
    """register-blocked NMS + candidate-list top-k (fast path) against the oracle, all tile-edge cases"""
    s = _scores(r * 13 + shape[2], *shape, plateaus=plateaus)
    _check_detect_fast(dev, s, r, 4, 0.005, 100)


@pytest.mark.parametrize("thr,k", [(0.0, 300), (-1.0, 128), (-1.0, 4000), (0.5, 4096), (2.0, 16), (0.9, 7)])
@pytest.mark.parametrize("plateaus", [False, True])
def test_detect_fast_thresholds(dev, thr, k, plateaus):
    """few / no candidates, all candidates kept (row-major order), zero-score pixels among the top k
    (negative threshold: the full-map kernel takes over for that image)"""
    s = _scores(17, 3, 72, 88, plateaus=plateaus)
    _check_detect_fast(dev, s, 2, 4, thr, k)


def test_detect_fast_no_border(dev):
    _check_detect_fast(dev, _scores(5, 1, 80, 96), 3, 0, 0.1, 64)


def test_detect_fast_list_overflow(dev):
    """radius 1 on white noise leaves > 65536 maxima per megapixel image: the list overflows and the
    full-map kernel must produce the result"""
    s = _scores(23, 1, 1024, 1024)
    _check_detect_fast(dev, s, 1, 4, 0.0, 2048)


def test_detect_fast_config2_shape(dev):
    s = _scores(29, 2, 1024, 1024) ** 4
    _check_detect_fast(dev, s, 3, 4, -1.0, 2048)
    _check_detect_fast(dev, s, 4, 4, 0.005, 2048)


def test_sample_descriptors(dev):
    from glue_factory_b200 import ops

    g = torch.Generator().manual_seed(2)
    dense = torch.nn.functional.normalize(torch.randn(2, 256, 15, 19, generator=g), dim=1)
    kp = torch.stack([torch.randint(0, 19 * 8, (2, 77), generator=g), torch.randint(0, 15 * 8, (2, 77), generator=g)], -1).float()
    want = osp.sample_descriptors(kp, dense, 8).transpose(1, 2)
    got = ops.sp_sample_descriptors(dense.permute(0, 2, 3, 1).contiguous().to(dev), (kp + 0.5).to(dev), 8).cpu()
    # tolerance: 1e-3 relative (L2 over the 256-vector), achieved ~1e-6
    rel = (got - want).norm(dim=-1) / want.norm(dim=-1)
    assert rel.max().item() < 1e-5, rel.max().item()


@pytest.mark.parametrize("with_size", [True, False])
def test_posenc(dev, lg_weights, with_size):
    from glue_factory_b200 import ops

    g = torch.Generator().manual_seed(4)
    kp = torch.rand(3, 50, 2, generator=g) * torch.tensor([640.0, 480.0])
    size = torch.tensor([[640.0, 480.0]] * 3) if with_size else None
    cos, sin = olg.positional_encoding(lg_weights, olg.normalize_keypoints(kp, size))
    cs = torch.empty(150, 32, 2, device=dev)
    ops.lg_posenc(kp.to(dev), None if size is None else size.to(dev), lg_weights["posenc.Wr.weight"].to(dev), cs)
    got = cs.cpu().view(3, 50, 32, 2)
    np.testing.assert_allclose(got[..., 0].numpy(), cos[:, 0, :, 0::2].numpy(), atol=2e-6)
    np.testing.assert_allclose(got[..., 1].numpy(), sin[:, 0, :, 0::2].numpy(), atol=2e-6)


def test_ingest_and_rowdot(dev):
    from glue_factory_b200 import ops

    g = torch.Generator().manual_seed(6)
    d = torch.randn(70, 256, generator=g)
    x32 = torch.zeros(100, 256, device=dev)
    xm = torch.zeros(100, 512, dtype=torch.bfloat16, device=dev)
    ops.lg_ingest(d.to(dev), x32, xm, 512, row0=30)
    assert torch.equal(x32[30:].cpu(), d) and x32[:30].abs().max().item() == 0
    assert torch.equal(xm[30:, :256].cpu(), d.to(torch.bfloat16)) and xm[:, 256:].abs().max().item() == 0
    w, b = torch.randn(256, generator=g), torch.randn(1, generator=g)
    out = torch.zeros(100, device=dev)
    ops.lg_rowdot(x32, 100, w.to(dev), b.to(dev), out)
    np.testing.assert_allclose(out.cpu().numpy(), (x32.cpu() @ w + b).numpy(), rtol=1e-5, atol=1e-5)
    ops.lg_rowdot(x32, 100, w.to(dev), b.to(dev), out, sigmoid=True)
    np.testing.assert_allclose(out.cpu().numpy(), torch.sigmoid(x32.cpu() @ w + b).numpy(), rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("B,M,N", [(1, 64, 256), (2, 97, 81), (1, 300, 777), (2, 1, 5)])
def test_log_assignment_and_filter(dev, B, M, N):
    from glue_factory_b200 import ops

    g = torch.Generator().manual_seed(B * 1000 + M)
    sim = torch.randn(B, M, N, generator=g) * 4
    z0, z1 = torch.randn(B, M, 1, generator=g) * 3, torch.randn(B, N, 1, generator=g) * 3
    certainties = torch.nn.functional.logsigmoid(z0) + torch.nn.functional.logsigmoid(z1).transpose(1, 2)
    want = sim.new_zeros(B, M + 1, N + 1)
    want[:, :M, :N] = torch.log_softmax(sim, 2) + torch.log_softmax(sim, 1) + certainties
    want[:, :M, N] = torch.nn.functional.logsigmoid(-z0.squeeze(-1))
    want[:, M, :N] = torch.nn.functional.logsigmoid(-z1.squeeze(-1))
    scores, m0, m1, ms0, ms1 = ops.lg_assignment(sim.to(dev), z0.squeeze(-1).contiguous().to(dev),
                                                 z1.squeeze(-1).contiguous().to(dev), 0.1)
    # tolerance on the log-assignment: 1e-3 relative (stated by north_star); achieved ~1e-6
    np.testing.assert_allclose(scores.cpu().numpy(), want.numpy(), rtol=1e-5, atol=2e-5)
    # filter on the kernel's own scores must equal the oracle's filter on the same matrix: bit-exact
    w0, w1, ws0, ws1 = olg.filter_matches(scores.cpu(), 0.1)
    assert torch.equal(m0.cpu(), w0) and torch.equal(m1.cpu(), w1)
    np.testing.assert_allclose(ms0.cpu().numpy(), ws0.numpy(), rtol=1e-5)
    np.testing.assert_allclose(ms1.cpu().numpy(), ws1.numpy(), rtol=1e-5)
    assert m0.dtype == torch.int64


@pytest.mark.parametrize("n0,n1,with_conf", [(160, 160, True), (1500, 2300, True), (1024, 1, False), (33, 4096, False)])
def test_prune_compaction_bit_exact(dev, n0, n1, with_conf):
    """get_pruning_mask + index_select compaction + prune counters + scatter back (lightglue.py:478-511)."""
    from glue_factory_b200 import ops

    g = torch.Generator().manual_seed(n0 + n1)
    t = n0 + n1
    match, conf = torch.rand(t, generator=g), torch.rand(t, generator=g)
    src = {"x32": torch.randn(t, 256, generator=g), "xm": torch.randn(t, 512, generator=g).to(torch.bfloat16),
           "cs": torch.randn(t, 32, 2, generator=g)}
    ind = torch.cat([torch.randperm(n0, generator=g), torch.randperm(n1, generator=g)]).to(torch.int32)
    keep = match > 0.3
    if with_conf:
        keep = keep | (conf <= 0.6)
    k0, k1 = torch.where(keep[:n0])[0], torch.where(keep[n0:])[0] + n0
    sel = torch.cat([k0, k1])
    d = {k: v.to(dev) for k, v in src.items()}
    dst = {k: torch.zeros_like(v) for k, v in d.items()}
    pos = torch.empty(t, dtype=torch.int32, device=dev)
    counts = torch.zeros(4, dtype=torch.int32, device=dev)
    ind_dst = torch.zeros(t, dtype=torch.int32, device=dev)
    pr0 = torch.ones(1, n0, dtype=torch.int64, device=dev)
    pr1 = torch.ones(1, n1, dtype=torch.int64, device=dev)
    ops.lg_prune(match.to(dev), conf.to(dev) if with_conf else None, n0, n1, 0.3, 0.6, pos, counts, d, dst, ind.to(dev),
                 ind_dst, pr0, pr1)
    assert counts[:2].tolist() == [k0.numel(), k1.numel()]
    n = sel.numel()
    assert torch.equal(dst["x32"][:n].cpu(), src["x32"][sel])
    assert torch.equal(dst["xm"][:n, :256].cpu(), src["xm"][sel, :256])
    assert torch.equal(dst["cs"][:n].cpu(), src["cs"][sel])
    assert torch.equal(ind_dst[:n].cpu(), ind[sel])
    want0 = torch.ones(n0, dtype=torch.int64)
    want0[ind[k0].long()] += 1
    assert torch.equal(pr0[0].cpu(), want0)
    cnt = torch.zeros(1, dtype=torch.int32, device=dev)
    ops.lg_count_below(conf.to(dev), t, 0.25, cnt)
    assert cnt.item() == (conf < 0.25).sum().item()
    # scatter back
    m, nn = k0.numel(), k1.numel()
    m0 = torch.randint(-1, nn, (1, m), generator=g)
    m1 = torch.randint(-1, m, (1, nn), generator=g)
    ms0, ms1 = torch.rand(1, m, generator=g), torch.rand(1, nn, generator=g)
    o0 = torch.full((1, n0), -1, dtype=torch.int64, device=dev)
    o1 = torch.full((1, n1), -1, dtype=torch.int64, device=dev)
    s0, s1 = torch.zeros(1, n0, device=dev), torch.zeros(1, n1, device=dev)
    ops.lg_scatter_matches(m0.to(dev), m1.to(dev), ms0.to(dev), ms1.to(dev), ind_dst, m, nn, o0, o1, s0, s1)
    i0, i1 = ind[k0].long()[None], ind[k1].long()[None]
    w0 = torch.full((1, n0), -1, dtype=torch.int64)
    w0[:, i0[0]] = torch.where(m0 == -1, -1, i1.gather(1, m0.clamp(min=0)))
    w1 = torch.full((1, n1), -1, dtype=torch.int64)
    w1[:, i1[0]] = torch.where(m1 == -1, -1, i0.gather(1, m1.clamp(min=0)))
    assert torch.equal(o0.cpu(), w0) and torch.equal(o1.cpu(), w1)
    ws0 = torch.zeros(1, n0)
    ws0[:, i0[0]] = ms0
    assert torch.equal(s0.cpu(), ws0)
