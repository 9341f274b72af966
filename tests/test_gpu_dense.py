"""GPU parity tests of the dense tcgen05 kernels (GEMM epilogues, implicit-GEMM conv, attention)
against plain PyTorch fp32 on the same bf16-rounded operands.  Run on the B200 box: pytest -m gpu."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    return torch.device("cuda:0")


def _rand(shape, dev, seed, scale=1.0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dev)


def _bf(t):
    return t.to(torch.bfloat16)


def _close(got, want, rtol, atol, what=""):
    got, want = got.float(), want.float()
    err = (got - want).abs()
    tol = atol + rtol * want.abs()
    bad = (err > tol).sum().item()
    assert bad == 0, f"{what}: {bad}/{err.numel()} elements off, max err {err.max().item():.4g}, max ref {want.abs().max().item():.4g}"


@pytest.mark.parametrize("M", [128, 300, 4096 + 17])
def test_linear_bias(dev, M):
    from glue_factory_b200 import ops

    a, w, b = _bf(_rand((M, 256), dev, 1)), _bf(_rand((256, 256), dev, 2, 1 / 16)), _rand((256,), dev, 3)
    out = torch.zeros(M, 512, dtype=torch.bfloat16, device=dev)
    ops.gemm(a, w, M=M, N=256, K=256, lda=256, block_n=256, epilogue=ops.EPI_BIAS_BF16, bias=b, out16=out,
             out16_col0=256, ld16=512)
    torch.cuda.synchronize()
    want = a.float() @ w.float().t() + b
    _close(out[:, 256:], want, 1e-2, 1e-2, "linear")
    assert out[:, :256].abs().max().item() == 0  # the left half is untouched


def test_linear_strided_a_and_n512(dev):
    from glue_factory_b200 import ops

    M = 777
    xm = _bf(_rand((M, 512), dev, 4))
    w, b = _bf(_rand((512, 256), dev, 5, 1 / 16)), _rand((512,), dev, 6)
    out = torch.zeros(M, 768, dtype=torch.bfloat16, device=dev)
    ops.gemm(xm, w, M=M, N=512, K=256, lda=512, block_n=256, epilogue=ops.EPI_BIAS_BF16, bias=b, out16=out, ld16=768)
    torch.cuda.synchronize()
    want = xm[:, :256].float() @ w.float().t() + b
    _close(out[:, :512], want, 1e-2, 1e-2, "strided A")


def test_qkv_rotary(dev):
    from glue_factory_b200 import ops

    M = 500
    a, w, b = _bf(_rand((M, 256), dev, 7)), _bf(_rand((768, 256), dev, 8, 1 / 16)), _rand((768,), dev, 9)
    ang = _rand((M, 32), dev, 10, 2.0)
    cs = torch.stack([ang.cos(), ang.sin()], -1).contiguous()
    out = torch.empty(M, 768, dtype=torch.bfloat16, device=dev)
    ops.gemm(a, w, M=M, N=768, K=256, lda=256, block_n=256, epilogue=ops.EPI_QKV_ROTARY, bias=b, cossin=cs,
             rot_cols=512, out16=out, ld16=768)
    torch.cuda.synchronize()
    y = a.float() @ w.float().t() + b
    qk = y[:, :512].view(M, 8, 32, 2)
    c, s = ang.cos()[:, None, :], ang.sin()[:, None, :]
    rot = torch.stack([qk[..., 0] * c - qk[..., 1] * s, qk[..., 1] * c + qk[..., 0] * s], -1).reshape(M, 512)
    _close(out[:, :512], rot, 1e-2, 2e-2, "rotary q,k")
    _close(out[:, 512:], y[:, 512:], 1e-2, 2e-2, "v")


def test_ffn_layernorm_gelu_and_residual(dev):
    from glue_factory_b200 import ops

    M = 391
    xm = _bf(_rand((M, 512), dev, 11))
    w1, b1 = _bf(_rand((512, 512), dev, 12, 1 / 22)), _rand((512,), dev, 13, 0.1)
    g, be = 1 + _rand((512,), dev, 14, 0.1), _rand((512,), dev, 15, 0.1)
    h = torch.empty(M, 512, dtype=torch.bfloat16, device=dev)
    ops.gemm(xm, w1, M=M, N=512, K=512, lda=512, block_n=512, epilogue=ops.EPI_LN_GELU, bias=b1, ln_gamma=g,
             ln_beta=be, ln_eps=1e-5, out16=h, ld16=512)
    torch.cuda.synchronize()
    want = F.gelu(F.layer_norm(xm.float() @ w1.float().t() + b1, (512,), g, be, 1e-5))
    _close(h, want, 1e-2, 2e-2, "LN+GELU")

    w2, b2 = _bf(_rand((256, 512), dev, 16, 1 / 22)), _rand((256,), dev, 17, 0.1)
    x32 = _rand((M, 256), dev, 18)
    x_before = x32.clone()
    xm2 = torch.zeros(M, 512, dtype=torch.bfloat16, device=dev)
    ops.gemm(h, w2, M=M, N=256, K=512, lda=512, block_n=256, epilogue=ops.EPI_RESID, bias=b2, resid=x32, out32=x32,
             ld32=256, out16=xm2, ld16=512)
    torch.cuda.synchronize()
    want = x_before + h.float() @ w2.float().t() + b2
    _close(x32, want, 1e-4, 1e-4, "residual fp32")
    _close(xm2[:, :256], want, 1e-2, 1e-2, "residual bf16 copy")


@pytest.mark.parametrize("B,M,N", [(1, 128, 256), (3, 200, 333), (2, 96, 80)])
def test_similarity_batched(dev, B, M, N):
    from glue_factory_b200 import ops

    md = _bf(_rand((B * (M + N), 256), dev, 19, 0.5))
    sim = torch.full((B, M, N), float("nan"), device=dev)
    ops.gemm(md, md, M=M, N=N, K=256, lda=256, ldb=256, a_rows=B * (M + N), b_row0=B * M, b_rows=B * N, batch=B,
             a_batch_rows=M, b_batch_rows=N, block_n=256 if N > 128 else 128, epilogue=ops.EPI_F32, alpha=1.0,
             out32=sim, ld32=N, out32_batch_stride=M * N)
    torch.cuda.synchronize()
    a = md[: B * M].float().view(B, M, 256)
    b = md[B * M:].float().view(B, N, 256)
    _close(sim, a @ b.transpose(1, 2), 1e-4, 1e-3, "similarity")


@pytest.mark.parametrize("cin,cout,H,W,pool", [(64, 64, 40, 48, 1), (64, 64, 37, 53, 0), (64, 128, 24, 40, 0),
                                               (64, 128, 37, 53, 1), (128, 128, 19, 21, 1), (128, 512, 16, 24, 0)])
def test_conv3x3_relu_bn_pool(dev, cin, cout, H, W, pool):
    from glue_factory_b200 import ops

    B = 2
    x = _bf(_rand((B, H, W, cin), dev, 20))
    w = _rand((cout, cin, 3, 3), dev, 21, 1 / math.sqrt(9 * cin))
    wp = _bf(w.permute(0, 2, 3, 1).reshape(cout, -1)).contiguous()
    bias, scale, shift = _rand((cout,), dev, 22, 0.1), 1 + _rand((cout,), dev, 23, 0.2), _rand((cout,), dev, 24, 0.2)
    oh, ow = (H // 2, W // 2) if pool else (H, W)
    y = torch.zeros(B, oh, ow, cout, dtype=torch.bfloat16, device=dev)
    ops.gemm(x, wp, M=0, N=cout, K=9 * cin, lda=cin, block_n=min(cout, 256), epilogue=ops.EPI_CONV_RELU_BN, batch=B,
             conv=(H, W, cin), bias=bias, scale=scale, shift=shift, relu=1, pool=pool, out16=y, ld16=cout)
    torch.cuda.synchronize()
    w_r = wp.float().view(cout, 3, 3, cin).permute(0, 3, 1, 2)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w_r, bias, padding=1)
    ref = F.relu(ref) * scale[None, :, None, None] + shift[None, :, None, None]
    if pool:
        ref = F.max_pool2d(ref, 2, 2)
    _close(y.permute(0, 3, 1, 2), ref, 1e-2, 2e-2, "conv")


@pytest.mark.parametrize("C,H,W,pool,B", [(1, 48, 40, 1, 2), (1, 37, 53, 0, 2), (3, 64, 24, 1, 1), (1, 16, 8, 1, 3),
                                          (1, 200, 312, 1, 2)])
def test_conv_stem_fused(dev, C, H, W, pool, B):
    """conv 1->64 (+ReLU+BN) -> conv 64->64 (+ReLU+BN) -> pool in one kernel against fp32 PyTorch (with the
    intermediate activation and the second weights rounded to bf16 as the kernel holds them) and against the
    two-kernel CUDA path (conv1a, conv3x3)."""
    from glue_factory_b200 import ops

    img = torch.rand(B, C, H, W, device=dev, generator=torch.Generator(device=dev).manual_seed(40 + H))
    w1 = _rand((64, 9), dev, 41, 1 / 3.0)
    b1, s1, t1 = _rand((64,), dev, 42, 0.1), 1 + _rand((64,), dev, 43, 0.2), _rand((64,), dev, 44, 0.2)
    w2 = _rand((64, 64, 3, 3), dev, 45, 1 / math.sqrt(9 * 64))
    w2p = _bf(w2.permute(0, 2, 3, 1).reshape(64, -1)).contiguous()
    b2, s2, t2 = _rand((64,), dev, 46, 0.1), 1 + _rand((64,), dev, 47, 0.2), _rand((64,), dev, 48, 0.2)
    oh, ow = (H // 2, W // 2) if pool else (H, W)
    y = torch.zeros(B, oh, ow, 64, dtype=torch.bfloat16, device=dev)
    ops.sp_conv_stem(img, w1, torch.stack([b1, s1, t1, b2, s2, t2]).cpu(), w2p, pool, y)
    # two-kernel path
    mid = torch.empty(B, H, W, 64, dtype=torch.bfloat16, device=dev)
    ops.sp_conv1a(img, w1, b1, s1, t1, mid)
    y2 = torch.zeros_like(y)
    ops.gemm(mid, w2p, M=0, N=64, K=576, lda=64, block_n=64, epilogue=ops.EPI_CONV_RELU_BN, batch=B, conv=(H, W, 64),
             bias=b2, scale=s2, shift=t2, relu=1, pool=pool, out16=y2, ld16=64)
    torch.cuda.synchronize()
    gray = img if C == 1 else (img * torch.tensor([0.299, 0.587, 0.114], device=dev).view(1, 3, 1, 1)).sum(1, keepdim=True)
    a = F.conv2d(gray, w1.view(64, 1, 3, 3), b1, padding=1)
    a = (F.relu(a) * s1[None, :, None, None] + t1[None, :, None, None]).to(torch.bfloat16).float()
    w_r = w2p.float().view(64, 3, 3, 64).permute(0, 3, 1, 2)
    ref = F.conv2d(a, w_r, b2, padding=1)
    ref = F.relu(ref) * s2[None, :, None, None] + t2[None, :, None, None]
    if pool:
        ref = F.max_pool2d(ref, 2, 2)
    _close(y.permute(0, 3, 1, 2), ref, 1e-2, 2e-2, "fused stem vs fp32 reference")
    # the two CUDA paths differ only where an intermediate value sat on a bf16 rounding boundary
    diff = (y.float() - y2.float()).abs()
    assert diff.max().item() <= 0.05 and (diff > 0.02).float().mean().item() < 1e-3, diff.max().item()


def test_heads_desc_l2_and_score_softmax(dev):
    from glue_factory_b200 import ops

    B, h, w = 2, 12, 20
    cells = B * h * w
    heads = _bf(_rand((cells, 512), dev, 25))
    wd, bd = _bf(_rand((256, 256), dev, 26, 1 / 16)), _rand((256,), dev, 27, 0.1)
    sd, td = 1 + _rand((256,), dev, 28, 0.1), _rand((256,), dev, 29, 0.1)
    dense = torch.empty(B, h, w, 256, device=dev)
    ops.gemm(heads, wd, M=cells, N=256, K=256, lda=512, a_col0=256, block_n=256, epilogue=ops.EPI_CONV_BN_L2,
             bias=bd, scale=sd, shift=td, out32=dense, ld32=256)
    ws, bs = _bf(_rand((65, 256), dev, 30, 1 / 8)), _rand((65,), dev, 31, 0.1)
    ss, ts = 1 + _rand((65,), dev, 32, 0.1), _rand((65,), dev, 33, 0.1)
    heat = torch.empty(B, h * 8, w * 8, device=dev)
    ops.gemm(heads, ws, M=cells, N=65, K=256, lda=512, a_col0=0, block_n=80, epilogue=ops.EPI_CONV_BN_SOFTMAX_D2S,
             bias=bs, scale=ss, shift=ts, out32=heat, cells=(h, w))
    torch.cuda.synchronize()
    ref_d = F.normalize((heads[:, 256:].float() @ wd.float().t() + bd) * sd + td, dim=1)
    _close(dense.view(cells, 256), ref_d, 1e-4, 1e-5, "descriptor head")
    logits = (heads[:, :256].float() @ ws.float().t() + bs) * ss + ts
    prob = torch.softmax(logits, 1)[:, :64].view(B, h, w, 8, 8).permute(0, 1, 3, 2, 4).reshape(B, h * 8, w * 8)
    _close(heat, prob, 1e-4, 1e-7, "score head")


def _attn_ref(q, k, v, scale):
    s = (q.float() @ k.float().t()) * scale
    return torch.softmax(s, -1) @ v.float()


@pytest.mark.parametrize("layout", ["narrow", "wide", "auto"])
@pytest.mark.parametrize("lens", [[(128, 128)], [(200, 333), (77, 130)], [(2048, 2048)], [(129, 64), (256, 257), (300, 1)],
                                  [(513, 700), (1, 5), (640, 384)], [(1100, 900)] * 16 + [(385, 70), (40, 0), (900, 64)]])
def test_attention(dev, lens, layout, monkeypatch):
    """Both stream layouts of the kernel (256-query items with key-parity streams and a merge; 512-query items with one
    stream per query tile) and the launcher's own choice; ragged lengths, several items per CTA, a sequence without keys
    (zeros, as the masked softmax of the reference over an empty key set never occurs: the kernel defines it)."""
    from glue_factory_b200 import ops

    if layout != "auto":
        monkeypatch.setenv("B200GLUE_ATT_LAYOUT", layout)
    else:
        monkeypatch.delenv("B200GLUE_ATT_LAYOUT", raising=False)
    H = 4
    total_q = sum(a for a, _ in lens)
    total_kv = max(sum(b for _, b in lens), 1)
    qb = _bf(_rand((total_q, 768), dev, 34))
    kvb = _bf(_rand((total_kv, 768), dev, 35))
    entries, qo, ko = [], 0, 0
    for a, b in lens:
        entries.append((qo, a, ko, b))
        qo += a
        ko += b
    seqs = ops.make_attn_seqs(entries, dev)
    out = torch.full((total_q, 256), 7.0, dtype=torch.bfloat16, device=dev)
    ops.attention(qb, kvb, kvb, seqs, num_seqs=len(lens), max_q_len=max(a for a, _ in lens), heads=H, ldq=768,
                  ldk=768, ldv=768, q_col0=0, k_col0=256, v_col0=512, scale=0.125, out=out, ld_out=256)
    torch.cuda.synchronize()
    for (q0, ql, k0, kl) in entries:
        if kl == 0:
            assert torch.count_nonzero(out[q0:q0 + ql]) == 0
            continue
        q4 = qb[q0:q0 + ql, :256].float().view(ql, H, 64).transpose(0, 1)
        k4 = kvb[k0:k0 + kl, 256:512].float().view(kl, H, 64).transpose(0, 1)
        v4 = kvb[k0:k0 + kl, 512:768].float().view(kl, H, 64).transpose(0, 1)
        ref = (torch.softmax(q4 @ k4.transpose(1, 2) * 0.125, -1) @ v4).transpose(0, 1).reshape(ql, 256)
        _close(out[q0:q0 + ql], ref, 2e-2, 2e-2, f"attention rows {q0}..{q0 + ql}")


def test_attention_live_lengths(dev, monkeypatch):
    """`len_override`: lengths read on the device (adaptive width); rows beyond the live length stay untouched."""
    from glue_factory_b200 import ops

    H, n = 4, 1024
    live = [700, 0, 1024, 513, 130, 1]
    x = _bf(_rand((len(live) * n, 768), dev, 36))
    seqs = ops.make_attn_seqs([(i * n, n, i * n, n, i, i) for i in range(len(live))], dev)
    lens = torch.tensor(live, dtype=torch.int32, device=dev)
    outs = {}
    for layout in ("narrow", "wide"):
        monkeypatch.setenv("B200GLUE_ATT_LAYOUT", layout)
        out = torch.full((len(live) * n, 256), 7.0, dtype=torch.bfloat16, device=dev)
        ops.attention(x, x, x, seqs, num_seqs=len(live), max_q_len=n, heads=H, ldq=768, ldk=768, ldv=768, q_col0=0,
                      k_col0=256, v_col0=512, scale=0.125, out=out, ld_out=256, len_override=lens)
        torch.cuda.synchronize()
        outs[layout] = out
        for i, m in enumerate(live):
            assert bool((out[i * n + m:(i + 1) * n] == 7.0).all()), (layout, i)
            if m == 0:
                continue
            q4 = x[i * n:i * n + m, :256].float().view(m, H, 64).transpose(0, 1)
            k4 = x[i * n:i * n + m, 256:512].float().view(m, H, 64).transpose(0, 1)
            v4 = x[i * n:i * n + m, 512:768].float().view(m, H, 64).transpose(0, 1)
            ref = (torch.softmax(q4 @ k4.transpose(1, 2) * 0.125, -1) @ v4).transpose(0, 1).reshape(m, 256)
            _close(out[i * n:i * n + m], ref, 2e-2, 2e-2, f"{layout} live rows of sequence {i}")


def test_uint8_images_equal_the_float_contract(dev, sp_weights, tmp_path):
    """uint8 pixels converted in the first kernel's load (value / 255) give bit-identical results to the float32
    images the reference's loader produces (utils/image.py:112-121), gray and RGB, fused stem and stand-alone conv1a."""
    from glue_factory_b200 import ops
    from gluefactory.models import get_model

    torch.save(sp_weights, tmp_path / "sp.pth")
    model = get_model("extractors.superpoint_open")({"weights": str(tmp_path / "sp.pth"), "max_num_keypoints": 128,
                                                      "detection_threshold": -1, "nms_radius": 3}).to(dev).eval()
    g = torch.Generator().manual_seed(1)
    for c in (1, 3):
        u8 = torch.randint(0, 256, (2, c, 96, 160), generator=g, dtype=torch.uint8)
        u8[0, 0, 40, :] = torch.arange(160, dtype=torch.uint8)          # every value 0..255 occurs
        u8[0, 0, 41, :96] = torch.arange(160, 256, dtype=torch.uint8)
        f32 = torch.from_numpy((u8.numpy() / 255.0).astype("float32"))
        with torch.no_grad():
            a = model({"image": u8.to(dev)})
            b = model({"image": f32.to(dev)})
        for k in ("heatmap", "keypoints", "keypoint_scores", "descriptors"):
            assert torch.equal(a[k], b[k]), (c, k)
        P = model._pack()
        o1 = torch.empty(2, 96, 160, 64, dtype=torch.bfloat16, device=dev)
        o2 = torch.empty_like(o1)
        ops.sp_conv1a(u8.to(dev), P["c1a_w"], P["c1a_b"], P["c1a_s"], P["c1a_t"], o1)
        ops.sp_conv1a(f32.to(dev), P["c1a_w"], P["c1a_b"], P["c1a_s"], P["c1a_t"], o2)
        assert torch.equal(o1, o2)
