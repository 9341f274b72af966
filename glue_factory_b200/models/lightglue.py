"""LightGlue matcher on the B200 kernels, behind the Glue Factory model API.

Drop-in for `gluefactory.models.matchers.lightglue.LightGlue`
(reference: gluefactory/models/matchers/lightglue.py:306-525): same `default_conf`, same
state-dict schema, same `forward(data)` keys and output dtypes.  The nn.Linear / nn.LayerNorm
children only own the parameters; `forward` launches libb200glue.so kernels:

  ingest + Fourier encoding -> per layer { QKV GEMM with rotary epilogue, flash self-attention,
  out_proj GEMM, concat-FFN GEMMs with LayerNorm+GELU and residual epilogues, shared qk|v GEMM,
  two cross-attention directions, to_out GEMM, FFN } -> final_proj GEMM, similarity GEMM,
  fused dual-softmax log-assignment + mutual-NN filter.

Both images' tokens share every GEMM launch (the weights are shared, lightglue.py:224-238), laid
out as [image0: B*M rows | image1: B*N rows].  No PyTorch fallback: CPU tensors raise.
"""
import os
import warnings
from pathlib import Path

import numpy as np
import torch
from torch import nn

from gluefactory.models.utils.losses import NLLLoss
from gluefactory.settings import DATA_PATH

from .. import ops
from .._conf import get_conf_backend

_, OmegaConf = get_conf_backend()


# ---- parameter containers with the reference's attribute names ----------------------------------
class LearnableFourierPositionalEncoding(nn.Module):
    def __init__(self, M: int, dim: int, F_dim: int = None, gamma: float = 1.0):
        super().__init__()
        F_dim = F_dim if F_dim is not None else dim
        self.gamma = gamma
        self.Wr = nn.Linear(M, F_dim // 2, bias=False)
        nn.init.normal_(self.Wr.weight.data, mean=0, std=self.gamma**-2)


class TokenConfidence(nn.Module):
    def __init__(self, dim: int):
        super().__init__()
        self.token = nn.Sequential(nn.Linear(dim, 1), nn.Sigmoid())


def _ffn(dim):
    return nn.Sequential(nn.Linear(2 * dim, 2 * dim), nn.LayerNorm(2 * dim, elementwise_affine=True), nn.GELU(),
                         nn.Linear(2 * dim, dim))


class SelfBlock(nn.Module):
    def __init__(self, embed_dim: int, num_heads: int, flash: bool = False, bias: bool = True):
        super().__init__()
        self.Wqkv = nn.Linear(embed_dim, 3 * embed_dim, bias=bias)
        self.out_proj = nn.Linear(embed_dim, embed_dim, bias=bias)
        self.ffn = _ffn(embed_dim)


class CrossBlock(nn.Module):
    def __init__(self, embed_dim: int, num_heads: int, flash: bool = False, bias: bool = True):
        super().__init__()
        self.to_qk = nn.Linear(embed_dim, embed_dim, bias=bias)
        self.to_v = nn.Linear(embed_dim, embed_dim, bias=bias)
        self.to_out = nn.Linear(embed_dim, embed_dim, bias=bias)
        self.ffn = _ffn(embed_dim)


class TransformerLayer(nn.Module):
    def __init__(self, *args, **kwargs):
        super().__init__()
        self.self_attn = SelfBlock(*args, **kwargs)
        self.cross_attn = CrossBlock(*args, **kwargs)


class MatchAssignment(nn.Module):
    def __init__(self, dim: int):
        super().__init__()
        self.dim = dim
        self.matchability = nn.Linear(dim, 1, bias=True)
        self.final_proj = nn.Linear(dim, dim, bias=True)


def _bf16(t):
    return t.detach().to(torch.bfloat16).contiguous()


def _f32(t):
    return t.detach().float().contiguous()


def _split_b(w):
    """fp32 [rows, K] -> bf16 [rows, 3K] = [hi | lo | hi]: the weight side of the two-term split GEMM
    (pairs with ops.split_hilo's [hi | hi | lo] on the activation side: x_hi.w_hi + x_hi.w_lo + x_lo.w_hi)."""
    w = w.detach().float()
    hi = w.to(torch.bfloat16)
    lo = (w - hi.float()).to(torch.bfloat16)
    return torch.cat([hi, lo, hi], 1).contiguous()


def _fold_out_proj(w1, b1, wo, bo):
    """cat[x, out_proj(ctx)] . W1^T + b1  ==  cat[x, ctx] . [W1a | W1b Wo]^T + (b1 + W1b bo): the message projection
    (out_proj / to_out, lightglue.py:155,211-212) is a linear map feeding the first FFN layer (:138-143,157) with nothing
    in between, so the two weights are multiplied once at packing time (fp64 product, one bf16 rounding) and the
    256 x 256 GEMM per block disappears from the forward pass."""
    w1, wo = w1.detach().double(), wo.detach().double()
    d = wo.shape[0]
    w = torch.cat([w1[:, :d], w1[:, d:] @ wo], 1)
    b = b1.detach().double() + w1[:, d:] @ bo.detach().double()
    return _bf16(w.float()), _f32(b.float())


# B200GLUE_FOLD_OUT_PROJ=0 keeps out_proj / to_out as their own GEMMs (A/B measurement)
FOLD_OUT_PROJ = os.environ.get("B200GLUE_FOLD_OUT_PROJ", "1") != "0"
# B200GLUE_ASSIGN_BF16=1: plain bf16 final projection / similarity (A/B measurement of the fp32-grade path)
ASSIGN_BF16 = os.environ.get("B200GLUE_ASSIGN_BF16", "0") == "1"


class LightGlue(nn.Module):
    default_conf = {
        "name": "lightglue",
        "input_dim": 256,
        "add_scale_ori": False,
        "descriptor_dim": 256,
        "n_layers": 9,
        "num_heads": 4,
        "flash": False,
        "mp": False,
        "depth_confidence": -1,
        "width_confidence": -1,
        "filter_threshold": 0.0,
        "checkpointed": False,
        "weights": None,
        "weights_from_version": "v0.1_arxiv",
        "loss": {"gamma": 1.0, "fn": "nll", "nll_balancing": 0.5},
    }
    required_data_keys = ["keypoints0", "keypoints1", "descriptors0", "descriptors1"]
    url = "https://github.com/cvg/LightGlue/releases/download/{}/{}_lightglue.pth"

    def __init__(self, conf) -> None:
        super().__init__()
        self.conf = conf = OmegaConf.merge(self.default_conf, conf)
        if conf.input_dim != 256 or conf.descriptor_dim != 256 or conf.num_heads != 4 or conf.add_scale_ori:
            raise NotImplementedError("the B200 kernels are built for 256-d descriptors, 4 heads, no scale/orientation")
        self.input_proj = nn.Identity()
        head_dim = conf.descriptor_dim // conf.num_heads
        self.posenc = LearnableFourierPositionalEncoding(2, head_dim, head_dim)
        h, n, d = conf.num_heads, conf.n_layers, conf.descriptor_dim
        self.transformers = nn.ModuleList([TransformerLayer(d, h, conf.flash) for _ in range(n)])
        self.log_assignment = nn.ModuleList([MatchAssignment(d) for _ in range(n)])
        self.token_confidence = nn.ModuleList([TokenConfidence(d) for _ in range(n - 1)])
        self.loss_fn = NLLLoss(conf.loss)
        self.confidence_thresholds = [self.confidence_threshold(i) for i in range(n)]

        state_dict = None
        if conf.weights is not None:
            if Path(conf.weights).exists():
                state_dict = torch.load(conf.weights, map_location="cpu")
            elif (Path(DATA_PATH) / conf.weights).exists():
                state_dict = torch.load(str(DATA_PATH / conf.weights), map_location="cpu")
            else:
                fname = f"{conf.weights}_{conf.weights_from_version}".replace(".", "-") + ".pth"
                state_dict = torch.hub.load_state_dict_from_url(
                    self.url.format(conf.weights_from_version, conf.weights), file_name=fname
                )
        if state_dict:
            for i in range(n):  # legacy key names of the official checkpoints
                state_dict = {k.replace(f"self_attn.{i}", f"transformers.{i}.self_attn"): v for k, v in state_dict.items()}
                state_dict = {k.replace(f"cross_attn.{i}", f"transformers.{i}.cross_attn"): v for k, v in state_dict.items()}
            self.load_state_dict(state_dict, strict=False)
        self._packed = None
        self._packed_key = None
        self._seq_cache = {}
        self.layer_tap = None  # set to a list to collect the residual stream after every layer (tests)

    def compile(self, mode="reduce-overhead"):
        warnings.warn("compile() is a no-op: the layers already run as hand-written kernels", stacklevel=2)

    def confidence_threshold(self, layer_index: int) -> float:
        """(reference: lightglue.py:527-530)"""
        return float(np.clip(0.8 + 0.1 * np.exp(-4.0 * layer_index / self.conf.n_layers), 0, 1))

    # -- weight repacking --------------------------------------------------------------------------
    def _pack_key(self):
        return tuple((p.data_ptr(), p._version) for p in self.parameters())

    def _pack(self):
        key = self._pack_key()
        if self._packed is not None and key == self._packed_key:
            return self._packed
        L = []
        # Wqkv rows are packed (head, dim, {q,k,v}) in the reference (lightglue.py:151); regroup to [q | k | v]
        perm = torch.arange(768).view(4, 64, 3).permute(2, 0, 1).reshape(-1)
        for t in self.transformers:
            sa, ca = t.self_attn, t.cross_attn
            perm_d = perm.to(sa.Wqkv.weight.device)
            L.append(dict(
                wqkv=_bf16(sa.Wqkv.weight[perm_d]), bqkv=_f32(sa.Wqkv.bias[perm_d]),
                wo=_bf16(sa.out_proj.weight), bo=_f32(sa.out_proj.bias),
                s_w1=_bf16(sa.ffn[0].weight), s_b1=_f32(sa.ffn[0].bias), s_g=_f32(sa.ffn[1].weight),
                s_be=_f32(sa.ffn[1].bias), s_eps=float(sa.ffn[1].eps), s_w2=_bf16(sa.ffn[3].weight),
                s_b2=_f32(sa.ffn[3].bias),
                wc=_bf16(torch.cat([ca.to_qk.weight, ca.to_v.weight], 0)),
                bc=_f32(torch.cat([ca.to_qk.bias, ca.to_v.bias], 0)),
                wco=_bf16(ca.to_out.weight), bco=_f32(ca.to_out.bias),
                c_w1=_bf16(ca.ffn[0].weight), c_b1=_f32(ca.ffn[0].bias), c_g=_f32(ca.ffn[1].weight),
                c_be=_f32(ca.ffn[1].bias), c_eps=float(ca.ffn[1].eps), c_w2=_bf16(ca.ffn[3].weight),
                c_b2=_f32(ca.ffn[3].bias),
            ))
            if FOLD_OUT_PROJ:
                L[-1]["s_w1"], L[-1]["s_b1"] = _fold_out_proj(sa.ffn[0].weight, sa.ffn[0].bias, sa.out_proj.weight,
                                                             sa.out_proj.bias)
                L[-1]["c_w1"], L[-1]["c_b1"] = _fold_out_proj(ca.ffn[0].weight, ca.ffn[0].bias, ca.to_out.weight,
                                                             ca.to_out.bias)
        A = []
        for a in self.log_assignment:
            s = float(a.dim) ** -0.25  # 0.25 for dim 256: folded into the projection (exact power of two)
            A.append(dict(wf=_bf16(a.final_proj.weight * s), wf_split=_split_b(a.final_proj.weight * s),
                          bf=_f32(a.final_proj.bias * s),
                          wm=_f32(a.matchability.weight.reshape(-1)), bm=_f32(a.matchability.bias)))
        T = [dict(w=_f32(t.token[0].weight.reshape(-1)), b=_f32(t.token[0].bias)) for t in self.token_confidence]
        self._packed = dict(layers=L, assign=A, token=T, wr=_f32(self.posenc.Wr.weight))
        self._packed_key = key
        return self._packed

    # -- building blocks ---------------------------------------------------------------------------
    @staticmethod
    def _ffn_block(ws, T, w1, b1, g, be, eps, w2, b2, tile_live=None):
        ops.gemm(ws["xm"], w1, M=T, N=512, K=512, lda=512, block_n=512, epilogue=ops.EPI_LN_GELU, bias=b1,
                 ln_gamma=g, ln_beta=be, ln_eps=eps, out16=ws["h"], ld16=512, tile_live=tile_live)
        ops.gemm(ws["h"], w2, M=T, N=256, K=512, lda=512, block_n=256, epilogue=ops.EPI_RESID, bias=b2,
                 resid=ws["x32"], out32=ws["x32"], ld32=256, out16=ws["xm"], ld16=512, tile_live=tile_live)

    def _layer(self, P, ws, T, seqs_self, seqs_cross, nseq, max_len, work_self=0.0, work_cross=0.0, lens=None,
               tile_live=None):
        """One TransformerLayer (lightglue.py:218-247).  `lens` / `tile_live` (adaptive mode): device-side live
        sequence lengths for the attention kernels and live 128-row tiles for the GEMMs."""
        tl = tile_live
        # self attention (lightglue.py:144-157)
        ops.gemm(ws["xm"], P["wqkv"], M=T, N=768, K=256, lda=512, block_n=256, epilogue=ops.EPI_QKV_ROTARY,
                 bias=P["bqkv"], cossin=ws["cs"], rot_cols=512, out16=ws["qkv"], ld16=768, tile_live=tl)
        # with the message projection folded into the FFN weights the attention writes its context straight into the
        # right half of XM = cat[x, .]
        fold = FOLD_OUT_PROJ
        ctx, ld_ctx, col_ctx = (ws["xm"], 512, 256) if fold else (ws["ctx"], 256, 0)
        ops.attention(ws["qkv"], ws["qkv"], ws["qkv"], seqs_self, num_seqs=nseq, max_q_len=max_len, heads=4,
                      ldq=768, ldk=768, ldv=768, q_col0=0, k_col0=256, v_col0=512, scale=0.125, out=ctx,
                      ld_out=ld_ctx, out_col0=col_ctx, attn_work=work_self, len_override=lens)
        if not fold:
            ops.gemm(ws["ctx"], P["wo"], M=T, N=256, K=256, lda=256, block_n=256, epilogue=ops.EPI_BIAS_BF16,
                     bias=P["bo"], out16=ws["xm"], out16_col0=256, ld16=512, tile_live=tl)
        self._ffn_block(ws, T, P["s_w1"], P["s_b1"], P["s_g"], P["s_be"], P["s_eps"], P["s_w2"], P["s_b2"], tl)
        # bidirectional cross attention (lightglue.py:186-215)
        ops.gemm(ws["xm"], P["wc"], M=T, N=512, K=256, lda=512, block_n=256, epilogue=ops.EPI_BIAS_BF16,
                 bias=P["bc"], out16=ws["qkv"], ld16=768, tile_live=tl)
        ops.attention(ws["qkv"], ws["qkv"], ws["qkv"], seqs_cross, num_seqs=nseq, max_q_len=max_len, heads=4,
                      ldq=768, ldk=768, ldv=768, q_col0=0, k_col0=0, v_col0=256, scale=0.125, out=ctx,
                      ld_out=ld_ctx, out_col0=col_ctx, attn_work=work_cross, len_override=lens)
        if not fold:
            ops.gemm(ws["ctx"], P["wco"], M=T, N=256, K=256, lda=256, block_n=256, epilogue=ops.EPI_BIAS_BF16,
                     bias=P["bco"], out16=ws["xm"], out16_col0=256, ld16=512, tile_live=tl)
        self._ffn_block(ws, T, P["c_w1"], P["c_b1"], P["c_g"], P["c_be"], P["c_eps"], P["c_w2"], P["c_b2"], tl)

    def _project(self, A, ws, T, tile_live=None):
        """final_proj (fp32-grade two-term split GEMM, scaled by dim^-1/4) and matchability logits of every row:
        (md32 [T,256], z [T])  (lightglue.py:270-275,283-284)"""
        dev = ws["x32"].device
        z = torch.empty(T, dtype=torch.float32, device=dev)
        ops.lg_rowdot(ws["x32"], T, A["wm"], A["bm"], z)
        xs = torch.empty(T, 768, dtype=torch.bfloat16, device=dev)
        ops.split_hilo(ws["x32"], T, 256, T, xs)
        md32 = torch.empty(T, 256, dtype=torch.float32, device=dev)
        ops.gemm(xs, A["wf_split"], M=T, N=256, K=768, lda=768, block_n=256, epilogue=ops.EPI_F32, bias=A["bf"],
                 alpha=1.0, out32=md32, ld32=256, out32_batch_stride=0, tile_live=tile_live)
        return md32, z

    @staticmethod
    def _similarity(md32, B, M, N, T):
        """sim[b] = md0_b . md1_b^T through the two-term split (K = 768), fp32 [B,M,N] (lightglue.py:276)"""
        dev = md32.device
        md = torch.empty(T, 768, dtype=torch.bfloat16, device=dev)
        ops.split_hilo(md32, T, 256, B * M, md)
        sim = torch.empty(B, M, N, dtype=torch.float32, device=dev)
        ops.gemm(md, md, M=M, N=N, K=768, lda=768, ldb=768, a_rows=T, b_row0=B * M, b_rows=B * N, batch=B,
                 a_batch_rows=M, b_batch_rows=N, block_n=256 if N > 128 else 128, epilogue=ops.EPI_F32, alpha=1.0,
                 out32=sim, ld32=N, out32_batch_stride=M * N)
        return sim

    def _assign(self, A, ws, B, M, N, T):
        """MatchAssignment.forward + filter_matches on the current descriptors (lightglue.py:265-303).

        north_star asks log_assignment within 1e-3 relative of the fp32 reference.  The final projection and the
        similarity therefore run as two-term bf16 split GEMMs (K = 3 x 256: hi.hi + hi.lo + lo.hi, fp32
        accumulate) on the fp32 residual stream, i.e. with ~2^-16 relative operand error instead of 2^-9."""
        dev = ws["x32"].device
        if ASSIGN_BF16:
            z = torch.empty(T, dtype=torch.float32, device=dev)
            ops.lg_rowdot(ws["x32"], T, A["wm"], A["bm"], z)
            sim = torch.empty(B, M, N, dtype=torch.float32, device=dev)
            bn = 256 if N > 128 else 128
            ops.gemm(ws["xm"], A["wf"], M=T, N=256, K=256, lda=512, block_n=256, epilogue=ops.EPI_BIAS_BF16,
                     bias=A["bf"], out16=ws["md"], ld16=256)
            ops.gemm(ws["md"], ws["md"], M=M, N=N, K=256, lda=256, ldb=256, a_rows=T, b_row0=B * M, b_rows=B * N,
                     batch=B, a_batch_rows=M, b_batch_rows=N, block_n=bn, epilogue=ops.EPI_F32,
                     alpha=1.0, out32=sim, ld32=N, out32_batch_stride=M * N)
        else:
            md32, z = self._project(A, ws, T)
            sim = self._similarity(md32, B, M, N, T)
        return ops.lg_assignment(sim, z[: B * M], z[B * M:], self.conf.filter_threshold) + (sim,)

    def assign_from_descriptors(self, desc0, desc1, layer=None):
        """Final stage alone, teacher-forced: fp32 descriptors of the last layer (B,M,256), (B,N,256) ->
        (log_assignment, matches0, matches1, matching_scores0, matching_scores1).  Used by the parity tests."""
        B, M, _ = desc0.shape
        N = desc1.shape[1]
        T = B * (M + N)
        dev = desc0.device
        with torch.cuda.device(dev):
            P = self._pack()
            ws = self._workspace(T, dev)
            ops.lg_ingest(desc0.float().contiguous(), ws["x32"], ws["xm"], 512, row0=0)
            ops.lg_ingest(desc1.float().contiguous(), ws["x32"], ws["xm"], 512, row0=B * M)
            layer = self.conf.n_layers - 1 if layer is None else layer
            return self._assign(P["assign"][layer], ws, B, M, N, T)[:5]

    def _seqs(self, B, M, N, dev):
        """Attention problem lists of a (B, M, N) batch: self (each image with itself) and cross (both directions),
        with len_override slots (b, B + b) for the adaptive mode.  Built once per shape and kept on the device."""
        key = (B, M, N, str(dev))
        if key not in self._seq_cache:
            off1 = B * M
            if len(self._seq_cache) > 16:
                self._seq_cache.clear()
            self._seq_cache[key] = (
                ops.make_attn_seqs([(b * M, M, b * M, M, b, b) for b in range(B)] +
                                   [(off1 + b * N, N, off1 + b * N, N, B + b, B + b) for b in range(B)], dev),
                ops.make_attn_seqs([(b * M, M, off1 + b * N, N, b, B + b) for b in range(B)] +
                                   [(off1 + b * N, N, b * M, M, B + b, b) for b in range(B)], dev))
        return self._seq_cache[key]

    def _workspace(self, T, dev, zero=False):
        """`zero`: adaptive mode keeps dead rows in the buffers (pruned tokens, skipped tiles); they must hold finite
        values because the attention kernel multiplies masked keys' V rows by exact zeros."""
        bf = torch.bfloat16
        new = torch.zeros if zero else torch.empty
        return dict(
            x32=new(T, 256, dtype=torch.float32, device=dev),
            xm=new(T, 512, dtype=bf, device=dev),
            qkv=new(T, 768, dtype=bf, device=dev),
            ctx=new(T, 256, dtype=bf, device=dev),
            h=new(T, 512, dtype=bf, device=dev),
            md=new(T, 256, dtype=bf, device=dev),
            cs=new(T, 32, 2, dtype=torch.float32, device=dev),
        )

    def forward(self, data: dict) -> dict:
        for key in self.required_data_keys:
            assert key in data, f"Missing key {key} in data"
        if not data["keypoints0"].is_cuda:
            raise RuntimeError("glue_factory_b200 LightGlue runs on CUDA only (no CPU fallback)")
        # kernels launch on the current device's stream: make the tensors' device current for the whole pass
        with torch.cuda.device(data["keypoints0"].device):
            return self._forward(data)

    def _forward(self, data: dict) -> dict:
        kpts0, kpts1 = data["keypoints0"], data["keypoints1"]
        B, M, _ = kpts0.shape
        N = kpts1.shape[1]
        dev = kpts0.device
        size0 = size1 = None
        if "view0" in data.keys() and "view1" in data.keys():
            size0 = data["view0"].get("image_size")
            size1 = data["view1"].get("image_size")
        desc0 = data["descriptors0"].float().contiguous()
        desc1 = data["descriptors1"].float().contiguous()
        assert desc0.shape[-1] == self.conf.input_dim
        assert desc1.shape[-1] == self.conf.input_dim
        do_early_stop = self.conf.depth_confidence > 0 and not self.training
        do_point_pruning = self.conf.width_confidence > 0 and not self.training
        if do_early_stop or do_point_pruning:
            from .lightglue_adaptive import adaptive_forward

            return adaptive_forward(self, data, kpts0, kpts1, desc0, desc1, size0, size1)

        P = self._pack()
        T = B * (M + N)
        ws = self._workspace(T, dev)
        ops.lg_ingest(desc0, ws["x32"], ws["xm"], 512, row0=0)
        ops.lg_ingest(desc1, ws["x32"], ws["xm"], 512, row0=B * M)
        ops.lg_posenc(kpts0.float().contiguous(), _size(size0, dev, B), P["wr"], ws["cs"], row0=0)
        ops.lg_posenc(kpts1.float().contiguous(), _size(size1, dev, B), P["wr"], ws["cs"], row0=B * M)
        seqs_self, seqs_cross = self._seqs(B, M, N, dev)
        for i in range(self.conf.n_layers):
            # algorithmic FLOPs: self 1024 n^2 per image, cross 1536 M N per pair (sim counted once)
            self._layer(P["layers"][i], ws, T, seqs_self, seqs_cross, 2 * B, max(M, N),
                        work_self=1024.0 * B * (M * M + N * N), work_cross=1536.0 * B * M * N)
            if self.layer_tap is not None:  # parity tests: per-layer descriptors (fp32 residual stream)
                self.layer_tap.append(ws["x32"].clone())
        scores, m0, m1, ms0, ms1, _ = self._assign(P["assign"][self.conf.n_layers - 1], ws, B, M, N, T)
        x = ws["x32"]
        return {
            "matches0": m0,
            "matches1": m1,
            "matching_scores0": ms0,
            "matching_scores1": ms1,
            "ref_descriptors0": x[: B * M].view(B, 1, M, 256),
            "ref_descriptors1": x[B * M:].view(B, 1, N, 256),
            "log_assignment": scores,
            "prune0": torch.full_like(ms0, self.conf.n_layers),
            "prune1": torch.full_like(ms1, self.conf.n_layers),
        }

    def loss(self, pred, data):
        raise NotImplementedError("training is outside the forward-only scope of the B200 path")


def _size(size, dev, B):
    """image_size as the kernels index it: float32 [B, 2] = (w, h).  The reference broadcasts
    (`shift[..., None, :]`, lightglue.py:29-32), so a (2,), (1,2) or list size is accepted for any B."""
    if size is None:
        return None
    if not isinstance(size, torch.Tensor):
        size = torch.tensor(size)
    size = size.to(device=dev, dtype=torch.float32)
    if size.numel() not in (2, 2 * B) or size.shape[-1] != 2:
        raise ValueError(f"image_size of shape {tuple(size.shape)} does not broadcast to ({B}, 2)")
    return size.reshape(-1, 2).expand(B, 2).contiguous()


__main_model__ = LightGlue
