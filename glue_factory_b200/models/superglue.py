"""SuperGlue matcher (the `gluefactory_nonfree.superglue` model) on the B200 kernels.

Drop-in for `gluefactory_nonfree.superglue.SuperGlue` (reference: gluefactory_nonfree/superglue.py:204-307,
studied for behaviour only): same `default_conf`, same state-dict keys / Conv1d shapes, same `forward(data)`
contract.  The nn.Conv1d / nn.BatchNorm1d children only own the parameters; `_forward` launches
libb200glue.so kernels on token-major tensors [image 0: B*M rows | image 1: B*N rows]:

  keypoint encoder (CUDA-core head 3->32->64, then tcgen05 GEMMs 64->128->256->256, eval-mode BatchNorm1d
  folded, ReLU in the epilogue, descriptor added in the residual epilogue)
  -> 18 x { fused q|k|v GEMM, flash attention (self or cross), merge GEMM, MLP GEMMs with ReLU and residual
  epilogues } -> final_proj -> two-term bf16 split -> K=768 similarity GEMM (near-fp32 cost) ->
  persistent log-domain Sinkhorn kernel -> arg-max + mutual-NN filter.

The reference interleaves attention heads as channel = dim * 4 + head (`view(b, 64, 4, N)`, :123-126); the
kernels want a head as 64 contiguous columns, so projection rows / merge columns are permuted when packing.
There is no PyTorch fallback: CPU tensors raise.
"""
import logging
from copy import deepcopy

import torch
from torch import nn

from gluefactory.models.base_model import BaseModel

from .. import ops
from .lightglue import _size


def _pointwise_stack(widths, norm=True):
    """Stack of 1x1 Conv1d layers held in an nn.Sequential whose child indices give the checkpoint key names:
    a hidden layer takes the slots (conv, [BatchNorm1d], ReLU), the output layer a single conv slot."""
    table = {}
    pairs = list(zip(widths[:-1], widths[1:]))
    for li, (c_in, c_out) in enumerate(pairs):
        table[len(table)] = nn.Conv1d(c_in, c_out, 1)
        if li + 1 < len(pairs):
            if norm:
                table[len(table)] = nn.BatchNorm1d(c_out)
            table[len(table)] = nn.ReLU()
    return nn.Sequential(*(table[i] for i in range(len(table))))


class KeypointEncoder(nn.Module):
    def __init__(self, feature_dim, layers, use_scores=True):
        super().__init__()
        self.use_scores = use_scores
        self.encoder = _pointwise_stack([3 if use_scores else 2, *layers, feature_dim])
        nn.init.constant_(self.encoder[-1].bias, 0.0)


class MultiHeadedAttention(nn.Module):
    def __init__(self, h, d_model):
        super().__init__()
        self.dim, self.h = d_model // h, h
        self.merge = nn.Conv1d(d_model, d_model, kernel_size=1)
        self.proj = nn.ModuleList([deepcopy(self.merge) for _ in range(3)])


class AttentionalPropagation(nn.Module):
    def __init__(self, num_dim, num_heads):
        super().__init__()
        self.attn = MultiHeadedAttention(num_heads, num_dim)
        self.mlp = _pointwise_stack([2 * num_dim, 2 * num_dim, num_dim])
        nn.init.constant_(self.mlp[-1].bias, 0.0)


class AttentionalGNN(nn.Module):
    def __init__(self, feature_dim, layer_names):
        super().__init__()
        self.layers = nn.ModuleList([AttentionalPropagation(feature_dim, 4) for _ in range(len(layer_names))])
        self.names = layer_names


def _fold(conv, bn=None):
    """Conv1d(k=1) [+ eval-mode BatchNorm1d] -> (W [out,in] fp32, b [out] fp32)."""
    w = conv.weight.detach().float().squeeze(-1)
    b = conv.bias.detach().float()
    if bn is not None:
        s = bn.weight.detach().float() / torch.sqrt(bn.running_var.detach().float() + bn.eps)
        w = w * s[:, None]
        b = (b - bn.running_mean.detach().float()) * s + bn.bias.detach().float()
    return w.contiguous(), b.contiguous()


def _bf16(t):
    return t.to(torch.bfloat16).contiguous()


class SuperGlue(BaseModel):
    default_conf = {
        "descriptor_dim": 256,
        "weights": "outdoor",
        "keypoint_encoder": [32, 64, 128, 256],
        "GNN_layers": ["self", "cross"] * 9,
        "num_sinkhorn_iterations": 50,
        "filter_threshold": 0.2,
        "use_scores": True,
        "loss": {"nll_balancing": 0.5},
    }
    required_data_keys = ["view0", "view1", "keypoints0", "keypoints1", "descriptors0", "descriptors1",
                          "keypoint_scores0", "keypoint_scores1"]
    checkpoint_url = "https://github.com/magicleap/SuperGluePretrainedNetwork/raw/master/models/weights/superglue_{}.pth"

    def _init(self, conf):
        if conf.descriptor_dim != 256 or list(conf.keypoint_encoder) != [32, 64, 128, 256] or not conf.use_scores:
            raise NotImplementedError("the B200 kernels are built for 256-d descriptors and the default keypoint encoder")
        self.kenc = KeypointEncoder(conf.descriptor_dim, conf.keypoint_encoder, conf.use_scores)
        self.gnn = AttentionalGNN(conf.descriptor_dim, list(conf.GNN_layers))
        self.final_proj = nn.Conv1d(conf.descriptor_dim, conf.descriptor_dim, kernel_size=1, bias=True)
        self.register_parameter("bin_score", torch.nn.Parameter(torch.tensor(1.0)))
        if conf.weights:
            assert conf.weights in ["indoor", "outdoor"]
            self.load_state_dict(torch.hub.load_state_dict_from_url(self.checkpoint_url.format(conf.weights)))
            logging.info(f"Loading SuperGlue trained for {conf.weights}.")
        self._packed = None
        self._packed_key = None

    # -- weight repacking (lazy, redone when parameters / buffers move or change) ------------------
    def _pack(self):
        key = tuple((p.data_ptr(), p._version) for p in list(self.parameters()) + list(self.buffers()))
        if self._packed is not None and key == self._packed_key:
            return self._packed
        enc = self.kenc.encoder
        P = {}
        P["k1_w"], P["k1_b"] = _fold(enc[0], enc[1])
        P["k2_w"], P["k2_b"] = _fold(enc[3], enc[4])
        w, P["k3_b"] = _fold(enc[6], enc[7])
        P["k3_w"] = _bf16(w)
        w, P["k4_b"] = _fold(enc[9], enc[10])
        P["k4_w"] = _bf16(w)
        w, P["k5_b"] = _fold(enc[12])
        P["k5_w"] = _bf16(w)
        dev = self.bin_score.device
        # kernel column h*64+d  <-  reference channel d*4+h
        perm = (torch.arange(64)[None, :] * 4 + torch.arange(4)[:, None]).reshape(-1).to(dev)
        L = []
        for layer in self.gnn.layers:
            a = layer.attn
            ws, bs = zip(*(_fold(pr) for pr in a.proj))
            wq = torch.cat([w_[perm] for w_ in ws], 0)
            bq = torch.cat([b_[perm] for b_ in bs], 0)
            wm, bm = _fold(a.merge)
            w0, b0 = _fold(layer.mlp[0], layer.mlp[1])
            w3, b3 = _fold(layer.mlp[3])
            # the merge convolution feeds the first MLP layer directly (superglue.py:131-140: mlp(cat[x, merge(ctx)])):
            # the two linear maps are multiplied once here (fp64 product, one bf16 rounding), so the forward pass
            # has no merge GEMM and the attention writes its context straight into the right half of cat[x, .]
            w0d, wmd = w0.double(), wm[:, perm].double()
            w0 = torch.cat([w0d[:, :256], w0d[:, 256:] @ wmd], 1).float()
            b0 = (b0.double() + w0d[:, 256:] @ bm.double()).float().contiguous()
            L.append(dict(wqkv=_bf16(wq), bqkv=bq.contiguous(), w0=_bf16(w0), b0=b0, w3=_bf16(w3), b3=b3))
        P["layers"] = L
        w, P["f_b"] = _fold(self.final_proj)
        P["f_w"] = _bf16(w)
        P["bin"] = self.bin_score.detach().float().reshape(1).contiguous()
        self._packed, self._packed_key = P, key
        return P

    @ops.on_device("keypoints0")
    def _forward(self, data):
        kpts0, kpts1 = data["keypoints0"], data["keypoints1"]
        if kpts0.shape[1] == 0 or kpts1.shape[1] == 0:  # no keypoints
            shape0, shape1 = kpts0.shape[:-1], kpts1.shape[:-1]
            return {
                "matches0": kpts0.new_full(shape0, -1, dtype=torch.int),
                "matches1": kpts1.new_full(shape1, -1, dtype=torch.int),
                "matching_scores0": kpts0.new_zeros(shape0),
                "matching_scores1": kpts1.new_zeros(shape1),
            }
        if not kpts0.is_cuda:
            raise RuntimeError("glue_factory_b200 SuperGlue runs on CUDA only (no CPU fallback)")
        dev = kpts0.device
        B, M, _ = kpts0.shape
        N = kpts1.shape[1]
        T = B * (M + N)
        P = self._pack()

        def size_of(view):
            size = view.get("image_size")
            if size is None:
                _, _, h, w = view["image"].shape
                size = [float(w), float(h)]
            return _size(size, dev, B)  # broadcast to [B, 2] as the kernel indexes it

        size0, size1 = size_of(data["view0"]), size_of(data["view1"])
        bf = torch.bfloat16
        x32 = torch.empty(T, 256, dtype=torch.float32, device=dev)
        xm = torch.empty(T, 512, dtype=bf, device=dev)
        qkv = torch.empty(T, 768, dtype=bf, device=dev)
        ctx = torch.empty(T, 256, dtype=bf, device=dev)
        hbuf = torch.empty(T, 512, dtype=bf, device=dev)
        off1 = B * M
        # descriptors -> residual stream; keypoint encoder added on top (superglue.py:89-102,287-288)
        ops.lg_ingest(data["descriptors0"].float().contiguous(), x32, xm, 512, row0=0)
        ops.lg_ingest(data["descriptors1"].float().contiguous(), x32, xm, 512, row0=off1)
        e64 = torch.empty(T, 64, dtype=bf, device=dev)
        ops.sg_kenc_head(kpts0.float().contiguous(), data["keypoint_scores0"].float().contiguous(), size0,
                         P["k1_w"], P["k1_b"], P["k2_w"], P["k2_b"], e64, row0=0)
        ops.sg_kenc_head(kpts1.float().contiguous(), data["keypoint_scores1"].float().contiguous(), size1,
                         P["k1_w"], P["k1_b"], P["k2_w"], P["k2_b"], e64, row0=off1)
        ops.gemm(e64, P["k3_w"], M=T, N=128, K=64, lda=64, block_n=128, epilogue=ops.EPI_BIAS_BF16, bias=P["k3_b"],
                 relu=1, out16=ctx, ld16=256)
        ops.gemm(ctx, P["k4_w"], M=T, N=256, K=128, lda=256, block_n=256, epilogue=ops.EPI_BIAS_BF16, bias=P["k4_b"],
                 relu=1, out16=hbuf, ld16=512)
        ops.gemm(hbuf, P["k5_w"], M=T, N=256, K=256, lda=512, block_n=256, epilogue=ops.EPI_RESID, bias=P["k5_b"],
                 resid=x32, out32=x32, ld32=256, out16=xm, ld16=512)

        seqs_self = ops.make_attn_seqs(
            [(b * M, M, b * M, M) for b in range(B)] + [(off1 + b * N, N, off1 + b * N, N) for b in range(B)], dev)
        seqs_cross = ops.make_attn_seqs(
            [(b * M, M, off1 + b * N, N) for b in range(B)] + [(off1 + b * N, N, b * M, M) for b in range(B)], dev)
        for Lw, name in zip(P["layers"], self.gnn.names):
            if name not in ("self", "cross"):
                raise ValueError(name)
            is_self = name == "self"
            ops.gemm(xm, Lw["wqkv"], M=T, N=768, K=256, lda=512, block_n=256, epilogue=ops.EPI_BIAS_BF16,
                     bias=Lw["bqkv"], out16=qkv, ld16=768)
            ops.attention(qkv, qkv, qkv, seqs_self if is_self else seqs_cross, num_seqs=2 * B, max_q_len=max(M, N),
                          heads=4, ldq=768, ldk=768, ldv=768, q_col0=0, k_col0=256, v_col0=512, scale=0.125, out=xm,
                          ld_out=512, out_col0=256, attn_work=4.0 * 256 * B * ((M * M + N * N) if is_self else 2 * M * N))
            ops.gemm(xm, Lw["w0"], M=T, N=512, K=512, lda=512, block_n=256, epilogue=ops.EPI_BIAS_BF16, bias=Lw["b0"],
                     relu=1, out16=hbuf, ld16=512)
            ops.gemm(hbuf, Lw["w3"], M=T, N=256, K=512, lda=512, block_n=256, epilogue=ops.EPI_RESID, bias=Lw["b3"],
                     resid=x32, out32=x32, ld32=256, out16=xm, ld16=512)

        # final projection in fp32, two-term bf16 split, cost = <m0, m1> / sqrt(256) (superglue.py:282-285)
        md32 = torch.empty(T, 256, dtype=torch.float32, device=dev)
        ops.gemm(xm, P["f_w"], M=T, N=256, K=256, lda=512, block_n=256, epilogue=ops.EPI_F32, bias=P["f_b"], alpha=1.0,
                 out32=md32, ld32=256, out32_batch_stride=0)
        md = torch.empty(T, 768, dtype=bf, device=dev)
        ops.sg_split_hilo(md32, T, off1, md)
        cost = torch.empty(B, M, N, dtype=torch.float32, device=dev)
        ops.gemm(md, md, M=M, N=N, K=768, lda=768, ldb=768, a_rows=T, b_row0=off1, b_rows=B * N, batch=B,
                 a_batch_rows=M, b_batch_rows=N, block_n=256 if N > 128 else 128, epilogue=ops.EPI_F32,
                 alpha=1.0 / 16.0, out32=cost, ld32=N, out32_batch_stride=M * N)
        scores, m0, m1, ms0, ms1 = ops.sg_sinkhorn_assignment(cost, P["bin"], self.conf.num_sinkhorn_iterations,
                                                              self.conf.filter_threshold)
        return {
            "sinkhorn_cost": cost,
            "log_assignment": scores,
            "matches0": m0,
            "matches1": m1,
            "matching_scores0": ms0,
            "matching_scores1": ms1,
        }

    def loss(self, pred, data):
        raise NotImplementedError("training is outside the forward-only scope of the B200 path")

    def metrics(self, pred, data):
        raise NotImplementedError
