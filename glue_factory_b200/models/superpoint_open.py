"""SuperPoint-open extractor on the B200 kernels, behind the Glue Factory BaseModel API.

Drop-in for `gluefactory.models.extractors.superpoint_open.SuperPoint`
(reference: gluefactory/models/extractors/superpoint_open.py:78-220): same `default_conf`,
same state-dict schema (84 tensors), same `forward(data)` contract.  The nn.Conv2d /
nn.BatchNorm2d children exist only to own the parameters; `_forward` launches
libb200glue.so kernels:

  conv1a (CUDA cores) -> 7 implicit-GEMM convs + fused heads on tcgen05 (ReLU, BN affine and 2x2
  max-pool fused in the epilogues; detector 1x1 + softmax + depth-to-space and descriptor 1x1 +
  L2-norm fused as well) -> NMS + border kernel -> per-image radix-select top-k -> bilinear
  descriptor sampling.

There is no PyTorch fallback: CPU tensors raise.
"""
import os
from collections import OrderedDict
from pathlib import Path
from types import SimpleNamespace

import torch
from torch import nn

from gluefactory.models.base_model import BaseModel

from .. import ops


class VGGBlock(nn.Sequential):
    """Parameter container with the reference's child names (conv, activation, bn)."""

    def __init__(self, c_in, c_out, kernel_size, relu=True):
        super().__init__(
            OrderedDict(
                conv=nn.Conv2d(c_in, c_out, kernel_size, stride=1, padding=(kernel_size - 1) // 2),
                activation=nn.ReLU(inplace=True) if relu else nn.Identity(),
                bn=nn.BatchNorm2d(c_out, eps=0.001),
            )
        )


def _bn_affine(bn):
    scale = bn.weight.float() / torch.sqrt(bn.running_var.float() + bn.eps)
    return scale.contiguous(), (bn.bias.float() - bn.running_mean.float() * scale).contiguous()


def _pack_conv(conv):
    w = conv.weight.detach()
    return w.permute(0, 2, 3, 1).reshape(w.shape[0], -1).to(torch.bfloat16).contiguous()


# B200GLUE_FUSED_STEM=0 keeps conv1a / conv1b as two kernels (A/B comparison of the same CUDA path)
FUSED_STEM = os.environ.get("B200GLUE_FUSED_STEM", "1") != "0"


def sp_dense_forward(P, image):
    """Shared dense path of both SuperPoint variants on packed weights `P` (see SuperPoint._pack):
    image (B,1|3,H,W) fp32 -> heat map (B,8h,8w) fp32, dense descriptors NHWC (B,h,w,256) fp32."""
    if not image.is_cuda:
        raise RuntimeError("glue_factory_b200 SuperPoint runs on CUDA only (no CPU fallback)")
    # uint8 pixels (as read from disk) are converted in the first kernel's load: value / 255, bit-identical to the
    # reference's host conversion (utils/image.py:112-121); anything else follows the float [0, 1] contract
    image = image.contiguous() if image.dtype == torch.uint8 else image.float().contiguous()
    B, _, H, W = image.shape
    dev = image.device
    L0 = P["convs"][0]
    if FUSED_STEM and L0["cin"] == 64 and L0["cout"] == 64 and (not L0["pool"] or (H % 2 == 0 and W % 2 == 0)):
        # first VGG block in one kernel: the full-resolution 64-channel activation stays on chip
        h, w = (H // 2, W // 2) if L0["pool"] else (H, W)
        x = torch.empty(B, h, w, 64, dtype=torch.bfloat16, device=dev)
        if "stem_par" not in P:  # per-channel operands of both convolutions, on the host (kernel parameters)
            P["stem_par"] = torch.stack([P["c1a_b"], P["c1a_s"], P["c1a_t"], L0["b"], L0["s"], L0["t"]]).float().cpu()
        ops.sp_conv_stem(image, P["c1a_w"], P["stem_par"], L0["w"], L0["pool"], x)
        rest = P["convs"][1:]
    else:
        x = torch.empty(B, H, W, 64, dtype=torch.bfloat16, device=dev)
        ops.sp_conv1a(image, P["c1a_w"], P["c1a_b"], P["c1a_s"], P["c1a_t"], x)
        h, w = H, W
        rest = P["convs"]
    for L in rest:
        oh, ow = (h // 2, w // 2) if L["pool"] else (h, w)
        y = torch.empty(B, oh, ow, L["cout"], dtype=torch.bfloat16, device=dev)
        ops.gemm(x, L["w"], M=0, N=L["cout"], K=9 * L["cin"], lda=L["cin"], block_n=min(L["cout"], 256),
                 epilogue=ops.EPI_CONV_RELU_BN, batch=B, conv=(h, w, L["cin"]), bias=L["b"], scale=L["s"],
                 shift=L["t"], relu=1, pool=L["pool"], out16=y, ld16=L["cout"])
        x, h, w = y, oh, ow
    heads = torch.empty(B, h, w, 512, dtype=torch.bfloat16, device=dev)
    ops.gemm(x, P["head_w"], M=0, N=512, K=9 * 128, lda=128, block_n=256, epilogue=ops.EPI_CONV_RELU_BN,
             batch=B, conv=(h, w, 128), bias=P["head_b"], scale=P["head_s"], shift=P["head_t"], relu=1, pool=0,
             out16=heads, ld16=512)
    cells = B * h * w
    heat = torch.empty(B, h * 8, w * 8, dtype=torch.float32, device=dev)
    ops.gemm(heads, P["det_w"], M=cells, N=65, K=256, lda=512, a_col0=0, block_n=80,
             epilogue=ops.EPI_CONV_BN_SOFTMAX_D2S, bias=P["det_b"], scale=P["det_s"], shift=P["det_t"],
             out32=heat, cells=(h, w))
    dense = torch.empty(B, h, w, 256, dtype=torch.float32, device=dev)
    ops.gemm(heads, P["desc_w"], M=cells, N=256, K=256, lda=512, a_col0=256, block_n=256,
             epilogue=ops.EPI_CONV_BN_L2, bias=P["desc_b"], scale=P["desc_s"], shift=P["desc_t"],
             out32=dense, ld32=256)
    return heat, dense


class SuperPoint(BaseModel):
    default_conf = {
        "descriptor_dim": 256,
        "nms_radius": 4,
        "max_num_keypoints": None,
        "force_num_keypoints": False,
        "detection_threshold": 0.005,
        "remove_borders": 4,
        "channels": [64, 64, 128, 128, 256],
        "dense_outputs": None,
        "weights": None,
    }
    checkpoint_url = "https://github.com/rpautrat/SuperPoint/raw/master/weights/superpoint_v6_from_tf.pth"

    def _init(self, conf):
        self.conf = SimpleNamespace(**conf)
        if list(self.conf.channels) != [64, 64, 128, 128, 256] or self.conf.descriptor_dim != 256:
            raise NotImplementedError("the B200 kernels are built for channels [64,64,128,128,256], descriptor_dim 256")
        self.stride = 2 ** (len(self.conf.channels) - 2)
        chans = [1, *self.conf.channels[:-1]]
        stages = []
        for i, c in enumerate(chans[1:], 1):
            layers = [VGGBlock(chans[i - 1], c, 3), VGGBlock(c, c, 3)]
            if i < len(chans) - 1:
                layers.append(nn.MaxPool2d(kernel_size=2, stride=2))
            stages.append(nn.Sequential(*layers))
        self.backbone = nn.Sequential(*stages)
        c = self.conf.channels[-1]
        self.detector = nn.Sequential(VGGBlock(chans[-1], c, 3), VGGBlock(c, self.stride**2 + 1, 1, relu=False))
        self.descriptor = nn.Sequential(VGGBlock(chans[-1], c, 3), VGGBlock(c, self.conf.descriptor_dim, 1, relu=False))
        if conf.weights is not None and Path(conf.weights).exists():
            state_dict = torch.load(conf.weights, map_location="cpu")
        else:
            state_dict = torch.hub.load_state_dict_from_url(self.checkpoint_url)
        self.load_state_dict(state_dict)
        self._packed = None
        self._packed_key = None

    # -- weight repacking (lazy, redone when parameters move or change) ---------------------------
    def _pack_key(self):
        return tuple((p.data_ptr(), p._version) for p in list(self.parameters()) + list(self.buffers()))

    def _pack(self):
        key = self._pack_key()
        if self._packed is not None and key == self._packed_key:
            return self._packed
        P = {}
        b0 = self.backbone[0][0]
        P["c1a_w"] = b0.conv.weight.detach().reshape(64, 9).float().contiguous()
        P["c1a_b"] = b0.conv.bias.detach().float().contiguous()
        P["c1a_s"], P["c1a_t"] = _bn_affine(b0.bn)
        convs = []
        for si in range(4):
            for bi in range(2):
                if si == 0 and bi == 0:
                    continue
                blk = self.backbone[si][bi]
                s, t = _bn_affine(blk.bn)
                convs.append(dict(w=_pack_conv(blk.conv), b=blk.conv.bias.detach().float().contiguous(), s=s, t=t,
                                  cin=blk.conv.in_channels, cout=blk.conv.out_channels, pool=int(bi == 1 and si < 3)))
        P["convs"] = convs
        det0, desc0 = self.detector[0], self.descriptor[0]
        s0, t0 = _bn_affine(det0.bn)
        s1, t1 = _bn_affine(desc0.bn)
        P["head_w"] = torch.cat([_pack_conv(det0.conv), _pack_conv(desc0.conv)], 0).contiguous()
        P["head_b"] = torch.cat([det0.conv.bias.detach().float(), desc0.conv.bias.detach().float()]).contiguous()
        P["head_s"], P["head_t"] = torch.cat([s0, s1]).contiguous(), torch.cat([t0, t1]).contiguous()
        det1, desc1 = self.detector[1], self.descriptor[1]
        P["det_w"] = det1.conv.weight.detach().reshape(65, 256).to(torch.bfloat16).contiguous()
        P["det_b"] = det1.conv.bias.detach().float().contiguous()
        P["det_s"], P["det_t"] = _bn_affine(det1.bn)
        P["desc_w"] = desc1.conv.weight.detach().reshape(256, 256).to(torch.bfloat16).contiguous()
        P["desc_b"] = desc1.conv.bias.detach().float().contiguous()
        P["desc_s"], P["desc_t"] = _bn_affine(desc1.bn)
        self._packed, self._packed_key = P, key
        return P

    # -- stages (also used one by one by the teacher-forced tests) ---------------------------------
    def dense_forward(self, image):
        """image (B,1|3,H,W) fp32 -> heat map (B,8h,8w) fp32, dense descriptors NHWC (B,h,w,256) fp32."""
        return sp_dense_forward(self._pack(), image)

    def detect(self, heat):
        """heat map -> (keypoints (B,K,2) incl. +0.5, scores (B,K), counts (B,) int32 on device)."""
        c = self.conf
        k = c.max_num_keypoints
        if k is not None and 1 <= c.nms_radius <= 4:
            _, kpts, scores, _, counts = ops.sp_detect_fast(heat, c.nms_radius, c.remove_borders,
                                                            c.detection_threshold, k)
            return kpts, scores, counts
        nms = ops.sp_nms_border(heat, c.nms_radius, c.remove_borders)
        if k is None:
            k = int(ops.sp_count_candidates(nms, c.detection_threshold).max().item())  # host sync, as in the reference
            k = max(k, 1)
        kpts, scores, _, counts = ops.sp_select_topk(nms, c.detection_threshold, k)
        return kpts, scores, counts

    @ops.on_device("image")
    def _forward(self, data):
        image = data["image"]
        heat, dense = self.dense_forward(image)
        B = image.shape[0]
        kpts, scores, counts = self.detect(heat)
        k = kpts.shape[1]
        interior = (heat.shape[1] - 2 * (self.conf.remove_borders or 0)) * (heat.shape[2] - 2 * (self.conf.remove_borders or 0))
        guaranteed = self.conf.max_num_keypoints is not None and self.conf.detection_threshold < 0 and interior >= k
        if not guaranteed:
            # data-dependent keypoint counts: the reference synchronises here too (torch.where)
            cnt = counts.tolist()
            if self.conf.force_num_keypoints:
                kpts, scores = self._pad(kpts, scores, cnt, data, image)
            elif len(set(cnt)) != 1:
                raise RuntimeError(f"stack expects each tensor to be equal size, but got keypoint counts {cnt}")
            elif cnt[0] < k:
                kpts, scores = kpts[:, : cnt[0]].contiguous(), scores[:, : cnt[0]].contiguous()
        desc = ops.sp_sample_descriptors(dense, kpts, self.stride)
        pred = {"keypoints": kpts, "keypoint_scores": scores, "descriptors": desc, "heatmap": heat}
        if self.conf.dense_outputs:
            pred["dense_descriptors"] = dense.permute(0, 3, 1, 2)
        return pred

    def _pad(self, kpts, scores, cnt, data, image):
        """force_num_keypoints: random in-extent padding drawn from the torch RNG, scores 0
        (reference: superpoint_open.py:181-194, models/utils/misc.py:19-70); host-side torch by design."""
        from gluefactory.models.utils.misc import pad_and_stack

        kp_list = [kpts[i, : cnt[i]] - 0.5 for i in range(len(cnt))]
        sc_list = [scores[i, : cnt[i]] for i in range(len(cnt))]
        bound = data.get("image_size", torch.tensor(image.shape[-2:])).min().item()
        kp = pad_and_stack(kp_list, self.conf.max_num_keypoints, -2, mode="random_c", bounds=(0, bound))
        sc = pad_and_stack(sc_list, self.conf.max_num_keypoints, -1, mode="zeros")
        return (kp + 0.5).contiguous(), sc.contiguous()

    def loss(self, pred, data):
        raise NotImplementedError
