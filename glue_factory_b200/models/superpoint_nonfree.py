"""MagicLeap SuperPoint extractor (the `gluefactory_nonfree.superpoint` model) on the B200 kernels.

Drop-in for `gluefactory_nonfree.superpoint.SuperPoint` (reference: gluefactory_nonfree/superpoint.py:152-356,
studied for behaviour only): same `default_conf`, same state-dict keys (conv1a ... convDb), same
`forward(data)` contract.  Same VGG topology as SuperPoint-open without batch norm, so the dense path
is the shared `sp_dense_forward` with an identity affine in the conv epilogues; detection uses the NMS /
top-k kernels with the per-image bottom / right border, descriptors are sampled with the legacy
(align_corners=True) or the corrected grid.

There is no PyTorch fallback: CPU tensors raise.
"""
from pathlib import Path

import torch
from torch import nn

from gluefactory.models.base_model import BaseModel

from .. import ops
from .superpoint_open import _pack_conv, sp_dense_forward


class SuperPoint(BaseModel):
    default_conf = {
        "has_detector": True,
        "has_descriptor": True,
        "descriptor_dim": 256,
        "sparse_outputs": True,
        "dense_outputs": False,
        "nms_radius": 4,
        "refinement_radius": 0,
        "detection_threshold": 0.005,
        "max_num_keypoints": -1,
        "max_num_keypoints_val": None,
        "force_num_keypoints": False,
        "randomize_keypoints_training": False,
        "remove_borders": 4,
        "legacy_sampling": True,
    }
    required_data_keys = ["image"]
    checkpoint_url = "https://github.com/magicleap/SuperGluePretrainedNetwork/raw/master/models/weights/superpoint_v1.pth"

    def _init(self, conf):
        if conf.descriptor_dim != 256 or not (conf.has_detector and conf.has_descriptor):
            raise NotImplementedError("the B200 kernels are built for the full model with descriptor_dim 256")
        if conf.refinement_radius > 0:
            raise NotImplementedError("soft-argmax refinement is outside the B200 path")
        # checkpoint layer names -> (in, out, kernel): the VGG topology of the kernels (two 64-wide and two 128-wide
        # stages, 256-wide heads)
        layout = {"conv1a": (1, 64, 3), "conv1b": (64, 64, 3), "conv2a": (64, 64, 3), "conv2b": (64, 64, 3),
                  "conv3a": (64, 128, 3), "conv3b": (128, 128, 3), "conv4a": (128, 128, 3), "conv4b": (128, 128, 3),
                  "convPa": (128, 256, 3), "convPb": (256, 65, 1), "convDa": (128, 256, 3),
                  "convDb": (256, conf.descriptor_dim, 1)}
        for name, (ci, co, k) in layout.items():
            setattr(self, name, nn.Conv2d(ci, co, kernel_size=k, stride=1, padding=(k - 1) // 2))
        # the reference always fetches the MagicLeap checkpoint (superpoint.py:198-200); an optional local
        # `weights` path is accepted in addition so that the model can be built offline
        path = conf.get("weights", None) if "weights" in conf else None
        if path is not None and Path(str(path)).exists():
            state_dict = torch.load(str(path), map_location="cpu")
        else:
            state_dict = torch.hub.load_state_dict_from_url(str(self.checkpoint_url))
        self.load_state_dict(state_dict, strict=False)
        self._packed = None
        self._packed_key = None

    def _pack(self):
        key = tuple((p.data_ptr(), p._version) for p in self.parameters())
        if self._packed is not None and key == self._packed_key:
            return self._packed
        dev = self.conv1a.weight.device

        def ident(c):
            return torch.ones(c, device=dev), torch.zeros(c, device=dev)

        def f32(t):
            return t.detach().float().contiguous()

        P = {"c1a_w": f32(self.conv1a.weight).reshape(64, 9).contiguous(), "c1a_b": f32(self.conv1a.bias)}
        P["c1a_s"], P["c1a_t"] = ident(64)
        convs = []
        for name, pool in (("conv1b", 1), ("conv2a", 0), ("conv2b", 1), ("conv3a", 0), ("conv3b", 1), ("conv4a", 0),
                           ("conv4b", 0)):
            conv = getattr(self, name)
            s, t = ident(conv.out_channels)
            convs.append(dict(w=_pack_conv(conv), b=f32(conv.bias), s=s, t=t, cin=conv.in_channels,
                              cout=conv.out_channels, pool=pool))
        P["convs"] = convs
        P["head_w"] = torch.cat([_pack_conv(self.convPa), _pack_conv(self.convDa)], 0).contiguous()
        P["head_b"] = torch.cat([f32(self.convPa.bias), f32(self.convDa.bias)]).contiguous()
        P["head_s"], P["head_t"] = ident(512)
        P["det_w"] = self.convPb.weight.detach().reshape(65, 256).to(torch.bfloat16).contiguous()
        P["det_b"] = f32(self.convPb.bias)
        P["det_s"], P["det_t"] = ident(65)
        P["desc_w"] = self.convDb.weight.detach().reshape(256, 256).to(torch.bfloat16).contiguous()
        P["desc_b"] = f32(self.convDb.bias)
        P["desc_s"], P["desc_t"] = ident(256)
        self._packed, self._packed_key = P, key
        return P

    def dense_forward(self, image):
        return sp_dense_forward(self._pack(), image)

    @ops.on_device("image")
    def _forward(self, data):
        image = data["image"]
        c = self.conf
        heat, dense = self.dense_forward(image)
        pred = {}
        if not c.sparse_outputs:
            return {"keypoint_scores": heat, "descriptors": dense.permute(0, 3, 1, 2)}
        B, H, W = heat.shape
        dev = heat.device
        max_kps = c.max_num_keypoints
        if not self.training and c.max_num_keypoints_val is not None:
            max_kps = c.max_num_keypoints_val
        if c.randomize_keypoints_training and self.training:
            raise NotImplementedError("training-time keypoint sampling is outside the forward-only B200 path")
        image_hw = None
        if c.remove_borders and "image_size" in data:
            # bottom / right border from the per-image (w, h), truncated to int as the reference does
            image_hw = data["image_size"].to(device=dev).flip(-1).to(torch.int32).contiguous()
        pad = c.remove_borders or 0
        fast = 1 <= c.nms_radius <= 4
        if not fast and image_hw is not None:
            raise NotImplementedError("per-image borders need nms_radius in [1, 4] on the B200 path")
        if max_kps is not None and max_kps > 0 and fast:
            k = int(max_kps)
            _, kpts, scores, _, counts = ops.sp_detect_fast(heat, c.nms_radius, pad, c.detection_threshold, k, image_hw)
        else:
            if fast:
                nms, _, _, _, _ = ops.sp_detect_fast(heat, c.nms_radius, pad, c.detection_threshold, 1, image_hw)
            else:
                nms = ops.sp_nms_border(heat, c.nms_radius, pad)
            k = int(ops.sp_count_candidates(nms, c.detection_threshold).max().item())  # host sync, as torch.where is
            if max_kps is not None and max_kps > 0:
                k = min(k, int(max_kps))
            k = max(k, 1)
            kpts, scores, _, counts = ops.sp_select_topk(nms, c.detection_threshold, k)
        interior = (H - 2 * pad) * (W - 2 * pad)
        guaranteed = (max_kps is not None and max_kps > 0 and c.detection_threshold < 0 and interior >= k
                      and image_hw is None)
        if not guaranteed:
            cnt = counts.tolist()  # data-dependent keypoint counts: the reference synchronises here too
            if c.force_num_keypoints:
                from gluefactory.models.utils.misc import pad_and_stack

                kp_list = [kpts[i, : cnt[i]] - 0.5 for i in range(B)]
                sc_list = [scores[i, : cnt[i]] for i in range(B)]
                bound = data.get("image_size", torch.tensor(image.shape[-2:])).min().item()
                kpts = (pad_and_stack(kp_list, max_kps, -2, mode="random_c", bounds=(0, bound)) + 0.5).contiguous()
                scores = pad_and_stack(sc_list, max_kps, -1, mode="zeros").contiguous()
            elif len(set(cnt)) != 1:
                raise RuntimeError(f"stack expects each tensor to be equal size, but got keypoint counts {cnt}")
            elif cnt[0] < k:
                kpts, scores = kpts[:, : cnt[0]].contiguous(), scores[:, : cnt[0]].contiguous()
        desc = ops.sp_sample_descriptors(dense, kpts, 8, mode=2 if c.legacy_sampling else 1)
        pred = {"keypoints": kpts, "keypoint_scores": scores, "descriptors": desc}
        if c.dense_outputs:
            pred["dense_descriptors"] = dense.permute(0, 3, 1, 2)
        return pred

    def loss(self, pred, data):
        raise NotImplementedError

    def metrics(self, pred, data):
        raise NotImplementedError
