"""Runs the rows either side of the path once at evaluation sizes (for an ncu capture of their kernels):
ImagePreprocessor 1200x1600 RGB uint8 -> 1024 long side (+ square pad), export packing of a 32-pair batch at 2048
keypoints, match metrics + DLT on the same batch.  python scripts/prof_io.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from glue_factory_b200.export import RecordPacker  # noqa: E402
from gluefactory.eval.utils import eval_homography_dlt, eval_matches_epipolar, eval_matches_homography  # noqa: E402
from gluefactory.utils.image import ImagePreprocessor  # noqa: E402

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
u8 = torch.randint(0, 256, (8, 1200, 1600, 3), generator=g, dtype=torch.uint8).to(dev)
pre = ImagePreprocessor(dict(resize=1024, side="long", square_pad=True))
B, K = 32, 2048
kp0, kp1 = torch.rand(B, K, 2, device=dev) * 1024, torch.rand(B, K, 2, device=dev) * 1024
m0 = torch.randint(-1, K, (B, K), device=dev)
pred = {"keypoints0": kp0, "keypoints1": kp1, "keypoint_scores0": torch.rand(B, K, device=dev),
        "keypoint_scores1": torch.rand(B, K, device=dev), "matches0": m0, "matches1": m0.clone(),
        "matching_scores0": torch.rand(B, K, device=dev), "matching_scores1": torch.rand(B, K, device=dev)}
sc = torch.rand(B, 2, device=dev) + 0.5
data = {"view0": {"scales": sc, "image_size": torch.tensor([[1024.0, 1024.0]] * B, device=dev),
                  "camera": torch.tensor([[1024.0, 1024.0, 800.0, 800.0, 512.0, 512.0]] * B, device=dev)},
        "view1": {"scales": sc, "camera": torch.tensor([[1024.0, 1024.0, 800.0, 800.0, 512.0, 512.0]] * B, device=dev)},
        "H_0to1": torch.eye(3, device=dev).repeat(B, 1, 1), "T_0to1": torch.tensor([[1.0, 0, 0, 0, 1, 0, 0, 0, 1, 0.1, 0, 0]] * B, device=dev)}
packer = RecordPacker()
for _ in range(3):
    out = pre(u8)
    rec = packer.pack(pred, data)
    eval_matches_homography(data, pred)
    eval_matches_epipolar(data, pred)
    eval_homography_dlt(data, pred)
torch.cuda.synchronize()
print("ok", out["image"].shape, rec.shape)
