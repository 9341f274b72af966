"""Fused first VGG block alone at the bench shape:  python scripts/bench_stem.py [images] [H] [W]"""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from glue_factory_b200 import _lib, ops  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
H = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
W = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
dev = torch.device("cuda:0")
torch.manual_seed(0)
img = torch.rand(B, 1, H, W, device=dev)
w1 = torch.randn(64, 9, device=dev) / 3
w2 = (torch.randn(64, 576, device=dev) / 24).to(torch.bfloat16)
v = [torch.randn(64, device=dev) * 0.1 for _ in range(6)]
hp = torch.stack([v[0], 1 + v[1], v[2], v[3], 1 + v[4], v[5]]).cpu()
out = torch.empty(B, H // 2, W // 2, 64, dtype=torch.bfloat16, device=dev)


def run():
    ops.sp_conv_stem(img, w1, hp, w2, 1, out)


for _ in range(2):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
tiles = B * ((H + 15) // 16) * ((W + 7) // 8)
print(f"stem {ms:.3f} ms/launch, {ms * 1e-3 * 1.9e9 / (tiles / 148):.0f} clk per tile per SM")
h = _lib.lib()
if hasattr(h, "b200_debug_stem_trace"):
    import numpy as np

    buf = np.zeros((3, 8, 8), dtype=np.int64)
    h.b200_debug_stem_trace(ctypes.c_void_p(buf.ctypes.data))
    t0 = buf[0, 0, 0]
    print("issuer [a_full(i+1) seen, stem issued, patch_full(i) seen, acc_empty seen, main issued] | "
          "worker [start, im2col done, stem result seen, patch slot free, patch written, next pixels requested] | epilogue [acc seen, stores done]")
    for j in range(8):
        print(j + 10, (buf[0, j, :5] - t0).tolist(), (buf[1, j, :8] - t0).tolist(), (buf[2, j, :2] - t0).tolist())
