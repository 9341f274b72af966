"""Timing of the pre-processing kernels at evaluation size (8 RGB uint8 images 1200x1600 -> 1024 long side, square pad)
and of the assignment stage at the bench shape.  python scripts/bench_io.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from glue_factory_b200 import ops  # noqa: E402
from gluefactory.utils.image import ImagePreprocessor  # noqa: E402

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
u8 = torch.randint(0, 256, (8, 1200, 1600, 3), generator=g, dtype=torch.uint8).to(dev)
pre = ImagePreprocessor(dict(resize=1024, side="long", square_pad=True))
sim = torch.randn(32, 2048, 2048, device=dev)
z = torch.randn(32, 2048, device=dev)
for _ in range(3):
    pre(u8)
    ops.lg_assignment(sim, z, z, 0.1)
torch.cuda.synchronize()
ops.reset_stats(profile=True)
for _ in range(10):
    pre(u8)
    ops.lg_assignment(sim, z, z, 0.1)
torch.cuda.synchronize()
per = {}
for tag, work, a, b in ops.profile_records():
    t = per.setdefault(tag, [0.0, 0.0])
    t[0] += a.elapsed_time(b) / 10
    t[1] = work
for tag, (ms, work) in per.items():
    print(f"{tag:16s} {ms:.3f} ms  {work / ms / 1e6:.0f} GB/s algorithmic")
