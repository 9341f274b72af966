"""Per-launch attention times inside the benchmark step (self / cross, layer by layer).  GPU box only."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from glue_factory_b200 import ops  # noqa: E402
from gluefactory.models import get_model  # noqa: E402

dev = torch.device("cuda:0")
sp_path, lg_path = bench.weight_files()
pipe = get_model("two_view_pipeline")({
    "extractor": {**bench.EXTRACTOR_CONF, "weights": sp_path},
    "matcher": {**bench.MATCHER_CONF, "weights": lg_path},
}).to(dev).eval()
im0, im1 = bench.make_pairs(0, bench.PAIRS)
size = torch.tensor([[1024.0, 1024.0]] * bench.PAIRS).to(dev)
data = {"view0": {"image": im0.to(dev), "image_size": size}, "view1": {"image": im1.to(dev), "image_size": size}}
with torch.no_grad():
    for _ in range(3):
        pipe(data)
    torch.cuda.synchronize()
    ops.reset_stats(profile=True, only=["attention"])
    for _ in range(3):
        pipe(data)
    torch.cuda.synchronize()
rec = ops.profile_records()
n = len(rec) // 3
for i in range(n):
    ms = sum(rec[k * n + i][2].elapsed_time(rec[k * n + i][3]) for k in range(3)) / 3
    print(f"layer {i // 2} {'self ' if i % 2 == 0 else 'cross'} {ms:.3f} ms")
