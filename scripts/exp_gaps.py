"""Experiment: how much of the step is launch gaps?  Times the resident step (a) plain, (b) with the per-launch
profiling events of bench.py, (c) replayed from a CUDA graph.  Run on the GPU box: python scripts/exp_gaps.py"""
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


def step():
    with torch.no_grad():
        return pipe(data)


def timed(fn, n=10):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for _ in range(5):
    step()
ops.reset_stats(profile=False)
print("plain        ms/step", timed(step))
ops.reset_stats(profile=True)
print("with events  ms/step", timed(step, 5))
ops.reset_stats(profile=False)
print("plain again  ms/step", timed(step))
if "--graph" in sys.argv:
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        step()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            out = step()
    torch.cuda.synchronize()
    for _ in range(3):
        g.replay()
    print("graph replay ms/step", timed(g.replay))
