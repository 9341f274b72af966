"""Latency of one forward of the pipeline at small batches, eager (one launch per kernel through ctypes) against a
CUDA-graph replay (glue_factory_b200/graph.py): 1024 x 1024 synthetic pairs, 2048 keypoints.
    python scripts/bench_latency.py [batches ...]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from glue_factory_b200.graph import GraphedForward  # noqa: E402

dev = torch.device("cuda:0")
pipe = bench.build_pipe("sweep", dev)  # the headline configuration: SuperPoint-open + LightGlue, 2048 keypoints
batches = [int(a) for a in sys.argv[1:]] or [1, 2, 4, 8]
gen = torch.Generator(device=dev).manual_seed(0)
for B in batches:
    data = {"view0": {"image": torch.rand(B, 1, 1024, 1024, device=dev, generator=gen)},
            "view1": {"image": torch.rand(B, 1, 1024, 1024, device=dev, generator=gen)}}

    def timed(fn, reps=20):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        host_enqueue = (time.perf_counter() - t0) / reps * 1e3
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps, host_enqueue

    with torch.no_grad():
        eager_ms, eager_host = timed(lambda: pipe(data))
        g = GraphedForward(pipe, data)
        graph_ms, graph_host = timed(lambda: g(data))
    print(f"batch {B}: eager {eager_ms:.3f} ms per forward (host enqueue {eager_host:.3f} ms), "
          f"graph replay {graph_ms:.3f} ms (host {graph_host:.3f} ms) -> {B / graph_ms * 1e3:.0f} pairs/s")
