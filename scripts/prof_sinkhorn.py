"""Sinkhorn kernel alone (for ncu): python scripts/prof_sinkhorn.py [pairs] [iters]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from glue_factory_b200 import ops  # noqa: E402

pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 32
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 100
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
cost = (torch.randn(pairs, 2048, 2048, generator=g) * 5).to(dev)
alpha = torch.ones(1, device=dev)
for _ in range(2):
    out = ops.sg_sinkhorn_assignment(cost, alpha, iters, 0.2)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
out = ops.sg_sinkhorn_assignment(cost, alpha, iters, 0.2)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
print(f"pairs {pairs} iters {iters}: {ms:.3f} ms, {4.0 * pairs * 2049 * 2049 * iters / ms / 1e6:.1f} GB/s algorithmic")
