"""NMS + top-k fast path alone at the bench shape (64 heat maps of 1024x1024, radius 3, 2048 keypoints), with a
bit-exactness check of the NMS map against the max-pool formulation (torch ops on the GPU, checker only).
Run on the GPU box: python scripts/bench_nms.py [reps]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from glue_factory_b200 import ops  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
dev = torch.device("cuda:0")
torch.manual_seed(0)
B, H, W, R, K = 64, 1024, 1024, 3, 2048
# peaked heat map like a 65-way softmax output: mostly tiny values, plateaus of exact ties included
scores = torch.rand(B, H, W, device=dev) ** 8
scores[:, ::37, ::41] = 0.5


def simple_nms(s, radius):  # superpoint_open.py:34-49
    def mp(x):
        return torch.nn.functional.max_pool2d(x, kernel_size=radius * 2 + 1, stride=1, padding=radius)

    zeros = torch.zeros_like(s)
    max_mask = s == mp(s)
    for _ in range(2):
        supp_mask = mp(max_mask.float()) > 0
        supp_scores = torch.where(supp_mask, zeros, s)
        new_max_mask = supp_scores == mp(supp_scores)
        max_mask = max_mask | (new_max_mask & (~supp_mask))
    return torch.where(max_mask, s, zeros)


nms, kpts, ksc, idx, counts = ops.sp_detect_fast(scores, R, 4, -1.0, K)
ref = simple_nms(scores[:4, None], R)[:, 0]
ref[:, :4] = -1
ref[:, -4:] = -1
ref[:, :, :4] = -1
ref[:, :, -4:] = -1
print("nms map bit-exact:", bool(torch.equal(nms[:4], ref)))
flat = ref.reshape(4, -1)
top = torch.topk(flat, K, dim=1)
print("top-k scores equal:", bool(torch.equal(top.values, ksc[:4])))
for _ in range(3):
    ops.sp_detect_fast(scores, R, 4, -1.0, K)
torch.cuda.synchronize()
ops.reset_stats(profile=True)
for _ in range(reps):
    ops.sp_detect_fast(scores, R, 4, -1.0, K)
torch.cuda.synchronize()
per = {}
for tag, work, a, b in ops.profile_records():
    per[tag] = per.get(tag, 0.0) + a.elapsed_time(b)
for tag, ms in per.items():
    print(f"{tag:16s} {ms / reps:.3f} ms per call ({B} maps)")
