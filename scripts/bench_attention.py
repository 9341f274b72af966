"""Attention kernel alone at the bench shape (64 sequences x 2048 tokens x 4 heads x 64):
    python scripts/bench_attention.py [seqs] [tokens] [reps]
Prints ms per launch and the clocks per (CTA, key tile) it implies at 148 SMs."""
import sys

import os

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from glue_factory_b200 import ops

nseq = int(sys.argv[1]) if len(sys.argv) > 1 else 64
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 20
dev = torch.device("cuda:0")
torch.manual_seed(0)
qkv = (torch.randn(nseq * n, 768, device=dev) * 0.5).to(torch.bfloat16)
out = torch.empty(nseq * n, 256, device=dev, dtype=torch.bfloat16)
seqs = ops.make_attn_seqs([(i * n, n, i * n, n) for i in range(nseq)], dev)


def run():
    ops.attention(qkv, qkv, qkv, seqs, num_seqs=nseq, max_q_len=n, heads=4, ldq=768, ldk=768, ldv=768, q_col0=0,
                  k_col0=256, v_col0=512, scale=0.125, out=out, ld_out=256)


for _ in range(3):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
ctas = nseq * 4 * ((n + 255) // 256)
waves = -(-ctas // 148)
print(f"attention {ms:.3f} ms/launch  {4.0 * nseq * 4 * n * n * 64 / ms / 1e9:.1f} TFLOP/s  "
      f"~{ms * 1e-3 * 1.9e9 / waves / ((n + 127) // 128):.0f} clk per (CTA, key tile) at 1.9 GHz")

import ctypes

from glue_factory_b200 import _lib

h = _lib.lib()
if hasattr(h, "b200_debug_attention_trace"):
    import numpy as np

    buf = np.zeros((4, 8, 8), dtype=np.int64)
    h.b200_debug_attention_trace(ctypes.c_void_p(buf.ctypes.data))
    t0 = buf[0, 0, 0]
    print("per stream, items 3..10 of CTA 5: [item start, first S seen, last S seen, loop done, pair met, O ready, merge done]")
    for st in range(4):
        for j in range(8):
            print(st, j + 3, (buf[st, j, :7] - t0).tolist())
