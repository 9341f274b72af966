"""Stand-alone timing of the LightGlue GEMM shapes of the benchmark step (T = 32 pairs x 4096 tokens).
Run on the GPU box: python scripts/bench_gemm.py [iters]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from glue_factory_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
T = 32 * 4096
bf = torch.bfloat16
g = torch.Generator(device=dev).manual_seed(1)


def rnd(*shape, dtype=bf, s=1.0):
    return (torch.randn(*shape, device=dev, generator=g) * s).to(dtype)


xm, h, qkv, ctx, md = rnd(T, 512), rnd(T, 512), rnd(T, 768), rnd(T, 256), rnd(T, 256)
x32 = rnd(T, 256, dtype=torch.float32)
cs = rnd(T, 32, 2, dtype=torch.float32)
w = {n: rnd(*s, s=0.05) for n, s in dict(qkv=(768, 256), o=(256, 256), c=(512, 256), w1=(512, 512), w2=(256, 512)).items()}
b = {n: rnd(s, dtype=torch.float32) for n, s in dict(qkv=768, o=256, c=512, w1=512, w2=256).items()}
gam, bet = rnd(512, dtype=torch.float32), rnd(512, dtype=torch.float32)

cases = {
    "qkv_rotary n768 k256": (lambda: ops.gemm(xm, w["qkv"], M=T, N=768, K=256, lda=512, block_n=256, epilogue=ops.EPI_QKV_ROTARY,
                                              bias=b["qkv"], cossin=cs, rot_cols=512, out16=qkv, ld16=768), 2.0 * T * 768 * 256,
                             T * (512 + 1536 + 256)),
    "proj n256 k256": (lambda: ops.gemm(ctx, w["o"], M=T, N=256, K=256, lda=256, block_n=256, epilogue=ops.EPI_BIAS_BF16,
                                        bias=b["o"], out16=xm, out16_col0=256, ld16=512), 2.0 * T * 256 * 256, T * 1024),
    "cross n512 k256": (lambda: ops.gemm(xm, w["c"], M=T, N=512, K=256, lda=512, block_n=256, epilogue=ops.EPI_BIAS_BF16,
                                         bias=b["c"], out16=qkv, ld16=768), 2.0 * T * 512 * 256, T * 1536),
    "ffn1 ln_gelu n512 k512": (lambda: ops.gemm(xm, w["w1"], M=T, N=512, K=512, lda=512, block_n=512, epilogue=ops.EPI_LN_GELU,
                                                bias=b["w1"], ln_gamma=gam, ln_beta=bet, ln_eps=1e-5, out16=h, ld16=512),
                               2.0 * T * 512 * 512, T * 2048),
    "ffn2 resid n256 k512": (lambda: ops.gemm(h, w["w2"], M=T, N=256, K=512, lda=512, block_n=256, epilogue=ops.EPI_RESID,
                                              bias=b["w2"], resid=x32, out32=x32, ld32=256, out16=xm, ld16=512),
                             2.0 * T * 256 * 512, T * (1024 + 1024 + 1024 + 512)),
}
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 30
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for name, (fn, flops, bytes_) in cases.items():
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.zero_()  # the step's working set exceeds L2 between two uses of the same tensor
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    ms = tot / iters
    print(f"{name:26s} {1e3 * ms:7.1f} us  {flops / ms / 1e9:7.1f} TFLOP/s  {bytes_ / ms / 1e6:7.1f} GB/s (algorithmic)")
