"""Time the max-pool kernels on the ResNet-50 stem shape (256 x 112 x 112 x 64 -> 56 x 56, 3x3/2 SAME), CUDA events, median of 20
(the 411 MB input is larger than L2). usage: python scripts/pool_bench.py"""
import sys
import torch
sys.path.insert(0, ".")
from azure_hc_intel_tf_b200.models.builder import BuildConfig, ConvNetBuilder
from azure_hc_intel_tf_b200 import ops

cfg = BuildConfig(kernels="b200")
b = ConvNetBuilder(cfg, 64, (112, 112))
pool = b.mpool(3, 2, "SAME")
x = torch.randn(256, 112, 112, 64, device="cuda").to(torch.bfloat16).permute(0, 3, 1, 2).requires_grad_(True)
y = pool(x)
dy = torch.randn_like(y)


def med(fn, n=20):
    for _ in range(3):
        fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for a, c in ev:
        a.record(); fn(); c.record()
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(c) for a, c in ev)
    return t[n // 2] * 1e3


with torch.no_grad():
    f = med(lambda: pool(x))
bw = med(lambda: torch.autograd.grad(y, x, dy, retain_graph=True))
print(f"maxpool 3x3/2 256x112x112x64: fwd {f:.1f} us (roofline {(411 + 103 + 51) / 6.454:.0f}), bwd {bw:.1f} us (roofline {(103 + 51 + 411) / 6.454:.0f})")
