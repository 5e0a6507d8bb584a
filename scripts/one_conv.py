"""Runs one convolution layer through the tcgen05 kernels a few times (ncu capture target).
usage: one_conv.py cin h cout k stride [batch]"""
import os
import sys
import torch
sys.path.insert(0, ".")
from azure_hc_intel_tf_b200.ops import conv_umma
cin, h, cout, k, stride = map(int, sys.argv[1:6])
n = int(sys.argv[6]) if len(sys.argv) > 6 else 256
WS = os.environ.get("NOSTATS") != "1"
class M: pass
m = M(); m.cin, m.cout, m.k, m.stride = cin, cout, (k, k), (stride, stride)
pd = (k - 1) // 2
m.pad = ((pd, pd), (pd, pd)); m.out_hw = ((h + 2 * pd - k) // stride + 1,) * 2
x = torch.randn(n, h, h, cin, device="cuda").bfloat16().permute(0, 3, 1, 2)
w = torch.randn(cout, k, k, cin, device="cuda").bfloat16().permute(0, 3, 1, 2)
dy = torch.randn(n, m.out_hw[0], m.out_hw[1], cout, device="cuda").bfloat16().permute(0, 3, 1, 2)
ONLY = os.environ.get("ONLY", "")          # fprop | dgrad | wgrad: run just that kernel (ncu capture target)
for _ in range(3):
    if ONLY in ("", "fprop"):
        y, st = conv_umma.fprop(x, w, m, want_stats=WS)
    if ONLY in ("", "dgrad") and conv_umma.supports_dgrad(dy, w, m, x.shape):
        dx = conv_umma.dgrad(dy, w, m, x.shape)
    if ONLY in ("", "wgrad"):
        dw = conv_umma.wgrad(x, dy, m)
torch.cuda.synchronize()
print("ok")
if len(sys.argv) > 7 and sys.argv[7] == "prof":
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as pr:
        for _ in range(5):
            y, st = conv_umma.fprop(x, w, m, want_stats=WS)
            if conv_umma.supports_dgrad(dy, w, m, x.shape):
                dx = conv_umma.dgrad(dy, w, m, x.shape)
            dw = conv_umma.wgrad(x, dy, m)
        torch.cuda.synchronize()
    for e in pr.key_averages():
        if e.device_time_total > 0:
            print(f"{e.key[:70]:70s} n={e.count:3d} avg={e.device_time_total / e.count:8.1f} us")
