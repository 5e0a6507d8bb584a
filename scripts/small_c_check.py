"""Numerics of the tcgen05 conv kernels on layers with fewer than 64 channels (run with B200_CONV_MIN_C=8)."""
import sys
import torch
import torch.nn.functional as F
sys.path.insert(0, ".")
from azure_hc_intel_tf_b200.ops import conv_umma

class M: pass
def run(n, h, cin, cout, k, stride, pd):
    m = M(); m.cin, m.cout, m.k, m.stride = cin, cout, (k, k), (stride, stride)
    m.pad = ((pd, pd), (pd, pd)); m.out_hw = ((h + 2 * pd - k) // stride + 1,) * 2
    g = torch.Generator().manual_seed(1)
    x = torch.randn(n, h, h, cin, generator=g).cuda().bfloat16().permute(0, 3, 1, 2)
    w = (torch.randn(cout, k, k, cin, generator=g) / (cin * k * k) ** 0.5).cuda().bfloat16().permute(0, 3, 1, 2)
    dy = torch.randn(n, m.out_hw[0], m.out_hw[1], cout, generator=g).cuda().bfloat16().permute(0, 3, 1, 2)
    xr = x.float().requires_grad_(True); wr = w.float().requires_grad_(True)
    yr = F.conv2d(xr, wr, None, stride, pd)
    yr.backward(dy.float())
    def rel(a, b): return ((a.float() - b.float()).norm() / b.float().norm()).item()
    out = []
    try:
        y, st = conv_umma.fprop(x, w, m, want_stats=True)
        out.append(f"fprop {rel(y, yr):.2e}")
    except Exception as e:
        out.append(f"fprop ERR {str(e)[:80]}")
    try:
        if conv_umma.supports_dgrad(dy, w, m, x.shape):
            dx = conv_umma.dgrad(dy, w, m, x.shape)
            out.append(f"dgrad {rel(dx, xr.grad):.2e}")
        else:
            out.append("dgrad unsupported")
    except Exception as e:
        out.append(f"dgrad ERR {str(e)[:80]}")
    try:
        dw = conv_umma.wgrad(x, dy, m)
        out.append(f"wgrad {rel(dw, wr.grad):.2e}")
    except Exception as e:
        out.append(f"wgrad ERR {str(e)[:80]}")
    torch.cuda.synchronize()
    print((n, h, cin, cout, k, stride, pd), " | ".join(out), flush=True)

for case in [(4, 35, 32, 32, 3, 1, 0), (4, 35, 32, 64, 3, 1, 1), (4, 35, 48, 64, 5, 1, 2), (4, 35, 64, 48, 1, 1, 0),
             (4, 35, 192, 32, 1, 1, 0), (4, 28, 16, 32, 5, 1, 2), (4, 28, 24, 64, 5, 1, 2), (4, 28, 192, 16, 1, 1, 0)]:
    run(*case)
