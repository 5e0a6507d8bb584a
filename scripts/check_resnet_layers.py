"""Per-layer numerics sweep: every conv geometry of a model through the tcgen05 kernels vs fp32 PyTorch."""
import sys
import torch
import torch.nn.functional as F
sys.path.insert(0, ".")
from azure_hc_intel_tf_b200 import models
from azure_hc_intel_tf_b200.models.builder import Conv
from azure_hc_intel_tf_b200.ops import conv_umma

model = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 8
net = models.create_model(model, kernels="b200")
shapes = {}
hooks = []
for mod in net.modules():
    if isinstance(mod, Conv):
        hooks.append(mod.register_forward_pre_hook(lambda m, inp: shapes.__setitem__(m, tuple(inp[0].shape))))
net_ref = net
convs_ref = [m for m in net_ref.modules() if isinstance(m, Conv)]
shapes_ref = {m: (batch, m.cin, m.in_hw[0], m.in_hw[1]) for m in convs_ref}
seen = set()
def rel(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()
worst = 0
for m in convs_ref:
    xs = shapes_ref[m]
    key = (xs, m.cout, m.k, m.stride, m.pad)
    if key in seen:
        continue
    seen.add(key)
    n, c, h, w = xs
    g = torch.Generator().manual_seed(len(seen))
    x = torch.randn(n, h, w, c, generator=g).cuda().bfloat16().permute(0, 3, 1, 2)
    wt = (torch.randn(m.cout, m.k[0], m.k[1], c, generator=g) / (c * m.k[0] * m.k[1]) ** 0.5).cuda().bfloat16().permute(0, 3, 1, 2)
    dy = (torch.randn(n, m.out_hw[0], m.out_hw[1], m.cout, generator=g) * 0.1).cuda().bfloat16().permute(0, 3, 1, 2)
    (pt, pb), (pl, pr) = m.pad
    xf, wf = x.float().requires_grad_(), wt.float().requires_grad_()
    ref = F.conv2d(F.pad(xf, (pl, pr, pt, pb)), wf, None, m.stride)
    ref.backward(dy.float())
    res = []
    if conv_umma.supports_fprop(x, wt, m):
        y, st = conv_umma.fprop(x, wt, m, want_stats=True)
        sums = conv_umma.stats_reduce(st)
        e = rel(y, ref)
        es = rel(sums[0], y.float().sum((0, 2, 3))) + rel(sums[1], (y.float() ** 2).sum((0, 2, 3)))
        res.append(f"fprop {e:.1e} stats {es:.1e}")
        worst = max(worst, e, es)
    else:
        res.append("fprop --")
    if conv_umma.supports_dgrad(dy, wt, m, x.shape):
        e = rel(conv_umma.dgrad(dy, wt, m, x.shape), xf.grad)
        res.append(f"dgrad {e:.1e}")
        worst = max(worst, e)
    else:
        res.append("dgrad --")
    if conv_umma.supports_wgrad(x, dy, m):
        e = rel(conv_umma.wgrad(x, dy, m), wf.grad)
        res.append(f"wgrad {e:.1e}")
        worst = max(worst, e)
    else:
        res.append("wgrad --")
    flag = "  <<<<<" if any("e-01" in r or "e+0" in r or "nan" in r for r in res) else ""
    print(f"{xs} -> {m.cout} k={m.k} s={m.stride} pad={m.pad}: " + " | ".join(res) + flag)
print("worst", worst)
