"""Per-shape timing of the BatchNorm kernels against the copy roofline. usage: bn_bench.py [hbm_GBps]"""
import sys, torch
sys.path.insert(0, ".")
from azure_hc_intel_tf_b200.ops import b200
bw = float(sys.argv[1]) if len(sys.argv) > 1 else 6555.0
class BN:
    pass
def timeit(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ts = []
    for _ in range(n):
        flush.zero_()
        a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b_.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b_) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]
shapes = [(3211264, 64, 1, False), (802816, 64, 6, False), (802816, 256, 4, True), (200704, 128, 8, False), (200704, 512, 5, True),
          (50176, 256, 12, False), (50176, 1024, 7, True), (12544, 512, 6, False), (12544, 2048, 4, True)]
tot = {"apply": 0.0, "reduce": 0.0, "bapply": 0.0, "ideal_apply": 0.0, "ideal_reduce": 0.0, "ideal_bapply": 0.0}
print(f"{'rows x C':>16s} cnt res | apply us (ideal) | bwd_reduce us (ideal) | bwd_apply us (ideal)")
for rows, c, cnt, res in shapes:
    n = 256; hw = rows // n; h = int(hw ** 0.5)
    y = torch.randn(n, h, h, c, device="cuda").bfloat16().permute(0, 3, 1, 2)
    r = torch.randn(n, h, h, c, device="cuda").bfloat16().permute(0, 3, 1, 2) if res else None
    dz = torch.randn(n, h, h, c, device="cuda").bfloat16().permute(0, 3, 1, 2)
    bn = BN(); bn.training = True; bn.decay = 0.9; bn.eps = 1e-5
    bn.gamma = torch.ones(c, device="cuda"); bn.beta = torch.zeros(c, device="cuda")
    bn.moving_mean = torch.zeros(c, device="cuda"); bn.moving_variance = torch.ones(c, device="cuda")
    z, stats, mask = b200.bn_forward(y, bn, r, True, None, want_mask=True)
    L = b200.lib(); P = b200._ptr; S = b200._stream
    nb = rows * c * 2
    dg = torch.empty(c, device="cuda"); db = torch.empty(c, device="cuda")
    dy = torch.empty_like(dz)
    t_apply = timeit(lambda: L.b200_bn_apply(P(y), P(r), P(z), P(mask), rows, c, P(stats[2]), P(stats[3]), 1, S()))
    t_red = timeit(lambda: L.b200_bn_bwd_reduce(P(dz), P(y), P(None), P(mask), rows, c, P(stats[0]), P(stats[1]), P(stats[2]),
                                                P(stats[3]), 1, P(dg), P(db), P(b200.scratch(y.device)), S()))
    t_bap = timeit(lambda: L.b200_bn_bwd_apply(P(dz), P(y), P(None), P(mask), P(dy), P(None), rows, c, P(stats[0]), P(stats[1]),
                                               P(stats[2]), P(stats[3]), P(dg), P(db), 1, S()))
    i_apply = nb * (3 if res else 2) / bw / 1e3; i_red = nb * 2 / bw / 1e3; i_bap = nb * 3 / bw / 1e3
    print(f"{rows:8d} x {c:5d} {cnt:3d} {int(res):3d} | {t_apply:7.1f} ({i_apply:6.1f}) | {t_red:7.1f} ({i_red:6.1f}) | {t_bap:7.1f} ({i_bap:6.1f})")
    for k, t, i in (("apply", t_apply, i_apply), ("reduce", t_red, i_red), ("bapply", t_bap, i_bap)):
        tot[k] += cnt * t; tot["ideal_" + k] += cnt * i
print({k: round(v / 1e3, 3) for k, v in tot.items()})
