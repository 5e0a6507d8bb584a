"""Debug helper: fused allreduce+LARS vs the unfused reference, per tensor (run on >= 2 GPUs)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch


def worker(rank, world, sizes, delay=False, algo="auto", rounds=1):
    import torch.distributed as dist
    from azure_hc_intel_tf_b200 import hvd
    from azure_hc_intel_tf_b200.hvd.fusion import ALIGN_ELEMS
    ctx = hvd.symmetric_context()
    dev = hvd.device()
    torch.manual_seed(99)
    params = [(f"t{i}", torch.nn.Parameter(torch.randn(n, 1, device=dev) if i % 2 else torch.randn(n, device=dev)))
              for i, n in enumerate(sizes)]
    kw = dict(lr=0.05, momentum=0.9, weight_decay=1e-3, lowp=True, fusion_threshold=max(sizes) * 4 // 2, eta=0.01)
    base = hvd.LARS(params, **kw)
    opt = hvd.DistributedOptimizer(base, named_parameters=params)
    nb = len(base.plan.buckets)
    scale = 1.0 / world
    from azure_hc_intel_tf_b200.parallel import symm as symm_mod
    ctx.algo = symm_mod.ALGOS[algo]
    g = torch.Generator(device="cpu").manual_seed(rank)
    mine = torch.randn(base.total, generator=g).to(dev)
    mine[base.block_tensor.repeat_interleave(ALIGN_ELEMS) < 0] = 0.0
    every = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(every, mine)
    acc = every[0].clone()
    for r in range(1, world):
        acc += every[r]
    acc *= scale
    ref = hvd.LARS([(n, torch.nn.Parameter(p.detach().clone())) for n, p in params], **kw)
    ref.master.copy_(base.master); ref.mom.copy_(base.mom); ref.grad.copy_(acc)
    ref.compute_trust()
    ref._update_range_torch(0, ref.total, ref.lr, 0.9, 1e-3, 1.0)
    base.grad.copy_(mine)
    torch.cuda.synchronize(); dist.barrier()
    info = []
    for b in range(nb):
        lo, hi = base.bucket_range(b)
        la = base.lars_bucket_args(b, ctx.block)
        info.append((b, lo, hi, la))
        if delay and rank == b % world:
            torch.cuda._sleep(3_000_000)
        ctx.fused_allreduce_sgd(opt._symm_grad, base, lo, hi, scale, lars=la)
    torch.cuda.synchronize()
    out = [("buckets", info), ("ranges", base.ranges), ("flags", base.flags),
           ("grad_equal", bool(torch.allclose(base.grad, acc, atol=1e-6)))]
    for i, (s, n) in enumerate(base.ranges):
        out.append((i, n, float(base.trust[i]), float(ref.trust[i]),
                    float((base.mom[s:s + n] - ref.mom[s:s + n]).abs().max())))
    return out


if __name__ == "__main__":
    from azure_hc_intel_tf_b200.utils import testing
    sizes = [2_359_296, 393_216, 1_048_576 + 512, 2_359_296, 65_536, 1000, 589_824]
    for delay, algo in ((True, "auto"), (False, "twoshot"), (True, "twoshot")):
        outs = testing.run_ranks(worker, 2, "nvl", args=(sizes, delay, algo), timeout=300)
        print("delay", delay, "algo", algo)
        for o in outs[0][4:]:
            print(o)
