"""Worker bodies for the multi-process tests (they must live in an importable module for ``spawn``)."""
from __future__ import annotations

import copy

import torch
import torch.nn as nn


def w_basic(rank, world):
    from .. import hvd
    assert hvd.size() == world and hvd.rank() == rank and hvd.local_rank() == rank
    dev = hvd.device()
    t = torch.full((5,), float(rank + 1), device=dev)
    s = hvd.allreduce(t, op=hvd.Sum)
    a = hvd.allreduce(t)                       # average is the default, as in horovod.torch
    g = hvd.allgather(torch.full((rank + 1, 2), float(rank), device=dev))
    b = hvd.broadcast(torch.full((3,), float(rank), device=dev), root_rank=world - 1)
    h = hvd.allreduce_async_(torch.ones(4, device=dev), op=hvd.Sum)
    while not hvd.poll(h):
        pass
    z = hvd.synchronize(h)
    obj = hvd.broadcast_object({"r": rank}, 0)
    # the TensorFlow-flavoured entry points: ranks start with different weights and BN statistics, the hook aligns them
    bn = nn.BatchNorm2d(4).to(dev)
    with torch.no_grad():
        bn.weight.fill_(float(rank + 2))
        bn.running_mean.fill_(float(rank + 5))
    hook = hvd.BroadcastGlobalVariablesHook(0)
    hook.begin(bn)
    hook.begin(bn)                             # idempotent
    assert hook.done and float(bn.weight[0]) == 2.0 and float(bn.running_mean[0]) == 5.0, (bn.weight, bn.running_mean)
    return (s.cpu().tolist(), a.cpu().tolist(), g.shape[0], b.cpu().tolist(), z.cpu().tolist(), obj)


def _net(dev, kernels=None):
    torch.manual_seed(1234)
    net = nn.Sequential(nn.Conv2d(3, 8, 3, padding=1), nn.ReLU(), nn.Conv2d(8, 4, 3, padding=1), nn.Flatten(),
                        nn.Linear(4 * 8 * 8, 10))
    return net.to(dev).to(memory_format=torch.channels_last)


def w_dist_opt(rank, world, mode, steps):
    """Train a tiny BN-free net on this rank's shard; return final weights. ``mode``:
    'fused'  — hvd.MomentumSGD wrapped by DistributedOptimizer (flat fused path)
    'generic' — torch.optim.SGD wrapped (Horovod-over-library path)"""
    from .. import hvd
    dev = hvd.device()
    net = _net(dev)
    if rank != 0:                         # ranks start different; broadcast must fix that
        with torch.no_grad():
            for p in net.parameters():
                p.add_(1.0)
    named = list(net.named_parameters())
    if mode == "fused":
        base = hvd.MomentumSGD(named, lr=0.05, momentum=0.9, weight_decay=1e-3, fusion_threshold=8192)
    elif mode == "lars":
        base = hvd.LARS(named, lr=0.05, momentum=0.9, weight_decay=1e-3, eta=0.01, fusion_threshold=8192)
    elif mode in ("rmsprop", "adam"):
        base = _adaptive(hvd, mode, named, fusion_threshold=8192)
    else:
        base = torch.optim.SGD(net.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-3)
    opt = hvd.DistributedOptimizer(base, named_parameters=named, fusion_threshold=8192)
    hvd.broadcast_parameters(net.state_dict(), root_rank=0)
    hvd.broadcast_optimizer_state(opt, root_rank=0)
    g = torch.Generator().manual_seed(7)
    full_x = torch.randn(steps, 4 * world, 3, 8, 8, generator=g)
    full_y = torch.randn(steps, 4 * world, 10, generator=g)
    for s in range(steps):
        x = full_x[s, rank * 4:(rank + 1) * 4].to(dev)
        y = full_y[s, rank * 4:(rank + 1) * 4].to(dev)
        opt.zero_grad()
        ((net(x) - y) ** 2).mean().backward()
        opt.step()
    return [p.detach().cpu().contiguous() for p in net.parameters()], len(opt.plan.buckets)


def single_process_reference(world, mode, steps):
    net = _net("cpu")
    if mode == "lars":
        from .. import hvd
        opt = hvd.LARS(list(net.named_parameters()), lr=0.05, momentum=0.9, weight_decay=1e-3, eta=0.01)
    elif mode in ("rmsprop", "adam"):
        from .. import hvd
        opt = _adaptive(hvd, mode, list(net.named_parameters()))
    else:
        opt = torch.optim.SGD(net.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-3)
    g = torch.Generator().manual_seed(7)
    full_x = torch.randn(steps, 4 * world, 3, 8, 8, generator=g)
    full_y = torch.randn(steps, 4 * world, 10, generator=g)
    for s in range(steps):
        opt.zero_grad()
        ((net(full_x[s]) - full_y[s]) ** 2).mean().backward()
        opt.step()
    return [p.detach().cpu().contiguous() for p in net.parameters()]


def _adaptive(hvd, mode, named, **kw):
    if mode == "rmsprop":
        return hvd.RMSProp(named, lr=0.01, momentum=0.9, weight_decay=1e-3, **kw)
    return hvd.Adam(named, lr=0.01, weight_decay=1e-3, **kw)


def w_plan_mismatch(rank, world):
    from .. import hvd
    dev = hvd.device()
    net = nn.Linear(4 + rank, 3).to(dev)           # ranks disagree on the model
    opt = hvd.DistributedOptimizer(hvd.MomentumSGD(list(net.named_parameters()), lr=0.1))
    net(torch.randn(2, 4 + rank, device=dev)).sum().backward()
    try:
        opt.step()
    except RuntimeError as e:
        return "mismatch" if "fusion plans differ" in str(e) else str(e)
    return "no error"


def w_cli(rank, world, argv):
    from ..cli import cnn_benchmarks
    from ..utils import flags
    p = flags.parse(argv)
    b = cnn_benchmarks.BenchmarkCNN(p)
    out = b.run()
    first = next(iter(b.model.parameters())).detach().float().cpu().sum().item()
    return out["images_per_sec"], first


# ------------------------------------------------------------------------------------ NVLink fabric (GPU)

def w_nvl_allreduce(rank, world, sizes, iters):
    """Own P2P allreduce kernels vs the arithmetic answer and vs NCCL, many sizes, repeated (flag-protocol stress)."""
    import torch.distributed as dist
    from .. import hvd
    ctx = hvd.symmetric_context()
    assert ctx is not None and hvd.fabric() == "nvl"
    dev = hvd.device()
    out = []
    for dtype, gran in ((torch.float32, 1024), (torch.bfloat16, 2048)):
        nmax = max(sizes)
        buf = ctx.alloc_like_flat(nmax, dtype, multicast=(dtype == torch.float32))   # NVLS for the large fp32 sizes
        for n in sizes:
            n = (n + gran - 1) // gran * gran
            g = torch.Generator(device="cpu").manual_seed(1000 * rank + n % 997)
            for it in range(iters):
                x = torch.randn(n, generator=g).to(dev).to(dtype)
                ref = x.clone().float()
                dist.all_reduce(ref)                                  # NCCL, fp32 accumulate
                buf.local[:n].copy_(x)
                torch.cuda.synchronize()
                dist.barrier()
                ctx.allreduce_range_(buf, 0, n, 0.5)
                torch.cuda.synchronize()
                got = buf.local[:n].float()
                tol = 1e-6 if dtype == torch.float32 else 2e-2
                ok = torch.allclose(got, ref * 0.5, atol=tol * max(1.0, ref.abs().max().item()), rtol=tol)
                # all ranks must hold bitwise identical results
                chk = got.double().sum().reshape(1)
                lo, hi = chk.clone(), chk.clone()
                dist.all_reduce(lo, op=dist.ReduceOp.MIN)
                dist.all_reduce(hi, op=dist.ReduceOp.MAX)
                out.append((str(dtype), n, bool(ok), bool(lo.item() == hi.item())))
    # back-to-back launches without host synchronisation in between (epochs must keep the ranks in step)
    buf = ctx.alloc_like_flat(1 << 20, torch.float32)
    buf.local.fill_(float(rank + 1))
    torch.cuda.synchronize()
    dist.barrier()
    for _ in range(200):
        ctx.allreduce_range_(buf, 0, 1 << 20, 1.0 / world)
    torch.cuda.synchronize()
    expect = sum(range(1, world + 1)) / world
    out.append(("chain", 1 << 20, bool(torch.allclose(buf.local, torch.full_like(buf.local, expect), rtol=1e-5)), True))
    # hvd.allreduce on an ordinary tensor (staged through symmetric scratch) and the broadcast kernel
    t = torch.arange(5000, device=dev, dtype=torch.float32) * (rank + 1)
    s = hvd.allreduce(t, op=hvd.Sum)
    out.append(("staged", 5000, bool(torch.allclose(s, torch.arange(5000, device=dev) * float(sum(range(1, world + 1))))), True))
    b = ctx.alloc_like_flat(4096, torch.float32)
    b.local.fill_(float(rank))
    torch.cuda.synchronize()
    dist.barrier()
    ctx.broadcast_(b, 0, 4096, world - 1)
    torch.cuda.synchronize()
    out.append(("bcast", 4096, bool((b.local == float(world - 1)).all().item()), True))
    out.append(("nvls_available", 0, True, True) if ctx.nvls_ok else ("nvls_unavailable_fell_back_to_p2p", 0, True, True))
    return out


def w_nvl_flag_stress(rank, world, iters, seed):
    """Flag-protocol stress: thousands of back-to-back collectives with no host synchronisation,
    message sizes hopping across the one-shot / two-shot / NVLS switch points, and a random device-side delay in
    front of every launch so the ranks arrive at the barriers in a different order each time. Every iteration each
    rank overwrites its buffer with a fresh value, so a peer read that is one iteration stale (a barrier that let
    someone through early) or a write that lands late shows up as a wrong sum. Mismatches are counted on the device."""
    import random
    import torch.distributed as dist
    from .. import hvd
    ctx = hvd.symmetric_context()
    assert ctx is not None and hvd.fabric() == "nvl"
    dev = hvd.device()
    one_shot = ctx.one_shot_max_bytes // 4 if hasattr(ctx, "one_shot_max_bytes") else 32 * 1024
    nvls = max(ctx.nvls_min_bytes // 4, 1 << 20) if hasattr(ctx, "nvls_min_bytes") else 2 << 20
    sizes = [1024, one_shot, one_shot + 1024, 256 * 1024, nvls, nvls + 1024]
    nmax = max(sizes)
    buf = ctx.alloc_like_flat(nmax, torch.float32, multicast=True)
    bad = torch.zeros((), dtype=torch.int64, device=dev)
    rng = random.Random(seed * 131 + rank)          # per-rank delays
    pick = random.Random(seed)                      # same size sequence on every rank
    torch.cuda.synchronize()
    dist.barrier()
    for it in range(iters):
        n = pick.choice(sizes)
        torch.cuda._sleep(rng.randrange(0, 60_000))
        buf.local[:n].fill_(float(rank + it % 97))
        ctx.allreduce_range_(buf, 0, n, 1.0 / world)
        expect = (it % 97) + (world - 1) / 2.0
        bad += (buf.local[:n] != expect).sum()
    torch.cuda.synchronize()
    return int(bad.item())


def w_nvl_fused_training(rank, world, steps, lars):
    """DistributedOptimizer on the nvl fabric (fused allreduce+SGD kernel) vs a single-process run on the
    concatenated batch; also returns whether ranks ended bit-identical."""
    from .. import hvd
    dev = hvd.device()
    net = _net(dev)
    named = list(net.named_parameters())
    if lars in ("rmsprop", "adam"):        # own allreduce kernel + flat RMSProp / Adam kernel per bucket
        inner = _adaptive(hvd, lars, named, fusion_threshold=8192)
    else:
        mk = hvd.LARS if lars else hvd.MomentumSGD
        kw = dict(lr=0.05, momentum=0.9, weight_decay=1e-3, fusion_threshold=8192)
        if lars:
            kw["eta"] = 0.01
        inner = mk(named, **kw)
    opt = hvd.DistributedOptimizer(inner, named_parameters=named)
    hvd.broadcast_parameters(net.state_dict(), root_rank=0)
    hvd.broadcast_optimizer_state(opt, root_rank=0)
    g = torch.Generator().manual_seed(7)
    full_x = torch.randn(steps, 4 * world, 3, 8, 8, generator=g)
    full_y = torch.randn(steps, 4 * world, 10, generator=g)
    for s in range(steps):
        x = full_x[s, rank * 4:(rank + 1) * 4].to(dev)
        y = full_y[s, rank * 4:(rank + 1) * 4].to(dev)
        opt.zero_grad()
        ((net(x) - y) ** 2).mean().backward()
        opt.step()
    torch.cuda.synchronize()
    return [p.detach().cpu().contiguous() for p in net.parameters()], len(opt.plan.buckets), opt._symm is not None


def w_nvl_fused_buckets(rank, world, sizes, algo, lars, rounds):
    """Fused allreduce + update kernels on buckets far above the one-shot range (the sizes real models produce), with
    one rank delayed on the device before every launch and no host synchronisation between buckets. Checks, per round:
    every rank ends with bit-identical master / momentum / bf16 weights, and they match the unfused arithmetic
    (gradients gathered with NCCL, summed in rank order in fp32, then the plain PyTorch update rule). ``sizes``:
    tensor sizes in elements; the plan cuts them into buckets of ``sizes[0]`` elements' worth of bytes."""
    import torch.distributed as dist
    from .. import hvd
    from ..parallel import symm as symm_mod
    ctx = hvd.symmetric_context()
    assert ctx is not None and hvd.fabric() == "nvl"
    dev = hvd.device()
    torch.manual_seed(99)
    params = [(f"t{i}", torch.nn.Parameter(torch.randn(n, 1, device=dev) if i % 2 else torch.randn(n, device=dev)))
              for i, n in enumerate(sizes)]
    mk = hvd.LARS if lars else hvd.MomentumSGD
    kw = dict(lr=0.05, momentum=0.9, weight_decay=1e-3, lowp=True, fusion_threshold=max(sizes) * 4 // 2)
    if lars:
        kw["eta"] = 0.01
    base = mk(params, **kw)
    opt = hvd.DistributedOptimizer(base, named_parameters=params)
    assert opt._symm is not None
    ctx.algo = symm_mod.ALGOS[algo]
    if algo == "nvls" and not getattr(opt._symm_grad, "mc_ptr", 0):
        return "skip: no multicast mapping"
    nb = len(base.plan.buckets)
    assert nb >= 2, nb
    scale = 1.0 / world
    bad = []
    for rnd in range(rounds):
        g = torch.Generator(device="cpu").manual_seed(1000 * rnd + rank)
        mine = torch.randn(base.total, generator=g).to(dev)
        # elements that belong to no tensor (alignment padding, the tail of a tensor's last block) never receive a
        # gradient in training: keep them zero here too
        live = torch.zeros(base.total, dtype=torch.bool, device=dev)
        for s0, n0 in base.ranges:
            live[s0:s0 + n0] = True
        mine[~live] = 0.0
        # reference: what the update must produce
        every = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(every, mine)
        acc = every[0].clone()
        for r in range(1, world):
            acc += every[r]
        acc *= scale
        ref = mk([(n, torch.nn.Parameter(p.detach().clone())) for n, p in params], **kw)
        ref.master.copy_(base.master)
        ref.mom.copy_(base.mom)
        ref.grad.copy_(acc)
        ref.compute_trust()
        ref._update_range_torch(0, ref.total, ref.lr, 0.9, 1e-3, 1.0)
        base.grad.copy_(mine)
        torch.cuda.synchronize()
        dist.barrier()
        for b in range(nb):
            if rank == (rnd + b) % world:
                torch.cuda._sleep(3_000_000)            # ~1.5 ms: this rank reaches the kernel late
            lo, hi = base.bucket_range(b)
            la = base.lars_bucket_args(b, ctx.block) if lars else None
            ctx.fused_allreduce_sgd(opt._symm_grad, base, lo, hi, scale, lars=la)
        torch.cuda.synchronize()
        for name, t, r_t, tol in (("master", base.master, ref.master, 2e-5), ("mom", base.mom, ref.mom, 2e-5),
                                  ("shadow", base.shadow.float(), ref.shadow.float(), 1e-2)):
            peers = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(peers, t.contiguous())
            if not all(torch.equal(peers[0], p) for p in peers[1:]):
                bad.append((rnd, name, "ranks differ"))
            if not torch.allclose(t, r_t, atol=tol, rtol=tol):
                where = [(i, n, round(float((t[s0:s0 + n] - r_t[s0:s0 + n]).abs().max()), 6))
                         for i, (s0, n) in enumerate(base.ranges)
                         if not torch.allclose(t[s0:s0 + n], r_t[s0:s0 + n], atol=tol, rtol=tol)]
                bad.append((rnd, name, "mismatch vs unfused", where,
                            [round(float(x), 6) for x in base.trust], [round(float(x), 6) for x in ref.trust]))
    ctx.algo = symm_mod.ALGO_AUTO
    return bad


def w_nvl_misc(rank, world):
    """All-gather, point-to-point probes, sub-block allreduce and the host-level hvd collectives on the nvl fabric."""
    from .. import hvd
    ctx = hvd.symmetric_context()
    dev = hvd.device()
    bad = []
    # all-gather kernel, plain and multicast-capable buffers
    for mc in (False, True):
        n = 4096 + 8
        buf = ctx.alloc_like_flat(n * world, torch.float32, multicast=mc)
        buf.local.fill_(-1.0)
        buf.local[rank * n:(rank + 1) * n] = torch.arange(n, device=dev, dtype=torch.float32) + 1000.0 * rank
        torch.cuda.synchronize()
        hvd.barrier()
        ctx.allgather_(buf, 0, n)
        torch.cuda.synchronize()
        want = torch.cat([torch.arange(n, device=dev, dtype=torch.float32) + 1000.0 * r for r in range(world)])
        if not torch.equal(buf.local, want):
            bad.append(("allgather", mc))
    # allreduce below one block (1 KiB and 16 bytes): the one-shot kernel masks the tail
    buf = ctx.alloc_like_flat(4096, torch.float32)
    for n in (4, 256, 1000, 1024 + 512):
        buf.local.fill_(7.0)
        buf.local[:n] = float(rank + 1)
        torch.cuda.synchronize()
        hvd.barrier()
        ctx.allreduce_range_(buf, 0, n, 1.0)
        torch.cuda.synchronize()
        tot = float(sum(range(1, world + 1)))
        if not (bool((buf.local[:n] == tot).all()) and bool((buf.local[n:] == 7.0).all())):
            bad.append(("small allreduce", n))
    # point-to-point: rank 0 pushes to / pulls from rank 1; ping-pong latency probe
    big = ctx.alloc_like_flat(1 << 20, torch.float32)
    big.local.fill_(float(rank))
    torch.cuda.synchronize()
    hvd.barrier()
    if rank == 0:
        big.local[:4096] = 42.0
        ctx.p2p_copy(big, 1, 0, 4096, pull=False)
        ctx.p2p_copy(big, 1, 8192, 4096, pull=True)
        torch.cuda.synchronize()
    hvd.barrier()
    if rank == 1 and not bool((big.local[:4096] == 42.0).all()):
        bad.append(("p2p push",))
    if rank == 0 and not bool((big.local[8192:8192 + 4096] == 1.0).all()):
        bad.append(("p2p pull",))
    out = torch.zeros(2, device=dev)
    for a, b in ((0, 1), (1, 0), (0, world - 1)):
        if a == b:
            continue
        ctx.p2p_pingpong(big, a, b, 256, 50, out)
        torch.cuda.synchronize()
        if rank == a and not (0.2 < float(out[0]) < 1000.0):
            bad.append(("pingpong", a, b, float(out[0])))
    # host-level API
    g = hvd.allgather(torch.full((rank + 1, 3), float(rank), device=dev))
    if g.shape[0] != sum(range(1, world + 1)) or float(g[-1, 0]) != float(world - 1):
        bad.append(("hvd.allgather",))
    t = torch.arange(10, device=dev, dtype=torch.int64) * (rank + 1)
    hvd.broadcast_(t, root_rank=world - 1)
    if not torch.equal(t, torch.arange(10, device=dev, dtype=torch.int64) * world):
        bad.append(("hvd.broadcast_ int64",))
    obj = hvd.broadcast_object({"rank": rank, "x": [1, 2, 3]}, 0)
    if obj != {"rank": 0, "x": [1, 2, 3]}:
        bad.append(("broadcast_object", obj))
    return bad
