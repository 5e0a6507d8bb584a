"""Optimizers on plan-ordered flat buffers, and the Horovod-style ``DistributedOptimizer``.

What Horovod does for the reference: every gradient is enqueued when
autograd produces it, a background thread negotiates which tensors are ready on all ranks, copies
them into a ≤128 MiB fusion buffer (HOROVOD_FUSION_THRESHOLD, run-tf-sing-ucx-openmpi.sh:105), runs
MPI_Allreduce, copies them out, and TF then applies momentum (``--optimizer=momentum``, :73).

What happens here instead:

* the fusion plan is static (hvd/fusion.py); parameters, gradients, momentum and the bf16 working copy
  of the weights are *views into flat buffers laid out in plan order*, so there is no copy-in/copy-out;
* when the last gradient of a bucket lands, one kernel launch handles the whole bucket. On fabric
  ``nvl`` that kernel reduces the peers' buffers over NVLink, applies 1/N, weight decay and the
  momentum/LARS update, and writes fp32 master + bf16 working weights (csrc/comm/fused_allreduce_sgd.cu).
  On ``nccl``/``gloo`` the allreduce is the library's and the update is a separate flat kernel;
* ``step()`` only waits for the side stream.
"""
from __future__ import annotations

import math
from typing import Iterable, List, Optional

import torch
import torch.distributed as dist

from .compression import Compression
from .fusion import ALIGN_ELEMS, DEFAULT_THRESHOLD, FLAG_DECAY, FLAG_LARS, FusionPlan


def _physical_view(flat: torch.Tensor, like: torch.Tensor) -> torch.Tensor:
    """View ``flat`` (1-D, numel == like.numel()) with like's logical shape *and* physical element order."""
    if like.dim() == 4 and like.is_contiguous(memory_format=torch.channels_last) and not like.is_contiguous():
        k, c, r, s = like.shape
        return flat.view(k, r, s, c).permute(0, 3, 1, 2)
    return flat.view(like.shape)


def _physical_flat(t: torch.Tensor) -> torch.Tensor:
    if t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last) and not t.is_contiguous():
        return t.permute(0, 2, 3, 1).reshape(-1)
    return t.reshape(-1)


class MomentumSGD:
    """Momentum SGD (TensorFlow ``MomentumOptimizer`` semantics: ``m = μ·m + g; w -= lr·m``) with optional
    L2 weight decay folded into the gradient and optional LARS, operating on flat buffers.

    ``params`` is an iterable of parameters or ``(name, parameter)`` pairs in *forward* order; the flat
    layout uses the reverse order, which is the order backward produces gradients in.
    ``lowp=True`` additionally maintains a bf16 copy of every parameter (``p.lp``) that the sm_100a
    kernels read as their weights.
    """

    def __init__(self, params: Iterable, lr: float, momentum: float = 0.9, weight_decay: float = 0.0,
                 lars: bool = False, lars_eta: float = 0.001, lars_eps: float = 1e-9, lowp: bool = False,
                 fusion_threshold: int = DEFAULT_THRESHOLD, grad_alloc=None):
        named = []
        for i, p in enumerate(params):
            if isinstance(p, tuple):
                named.append(p)
            else:
                named.append((f"param{i}", p))
        named = [(n, p) for n, p in named if p.requires_grad]
        if not named:
            raise ValueError("optimizer got no trainable parameters")
        self.names = [n for n, _ in reversed(named)]
        self.params: List[torch.nn.Parameter] = [p for _, p in reversed(named)]
        self.param_groups = [{"lr": float(lr), "momentum": float(momentum), "weight_decay": float(weight_decay),
                              "params": self.params}]
        self.lars, self.lars_eta, self.lars_eps = bool(lars), float(lars_eta), float(lars_eps)
        self.lowp = lowp
        dev = self.params[0].device
        self.device = dev
        flags = []
        for p in self.params:
            f = FLAG_DECAY if getattr(p, "decay", True) else 0
            if p.dim() > 1:
                f |= FLAG_LARS
            flags.append(f)
        self.flags = flags
        self.plan = FusionPlan([p.numel() for p in self.params], [4] * len(self.params), [0] * len(self.params),
                               flags, fusion_threshold, ALIGN_ELEMS)
        total = self.plan.total_elems[0]
        self.total = total
        self.master = torch.zeros(total, dtype=torch.float32, device=dev)
        self.mom = torch.zeros(total, dtype=torch.float32, device=dev)
        self.grad = grad_alloc(total) if grad_alloc is not None else torch.zeros(total, dtype=torch.float32, device=dev)
        self.shadow = torch.zeros(total, dtype=torch.bfloat16, device=dev) if lowp else None
        self.ranges = []
        # per-1024-element block: which tensor owns it (-1 = padding); per tensor: decay flag, trust ratio
        nblk = total // ALIGN_ELEMS
        blk = torch.full((nblk,), -1, dtype=torch.int32)
        with torch.no_grad():
            for i, p in enumerate(self.params):
                _, start, n = self.plan.tensor_range(i)
                self.ranges.append((start, n))
                blk[start // ALIGN_ELEMS: (start + n + ALIGN_ELEMS - 1) // ALIGN_ELEMS] = i
                self.master[start:start + n].copy_(_physical_flat(p.detach()))
                p.data = _physical_view(self.master[start:start + n], p)
                p.grad = _physical_view(self.grad[start:start + n], p)
                p.flat_index = i
                if lowp:
                    self.shadow[start:start + n].copy_(self.master[start:start + n])
                    p.lp = _physical_view(self.shadow[start:start + n], p)
        self.block_tensor = blk.to(dev)
        self.tensor_decay = torch.tensor([1.0 if f & FLAG_DECAY else 0.0 for f in flags], dtype=torch.float32, device=dev)
        self.tensor_lars = torch.tensor([1.0 if f & FLAG_LARS else 0.0 for f in flags], dtype=torch.float32, device=dev)
        self.trust = torch.ones(len(flags), dtype=torch.float32, device=dev)
        self.norms = torch.zeros(2 * len(flags), dtype=torch.float32, device=dev)
        self.steps = 0

    # -- torch.optim-like surface
    @property
    def lr(self):
        return self.param_groups[0]["lr"]

    def zero_grad(self, set_to_none: bool = False):
        self.grad.zero_()

    def bucket_range(self, b: int):
        bk = self.plan.buckets[b]
        return bk.start, bk.start + bk.numel

    def update_range(self, lo: int, hi: int, grad_scale: float = 1.0):
        """Apply the update to flat elements [lo, hi). ``self.grad`` already holds the reduced gradient."""
        g = self.param_groups[0]
        lr, mu, wd = g["lr"], g["momentum"], g["weight_decay"]
        if self.master.is_cuda:
            from ..ops import b200_optim
            b200_optim.sgd_update(self, lo, hi, lr, mu, wd, grad_scale)
            return
        self._update_range_torch(lo, hi, lr, mu, wd, grad_scale)

    def _update_range_torch(self, lo, hi, lr, mu, wd, grad_scale):
        """Reference semantics in plain PyTorch (CPU path and the oracle for the CUDA kernels)."""
        with torch.no_grad():
            w, m, g = self.master[lo:hi], self.mom[lo:hi], self.grad[lo:hi]
            t = self.block_tensor[lo // ALIGN_ELEMS: hi // ALIGN_ELEMS].long()
            valid = (t >= 0).float()
            tc = t.clamp(min=0)
            dec = (self.tensor_decay[tc] * valid).repeat_interleave(ALIGN_ELEMS)
            geff = g * grad_scale + wd * dec * w
            if self.lars:
                tr = torch.where(self.tensor_lars[tc] > 0, self.trust[tc], torch.ones_like(self.trust[tc]))
                geff = geff * tr.repeat_interleave(ALIGN_ELEMS)
            m.mul_(mu).add_(geff)
            w.add_(m, alpha=-lr)
            if self.shadow is not None:
                self.shadow[lo:hi].copy_(w)

    def compute_trust(self, grad_scale: float = 1.0):
        """LARS trust ratio per tensor from the (already reduced) flat gradient:
        ``η·‖w‖ / (‖g‖ + wd·‖w‖ + ε)`` for tensors flagged FLAG_LARS, else 1."""
        if not self.lars:
            return
        wd = self.param_groups[0]["weight_decay"]
        if self.master.is_cuda:
            from ..ops import b200_optim
            b200_optim.lars_trust(self, wd, grad_scale)
            return
        with torch.no_grad():
            for i, (s, n) in enumerate(self.ranges):
                if not (self.flags[i] & FLAG_LARS):
                    self.trust[i] = 1.0
                    continue
                wn = self.master[s:s + n].norm()
                gn = (self.grad[s:s + n] * grad_scale).norm()
                d = wd if self.flags[i] & FLAG_DECAY else 0.0
                self.trust[i] = torch.where((wn > 0) & (gn > 0), self.lars_eta * wn / (gn + d * wn + self.lars_eps),
                                            torch.ones(()))

    def step(self, closure=None):
        loss = closure() if closure is not None else None
        self.compute_trust()
        self.update_range(0, self.total)
        self.steps += 1
        return loss

    # -- state
    def state_dict(self):
        return {"master": self.master.detach().cpu().clone(), "momentum": self.mom.detach().cpu().clone(),
                "steps": self.steps, "param_groups": [{k: v for k, v in self.param_groups[0].items() if k != "params"}],
                "plan_hash": self.plan.hash}

    def load_state_dict(self, sd):
        if sd.get("plan_hash") not in (None, self.plan.hash):
            raise ValueError("optimizer state was saved with a different fusion plan")
        with torch.no_grad():
            self.master.copy_(sd["master"])
            self.mom.copy_(sd["momentum"])
            if self.shadow is not None:
                self.shadow.copy_(self.master)
        self.steps = sd.get("steps", 0)
        self.param_groups[0].update(sd["param_groups"][0])

    def momentum_of(self, p) -> torch.Tensor:
        s, n = self.ranges[p.flat_index]
        return _physical_view(self.mom[s:s + n], p)

    def broadcast_state(self, root_rank: int = 0):
        dist.broadcast(self.master, src=root_rank)
        dist.broadcast(self.mom, src=root_rank)
        if self.shadow is not None:
            self.shadow.copy_(self.master)

    def sync_shadow(self):
        if self.shadow is not None:
            self.shadow.copy_(self.master)


def LARS(params, lr, momentum=0.9, weight_decay=0.0, eta=0.001, **kw):
    return MomentumSGD(params, lr, momentum, weight_decay, lars=True, lars_eta=eta, **kw)


class DistributedOptimizer:
    """``hvd.DistributedOptimizer(optimizer, named_parameters=...)``.

    * ``optimizer`` is a :class:`MomentumSGD` → fused path (see module docstring).
    * ``optimizer`` is any ``torch.optim.Optimizer`` → gradients become views into fusion buckets, each
      bucket is allreduced on a side stream as soon as it is complete, and ``step()`` calls the wrapped
      optimizer. This is the Horovod-over-NCCL baseline shape.
    """

    def __init__(self, optimizer, named_parameters=None, compression=Compression.none,
                 backward_passes_per_step: int = 1, op: str = "average",
                 fusion_threshold: Optional[int] = None, timeline=None):
        from . import _state, size, rank, fabric, symmetric_context
        if not _state["init"]:
            raise ValueError("hvd.init() must be called before DistributedOptimizer")
        self._opt = optimizer
        self.world = size()
        self.rank = rank()
        self.fabric = fabric()
        self.op = op
        self.compression = compression
        self.passes = max(1, int(backward_passes_per_step))
        self.timeline = timeline
        self.fused = isinstance(optimizer, MomentumSGD)
        self._symm = symmetric_context()
        thr = fusion_threshold or DEFAULT_THRESHOLD
        if self.fused:
            self.params = optimizer.params
            self.plan = optimizer.plan
            self.grad = optimizer.grad
            if self._symm is not None and self.world > 1:
                # move the flat gradient buffer into NVLink symmetric memory; re-point p.grad
                self._symm_grad = self._symm.alloc_like_flat(optimizer.total, torch.float32, multicast=True)
                self._symm_grad.local.copy_(optimizer.grad)
                optimizer.grad = self._symm_grad.local
                self.grad = optimizer.grad
                for p, (s, n) in zip(self.params, optimizer.ranges):
                    p.grad = _physical_view(self.grad[s:s + n], p)
        else:
            if named_parameters is not None:
                named = [(n, p) for n, p in named_parameters if p.requires_grad]
            else:
                named = [(f"g{gi}.p{pi}", p) for gi, g in enumerate(optimizer.param_groups)
                         for pi, p in enumerate(g["params"]) if p.requires_grad]
            self.params = [p for _, p in reversed(named)]
            dev = self.params[0].device
            self.plan = FusionPlan([p.numel() for p in self.params], [4] * len(self.params),
                                   [0] * len(self.params), [1] * len(self.params), thr, ALIGN_ELEMS)
            self.grad = torch.zeros(self.plan.total_elems[0], dtype=torch.float32, device=dev)
            for i, p in enumerate(self.params):
                _, s, n = self.plan.tensor_range(i)
                p.grad = _physical_view(self.grad[s:s + n], p)
                p.flat_index = i
        self.tracker = self.plan.tracker()
        # Horovod's stall inspector: warn when a fused bucket has been in flight for too long (a rank that never
        # reaches the collective). HOROVOD_STALL_CHECK_TIME_SECONDS (default 60; 0 disables) /
        # HOROVOD_STALL_SHUTDOWN_TIME_SECONDS (default 0 = never abort), Horovod's own variable names.
        import os as _os
        warn_s = float(_os.environ.get("HOROVOD_STALL_CHECK_TIME_SECONDS", "60"))
        abort_s = float(_os.environ.get("HOROVOD_STALL_SHUTDOWN_TIME_SECONDS", "0"))
        self._stall = None
        if warn_s > 0 and self.world > 1:
            from .. import _native
            lib = _native.host()
            self._stall = (lib, lib.hvd_stall_start(self.tracker._h, min(5.0, warn_s / 2), warn_s, abort_s))
        self._cuda = self.grad.is_cuda
        self._comm_stream = torch.cuda.Stream(priority=-1) if self._cuda else None
        self._handles = []
        self._comm_used = False      # the side stream holds work the compute stream has not joined yet
        self._launched = set()
        self._counts = [0] * len(self.params)
        self._direct_seen = [False] * len(self.params)
        self._plan_checked = False
        self._hooks = []
        for i, p in enumerate(self.params):
            self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(i)))
            p.grad_ready = self._make_direct(i)
        self.tracker.begin_step()

    # -- plumbing
    def _make_hook(self, i):
        def hook(param):
            # autograd may still run the accumulate node (and this hook) with an undefined gradient after a
            # kernel wrote the gradient in place and signalled readiness itself: count that once, not twice
            if self._direct_seen[i]:
                self._direct_seen[i] = False
                return
            self._on_ready(i)
        return hook

    def _make_direct(self, i):
        def ready():
            self._direct_seen[i] = True
            self._on_ready(i)
        return ready

    def _on_ready(self, i):
        self._counts[i] += 1
        if self._counts[i] < self.passes:
            return
        for b in self.tracker.ready(i):
            self._launch(b)

    def _check_plan(self):
        if self._plan_checked or self.world == 1:
            self._plan_checked = True
            return
        h = torch.tensor([self.plan.hash & 0x7FFFFFFFFFFFFFFF], dtype=torch.int64,
                         device=self.grad.device if self.fabric != "gloo" else "cpu")
        lo, hi = h.clone(), h.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        if int(lo.item()) != int(hi.item()):
            raise RuntimeError("fusion plans differ between ranks (models or thresholds are not identical)")
        self._plan_checked = True

    def _range(self, b):
        bk = self.plan.buckets[b]
        return bk.start, bk.start + bk.numel

    def _launch(self, b):
        self._launched.add(b)
        lo, hi = self._range(b)
        scale = 1.0 / (self.world * self.passes) if self.op == "average" else 1.0 / self.passes
        t0 = self.timeline.now_us() if self.timeline is not None else 0.0
        if self.world == 1:
            if self.fused and not self._opt.lars:
                self._opt.update_range(lo, hi, scale)
            self.tracker.done(b)
            return
        if self._cuda:
            self._comm_stream.wait_stream(torch.cuda.current_stream())
            self._comm_used = True
        if self.fused and self._symm is not None:
            with torch.cuda.stream(self._comm_stream):
                if self._opt.lars:
                    self._symm.allreduce_range_(self._symm_grad, lo, hi, 1.0)
                else:
                    self._symm.fused_allreduce_sgd(self._symm_grad, self._opt, lo, hi, scale)
            self._handles.append((b, None, scale))
        else:
            chunk = self.grad[lo:hi]
            wire, cctx = self.compression.compress(chunk)
            if self._cuda:
                with torch.cuda.stream(self._comm_stream):
                    work = dist.all_reduce(wire, async_op=True)
            else:
                work = dist.all_reduce(wire, async_op=True)
            self._handles.append((b, (work, wire, cctx, chunk), scale))
        if self.timeline is not None:
            self.timeline.event(f"bucket{b}", "LAUNCH", 1, t0, self.timeline.now_us() - t0)

    def synchronize(self):
        """Launch whatever has not been launched (unused parameters), then wait for all buckets."""
        self._check_plan()
        for b in range(len(self.plan.buckets)):
            if b not in self._launched:
                self._launch(b)
        pending_update = []
        for b, h, scale in self._handles:
            if h is not None:
                work, wire, cctx, chunk = h
                if self._cuda:
                    with torch.cuda.stream(self._comm_stream):
                        work.wait()
                else:
                    work.wait()
                if wire is not chunk:
                    chunk.copy_(self.compression.decompress(wire, cctx))
                pending_update.append((b, scale))
            self.tracker.done(b)
        if self.fused:
            on_side = self._cuda and self._comm_used
            stream_ctx = torch.cuda.stream(self._comm_stream) if on_side else _null()
            with stream_ctx:
                if self._opt.lars:
                    scale = 1.0 / (self.world * self.passes) if self.op == "average" else 1.0 / self.passes
                    self._opt.compute_trust(scale)
                    self._opt.update_range(0, self._opt.total, scale)
                else:
                    for b, scale in pending_update:
                        lo, hi = self._range(b)
                        self._opt.update_range(lo, hi, scale)
        else:
            for b, scale in pending_update:
                lo, hi = self._range(b)
                if scale != 1.0:
                    if self._cuda:
                        with torch.cuda.stream(self._comm_stream):
                            self.grad[lo:hi].mul_(scale)
                    else:
                        self.grad[lo:hi].mul_(scale)
        if self._cuda and self._comm_used:
            # join the side stream (inside a CUDA-graph capture this is the join of the fork made in _launch)
            torch.cuda.current_stream().wait_stream(self._comm_stream)
            self._comm_used = False
        self._handles = []

    def step(self, closure=None):
        loss = closure() if closure is not None else None
        self.synchronize()
        if self.fused:
            self._opt.steps += 1
        else:
            self._opt.step()
        self._launched.clear()
        self._counts = [0] * len(self.params)
        self._direct_seen = [False] * len(self.params)
        self.tracker.begin_step()
        return loss

    def close(self):
        """Stop the stall inspector thread (also called on garbage collection)."""
        st = getattr(self, "_stall", None)
        if st is not None and st[1]:
            st[0].hvd_stall_stop(st[1])
        self._stall = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def zero_grad(self, set_to_none: bool = False):
        # gradients are views into the fusion buffer: never detach them, just clear the buffer
        if self._cuda and self._comm_used:
            torch.cuda.current_stream().wait_stream(self._comm_stream)
            self._comm_used = False
        self.grad.zero_()

    @property
    def param_groups(self):
        return self._opt.param_groups

    def state_dict(self):
        return self._opt.state_dict()

    def load_state_dict(self, sd):
        return self._opt.load_state_dict(sd)

    def __getattr__(self, name):
        return getattr(self.__dict__["_opt"], name)


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
