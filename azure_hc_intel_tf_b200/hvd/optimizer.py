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
  momentum/LARS update, and writes fp32 master + bf16 working weights (csrc/comm/allreduce.cu).
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
                 fusion_threshold: int = DEFAULT_THRESHOLD, grad_alloc=None, fusion_tail: int = 0,
                 rule: str = "momentum", hyper: Optional[dict] = None):
        if rule not in ("momentum", "rmsprop", "adam"):
            raise ValueError(f"unknown update rule {rule!r}")
        if rule != "momentum" and lars:
            raise ValueError("LARS is a momentum-SGD variant")
        # the update rule applied to the flat buffers: TF ApplyMomentum, ApplyRMSProp (rho, eps; ms starts at 1) or
        # ApplyAdam (beta1, beta2, eps) - tf_cnn_benchmarks' --optimizer=momentum|sgd|rmsprop|adam with its defaults
        self.rule = rule
        self.hyper = {"momentum": {}, "rmsprop": {"rho": 0.9, "eps": 1.0},
                      "adam": {"beta1": 0.9, "beta2": 0.999, "eps": 1e-8}}[rule]
        self.hyper.update(hyper or {})
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
                               flags, fusion_threshold, ALIGN_ELEMS, fusion_tail)
        total = self.plan.total_elems[0]
        self.total = total
        self.master = torch.zeros(total, dtype=torch.float32, device=dev)
        self.mom = torch.zeros(total, dtype=torch.float32, device=dev)       # Momentum | RMSProp momentum | Adam m
        self.state2 = None                                                    # RMSProp mean square | Adam v
        if rule == "rmsprop":
            self.state2 = torch.ones(total, dtype=torch.float32, device=dev)
        elif rule == "adam":
            self.state2 = torch.zeros(total, dtype=torch.float32, device=dev)
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
        # the learning rate also lives in device memory: the update kernels read it from there, so a schedule (warm-up,
        # decay) keeps working when the whole step is replayed from a CUDA graph
        self.lr_dev = torch.full((1,), float(lr), dtype=torch.float32, device=dev) if dev.type == "cuda" else None
        self._lr_dev_value = float(lr)
        self.lr_schedule = None
        self._lars_tab = None
        self.steps = 0
        self._push_lr()

    # -- torch.optim-like surface
    @property
    def lr(self):
        return self.param_groups[0]["lr"]

    def effective_lr(self) -> float:
        """The rate the update kernels apply in the step about to run: the base rate, for Adam times the bias
        correction sqrt(1 - beta2^t) / (1 - beta1^t) of step t = steps + 1."""
        lr = self.param_groups[0]["lr"]
        if self.rule == "adam":
            t = self.steps + 1
            lr = lr * math.sqrt(1.0 - self.hyper["beta2"] ** t) / (1.0 - self.hyper["beta1"] ** t)
        return lr

    def _push_lr(self):
        eff = float(self.effective_lr())
        if self.lr_dev is not None and eff != self._lr_dev_value:
            self.lr_dev.fill_(eff)
            self._lr_dev_value = eff

    def set_lr(self, lr: float):
        """Change the learning rate (host copy and the device scalar the kernels read)."""
        self.param_groups[0]["lr"] = float(lr)
        self._push_lr()

    def apply_schedule(self):
        """Set the rate for the step about to run: ``lr_schedule(step)`` (utils/lr_schedule.py) if one is set, and the
        device scalar the kernels read (which for Adam changes every step with the bias correction)."""
        if self.master.is_cuda and torch.cuda.is_current_stream_capturing():
            return      # a fill captured into the graph would reset the rate on every replay; the caller sets it outside
        if self.lr_schedule is not None:
            self.param_groups[0]["lr"] = float(self.lr_schedule(self.steps))
        self._push_lr()

    def lars_tables(self, block: int):
        """(per-block partial-norm buffer, first block of each tensor, blocks of each tensor), in units of ``block``
        elements, for the fused allreduce+LARS kernel (include/b200comm.h)."""
        if self._lars_tab is None or self._lars_tab[0] != block:
            lo = torch.tensor([s // block for s, _ in self.ranges], dtype=torch.int32, device=self.device)
            cnt = torch.tensor([(n + block - 1) // block for _, n in self.ranges], dtype=torch.int32, device=self.device)
            part = torch.zeros(2 * (self.total // block), dtype=torch.float32, device=self.device)
            self._lars_tab = (block, part, lo, cnt)
        return self._lars_tab[1:]

    def lars_bucket_args(self, b: int, block: int):
        """(t_lo, t_hi, upd_lo, upd_hi) for bucket b: the tensors whose last element lies in the bucket, and the block
        range they span (which starts in an earlier bucket when the first of them was split by the planner)."""
        lo, hi = self.bucket_range(b)
        ends = [i for i, (s, n) in enumerate(self.ranges) if lo < s + n <= hi]
        if not ends:
            return 0, 0, 0, 0
        t_lo, t_hi = ends[0], ends[-1] + 1
        s0 = self.ranges[t_lo][0]
        s1, n1 = self.ranges[t_hi - 1]
        return t_lo, t_hi, s0 // block, (s1 + n1 + block - 1) // block

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
            if self.rule == "momentum":
                b200_optim.sgd_update(self, lo, hi, lr, mu, wd, grad_scale)
            else:
                b200_optim.adaptive_update(self, lo, hi, lr, wd, grad_scale)
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
            if self.rule != "momentum":
                h, s2 = self.hyper, self.state2[lo:hi]
                live = valid.repeat_interleave(ALIGN_ELEMS) > 0       # padding between tensors stays untouched
                if self.rule == "rmsprop":
                    s2n = h["rho"] * s2 + (1.0 - h["rho"]) * geff * geff
                    mn = mu * m + lr * geff / torch.sqrt(s2n + h["eps"])
                    wn = w - mn
                else:
                    mn = h["beta1"] * m + (1.0 - h["beta1"]) * geff
                    s2n = h["beta2"] * s2 + (1.0 - h["beta2"]) * geff * geff
                    wn = w - self.effective_lr() * mn / (torch.sqrt(s2n) + h["eps"])
                s2.copy_(torch.where(live, s2n, s2))
                m.copy_(torch.where(live, mn, m))
                w.copy_(torch.where(live, wn, w))
                if self.shadow is not None:
                    self.shadow[lo:hi].copy_(w)
                return
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
        self.apply_schedule()
        self.compute_trust()
        self.update_range(0, self.total)
        self.steps += 1
        return loss

    # -- state
    def state_dict(self):
        extra = {"state2": self.state2.detach().cpu().clone()} if self.state2 is not None else {}
        return {"master": self.master.detach().cpu().clone(), "momentum": self.mom.detach().cpu().clone(), **extra,
                "rule": self.rule, "steps": self.steps, "param_groups": [{k: v for k, v in self.param_groups[0].items() if k != "params"}],
                "plan_hash": self.plan.hash}

    def load_state_dict(self, sd):
        if sd.get("plan_hash") not in (None, self.plan.hash):
            raise ValueError("optimizer state was saved with a different fusion plan")
        if sd.get("rule", "momentum") != self.rule:
            raise ValueError(f"optimizer state was saved by a {sd.get('rule')} optimizer, this one is {self.rule}")
        with torch.no_grad():
            self.master.copy_(sd["master"])
            self.mom.copy_(sd["momentum"])
            if self.state2 is not None:
                self.state2.copy_(sd["state2"])
            if self.shadow is not None:
                self.shadow.copy_(self.master)
        self.steps = sd.get("steps", 0)
        self.param_groups[0].update(sd["param_groups"][0])
        self._lr_dev_value = None
        self._push_lr()

    def momentum_of(self, p) -> torch.Tensor:
        s, n = self.ranges[p.flat_index]
        return _physical_view(self.mom[s:s + n], p)

    def slots_of(self, p) -> dict:
        """Optimizer slot variables of parameter ``p`` under their TensorFlow slot names (checkpoints, utils/tfnames.py)."""
        s, n = self.ranges[p.flat_index]
        first = _physical_view(self.mom[s:s + n], p)
        if self.rule == "momentum":
            return {"Momentum": first}
        second = _physical_view(self.state2[s:s + n], p)
        if self.rule == "rmsprop":
            return {"RMSProp": second, "RMSProp_1": first}        # TF: slot "rms" -> <var>/RMSProp, "momentum" -> /RMSProp_1
        return {"Adam": first, "Adam_1": second}

    def broadcast_state(self, root_rank: int = 0):
        from . import symmetric_context
        ctx = symmetric_context()
        for t in (self.master, self.mom) + ((self.state2,) if self.state2 is not None else ()):
            if ctx is not None and t.is_cuda:
                ctx.broadcast_tensor_(t, root_rank)      # own NVLink broadcast kernel, no NCCL
            else:
                dist.broadcast(t, src=root_rank)
        if self.shadow is not None:
            self.shadow.copy_(self.master)

    def sync_shadow(self):
        if self.shadow is not None:
            self.shadow.copy_(self.master)


def LARS(params, lr, momentum=0.9, weight_decay=0.0, eta=0.001, **kw):
    return MomentumSGD(params, lr, momentum, weight_decay, lars=True, lars_eta=eta, **kw)


def RMSProp(params, lr, momentum=0.9, weight_decay=0.0, rho=0.9, eps=1.0, **kw):
    """TF ``RMSPropOptimizer(lr, decay=0.9, momentum=0.9, epsilon=1.0)`` as tf_cnn_benchmarks builds it, on flat buffers."""
    return MomentumSGD(params, lr, momentum, weight_decay, rule="rmsprop", hyper={"rho": rho, "eps": eps}, **kw)


def Adam(params, lr, weight_decay=0.0, beta1=0.9, beta2=0.999, eps=1e-8, **kw):
    """TF ``AdamOptimizer(lr, 0.9, 0.999, 1e-8)`` on flat buffers."""
    return MomentumSGD(params, lr, 0.0, weight_decay, rule="adam", hyper={"beta1": beta1, "beta2": beta2, "eps": eps}, **kw)


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
        # on the NVLink fabric every collective of this process runs on the context's one communication stream (the
        # kernels share per-CTA flag slots and must never overlap each other); otherwise a private side stream
        if self._cuda and self._symm is not None:
            self._comm_stream = self._symm.stream
        else:
            self._comm_stream = torch.cuda.Stream(priority=-1) if self._cuda else None
        self._tl_events, self._tl_base = [], None
        self._ready = set()          # buckets whose gradients are complete but that wait for an earlier bucket
        self._next = 0               # buckets are launched in plan order: the same order on every rank by construction
        self._handles = []
        self._comm_used = False      # the side stream holds work the compute stream has not joined yet
        self._launched = set()
        self._counts = [0] * len(self.params)
        self._direct_seen = [False] * len(self.params)
        self._plan_checked = False
        self._hooks = []
        for i, p in enumerate(self.params):
            self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(i)))
            # The kernels' direct hand-off OVERWRITES the gradient view. With several backward passes per step the
            # passes must add up, so the hand-off is left off and autograd accumulates into the view itself.
            p.grad_ready = self._make_direct(i) if self.passes == 1 else None
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
            self._ready.add(b)
        while self._next in self._ready:
            self._ready.discard(self._next)
            self._launch(self._next)
            self._next += 1

    def _check_plan(self):
        if self._plan_checked or self.world == 1:
            self._plan_checked = True
            return
        h = torch.tensor([self.plan.hash & 0x7FFFFFFFFFFFFFFF], dtype=torch.int64,
                         device=self.grad.device if self.fabric != "gloo" else "cpu")
        if self._symm is not None and h.is_cuda:
            every = torch.cat(self._symm.allgather_tensor(h))        # own all-gather kernel
            lo, hi = every.min(), every.max()
        else:
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
        # device-timed timeline (HOROVOD_TIMELINE): per bucket an event when its last gradient is ready (compute
        # stream), when its collective starts and when it is done (communication stream). Not under graph capture.
        tl_dev = (self.timeline is not None and self._cuda and self.world > 1
                  and not torch.cuda.is_current_stream_capturing())
        ev = None
        if tl_dev:
            if not self._tl_events:
                self._tl_base = (torch.cuda.Event(enable_timing=True), self.timeline.now_us())
                self._tl_base[0].record()
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            ev[0].record()
        if self.fused and b == 0:
            self._opt.apply_schedule()       # before the first update of the step (the rate is a device scalar)
        if self.world == 1:
            if self.fused and not self._opt.lars:
                self._opt.update_range(lo, hi, scale)
            self.tracker.done(b)
            return
        if self._cuda:
            self._comm_stream.wait_stream(torch.cuda.current_stream())
            self._comm_used = True
        if ev is not None:
            ev[1].record(self._comm_stream)
        if self.fused and self._symm is not None:
            with torch.cuda.stream(self._comm_stream):
                if self._opt.rule != "momentum":
                    # RMSProp / Adam: own NVLink allreduce of the bucket, then the flat update kernel behind it
                    self._symm.allreduce_range_(self._symm_grad, lo, hi, scale)
                    self._opt.update_range(lo, hi, 1.0)
                else:
                    # one launch per bucket either way; LARS folds the norms of the tensors that end in this bucket
                    lars = self._opt.lars_bucket_args(b, self._symm.block) if self._opt.lars else None
                    self._symm.fused_allreduce_sgd(self._symm_grad, self._opt, lo, hi, scale, lars=lars)
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
        if ev is not None:
            ev[2].record(self._comm_stream)
            self._tl_events.append((b, hi - lo, ev))

    def synchronize(self):
        """Launch whatever has not been launched (unused parameters), then wait for all buckets."""
        self._check_plan()
        while self._next < len(self.plan.buckets):        # unused parameters: their buckets never became ready
            self._launch(self._next)
            self._next += 1
        self._ready.clear()
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
                if self._opt.lars and not (self._symm is not None and self.world > 1):
                    # library fabrics and single process: norms and update after the (library) allreduce
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
        self._flush_timeline()
        if self.fused:
            self._opt.steps += 1
        else:
            self._opt.step()
        self._launched.clear()
        self._next = 0
        self._counts = [0] * len(self.params)
        self._direct_seen = [False] * len(self.params)
        self.tracker.begin_step()
        return loss

    def _flush_timeline(self):
        """Write the device-side times of this step's buckets: QUEUE = last gradient ready → collective starts,
        ALLREDUCE_UPDATE = the fused kernel (or the library allreduce) on the communication stream. Times come from
        CUDA events, placed on the host clock through one base event per step."""
        if not self._tl_events:
            return
        torch.cuda.synchronize()
        base_ev, base_us = self._tl_base
        for b, numel, (ready, start, done) in self._tl_events:
            t_ready = base_us + base_ev.elapsed_time(ready) * 1e3
            t_start = base_us + base_ev.elapsed_time(start) * 1e3
            t_done = base_us + base_ev.elapsed_time(done) * 1e3
            name = f"bucket{b} ({numel * 4 >> 10} KiB)"
            self.timeline.event(name, "QUEUE", 2, t_ready, max(0.0, t_start - t_ready))
            self.timeline.event(name, "ALLREDUCE_UPDATE", 3, t_start, max(0.0, t_done - t_start))
        self._tl_events, self._tl_base = [], None

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
