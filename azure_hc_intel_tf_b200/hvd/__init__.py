"""``hvd`` — Horovod-style data-parallel API (torch flavour), B200-native underneath.

The reference selects Horovod with ``--variable_update=horovod --horovod_device=cpu`` and tunes it with
``HOROVOD_FUSION_THRESHOLD=134217728`` (benchmark-scripts/run-tf-sing-ucx-openmpi.sh:77-78,105). This
module keeps that API (``init/size/rank/local_rank/allreduce/broadcast_parameters/DistributedOptimizer``…)
and replaces the machinery:

* process group / rendezvous: ``torch.distributed`` (one process per GPU), not MPI;
* fabric ``nvl``: hand-written sm_100a kernels that read and write peer GPUs' symmetric buffers over
  NVLink/NVSwitch (parallel/symm.py + csrc/comm), fused with the optimizer update;
* fabric ``nccl``: stock ``ncclAllReduce`` — the measured baseline only;
* fabric ``gloo``: CPU plumbing path (the reference's ``sock`` fabric, run-tf-sing-ucx-openmpi.sh:94).
"""
from __future__ import annotations

import datetime
import os
from typing import Optional

import torch
import torch.distributed as dist

from .compression import Compression  # noqa: F401

Sum = "sum"
Average = "average"

_state = {"init": False, "fabric": None, "owns_pg": False, "device": None, "symm": None}


def _env_int(name, default):
    v = os.environ.get(name)
    return int(v) if v not in (None, "") else default


def init(fabric: str = "auto", device: Optional[str] = None, timeout_s: int = 600) -> None:
    """Initialise the data-parallel runtime. Reads RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR /
    MASTER_PORT (torchrun or launch/launcher.py set them). With WORLD_SIZE unset it runs single-process."""
    if _state["init"]:
        return
    world = _env_int("WORLD_SIZE", 1)
    rank = _env_int("RANK", 0)
    lrank = _env_int("LOCAL_RANK", rank)
    use_cuda = torch.cuda.is_available() and device != "cpu" and fabric not in ("gloo", "sock")
    if fabric == "auto":
        fabric = "nvl" if use_cuda else "gloo"
    if fabric in ("ib",):
        fabric = "nvl"
    if fabric == "sock":
        fabric = "gloo"
    if fabric in ("nvl", "nccl") and not use_cuda:
        raise RuntimeError(f"fabric {fabric} needs a CUDA device")
    if use_cuda:
        torch.cuda.set_device(lrank % torch.cuda.device_count())
        _state["device"] = torch.device("cuda", torch.cuda.current_device())
    else:
        _state["device"] = torch.device("cpu")
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        backend = "gloo" if fabric == "gloo" else "nccl"
        kw = {}
        if backend == "nccl":
            kw["device_id"] = _state["device"]
        dist.init_process_group(backend, rank=rank, world_size=world,
                                timeout=datetime.timedelta(seconds=timeout_s), **kw)
        _state["owns_pg"] = True
    elif world > 1 and fabric in ("nvl", "nccl") and dist.get_backend() != "nccl":
        raise RuntimeError(f"fabric {fabric} needs the existing process group to use the nccl backend")
    _state.update(init=True, fabric=fabric, rank=rank, world=world, local_rank=lrank,
                  local_size=_env_int("LOCAL_WORLD_SIZE", world))
    if fabric == "nvl" and world > 1:
        from ..parallel import symm
        _state["symm"] = symm.SymmetricContext.create(rank, world, _state["device"])


def shutdown(keep_process_group: bool = False) -> None:
    """Tear the runtime down. ``keep_process_group=True`` keeps the torch.distributed group alive so that a following
    ``init(fabric=...)`` in the same process can switch fabrics (bench.py measures the stock arm and the product arm
    back to back in one launch)."""
    if not _state["init"]:
        return
    if _state.get("symm") is not None:
        _state["symm"].close()
        _state["symm"] = None
    if _state["owns_pg"] and dist.is_initialized() and not keep_process_group:
        dist.destroy_process_group()
        _state["owns_pg"] = False
    _state.update(init=False, fabric=None)


def is_initialized() -> bool:
    return _state["init"]


def _need():
    if not _state["init"]:
        raise ValueError("hvd.init() has not been called")


def size() -> int:
    _need()
    return _state["world"]


def rank() -> int:
    _need()
    return _state["rank"]


def local_rank() -> int:
    _need()
    return _state["local_rank"]


def local_size() -> int:
    _need()
    return _state["local_size"]


def fabric() -> str:
    _need()
    return _state["fabric"]


def device() -> torch.device:
    _need()
    return _state["device"]


def symmetric_context():
    """The NVLink symmetric-memory context (fabric nvl, world > 1), else None."""
    return _state.get("symm")


def mpi_enabled() -> bool:
    return False


def nccl_built() -> bool:
    return torch.cuda.is_available() and dist.is_nccl_available()


def gloo_enabled() -> bool:
    return dist.is_gloo_available()


# ------------------------------------------------------------------------------------ collectives

class _Handle:
    def __init__(self, work, tensor, post=None):
        self.work, self.tensor, self.post = work, tensor, post
        self.done = work is None

    def wait(self):
        if not self.done:
            self.work.wait()
            self.done = True
        if self.post is not None:
            self.post(self.tensor)
            self.post = None
        return self.tensor


def allreduce_async_(tensor: torch.Tensor, average: Optional[bool] = None, name: Optional[str] = None,
                     op: Optional[str] = None) -> _Handle:
    """In-place asynchronous allreduce. ``op`` defaults to Average (Horovod's torch default)."""
    _need()
    if op is None:
        op = Average if (average is None or average) else Sum
    n = size()
    if n == 1:
        return _Handle(None, tensor)
    ctx = _state.get("symm")
    if ctx is not None and tensor.is_cuda and tensor.dtype in (torch.float32, torch.bfloat16):
        ctx.allreduce_(tensor, scale=(1.0 / n) if op == Average else 1.0)
        return _Handle(None, tensor)
    work = dist.all_reduce(tensor, op=dist.ReduceOp.SUM, async_op=True)
    post = (lambda t: t.div_(n)) if op == Average else None
    return _Handle(work, tensor, post)


def allreduce_(tensor, average=None, name=None, op=None):
    return synchronize(allreduce_async_(tensor, average, name, op))


def allreduce(tensor, average=None, name=None, op=None, compression=Compression.none):
    t, ctx = compression.compress(tensor.clone())
    out = allreduce_(t, average, name, op)
    return compression.decompress(out, ctx)


def allreduce_async(tensor, average=None, name=None, op=None):
    return allreduce_async_(tensor.clone(), average, name, op)


def allgather(tensor: torch.Tensor, name: Optional[str] = None) -> torch.Tensor:
    """Concatenate along dim 0; first dimensions may differ between ranks."""
    _need()
    n = size()
    if n == 1:
        return tensor.clone()
    ctx = _state.get("symm") if tensor.is_cuda else None
    mine = torch.tensor([tensor.shape[0]], dtype=torch.int64, device=tensor.device)
    if ctx is not None:
        sizes = ctx.allgather_tensor(mine)              # own all-gather kernel over NVLink
    else:
        sizes = [torch.zeros(1, dtype=torch.int64, device=tensor.device) for _ in range(n)]
        dist.all_gather(sizes, mine)
    sizes = [int(s.item()) for s in sizes]
    mx = max(sizes)
    pad = tensor.new_zeros((mx,) + tuple(tensor.shape[1:]))
    pad[: tensor.shape[0]] = tensor
    if ctx is not None:
        outs = ctx.allgather_tensor(pad)
    else:
        outs = [torch.empty_like(pad) for _ in range(n)]
        dist.all_gather(outs, pad)
    return torch.cat([o[:s] for o, s in zip(outs, sizes)], dim=0)


def broadcast_(tensor: torch.Tensor, root_rank: int = 0, name: Optional[str] = None) -> torch.Tensor:
    _need()
    if size() > 1:
        ctx = _state.get("symm")
        if ctx is not None and tensor.is_cuda:
            ctx.broadcast_tensor_(tensor, root_rank)     # own NVLink broadcast kernel (K-BC)
        else:
            dist.broadcast(tensor, src=root_rank)
    return tensor


def broadcast(tensor, root_rank=0, name=None):
    return broadcast_(tensor.clone(), root_rank, name)


def synchronize(handle: _Handle):
    return handle.wait()


def poll(handle: _Handle) -> bool:
    if handle.done or handle.work is None:
        return True
    return handle.work.is_completed()


def barrier() -> None:
    _need()
    if size() > 1:
        ctx = _state.get("symm")
        if ctx is not None:
            ctx.host_barrier()                           # flag barrier kernel over NVLink + stream wait
        else:
            dist.barrier()


def broadcast_parameters(params, root_rank: int = 0) -> None:
    """Broadcast a ``state_dict()`` or ``named_parameters()`` from root (Horovod's
    ``broadcast_global_variables(0)``, K14 in SURVEY.md §2.6). Tensors are packed into one flat buffer
    per dtype so the wire sees a handful of large messages instead of 161 small ones."""
    _need()
    if isinstance(params, dict):
        items = list(params.items())
    else:
        items = list(params)
    tensors = [t for _, t in items if isinstance(t, torch.Tensor)]
    if size() == 1 or not tensors:
        return
    by_key = {}
    for t in tensors:
        by_key.setdefault((t.dtype, t.device), []).append(t)
    ctx = _state.get("symm")
    for (_dt, _dev), ts in by_key.items():
        flat = torch.cat([t.detach().reshape(-1) for t in ts])
        if ctx is not None and flat.is_cuda:
            ctx.broadcast_tensor_(flat, root_rank)       # own NVLink broadcast kernel (K-BC), no NCCL
        else:
            dist.broadcast(flat, src=root_rank)
        off = 0
        with torch.no_grad():
            for t in ts:
                n = t.numel()
                # copy through a logical view so channels_last parameters keep their layout
                t.copy_(flat[off:off + n].view(t.shape) if t.is_contiguous() else
                        flat[off:off + n].view(t.shape))
                off += n


def broadcast_global_variables(model, root_rank: int = 0, optimizer=None) -> None:
    """``horovod.tensorflow.broadcast_global_variables(root_rank)``: TensorFlow's "global variables" are the model's
    parameters, its non-trainable state (BN moving statistics) and the optimizer slots; here they are named explicitly."""
    broadcast_parameters(model.state_dict(), root_rank)
    if optimizer is not None:
        broadcast_optimizer_state(optimizer, root_rank)


class BroadcastGlobalVariablesHook:
    """``hvd.BroadcastGlobalVariablesHook(0)``, the session hook tf_cnn_benchmarks installs under
    ``--variable_update=horovod`` (run-tf-sing-ucx-openmpi.sh:77): call ``begin(model, optimizer)`` once after the
    variables exist and before the first step."""

    def __init__(self, root_rank: int = 0):
        self.root_rank = root_rank
        self.done = False

    def begin(self, model, optimizer=None) -> None:
        if not self.done:
            broadcast_global_variables(model, self.root_rank, optimizer)
            self.done = True


def broadcast_optimizer_state(optimizer, root_rank: int = 0) -> None:
    _need()
    if size() == 1:
        return
    opt = getattr(optimizer, "_opt", optimizer)
    if hasattr(opt, "broadcast_state"):
        opt.broadcast_state(root_rank)
        return
    state = opt.state_dict()
    tensors = {}
    for pid, st in state.get("state", {}).items():
        for k, v in st.items():
            if isinstance(v, torch.Tensor):
                tensors[f"{pid}.{k}"] = v
    broadcast_parameters(tensors, root_rank)
    state["param_groups"] = broadcast_object(state.get("param_groups"), root_rank)
    opt.load_state_dict(state)


def broadcast_object(obj, root_rank: int = 0):
    _need()
    if size() == 1:
        return obj
    ctx = _state.get("symm")
    if ctx is not None:
        import pickle
        dev = _state["device"]
        raw = pickle.dumps(obj) if rank() == root_rank else b""
        n = broadcast_(torch.tensor([len(raw)], dtype=torch.int64, device=dev), root_rank)
        buf = torch.empty(int(n.item()), dtype=torch.uint8, device=dev)
        if rank() == root_rank:
            buf.copy_(torch.frombuffer(bytearray(raw), dtype=torch.uint8))
        broadcast_(buf, root_rank)
        return pickle.loads(bytes(buf.cpu().numpy()))
    objs = [obj]
    dist.broadcast_object_list(objs, src=root_rank)
    return objs[0]


from .optimizer import DistributedOptimizer, MomentumSGD, LARS, RMSProp, Adam  # noqa: E402,F401
