"""Multi-process test harness: run a function on N local ranks (gloo on CPU, nccl/nvl on GPUs) — the
analogue of the reference's single-node ``--oversubscribe`` mode (run-tf-sing-ucx-openmpi.sh:100)."""
from __future__ import annotations

import os
import socket
import traceback

import torch
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _entry(rank, world, port, fabric, fn, args, ret):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), LOCAL_WORLD_SIZE=str(world),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(1)
    from .. import hvd
    try:
        hvd.init(fabric=fabric)
        out = fn(rank, world, *args)
        ret[rank] = ("ok", out)
    except Exception:
        ret[rank] = ("err", traceback.format_exc())
    finally:
        try:
            hvd.shutdown()
        except Exception:
            pass


def run_ranks(fn, world: int = 2, fabric: str = "gloo", args=(), timeout: float = 300.0):
    """Run ``fn(rank, world, *args)`` on ``world`` spawned processes after ``hvd.init(fabric)``.
    Returns the list of per-rank return values (must be picklable); raises if any rank failed."""
    ctx = mp.get_context("spawn")
    mgr = ctx.Manager()
    ret = mgr.dict()
    port = _free_port()
    procs = [ctx.Process(target=_entry, args=(r, world, port, fabric, fn, args, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout)
    for p in procs:
        if p.is_alive():
            p.kill()
            raise TimeoutError("a rank did not finish")
    outs = []
    for r in range(world):
        st, val = ret.get(r, ("err", "rank exited without a result"))
        if st != "ok":
            raise RuntimeError(f"rank {r} failed:\n{val}")
        outs.append(val)
    return outs
