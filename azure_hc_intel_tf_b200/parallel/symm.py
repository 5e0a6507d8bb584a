"""Symmetric (peer-mapped) device memory and the NVLink collectives built on it.

Role in the reference stack: MPI_Init / communicator set-up and the UCX (`rc_x,sm,self`) or libfabric transport
selected by the launcher's fabric argument (benchmark-scripts/run-tf-sing-ucx-openmpi.sh:85-95) — here the
"transport" is direct load/store into every peer GPU's memory over NVLink 5 / NVSwitch.

Every rank allocates buffers of identical size, exports CUDA IPC handles, and all ranks exchange them once through
``torch.distributed`` (any backend). Kernels then receive the table of peer addresses (``b200_comm``) and read /
write the peers directly (csrc/comm/allreduce.cu, include/b200comm.h). ``torch.distributed`` is used for this
rendezvous only — never for a gradient.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from .. import _native

MAX_RANKS = 8


class CommStruct(C.Structure):
    _fields_ = [("buf", C.c_uint64 * MAX_RANKS), ("pad", C.c_uint64 * MAX_RANKS), ("rank", C.c_int32),
                ("world", C.c_int32), ("ctas", C.c_int32), ("one_shot_max_bytes", C.c_int32),
                ("mc", C.c_uint64), ("nvls_min_bytes", C.c_int64), ("algo", C.c_int32), ("reserved", C.c_int32)]


ALGO_AUTO, ALGO_ONE_SHOT, ALGO_TWO_SHOT, ALGO_NVLS = 0, 1, 2, 3
ALGOS = {"auto": ALGO_AUTO, "oneshot": ALGO_ONE_SHOT, "twoshot": ALGO_TWO_SHOT, "nvls": ALGO_NVLS}


class SgdArgs(C.Structure):
    _fields_ = [("master", C.c_void_p), ("mom", C.c_void_p), ("shadow", C.c_void_p), ("block_tensor", C.c_void_p),
                ("tensor_decay", C.c_void_p), ("tensor_lars", C.c_void_p), ("trust", C.c_void_p),
                ("lr", C.c_float), ("mu", C.c_float), ("wd", C.c_float), ("lr_dev", C.c_void_p),
                ("lars_block_partials", C.c_void_p), ("tensor_blk_lo", C.c_void_p), ("tensor_blk_cnt", C.c_void_p),
                ("lars_t_lo", C.c_int32), ("lars_t_hi", C.c_int32), ("lars_upd_lo", C.c_int64),
                ("lars_upd_hi", C.c_int64), ("lars_eta", C.c_float), ("lars_eps", C.c_float)]


class _DevArray:
    """Exposes a raw device allocation to torch through __cuda_array_interface__."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}


def _check(rc, what):
    if rc != 0:
        raise RuntimeError(f"{what} failed ({rc}): {_native.cuda().b200_last_error().decode()}")


def _declare(lib):
    if getattr(lib, "_b200comm_declared", False):
        return
    vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
    lib.b200_symm_alloc.argtypes, lib.b200_symm_alloc.restype = [i64, C.POINTER(vp)], i32
    lib.b200_symm_free.argtypes, lib.b200_symm_free.restype = [vp], i32
    lib.b200_symm_get_handle.argtypes, lib.b200_symm_get_handle.restype = [vp, vp], i32
    lib.b200_symm_open_handle.argtypes, lib.b200_symm_open_handle.restype = [vp, C.POINTER(vp)], i32
    lib.b200_symm_close_handle.argtypes, lib.b200_symm_close_handle.restype = [vp], i32
    lib.b200_comm_pad_bytes.argtypes, lib.b200_comm_pad_bytes.restype = [], i64
    lib.b200_vmm_supported.argtypes, lib.b200_vmm_supported.restype = [i32, C.POINTER(i32)], i32
    lib.b200_vmm_create.argtypes = [i64, i32, i32, i32, C.POINTER(vp), C.POINTER(i32), C.POINTER(i64)]
    lib.b200_vmm_create.restype = i32
    lib.b200_vmm_map_peer.argtypes, lib.b200_vmm_map_peer.restype = [vp, i32, i32], i32
    lib.b200_vmm_ptr.argtypes, lib.b200_vmm_ptr.restype = [vp, i32], C.c_uint64
    lib.b200_vmm_mc_create.argtypes, lib.b200_vmm_mc_create.restype = [vp, C.POINTER(i32)], i32
    lib.b200_vmm_mc_import.argtypes, lib.b200_vmm_mc_import.restype = [vp, i32], i32
    lib.b200_vmm_mc_add_device.argtypes, lib.b200_vmm_mc_add_device.restype = [vp], i32
    lib.b200_vmm_mc_bind.argtypes, lib.b200_vmm_mc_bind.restype = [vp], i32
    lib.b200_vmm_mc_ptr.argtypes, lib.b200_vmm_mc_ptr.restype = [vp], C.c_uint64
    lib.b200_vmm_free.argtypes, lib.b200_vmm_free.restype = [vp], i32
    lib.b200_fd_listen.argtypes, lib.b200_fd_listen.restype = [C.c_char_p], i32
    lib.b200_fd_send.argtypes, lib.b200_fd_send.restype = [i32, C.c_char_p, i32, i32], i32
    lib.b200_fd_recv.argtypes, lib.b200_fd_recv.restype = [i32, C.POINTER(i32)], i32
    lib.b200_fd_close.argtypes, lib.b200_fd_close.restype = [i32], i32
    lib.b200_comm_block_elems.argtypes, lib.b200_comm_block_elems.restype = [], i32
    lib.b200_allgather_f32.argtypes, lib.b200_allgather_f32.restype = [C.POINTER(CommStruct), i64, i64, vp], i32
    lib.b200_p2p_copy.argtypes, lib.b200_p2p_copy.restype = [C.POINTER(CommStruct), i32, i64, i64, i32, vp], i32
    lib.b200_p2p_pingpong.argtypes = [C.POINTER(CommStruct), i32, i32, i64, i64, i32, vp, vp]
    lib.b200_p2p_pingpong.restype = i32
    lib.b200_allreduce_f32.argtypes = [C.POINTER(CommStruct), i64, i64, C.c_float, C.POINTER(SgdArgs), vp]
    lib.b200_allreduce_f32.restype = i32
    lib.b200_allreduce_bf16.argtypes = [C.POINTER(CommStruct), i64, i64, C.c_float, vp]
    lib.b200_allreduce_bf16.restype = i32
    lib.b200_comm_barrier.argtypes, lib.b200_comm_barrier.restype = [C.POINTER(CommStruct), vp], i32
    lib.b200_broadcast_f32.argtypes = [C.POINTER(CommStruct), i64, i64, i32, vp]
    lib.b200_broadcast_f32.restype = i32
    lib._b200comm_declared = True


class SymmBuffer:
    """One symmetric allocation: ``local`` is this rank's tensor, ``peers[r]`` the device address of rank r's."""

    def __init__(self, ctx: "SymmetricContext", nbytes: int, dtype: torch.dtype):
        lib = ctx.lib
        self.ctx = ctx
        self.nbytes = nbytes
        ptr = C.c_void_p()
        _check(lib.b200_symm_alloc(nbytes, C.byref(ptr)), "symm_alloc")
        self.ptr = ptr.value
        handle = (C.c_ubyte * 64)()
        _check(lib.b200_symm_get_handle(C.c_void_p(self.ptr), handle), "symm_get_handle")
        mine = torch.tensor(list(handle), dtype=torch.uint8)
        gathered = ctx.all_gather_bytes(mine)
        self.peers: List[int] = []
        self._opened: List[int] = []
        for r in range(ctx.world):
            if r == ctx.rank:
                self.peers.append(self.ptr)
                continue
            h = (C.c_ubyte * 64)(*gathered[r].tolist())
            pp = C.c_void_p()
            _check(lib.b200_symm_open_handle(h, C.byref(pp)), f"symm_open_handle(rank {r})")
            self.peers.append(pp.value)
            self._opened.append(pp.value)
        self._keep = _DevArray(self.ptr, nbytes)
        self.local_bytes = torch.as_tensor(self._keep, device=ctx.device)
        self.local = self.local_bytes.view(dtype)

    def close(self):
        lib = self.ctx.lib
        for p in self._opened:
            lib.b200_symm_close_handle(C.c_void_p(p))
        self._opened = []
        if self.ptr:
            lib.b200_symm_free(C.c_void_p(self.ptr))
            self.ptr = 0


class VmmSymmBuffer:
    """Symmetric allocation on the CUDA virtual-memory-management API, optionally with a multicast (NVLS) mapping —
    this framework's own allocator (csrc/comm/symm_alloc.cpp): cuMemCreate per rank, POSIX file descriptors passed
    between the processes over abstract Unix-domain sockets, every peer mapped locally, and one multicast object that
    all ranks bind their allocation to. ``torch.distributed`` only provides the barriers between the steps."""

    def __init__(self, ctx: "SymmetricContext", nbytes: int, dtype: torch.dtype, multicast: bool):
        lib, world, rank = ctx.lib, ctx.world, ctx.rank
        self.ctx = ctx
        h, fd, padded = C.c_void_p(), C.c_int32(-1), C.c_int64(0)
        _check(lib.b200_vmm_create(nbytes, world, rank, int(multicast), C.byref(h), C.byref(fd), C.byref(padded)), "vmm_create")
        self._h = h
        self.nbytes = nbytes
        ctx.vmm_seq += 1
        name = lambda r: f"{ctx.sock_prefix}-{ctx.vmm_seq}-{r}".encode()
        sock = lib.b200_fd_listen(name(rank))
        if sock < 0:
            raise RuntimeError(f"fd_listen failed: {lib.b200_last_error().decode()}")
        try:
            dist.barrier()                                   # every rank's socket is bound
            for r in range(world):
                if r != rank:
                    _check(lib.b200_fd_send(sock, name(r), fd.value, rank), f"fd_send(rank {r})")
            for _ in range(world - 1):
                tag = C.c_int32(-1)
                pfd = lib.b200_fd_recv(sock, C.byref(tag))
                if pfd < 0:
                    raise RuntimeError(f"fd_recv failed: {lib.b200_last_error().decode()}")
                _check(lib.b200_vmm_map_peer(h, tag.value, pfd), f"vmm_map_peer(rank {tag.value})")
                lib.b200_fd_close(pfd)
            lib.b200_fd_close(fd.value)
            dist.barrier()                                   # all peer descriptors consumed
            self.mc_ptr = 0
            if multicast:
                if rank == 0:
                    mfd = C.c_int32(-1)
                    _check(lib.b200_vmm_mc_create(h, C.byref(mfd)), "vmm_mc_create")
                    for r in range(1, world):
                        _check(lib.b200_fd_send(sock, name(r), mfd.value, 1000), f"fd_send(multicast, rank {r})")
                    lib.b200_fd_close(mfd.value)
                else:
                    tag = C.c_int32(-1)
                    mfd = lib.b200_fd_recv(sock, C.byref(tag))
                    if mfd < 0 or tag.value != 1000:
                        raise RuntimeError(f"multicast descriptor not received: {lib.b200_last_error().decode()}")
                    _check(lib.b200_vmm_mc_import(h, mfd), "vmm_mc_import")
                    lib.b200_fd_close(mfd)
                _check(lib.b200_vmm_mc_add_device(h), "vmm_mc_add_device")
                dist.barrier()                               # every device is in the group before anyone binds
                _check(lib.b200_vmm_mc_bind(h), "vmm_mc_bind")
                dist.barrier()
                self.mc_ptr = int(lib.b200_vmm_mc_ptr(h))
        finally:
            lib.b200_fd_close(sock)
        self.peers = [int(lib.b200_vmm_ptr(h, r)) for r in range(world)]
        self.ptr = self.peers[rank]
        self._keep = _DevArray(self.ptr, nbytes)
        self.local_bytes = torch.as_tensor(self._keep, device=ctx.device)
        self.local = self.local_bytes.view(dtype)

    def close(self):
        if self._h is not None:
            self.ctx.lib.b200_vmm_free(self._h)
            self._h = None


class SymmetricContext:
    """Per-process handle on the NVLink fabric: the signal pad, the symmetric buffers and the collectives.

    Every kernel that uses the signal pad shares its per-CTA epoch slots, so two of them must never be in flight on
    one device at the same time. All collectives are therefore issued on ONE stream, ``self.stream`` (high priority);
    a call made from any other stream forks onto it and joins back (``_serial``), and ``DistributedOptimizer`` uses
    this same stream as its side stream. The fork/join pattern is also what a CUDA-graph capture needs."""

    def __init__(self, rank: int, world: int, device: torch.device):
        if world > MAX_RANKS:
            raise ValueError(f"the NVLink fabric supports up to {MAX_RANKS} ranks on one box, got {world}")
        self.rank, self.world, self.device = rank, world, device
        self.lib = _native.cuda()
        _declare(self.lib)
        self.block = int(self.lib.b200_comm_block_elems())
        # upper bound on CTAs per collective (128 threads each); small messages use fewer (one per four blocks)
        self.ctas = int(os.environ.get("B200_COMM_CTAS", "96"))
        # messages up to this many bytes use the one-shot kernel (one NVLink traversal); 0 disables it
        # (thresholds from the per-algorithm sweeps in profiles/osu/: at 8 GPUs one-shot and two-shot tie up to 256 KiB
        # and the in-switch reduction wins from 64 KiB on; at 2 GPUs everything up to 1 MiB is flat at the barrier cost)
        self.one_shot_max_bytes = int(os.environ.get("B200_COMM_ONE_SHOT_MAX", str(32 << 10)))
        # messages of at least this many bytes use the NVLS (in-switch reduction) kernel when the buffer has a
        # multicast mapping; -1 disables NVLS. With two ranks a direct peer read is the shorter path.
        default_nvls = (64 << 10) if world > 2 else -1
        self.nvls_min_bytes = int(os.environ.get("B200_COMM_NVLS_MIN", str(default_nvls)))
        self.algo = ALGOS[os.environ.get("B200_COMM_ALGO", "auto")]
        self.nvls_ok = None      # decided collectively at the first multicast allocation
        self.vmm_seq = 0
        pids = self.all_gather_bytes(torch.tensor([os.getpid() >> 24 & 255, os.getpid() >> 16 & 255, os.getpid() >> 8 & 255,
                                                   os.getpid() & 255], dtype=torch.uint8))
        pid0 = int.from_bytes(bytes(pids[0].tolist()), "big")
        self.sock_prefix = f"b200vmm-{os.environ.get('MASTER_PORT', '0')}-{pid0}"
        self.stream = torch.cuda.Stream(device=device, priority=-1)
        self.buffers: List[SymmBuffer] = []
        self.pad = SymmBuffer(self, int(self.lib.b200_comm_pad_bytes()), torch.uint8)
        self._scratch: Optional[SymmBuffer] = None
        self.launches = 0
        dist.barrier()

    @staticmethod
    def create(rank, world, device) -> "SymmetricContext":
        return SymmetricContext(rank, world, device)

    # -- rendezvous helper (works with nccl and gloo process groups)
    def all_gather_bytes(self, t: torch.Tensor) -> List[torch.Tensor]:
        backend = dist.get_backend()
        src = t.to(self.device) if backend == "nccl" else t
        outs = [torch.empty_like(src) for _ in range(self.world)]
        dist.all_gather(outs, src)
        return [o.cpu() for o in outs]

    def alloc(self, nbytes: int, dtype: torch.dtype, multicast: bool = False):
        """Collective. ``multicast=True`` asks for a buffer with an NVLS multicast mapping and falls back to a plain
        IPC buffer (on every rank alike) if the platform refuses."""
        if multicast and self.nvls_ok is not False and os.environ.get("B200_COMM_MULTICAST", "1") != "0":
            if self.nvls_ok is None:
                # one collective decision, up front: VMM + POSIX descriptors + multicast on every rank's device
                mc = C.c_int32(0)
                vmm = self.lib.b200_vmm_supported(self.device.index or 0, C.byref(mc))
                flag = torch.tensor([1 if (vmm == 1 and mc.value) else 0], device=self.device, dtype=torch.int32)
                dist.all_reduce(flag, op=dist.ReduceOp.MIN)
                self.nvls_ok = bool(flag.item())
                if not self.nvls_ok and self.rank == 0:
                    print("[b200 comm] NVLS multicast unavailable on this platform; using the P2P two-shot kernels")
            if self.nvls_ok:
                b = VmmSymmBuffer(self, nbytes, dtype, multicast=True)
                self.buffers.append(b)
                dist.barrier()
                return b
        b = SymmBuffer(self, nbytes, dtype)
        self.buffers.append(b)
        dist.barrier()           # every rank has mapped every peer before anyone launches a kernel on it
        return b

    def alloc_like_flat(self, numel: int, dtype: torch.dtype, multicast: bool = False):
        return self.alloc(numel * torch.empty((), dtype=dtype).element_size(), dtype, multicast)

    def comm_struct(self, buf: SymmBuffer, algo: Optional[int] = None) -> CommStruct:
        cs = CommStruct()
        for r in range(self.world):
            cs.buf[r] = buf.peers[r]
            cs.pad[r] = self.pad.peers[r]
        cs.rank, cs.world, cs.ctas, cs.one_shot_max_bytes = self.rank, self.world, self.ctas, self.one_shot_max_bytes
        cs.mc = getattr(buf, "mc_ptr", 0) or 0
        cs.nvls_min_bytes = self.nvls_min_bytes
        cs.algo = self.algo if algo is None else algo
        return cs

    @staticmethod
    def _stream():
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)

    @contextlib.contextmanager
    def _serial(self):
        """Run the body on the communication stream, ordered after the caller's stream, and make the caller's
        stream wait for it (no-op when the caller already is on the communication stream)."""
        cur = torch.cuda.current_stream()
        if cur.cuda_stream == self.stream.cuda_stream:
            yield
            return
        self.stream.wait_stream(cur)
        with torch.cuda.stream(self.stream):
            yield
        cur.wait_stream(self.stream)

    # -- collectives on symmetric buffers
    def allreduce_range_(self, buf: SymmBuffer, lo: int, hi: int, scale: float = 1.0, algo: Optional[int] = None):
        """In-place sum (times scale) of elements [lo, hi) of ``buf`` across ranks."""
        cs = self.comm_struct(buf, algo)
        with self._serial():
            if buf.local.dtype == torch.float32:
                _check(self.lib.b200_allreduce_f32(C.byref(cs), lo, hi - lo, scale, None, self._stream()), "allreduce_f32")
            elif buf.local.dtype == torch.bfloat16:
                _check(self.lib.b200_allreduce_bf16(C.byref(cs), lo, hi - lo, scale, self._stream()), "allreduce_bf16")
            else:
                raise TypeError(buf.local.dtype)
        self._count()

    def fused_allreduce_sgd(self, buf: SymmBuffer, opt, lo: int, hi: int, scale: float, lars=None):
        """One launch: allreduce [lo, hi) of the symmetric gradient buffer, 1/N scale, momentum-SGD update of
        ``opt``'s local master / momentum / bf16 weights. ``lars`` = (t_lo, t_hi, upd_lo_blk, upd_hi_blk) switches
        on the fused LARS half for the tensors that end in this bucket (see include/b200comm.h)."""
        g = opt.param_groups[0]
        a = SgdArgs()
        a.master, a.mom = opt.master.data_ptr(), opt.mom.data_ptr()
        a.shadow = opt.shadow.data_ptr() if opt.shadow is not None else None
        a.block_tensor, a.tensor_decay = opt.block_tensor.data_ptr(), opt.tensor_decay.data_ptr()
        a.tensor_lars = opt.tensor_lars.data_ptr()
        a.trust = opt.trust.data_ptr() if opt.lars else None
        a.lr, a.mu, a.wd = g["lr"], g["momentum"], g["weight_decay"]
        a.lr_dev = opt.lr_dev.data_ptr() if getattr(opt, "lr_dev", None) is not None else None
        if lars is not None:
            part, blk_lo, blk_cnt = opt.lars_tables(self.block)
            a.lars_block_partials, a.tensor_blk_lo, a.tensor_blk_cnt = part.data_ptr(), blk_lo.data_ptr(), blk_cnt.data_ptr()
            a.lars_t_lo, a.lars_t_hi, a.lars_upd_lo, a.lars_upd_hi = lars
            a.lars_eta, a.lars_eps = opt.lars_eta, opt.lars_eps
        cs = self.comm_struct(buf)
        with self._serial():
            _check(self.lib.b200_allreduce_f32(C.byref(cs), lo, hi - lo, scale, C.byref(a), self._stream()),
                   "fused allreduce+sgd")
        self._count()

    def barrier(self):
        """Device-side barrier across ranks on the communication stream (no host synchronisation)."""
        cs = self.comm_struct(self.pad)
        with self._serial():
            _check(self.lib.b200_comm_barrier(C.byref(cs), self._stream()), "comm_barrier")
        self._count()

    def host_barrier(self):
        """hvd.barrier(): every rank has reached this point when it returns (flag kernel + stream wait)."""
        self.barrier()
        torch.cuda.current_stream().synchronize()

    def broadcast_(self, buf: SymmBuffer, lo: int, hi: int, root: int):
        cs = self.comm_struct(buf)
        with self._serial():
            _check(self.lib.b200_broadcast_f32(C.byref(cs), lo, hi - lo, root, self._stream()), "broadcast_f32")
        self._count()

    def allgather_(self, buf: SymmBuffer, lo: int, n_per_rank: int):
        """Rank r's fp32 segment [lo + r*n, +n) of ``buf`` is replicated into every rank's ``buf``."""
        cs = self.comm_struct(buf)
        with self._serial():
            _check(self.lib.b200_allgather_f32(C.byref(cs), lo, n_per_rank, self._stream()), "allgather_f32")
        self._count()

    def p2p_copy(self, buf: SymmBuffer, peer: int, lo: int, n: int, pull: bool = False):
        """Bandwidth probe: move n floats to (or, ``pull``, from) one peer. Not serialised: it uses no flags."""
        cs = self.comm_struct(buf)
        _check(self.lib.b200_p2p_copy(C.byref(cs), peer, lo, n, int(pull), self._stream()), "p2p_copy")
        self._count()

    def p2p_pingpong(self, buf: SymmBuffer, a: int, b: int, n: int, iters: int, out_us: torch.Tensor):
        cs = self.comm_struct(buf)
        with self._serial():
            _check(self.lib.b200_p2p_pingpong(C.byref(cs), a, b, 0, n, iters, C.c_void_p(out_us.data_ptr()),
                                              self._stream()), "p2p_pingpong")
        self._count()

    def _staging(self, need_bytes: int) -> SymmBuffer:
        if self._scratch is None or self._scratch.nbytes < need_bytes:
            # sized collectively: every rank reaches this branch for the same call
            self._scratch = self.alloc(max(need_bytes, 64 << 20), torch.uint8, multicast=True)
        return self._scratch

    def allreduce_(self, tensor: torch.Tensor, scale: float = 1.0):
        """Allreduce an arbitrary CUDA tensor (fp32 / bf16) by staging it through a symmetric scratch buffer.
        Convenience path for hvd.allreduce(); the gradient path never stages."""
        flat = tensor.reshape(-1)
        esz = flat.element_size()
        gran = 1024 if flat.dtype == torch.float32 else 2048
        n = (flat.numel() + gran - 1) // gran * gran
        sc = self._staging(n * 4)
        view = sc.local_bytes[: n * esz].view(flat.dtype)
        view[: flat.numel()].copy_(flat)
        if n > flat.numel():
            view[flat.numel():].zero_()
        cs = self.comm_struct(sc)
        with self._serial():
            if flat.dtype == torch.float32:
                _check(self.lib.b200_allreduce_f32(C.byref(cs), 0, n, scale, None, self._stream()), "allreduce_f32")
            else:
                _check(self.lib.b200_allreduce_bf16(C.byref(cs), 0, n, scale, self._stream()), "allreduce_bf16")
        self._count()
        flat.copy_(view[: flat.numel()])
        return tensor

    def broadcast_tensor_(self, tensor: torch.Tensor, root: int):
        """Broadcast an ordinary CUDA tensor (any dtype, viewed as bytes) from ``root`` by staging it through
        symmetric scratch: the root writes every peer's copy with the broadcast kernel (Horovod's
        broadcast_global_variables role)."""
        flat = tensor.reshape(-1).view(torch.uint8) if tensor.is_contiguous() else tensor.contiguous().reshape(-1).view(torch.uint8)
        nb = flat.numel()
        n = (nb + 15) // 16 * 16
        sc = self._staging(n)
        view = sc.local_bytes[:n]
        if self.rank == root:
            view[:nb].copy_(flat)
        self.broadcast_(sc, 0, n // 4, root)            # offsets are in fp32 elements of the scratch allocation
        if tensor.is_contiguous():
            flat.copy_(view[:nb])
        else:
            tensor.copy_(view[:nb].view(tensor.dtype).view(tensor.shape))
        return tensor

    def allgather_tensor(self, tensor: torch.Tensor) -> List[torch.Tensor]:
        """All-gather equally sized CUDA tensors through symmetric scratch; returns one tensor per rank."""
        flat = tensor.contiguous().reshape(-1).view(torch.uint8)
        nb = flat.numel()
        seg = (nb + 15) // 16 * 16
        sc = self._staging(seg * self.world)
        sc.local_bytes[self.rank * seg: self.rank * seg + nb].copy_(flat)
        self.allgather_(sc, 0, seg // 4)
        return [sc.local_bytes[r * seg: r * seg + nb].clone().view(tensor.dtype).view(tensor.shape)
                for r in range(self.world)]

    def _count(self):
        self.launches += 1
        try:
            from ..ops import b200
            b200.count_launch(1)
        except Exception:
            pass

    def close(self):
        try:
            torch.cuda.synchronize()
        except Exception:
            pass
        for b in self.buffers:
            b.close()
        self.buffers = []
        self.pad.close()
