"""Symmetric memory for one NVSwitch domain.

Every rank allocates the same set of device buffers (gradients, parameters, flag pad) straight
from ``cudaMalloc`` through the native library, exports them as CUDA IPC handles, and maps every
peer's buffers, so kernels dereference peer pointers directly over NVLink. ``torch.distributed``
is used only as the bootstrap channel for the 64-byte handles.

Buffers allocated with ``multicast=True`` additionally get an NVLS view (csrc/nvls.cu): they are
physical allocations of the virtual-memory-management API, shared between the ranks' processes as POSIX
file descriptors over unix sockets, bound to one multicast object, so ``multimem.ld_reduce`` /
``multimem.st`` on ``mc_ptr`` reduce / broadcast inside the NVSwitch.

The reference has no communication layer: its network is the three numbers
``--bandwidth``, ``--internode_latency`` (run_sim.py:59-69) and ``FACTOR`` (infra/interference.py).
"""
from __future__ import annotations

import ctypes as C
import os
import socket
import tempfile
import threading
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch

from ..ops import native


class _CudaArray:
    """Minimal ``__cuda_array_interface__`` holder so torch can view raw device memory."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None,
        }


def tensor_from_ptr(ptr: int, nbytes: int, dtype: torch.dtype, device: torch.device) -> torch.Tensor:
    raw = torch.as_tensor(_CudaArray(ptr, nbytes), device=device)
    return raw.view(dtype)


def exchange_fds(rank: int, world: int, mine: List[int], group=None) -> Dict[int, List[int]]:
    """All-to-all hand-over of open file descriptors between the ranks of one host: every rank offers
    ``mine`` (the same list to every peer) and receives each peer's list. Descriptors cannot travel
    through torch.distributed, so each rank listens on a unix socket (paths are all-gathered) and sends
    them as SCM_RIGHTS ancillary data. Returns {peer rank: [fds]} — the caller closes them."""
    import torch.distributed as dist
    path = os.path.join(tempfile.gettempdir(), f"tb200-fd-{os.getpid()}-{rank}-{id(mine) & 0xffff:x}.sock")
    if os.path.exists(path):
        os.unlink(path)
    server = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    server.bind(path)
    server.listen(world)
    paths: List[Optional[str]] = [None] * world
    dist.all_gather_object(paths, path, group=group)          # also: everyone is listening from here on
    errors: List[BaseException] = []

    def serve() -> None:
        try:
            for _ in range(world - 1):
                conn, _addr = server.accept()
                with conn:
                    conn.recv(1)
                    socket.send_fds(conn, [b"f"], list(mine))
        except BaseException as e:                           # noqa: BLE001 - reported by the caller
            errors.append(e)

    t = threading.Thread(target=serve, daemon=True)
    t.start()
    got: Dict[int, List[int]] = {}
    try:
        for r in range(world):
            if r == rank:
                continue
            with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as c:
                c.connect(paths[r])
                c.sendall(b"?")
                _msg, fds, _flags, _addr = socket.recv_fds(c, 16, 8)
                got[r] = list(fds)
        t.join(timeout=60)
    finally:
        server.close()
        if os.path.exists(path):
            os.unlink(path)
    if errors:
        raise RuntimeError(f"file-descriptor exchange failed: {errors[0]!r}")
    return got


@dataclass
class SymmBuffer:
    name: str
    nbytes: int
    ptrs: List[int]                 # indexed by rank; ptrs[rank] is the local allocation
    local: torch.Tensor             # uint8 view of the local allocation
    opened: List[int] = field(default_factory=list)   # IPC-mapped pointers to close
    # VMM / NVLS buffers (multicast=True): allocation handles per rank, the multicast object and its view
    vmm_bytes: int = 0
    vmm_handles: Dict[int, int] = field(default_factory=dict)
    mc_handle: int = 0
    mc_ptr: int = 0

    def view(self, dtype: torch.dtype) -> torch.Tensor:
        return self.local.view(dtype)


class SymmetricWorld:
    """The set of ranks (1..8 GPUs of one box) that one job's workers / PS shards span."""

    def __init__(self, rank: int, world: int, device: torch.device, group=None):
        if not 1 <= world <= native.MAX_WORLD:
            raise ValueError(f"world size {world} outside 1..{native.MAX_WORLD}")
        self.rank, self.world, self.device, self.group = rank, world, device, group
        self.buffers: Dict[str, SymmBuffer] = {}
        self._emulated_peers: Optional[List["SymmetricWorld"]] = None

    # -- construction ---------------------------------------------------------------------------
    @classmethod
    def from_env(cls, device: Optional[torch.device] = None, group=None) -> "SymmetricWorld":
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            rank, world = dist.get_rank(group), dist.get_world_size(group)
        else:
            rank, world = 0, 1
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device())
        return cls(rank, world, device, group)

    @classmethod
    def emulated(cls, world: int, device: torch.device) -> List["SymmetricWorld"]:
        """W ranks that all live on one device (single-GPU tests of the multi-rank data path).

        Kernels of different emulated ranks must never wait on each other (they would run as
        separate launches on one GPU); callers pre-post the flags with :meth:`post_flags`.
        """
        ranks = [cls(r, world, device) for r in range(world)]
        for r in ranks:
            r._emulated_peers = ranks
        return ranks

    # -- allocation -----------------------------------------------------------------------------
    def alloc(self, name: str, nbytes: int, multicast: bool = False) -> SymmBuffer:
        """Collective: every rank must call with the same name and size, in the same order.
        ``multicast=True`` asks for an NVLS view (``mc_ptr``) where the box supports it; the buffer is
        then a VMM allocation shared by file descriptor instead of a cudaMalloc block shared by IPC."""
        if name in self.buffers:
            raise KeyError(f"symmetric buffer {name!r} already exists")
        if multicast and self.world > 1 and self._emulated_peers is None and self.nvls_available():
            return self._alloc_vmm(name, nbytes)
        lib = native.lib()
        with torch.cuda.device(self.device):
            out = C.c_void_p()
            native.check(lib.tb200_symm_alloc(nbytes, C.byref(out)))
        local_ptr = int(out.value)
        local = tensor_from_ptr(local_ptr, nbytes, torch.uint8, self.device)
        buf = SymmBuffer(name, nbytes, [0] * self.world, local)
        buf.ptrs[self.rank] = local_ptr
        self.buffers[name] = buf
        if self._emulated_peers is not None:
            # pointers are resolved lazily once every emulated rank has allocated `name`
            peers = self._emulated_peers
            if all(name in p.buffers for p in peers):
                for p in peers:
                    for q in peers:
                        p.buffers[name].ptrs[q.rank] = q.buffers[name].ptrs[q.rank]
        elif self.world > 1:
            self._exchange(buf)
        return buf

    # -- NVLS-capable buffers -------------------------------------------------------------------------
    def nvls_available(self) -> bool:
        """Collective: do all ranks' GPUs support multicast objects (and is it not switched off)?"""
        if os.environ.get("TB200_NVLS", "auto") == "0" or self.world < 2 or self._emulated_peers is not None:
            return False
        if not hasattr(self, "_nvls_ok"):
            import torch.distributed as dist
            mine = bool(native.nvls_supported(self.device.index or 0))
            votes: List[Optional[bool]] = [None] * self.world
            dist.all_gather_object(votes, mine, group=self.group)
            self._nvls_ok = all(votes)
        return self._nvls_ok

    def nvls_preferred(self) -> bool:
        """Collective: should a trainer on this world default to the NVLS kernels? From 4 ranks up, where the
        in-switch reduction saves (W-1)/W of the reads a rank would otherwise pull over NVLink (8 GPUs, 528 MB
        bucket: 1.18 ms against 1.65 ms through peer loads, profiles/COMM_r1.md); at 2 GPUs the switch has
        nothing to save and the peer load/store kernels measure faster. TB200_NVLS=1 / 0 forces it on / off."""
        env = os.environ.get("TB200_NVLS", "auto")
        if env == "1":
            return self.nvls_available()
        return env != "0" and self.world >= 4 and self.nvls_available()

    def _alloc_vmm(self, name: str, nbytes: int) -> SymmBuffer:
        import torch.distributed as dist
        lib = native.lib()
        u64, vp, ci = C.c_ulonglong, C.c_void_p, C.c_int
        with torch.cuda.device(self.device):
            rounded = C.c_size_t()
            native.check(lib.tb200_vmm_round(self.world, nbytes, C.byref(rounded)))
            size = int(rounded.value)
            phys, ptr, fd = u64(), vp(), ci(-1)
            native.check(lib.tb200_vmm_alloc(size, C.byref(phys), C.byref(ptr), C.byref(fd)))
            mc, mc_fd = u64(), ci(-1)
            if self.rank == 0:
                native.check(lib.tb200_mc_create(self.world, size, C.byref(mc), C.byref(mc_fd)))
            mine = [fd.value] + ([mc_fd.value] if self.rank == 0 else [])
            got = exchange_fds(self.rank, self.world, mine, self.group)
            local = tensor_from_ptr(int(ptr.value), nbytes, torch.uint8, self.device)
            buf = SymmBuffer(name, nbytes, [0] * self.world, local, vmm_bytes=size)
            buf.ptrs[self.rank] = int(ptr.value)
            buf.vmm_handles[self.rank] = int(phys.value)
            try:
                for r in range(self.world):
                    if r == self.rank:
                        continue
                    h, p = u64(), vp()
                    native.check(lib.tb200_vmm_import(got[r][0], size, C.byref(h), C.byref(p)))
                    buf.ptrs[r], buf.vmm_handles[r] = int(p.value), int(h.value)
                if self.rank != 0:
                    native.check(lib.tb200_mc_import(got[0][1], C.byref(mc)))
                buf.mc_handle = int(mc.value)
                native.check(lib.tb200_mc_add_device(mc))
                dist.barrier(group=self.group)          # every GPU is in the object before anyone binds
                mcp = vp()
                native.check(lib.tb200_mc_bind(mc, phys, size, C.byref(mcp)))
                buf.mc_ptr = int(mcp.value)
            finally:
                for fds in got.values():
                    for f in fds:
                        os.close(f)
                for f in mine:
                    os.close(f)
        self.buffers[name] = buf
        dist.barrier(group=self.group)
        return buf

    def _exchange(self, buf: SymmBuffer) -> None:
        import torch.distributed as dist
        lib = native.lib()
        handle = C.create_string_buffer(native.IPC_HANDLE_BYTES)
        native.check(lib.tb200_ipc_export(buf.ptrs[self.rank], handle))
        gathered: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(gathered, bytes(handle.raw), group=self.group)
        with torch.cuda.device(self.device):
            for r, h in enumerate(gathered):
                if r == self.rank:
                    continue
                out = C.c_void_p()
                native.check(lib.tb200_ipc_open(h, C.byref(out)))
                buf.ptrs[r] = int(out.value)
                buf.opened.append(int(out.value))
        dist.barrier(group=self.group)

    # -- native views ---------------------------------------------------------------------------
    def native_world(self, grad: Optional[str], param: Optional[str], flags: str) -> native.World:
        w = native.World()
        w.world, w.rank = self.world, self.rank
        for r in range(self.world):
            w.grad[r] = self.buffers[grad].ptrs[r] if grad else None
            w.param[r] = self.buffers[param].ptrs[r] if param else None
            w.flags[r] = self.buffers[flags].ptrs[r]
        w.grad_mc = (self.buffers[grad].mc_ptr or None) if grad else None
        w.param_mc = (self.buffers[param].mc_ptr or None) if param else None
        return w

    def post_flags(self, flags: str, first: int, count: int, value: int) -> None:
        """Host-side write of a run of flag words in the *local* pad."""
        native.check(native.lib().tb200_flags_fill(self.buffers[flags].ptrs[self.rank], first, count,
                                                   value, native.stream_ptr()))

    def barrier(self) -> None:
        if self.world > 1 and self._emulated_peers is None:
            import torch.distributed as dist
            dist.barrier(group=self.group)

    def close(self) -> None:
        lib = native.lib()
        torch.cuda.synchronize(self.device)
        self.barrier()
        for buf in self.buffers.values():
            for p in buf.opened:
                lib.tb200_ipc_close(p)
            buf.opened.clear()
            if buf.vmm_bytes:
                with torch.cuda.device(self.device):
                    if buf.mc_handle:
                        lib.tb200_mc_release(buf.mc_handle, buf.mc_ptr or None, buf.vmm_bytes)
                    for r, h in buf.vmm_handles.items():
                        if r != self.rank:
                            lib.tb200_vmm_release(h, buf.ptrs[r], buf.vmm_bytes)
        self.barrier()
        for buf in self.buffers.values():
            if buf.vmm_bytes:
                with torch.cuda.device(self.device):
                    lib.tb200_vmm_release(buf.vmm_handles[self.rank], buf.ptrs[self.rank], buf.vmm_bytes)
            else:
                lib.tb200_symm_free(buf.ptrs[self.rank])
        self.buffers.clear()
