"""Symmetric memory for one NVSwitch domain.

Every rank allocates the same set of device buffers (gradients, parameters, flag pad) straight
from ``cudaMalloc`` through the native library, exports them as CUDA IPC handles, and maps every
peer's buffers, so kernels dereference peer pointers directly over NVLink. ``torch.distributed``
is used only as the bootstrap channel for the 64-byte handles.

The reference has no communication layer: its network is the three numbers
``--bandwidth``, ``--internode_latency`` (run_sim.py:59-69) and ``FACTOR`` (infra/interference.py).
"""
from __future__ import annotations

import ctypes as C
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


@dataclass
class SymmBuffer:
    name: str
    nbytes: int
    ptrs: List[int]                 # indexed by rank; ptrs[rank] is the local allocation
    local: torch.Tensor             # uint8 view of the local allocation
    opened: List[int] = field(default_factory=list)   # IPC-mapped pointers to close

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
    def alloc(self, name: str, nbytes: int) -> SymmBuffer:
        """Collective: every rank must call with the same name and size, in the same order."""
        if name in self.buffers:
            raise KeyError(f"symmetric buffer {name!r} already exists")
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
        self.barrier()
        for buf in self.buffers.values():
            lib.tb200_symm_free(buf.ptrs[self.rank])
        self.buffers.clear()
