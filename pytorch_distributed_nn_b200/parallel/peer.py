"""Peer-mapped device memory: one arena per rank, visible to every other rank of the node over NVLink.

The reference's transport is MPI point-to-point over TCP between hosts. On one NVSwitch box the
transport is memory: every rank allocates an arena with ``cudaMalloc`` (through libpdnn_b200.so, so that the
allocation is a clean IPC unit rather than a slice of torch's caching allocator), exports it with CUDA IPC,
and maps every other rank's arena. Kernels then store gradients / weights / flags straight into peer memory.
``torch.distributed`` is used only once, to exchange the 64-byte handles.
"""
from __future__ import annotations

import ctypes as C
from typing import List

import torch

from ..ops import _lib


class _CudaBytes:
    """Minimal __cuda_array_interface__ holder so torch can view raw device memory without copying."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2, "strides": None,
        }


def tensor_from_ptr(ptr: int, nbytes: int, dtype: torch.dtype, device) -> torch.Tensor:
    t = torch.as_tensor(_CudaBytes(ptr, nbytes), device=device)
    return t.view(dtype)


class PeerArena:
    """A device allocation of ``nbytes`` on every rank plus the mapped pointers of all peers.

    ``local`` is this rank's base pointer; ``ptrs[r]`` is rank r's arena as seen from this rank.
    Sub-buffers are carved by byte offset (identical on every rank), see :meth:`carve`.
    """

    def __init__(self, nbytes: int, rank: int, world: int, device, group=None):
        self.nbytes = int((nbytes + 255) // 256 * 256)
        self.rank, self.world, self.device = rank, world, torch.device(device)
        self._cursor = 0
        out = C.c_void_p()
        _lib.call("pdnn_malloc", C.byref(out), self.nbytes)
        self.local = int(out.value)
        self.ptrs: List[int] = [0] * world
        self.ptrs[rank] = self.local
        self._opened = []
        if world > 1:
            import torch.distributed as dist
            handle = (C.c_ubyte * 64)()
            _lib.call("pdnn_ipc_get_handle", self.local, C.cast(handle, C.c_void_p).value)
            gathered = [None] * world
            dist.all_gather_object(gathered, bytes(handle), group=group)
            for r, hb in enumerate(gathered):
                if r == rank:
                    continue
                buf = (C.c_ubyte * 64).from_buffer_copy(hb)
                p = C.c_void_p()
                _lib.call("pdnn_ipc_open_handle", C.cast(buf, C.c_void_p).value, C.byref(p))
                self.ptrs[r] = int(p.value)
                self._opened.append(int(p.value))
            dist.barrier(group=group)

    def carve(self, nbytes: int, align: int = 256) -> int:
        """Reserve ``nbytes`` at the same offset on every rank; returns the byte offset."""
        off = (self._cursor + align - 1) // align * align
        if off + nbytes > self.nbytes:
            raise MemoryError(f"PeerArena exhausted: need {off + nbytes} of {self.nbytes} bytes")
        self._cursor = off + nbytes
        return off

    def ptr(self, rank: int, offset: int) -> int:
        return self.ptrs[rank] + offset

    def tensor(self, offset: int, nbytes: int, dtype: torch.dtype) -> torch.Tensor:
        """Torch view of a LOCAL sub-buffer."""
        return tensor_from_ptr(self.local + offset, nbytes, dtype, self.device)

    def close(self):
        for p in self._opened:
            try:
                _lib.call("pdnn_ipc_close", p)
            except Exception:
                pass
        self._opened = []
        if self.local:
            try:
                _lib.call("pdnn_free", self.local)
            except Exception:
                pass
            self.local = 0


class SymmArena(PeerArena):
    """Same interface as :class:`PeerArena`, but the allocation comes from torch's symmetric-memory allocator
    (CUDA VMM + fabric handles), which also binds the arenas of all ranks to one NVLS **multicast** address:
    a store to ``mc_ptr + offset`` lands in every rank's arena (the switch replicates it), and
    ``multimem.ld_reduce`` on it returns the sum over all ranks computed inside the switch.
    ``mc_ptr`` is 0 when the fabric has no multicast support; callers then use the unicast pointers."""

    def __init__(self, nbytes: int, rank: int, world: int, device, group=None):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        self.nbytes = int((nbytes + 255) // 256 * 256)
        self.rank, self.world, self.device = rank, world, torch.device(device)
        self._cursor = 0
        self._opened = []
        self._buf = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=self.device)
        self._buf.zero_()
        grp = group if group is not None else dist.group.WORLD
        self._hdl = symm_mem.rendezvous(self._buf, grp.group_name if hasattr(grp, "group_name") else grp)
        self.local = int(self._buf.data_ptr())
        self.ptrs = [int(p) for p in list(self._hdl.buffer_ptrs)[:world]]
        self.mc_ptr = int(getattr(self._hdl, "multicast_ptr", 0) or 0)
        torch.cuda.synchronize()
        dist.barrier(group=group)

    def mc(self, offset: int) -> int:
        return self.mc_ptr + offset if self.mc_ptr else 0

    def close(self):
        self._hdl = None
        self._buf = None
        self.local = 0


def make_arena(nbytes: int, rank: int, world: int, device, group=None, prefer_symmetric: bool = True):
    """Symmetric (multicast-capable) arena when torch provides one for this job, CUDA-IPC arena otherwise."""
    if world > 1 and prefer_symmetric:
        try:
            return SymmArena(nbytes, rank, world, device, group)
        except Exception as ex:  # noqa: BLE001 -- any failure of the optional path falls back to plain IPC
            import warnings
            warnings.warn(f"symmetric memory unavailable ({ex!r}); using CUDA IPC peer arenas without multicast")
    arena = PeerArena(nbytes, rank, world, device, group)
    arena.mc_ptr = 0
    arena.mc = lambda offset: 0
    return arena
