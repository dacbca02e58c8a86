"""Symmetric, peer-mapped device allocations.

Every rank allocates a block of the same size with ``cudaMalloc`` (through the
native library, not through torch's caching allocator, so the IPC handle
covers exactly this block), exports a CUDA-IPC handle, gathers all handles
over the bootstrap, and opens every peer's block.  Afterwards each rank holds
a table of ``size`` base pointers, and a kernel can store into
``peer_ptr(r) + offset`` directly over NVLink.

Reference counterpart (capability): handing a device pointer to a CUDA-aware
MPI and letting the library stage / IPC-map it per message
(``stencil2d/stencil2D.h:369-373``, ``test-benchmark/mpi-pingpong-gpu.cpp:52-53``).
Here the mapping is done once at set-up and kernels do the data movement.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Sequence

from .. import _lib
from .world import LocalGroup, World


class _CudaArray:
    """Adapter exposing raw device memory through ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, shape, typestr: str, strides=None, owner=None):
        self.__cuda_array_interface__ = {
            "shape": tuple(shape),
            "typestr": typestr,
            "data": (int(ptr), False),
            "version": 3,
            "strides": None if strides is None else tuple(strides),
        }
        self._owner = owner


_TYPESTR = {
    "float64": "<f8", "float32": "<f4", "int32": "<i4", "int64": "<i8",
    "uint8": "|u1", "int8": "|i1", "uint64": "<u8", "float16": "<f2", "int16": "<i2",
}


def as_tensor(ptr: int, shape, dtype, device: int, strides_bytes=None, owner=None):
    """View device memory at ``ptr`` as a torch tensor (no copy)."""
    import torch

    name = str(dtype).replace("torch.", "")
    if name == "uint64":  # torch has limited uint64 support: view as int64
        name = "int64"
    arr = _CudaArray(ptr, shape, _TYPESTR[name], strides_bytes, owner)
    return torch.as_tensor(arr, device=torch.device("cuda", device))


class SymmetricBuffer:
    """One rank's view of a symmetric allocation."""

    def __init__(self, rank: int, size: int, nbytes: int, device: int, local_ptr: int,
                 peer_ptrs: Sequence[int], opened: Sequence[int] = (), owns: bool = True,
                 block=None):
        self.rank, self.size, self.nbytes, self.device = rank, size, nbytes, device
        self.local_ptr = int(local_ptr)
        self.peer_ptrs = [int(p) for p in peer_ptrs]
        self._opened = list(opened)
        self._owns = owns
        self._freed = False
        self._block = block              # HeapBlock when the memory comes from the symmetric heap
        #: multicast address of the block (0: none): a store there lands in EVERY rank's copy
        self.mc_ptr = int(block.mc_ptr) if block is not None else 0

    # -- factories ---------------------------------------------------------
    @staticmethod
    def allocate(world: World, nbytes: int, multicast: bool = False) -> "SymmetricBuffer":
        """Collective over ``world``.  Multi-process worlds sub-allocate from the world's
        :class:`~cuda_mpi_scratch_b200.parallel.heap.SymmetricHeap` (one VMM mapping per arena and
        peer, optional NVLink multicast binding); ``B200COMM_HEAP=ipc`` or a driver without
        exportable allocations falls back to one ``cudaMalloc`` + CUDA-IPC mapping per buffer."""
        if world.size > 1:
            from .heap import SymmetricHeap
            if SymmetricHeap.enabled(world):
                try:
                    heap = SymmetricHeap.for_world(world)
                except _lib.CommError:
                    heap = None
                # all ranks take the same path
                if world.allreduce_max(0.0 if heap is not None else 1.0) == 0.0:
                    blk = heap.alloc(nbytes, multicast=multicast)
                    return SymmetricBuffer(world.rank, world.size, blk.nbytes, world.device, blk.local_ptr,
                                           blk.peer_ptrs, (), owns=False, block=blk)
        p = C.c_void_p()
        _lib.check(_lib.lib.b2c_malloc(C.byref(p), nbytes))
        if world.size == 1:
            return SymmetricBuffer(0, 1, nbytes, world.device, p.value, [p.value])
        handle = C.create_string_buffer(_lib_handle_bytes())
        _lib.check(_lib.lib.b2c_ipc_export(p, handle))
        me = handle.raw + os.getpid().to_bytes(8, "little") + int(world.device).to_bytes(4, "little")
        infos = world.allgather_bytes(me)
        peers, opened = [], []
        for r, info in enumerate(infos):
            if r == world.rank:
                peers.append(p.value)
                continue
            q = C.c_void_p()
            _lib.check(_lib.lib.b2c_ipc_open(info[: _lib_handle_bytes()], C.byref(q)))
            peers.append(q.value)
            opened.append(q.value)
        world.barrier()
        return SymmetricBuffer(world.rank, world.size, nbytes, world.device, p.value, peers, opened)

    @staticmethod
    def allocate_local(group: LocalGroup, nbytes: int) -> list["SymmetricBuffer"]:
        """All virtual ranks of a :class:`LocalGroup` at once (one process)."""
        prev = C.c_int()
        _lib.check(_lib.lib.b2c_get_device(C.byref(prev)))
        ptrs = []
        for dev in group.devices:
            _lib.check(_lib.lib.b2c_set_device(dev))
            p = C.c_void_p()
            _lib.check(_lib.lib.b2c_malloc(C.byref(p), nbytes))
            ptrs.append(p.value)
        _lib.check(_lib.lib.b2c_set_device(prev.value))
        return [SymmetricBuffer(r, group.size, nbytes, group.devices[r], ptrs[r], ptrs)
                for r in range(group.size)]

    # -- access --------------------------------------------------------------
    def peer_ptr(self, r: int) -> int:
        return self.peer_ptrs[r]

    def tensor(self, offset: int, shape, dtype, strides_bytes=None):
        """Torch view of my block at byte ``offset``."""
        return as_tensor(self.local_ptr + offset, shape, dtype, self.device, strides_bytes, self)

    def peer_tensor(self, r: int, offset: int, shape, dtype, strides_bytes=None):
        """Torch view of rank ``r``'s block (peer-mapped memory; debugging/tests)."""
        return as_tensor(self.peer_ptrs[r] + offset, shape, dtype, self.device, strides_bytes, self)

    def release_if_heap(self) -> None:
        """Give a heap block back (called by the plans' ``close()``).  The block is reused only
        after EVERY rank has released it, so this is safe from finalisers and in any order; buffers
        of the cudaMalloc / IPC fallback stay allocated until :meth:`free` or process exit, as
        peers may still hold mappings of them."""
        if self._block is not None:
            self.free()

    def free(self) -> None:
        if self._freed:
            return
        self._freed = True
        if self._block is not None:
            self._block.heap.free(self._block)
            return
        for q in self._opened:
            _lib.lib.b2c_ipc_close(C.c_void_p(q))
        if self._owns:
            _lib.lib.b2c_free(C.c_void_p(self.local_ptr))


def _lib_handle_bytes() -> int:
    return 64
