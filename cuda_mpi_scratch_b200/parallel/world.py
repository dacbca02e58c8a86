"""Process bootstrap: who am I, which GPU is mine, how do I reach my peers.

Capability parity with ``MPI_Init`` / ``MPI_Comm_rank`` / ``MPI_Comm_size`` /
``MPI_Get_processor_name`` / ``MPI_Wtime`` / ``MPI_Abort`` / ``MPI_Finalize``
(``mpi1.cpp:11-16``, ``mpi2.cpp:31-37``) plus the rank->GPU binding that the
reference does by hand before ``MPI_Init``
(``stencil2d/mpi-2d-stencil-subarray-cuda.cu:85-88``).

Three ways to get ranks:

* :class:`TorchWorld` -- one process per GPU, launched by ``torchrun`` (what
  ``bench.py`` uses).  ``torch.distributed`` is only the *plumbing*: it carries
  the 64-byte CUDA-IPC handles at set-up, NCCL baselines, and the final
  max-over-ranks of a timing.  No hot path calls it.
* :class:`LocalGroup` -- N virtual ranks inside one process (one or several
  GPUs).  Used by tests and by single-GPU emulation of multi-rank runs: ranks
  are stepped in lock-step in one stream, so no kernel ever waits for a
  kernel that has not been launched yet.
* The native host transport (``parallel/comm.py``) for MPI-style messaging.
"""
from __future__ import annotations

import ctypes as C
import os
import socket
import time
from typing import Sequence

from .. import _lib


def processor_name() -> str:
    """``MPI_Get_processor_name`` (``mpi1.cpp:14``)."""
    return socket.gethostname()


def wtime() -> float:
    """``MPI_Wtime`` (``test-benchmark/mpi-pingpong-gpu.cpp:51``): wall clock in s."""
    return time.perf_counter()


def bind_device(policy: str = "local_rank", rank: int = 0, node_count: int = 1) -> int:
    """Select this rank's GPU; returns the device index.

    ``local_rank`` reads LOCAL_RANK / OMPI_COMM_WORLD_LOCAL_RANK /
    MV2_COMM_WORLD_LOCAL_RANK / SLURM_LOCALID; ``rank_mod`` is the reference's
    ``task % device_count`` (``mpicuda2.cu:201``); ``round_robin`` is its
    ``(task / node_count) % device_count`` (``mpicuda2.cu:199``).
    ``NUM_GPU_DEVICES`` caps the device count as in the reference's BindDevice.
    """
    pol = {"local_rank": 0, "rank_mod": 1, "round_robin": 2}[policy]
    dev = C.c_int(-1)
    _lib.check(_lib.lib.b2c_bind_device(pol, rank, node_count, C.byref(dev)))
    try:
        import torch

        if torch.cuda.is_available():
            torch.cuda.set_device(dev.value)
    except ImportError:  # pragma: no cover
        pass
    return dev.value


class World:
    """Minimal bootstrap interface used by the symmetric heap and the models."""

    rank: int = 0
    size: int = 1
    device: int = 0

    def allgather_bytes(self, b: bytes) -> list[bytes]:
        raise NotImplementedError

    def barrier(self) -> None:
        raise NotImplementedError

    def allreduce_max(self, x: float) -> float:
        raise NotImplementedError

    def allreduce_sum(self, x: float) -> float:
        raise NotImplementedError

    def allreduce_max_vec(self, xs: Sequence[float]) -> list[float]:
        """Element-wise max over ranks of a list of numbers (per-batch timings)."""
        return [self.allreduce_max(float(x)) for x in xs]

    def abort(self, code: int = 1) -> None:
        """``MPI_Abort`` (``mpierr.h:41``)."""
        os._exit(code)

    def finalize(self) -> None:
        pass


class SingleWorld(World):
    """World of one rank (no launcher)."""

    def __init__(self, device: int | None = None):
        self.rank, self.size = 0, 1
        self.device = bind_device("local_rank", 0) if device is None else device
        if device is not None:
            _lib.check(_lib.lib.b2c_set_device(device))

    def allgather_bytes(self, b):
        return [b]

    def barrier(self):
        pass

    def allreduce_max(self, x):
        return x

    def allreduce_sum(self, x):
        return x


class TorchWorld(World):
    """One process per GPU under ``torchrun`` (RANK/WORLD_SIZE/LOCAL_RANK/MASTER_*)."""

    def __init__(self, backend: str | None = None):
        import torch
        import torch.distributed as dist

        self.rank = int(os.environ.get("RANK", "0"))
        self.size = int(os.environ.get("WORLD_SIZE", "1"))
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        self._owns_pg = False
        self._cuda = torch.cuda.is_available()
        if self._cuda:
            self.device = bind_device("local_rank", self.rank)
        else:
            self.device = -1
        if not dist.is_initialized():
            if backend is None:
                backend = "cpu:gloo,cuda:nccl" if self._cuda else "gloo"
            dist.init_process_group(backend=backend, rank=self.rank, world_size=self.size)
            self._owns_pg = True
        self._dist = dist
        self._torch = torch
        # object/byte exchange always goes over a CPU (gloo) group
        self._cpu_group = dist.new_group(backend="gloo")

    def allgather_bytes(self, b):
        out = [None] * self.size
        self._dist.all_gather_object(out, b, group=self._cpu_group)
        return out

    def barrier(self):
        self._dist.barrier(group=self._cpu_group)

    def _allreduce(self, x, op):
        t = self._torch.tensor([x], dtype=self._torch.float64)
        self._dist.all_reduce(t, op=op, group=self._cpu_group)
        return float(t.item())

    def allreduce_max(self, x):
        return self._allreduce(x, self._dist.ReduceOp.MAX)

    def allreduce_sum(self, x):
        return self._allreduce(x, self._dist.ReduceOp.SUM)

    def allreduce_max_vec(self, xs):
        t = self._torch.tensor([float(x) for x in xs], dtype=self._torch.float64)
        if t.numel():
            self._dist.all_reduce(t, op=self._dist.ReduceOp.MAX, group=self._cpu_group)
        return t.tolist()

    def finalize(self):
        if self._owns_pg and self._dist.is_initialized():
            self._dist.destroy_process_group()
            self._owns_pg = False


class CommWorld(World):
    """One process per GPU, bootstrapped by the library's OWN communicator (``parallel/comm.py``,
    the MPI-like API over the shared-memory fabric): what ``python -m cuda_mpi_scratch_b200.launch
    -n N prog.py`` starts.  No ``torch.distributed`` anywhere: the symmetric heap, the fused stencil,
    the fused dot and the device collectives get their handles, barriers and max-over-ranks from
    ``Comm.allgather`` / ``barrier`` / ``allreduce`` -- the same calls a user of the MPI-like API
    makes (``MPI_Init`` .. ``MPI_Allreduce``, ``mpi1.cpp:11-16``, ``mpi9.cpp:51-54``)."""

    def __init__(self, comm=None, device: int | None = None):
        import numpy as np
        from . import comm as cm
        self._np, self._cm = np, cm
        self.comm = comm if comm is not None else cm.init_process()
        self._owns = comm is None
        self.rank, self.size = self.comm.rank, self.comm.size
        try:
            import torch
            cuda = torch.cuda.is_available()
        except ImportError:  # pragma: no cover
            cuda = False
        self.device = (bind_device("local_rank", self.rank) if device is None else device) if cuda else -1

    def allgather_bytes(self, b: bytes) -> list[bytes]:
        np = self._np
        lens = np.zeros(self.size, dtype=np.int64)
        self.comm.allgather(np.array([len(b)], dtype=np.int64), lens)
        width = max(int(lens.max()), 1)
        mine = np.zeros(width, dtype=np.uint8)
        mine[:len(b)] = np.frombuffer(b, dtype=np.uint8)
        allb = np.zeros(self.size * width, dtype=np.uint8)
        self.comm.allgather(mine, allb)
        return [allb[r * width: r * width + int(lens[r])].tobytes() for r in range(self.size)]

    def barrier(self) -> None:
        self.comm.barrier()

    def _allreduce(self, x: float, op: int) -> float:
        np = self._np
        a, out = np.array([x], dtype=np.float64), np.zeros(1, dtype=np.float64)
        self.comm.allreduce(a, out, op)
        return float(out[0])

    def allreduce_max(self, x):
        return self._allreduce(x, self._cm.MAX)

    def allreduce_sum(self, x):
        return self._allreduce(x, self._cm.SUM)

    def allreduce_max_vec(self, xs):
        np = self._np
        a = np.array([float(x) for x in xs], dtype=np.float64)
        out = np.zeros_like(a)
        if a.size:
            self.comm.allreduce(a, out, self._cm.MAX)
        return out.tolist()

    def abort(self, code: int = 1) -> None:
        self.comm.abort(code)

    def finalize(self) -> None:
        if self._owns:
            self.comm.finalize()
            self._owns = False


def init(backend: str | None = None) -> World:
    """``MPI_Init`` analogue: returns the world for this process.  Under ``torchrun``
    (``WORLD_SIZE`` > 1): :class:`TorchWorld`; under the library's own launcher (``B200_RANK`` /
    ``B200_WORLD_SIZE``, ``python -m cuda_mpi_scratch_b200.launch``): :class:`CommWorld`, no
    ``torch.distributed`` involved; otherwise a world of one."""
    if backend != "torch" and int(os.environ.get("B200_WORLD_SIZE", "1")) > 1 and "WORLD_SIZE" not in os.environ:
        return CommWorld()
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        return TorchWorld(None if backend == "torch" else backend)
    return SingleWorld()


class LocalGroup:
    """N virtual ranks inside this process.

    ``devices[r]`` is the GPU of virtual rank ``r`` (all 0 for single-GPU
    emulation).  Not a :class:`World`: collective set-up is done for all ranks
    at once by the factory functions that take a LocalGroup.
    """

    def __init__(self, nranks: int, devices: Sequence[int] | None = None):
        if nranks < 1:
            raise ValueError("nranks must be >= 1")
        self.size = nranks
        self.devices = list(devices) if devices is not None else [0] * nranks
        if len(self.devices) != nranks:
            raise ValueError("devices must have one entry per rank")
        for a in set(self.devices):
            for b in set(self.devices):
                if a != b:
                    _lib.check(_lib.lib.b2c_enable_peer_access(a, b))
