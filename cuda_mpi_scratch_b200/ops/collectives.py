"""Collectives on CUDA tensors over peer-mapped symmetric memory (no host, no NCCL).

Capability parity with ``MPI_Reduce`` / ``MPI_Allreduce`` / ``MPI_Gather`` /
``MPI_Bcast`` / ``MPI_Scatter`` as used by the reference (``mpicuda2.cu:154,293``,
``mpi9.cpp:51-54``, ``mpi6.cpp:89-100``) -- there always on host scalars after a
device-to-host copy.  Here the operands stay on the GPUs: inputs are staged in a
symmetric workspace, kernels read/write the peers' staging directly and
synchronise with epoch flags.  Reductions add in rank order, so every rank gets
bit-identical results.

``allreduce`` picks one-shot (every rank reads all inputs) below
``twoshot_min_bytes`` and two-shot (reduce-scatter + broadcast of the reduced
chunks) above it.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _lib
from ..parallel.symmetric import SymmetricBuffer
from ..parallel.world import LocalGroup, World

_L = _lib.lib
_vp, _i32, _i64 = C.c_void_p, C.c_int, C.c_int64
_lib._proto("b2c_coll_workspace_bytes", [C.c_size_t, C.POINTER(C.c_size_t)])
_lib._proto("b2c_coll_create", [C.POINTER(_vp), _i32, _i32, C.POINTER(_vp), C.c_size_t])
_lib._proto("b2c_coll_destroy", [_vp])
_lib._proto("b2c_coll_set_multicast", [_vp, _vp, _i32])
_lib._proto("b2c_coll_allreduce_symmetric", [_vp, _vp, _i64, _i32, _vp])
_lib._proto("b2c_coll_spread_symmetric", [_vp, _vp, _vp, C.c_size_t, C.c_size_t, _vp])
_lib._proto("b2c_coll_allreduce", [_vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _vp])
_lib._proto("b2c_coll_move", [_vp, _i32, _vp, _vp, C.c_size_t, _i32, _i32, _vp])
_lib._proto("b2c_coll_barrier", [_vp, _i32, _vp])
_lib._proto("b2c_coll_check", [_vp, _vp])

_DT = {torch.int32: 5, torch.int64: 6, torch.float32: 8, torch.float64: 9}
OPS = {"sum": 0, "prod": 1, "max": 2, "min": 3}
BCAST, GATHER, ALLGATHER, SCATTER = range(4)


def workspace_bytes(max_bytes: int) -> int:
    n = C.c_size_t()
    _lib.check(_L.b2c_coll_workspace_bytes(max_bytes, C.byref(n)))
    return n.value


def _stream(device: int) -> int:
    return torch.cuda.current_stream(device).cuda_stream


class DeviceCollectives:
    """One rank's handle.  All calls are stream-ordered on the current torch stream."""

    def __init__(self, rank: int, nranks: int, symbuf: SymmetricBuffer, max_bytes: int,
                 world: World | None, twoshot_min_bytes: int = 512 << 10):
        self.rank, self.nranks, self.symbuf, self.max_bytes = rank, nranks, symbuf, max_bytes
        self.world, self.device = world, symbuf.device
        self.twoshot_min_bytes = twoshot_min_bytes
        ws = (C.c_void_p * nranks)(*[symbuf.peer_ptr(r) for r in range(nranks)])
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(_L.b2c_coll_create(C.byref(h), rank, nranks, ws, max_bytes))
        self._h = h
        self.kernel_launches = 0
        #: the workspace has an NVLink multicast mapping: in-switch allreduce ("nvls") and
        #: one-store broadcast are available
        self.multicast = bool(getattr(symbuf, "mc_ptr", 0))
        if self.multicast:
            _lib.check(_L.b2c_coll_set_multicast(h, C.c_void_p(symbuf.mc_ptr), 1))

    @classmethod
    def create(cls, world: World, max_bytes: int = 64 << 20, **kw) -> "DeviceCollectives":
        sym = SymmetricBuffer.allocate(world, workspace_bytes(max_bytes), multicast=True)
        return cls(world.rank, world.size, sym, max_bytes, world, **kw)

    @classmethod
    def create_group(cls, group: LocalGroup, max_bytes: int = 16 << 20, **kw) -> "CollectiveGroup":
        syms = SymmetricBuffer.allocate_local(group, workspace_bytes(max_bytes))
        return CollectiveGroup([cls(r, group.size, syms[r], max_bytes, None, **kw)
                                for r in range(group.size)])

    # ---- reductions ----------------------------------------------------------
    def _algo(self, t: torch.Tensor, algo: str) -> int:
        if algo == "auto":
            nbytes = t.numel() * t.element_size()
            # measured (profiles/r2/collectives_{2,4,8}gpu.json): with 8 ranks the in-switch path wins
            # from 1 KiB up (26 vs 31 us at 1 KiB, 80 vs 324 us at 16 MiB); with 2 ranks the switch
            # has nothing to save and one-shot / two-shot stay ahead
            # (4 ranks: one-shot stays ahead up to 1 MiB, 21-28 vs 22-29 us; from 4 MiB the switch wins)
            if self.multicast and self._nvls_op and ((self.nranks >= 8 and nbytes >= 1024) or
                                                     (self.nranks >= 4 and nbytes >= (4 << 20))):
                return 2
            return 1 if nbytes >= self.twoshot_min_bytes and self.nranks > 1 else 0
        if algo == "nvls" and not self.multicast:
            raise _lib.CommError(_lib.ERR_STATE, "no multicast mapping: the in-switch allreduce is not available")
        return {"oneshot": 0, "twoshot": 1, "nvls": 2}[algo]

    def allreduce(self, t: torch.Tensor, op: str = "sum", algo: str = "auto", out: torch.Tensor | None = None,
                  phases: int = 7, root: int = -1) -> torch.Tensor:
        """In-place (or into ``out``) all-reduce of a contiguous CUDA tensor."""
        if t.dtype not in _DT or not t.is_contiguous():
            raise TypeError("need a contiguous int32/int64/float32/float64 CUDA tensor")
        out = t if out is None else out
        self._nvls_op = op == "sum" and root < 0          # what the switch can reduce
        with torch.cuda.device(self.device):
            _lib.check(_L.b2c_coll_allreduce(self._h, t.data_ptr(), out.data_ptr(), t.numel(), _DT[t.dtype],
                                             OPS[op], root, self._algo(t, algo) if root < 0 else 0,
                                             phases, _stream(self.device)))
        self.kernel_launches += 1
        return out

    def symmetric_empty(self, numel: int, dtype: torch.dtype = torch.float32) -> torch.Tensor:
        """Collective: a tensor in a MULTICAST arena of the symmetric heap, at the same offset on
        every rank -- the operand type of :meth:`allreduce_symmetric` (no staging copies).  Keep the
        tensor for the life of the computation: it is freed with this handle."""
        if self.world is None or not self.multicast:
            raise _lib.CommError(_lib.ERR_STATE, "symmetric tensors need a multi-process world with multicast")
        sym = SymmetricBuffer.allocate(self.world, numel * torch.empty((), dtype=dtype).element_size(), multicast=True)
        if not sym.mc_ptr:
            raise _lib.CommError(_lib.ERR_STATE, "the heap could not bind this block to a multicast object")
        t = sym.tensor(0, (numel,), dtype)
        self._symm = getattr(self, "_symm", {})
        self._symm[t.data_ptr()] = sym
        return t

    def allreduce_symmetric(self, t: torch.Tensor) -> torch.Tensor:
        """In-place SUM over ranks of a tensor from :meth:`symmetric_empty` (or a leading slice of
        it): the NVSwitch reduces and redistributes, nothing is staged."""
        sym = getattr(self, "_symm", {}).get(t.data_ptr())
        if sym is None or t.dtype not in _DT or not t.is_contiguous():
            raise ValueError("allreduce_symmetric needs a (slice from the start of a) tensor made by symmetric_empty")
        with torch.cuda.device(self.device):
            _lib.check(_L.b2c_coll_allreduce_symmetric(self._h, C.c_void_p(sym.mc_ptr), t.numel(), _DT[t.dtype],
                                                       _stream(self.device)))
        self.kernel_launches += 2
        return t

    def _spread_symmetric(self, t: torch.Tensor, lo: int, hi: int) -> None:
        sym = getattr(self, "_symm", {}).get(t.data_ptr())
        if sym is None or not t.is_contiguous():
            raise ValueError("needs a (slice from the start of a) tensor made by symmetric_empty")
        with torch.cuda.device(self.device):
            _lib.check(_L.b2c_coll_spread_symmetric(self._h, C.c_void_p(sym.mc_ptr), C.c_void_p(t.data_ptr()),
                                                    lo, hi, _stream(self.device)))
        self.kernel_launches += 2

    def bcast_symmetric(self, t: torch.Tensor, root: int = 0) -> torch.Tensor:
        """``MPI_Bcast`` in place on a tensor from :meth:`symmetric_empty`: the root stores it ONCE
        to the multicast address and the switch writes every rank's copy -- one pass, no staging
        (``mpicuda2.cu:154`` broadcasts a host scalar)."""
        nb = t.numel() * t.element_size()
        if nb % 16:
            raise ValueError("bcast_symmetric moves whole 16-byte vectors: pad the tensor")
        self._spread_symmetric(t, 0, nb if self.rank == root else 0)
        return t

    def allgather_symmetric(self, t: torch.Tensor) -> torch.Tensor:
        """In-place allgather (``MPI_IN_PLACE``) on a tensor from :meth:`symmetric_empty` of
        ``nranks`` equal blocks: block ``rank`` holds this rank's contribution on entry, every block
        is filled on return.  Each rank stores its block once through the switch."""
        nb = t.numel() * t.element_size()
        if nb % self.nranks or (nb // self.nranks) % 16:
            raise ValueError("the tensor must split into nranks blocks of a multiple of 16 bytes")
        blk = nb // self.nranks
        self._spread_symmetric(t, self.rank * blk, (self.rank + 1) * blk)
        return t

    def reduce(self, t: torch.Tensor, root: int, op: str = "sum", out: torch.Tensor | None = None,
               phases: int = 7) -> torch.Tensor | None:
        """``MPI_Reduce``: the result is written on ``root`` only."""
        res = self.allreduce(t, op, "oneshot", out if out is not None else t, phases, root)
        return res if self.rank == root else None

    # ---- data movement ---------------------------------------------------------
    def _move(self, kind, send, recv, block_bytes, root, phases):
        with torch.cuda.device(self.device):
            _lib.check(_L.b2c_coll_move(self._h, kind, None if send is None else send.data_ptr(),
                                        None if recv is None else recv.data_ptr(), block_bytes, root,
                                        phases, _stream(self.device)))
        self.kernel_launches += 2

    def bcast(self, t: torch.Tensor, root: int = 0, phases: int = 3) -> torch.Tensor:
        self._move(BCAST, t, t, t.numel() * t.element_size(), root, phases)
        return t

    def gather(self, send: torch.Tensor, recv: torch.Tensor | None, root: int = 0, phases: int = 3):
        """Non-root ranks may pass ``recv=None`` (``mpi6.cpp:96-100``)."""
        self._move(GATHER, send, recv, send.numel() * send.element_size(), root, phases)
        return recv

    def allgather(self, send: torch.Tensor, recv: torch.Tensor, phases: int = 3) -> torch.Tensor:
        self._move(ALLGATHER, send, recv, send.numel() * send.element_size(), 0, phases)
        return recv

    def scatter(self, send: torch.Tensor | None, recv: torch.Tensor, root: int = 0, phases: int = 3):
        self._move(SCATTER, send, recv, recv.numel() * recv.element_size(), root, phases)
        return recv

    def barrier(self, phases: int = 3) -> None:
        with torch.cuda.device(self.device):
            _lib.check(_L.b2c_coll_barrier(self._h, phases, _stream(self.device)))
        self.kernel_launches += 1

    def check(self) -> None:
        """Synchronise and raise ``ERR_TIMEOUT`` if a peer never arrived."""
        with torch.cuda.device(self.device):
            _lib.check(_L.b2c_coll_check(self._h, _stream(self.device)))

    def close(self) -> None:
        if self._h:
            _L.b2c_coll_destroy(self._h)
            self._h = None
            self.symbuf.release_if_heap()
            for sym in getattr(self, "_symm", {}).values():
                sym.release_if_heap()


class CollectiveGroup:
    """Virtual ranks in one process: every collective is issued phase by phase over
    all ranks, so no kernel waits for one that has not been launched."""

    def __init__(self, ranks: list[DeviceCollectives]):
        self.ranks = ranks

    def _sync(self):
        for d in {r.device for r in self.ranks}:
            torch.cuda.synchronize(d)

    def allreduce(self, ts: list[torch.Tensor], op="sum", algo="auto") -> list[torch.Tensor]:
        for ph in (1, 2, 4):
            for r, t in zip(self.ranks, ts):
                r.allreduce(t, op, algo, phases=ph)
            self._sync()
        return ts

    def reduce(self, ts, root, op="sum"):
        for ph in (1, 2):
            for r, t in zip(self.ranks, ts):
                r.reduce(t, root, op, phases=ph)
            self._sync()
        return ts[root]

    def _two_phase(self, fn):
        for ph in (1, 2):
            fn(ph)
            self._sync()

    def bcast(self, ts, root=0):
        self._two_phase(lambda ph: [r.bcast(t, root, ph) for r, t in zip(self.ranks, ts)])
        return ts

    def gather(self, sends, recv, root=0):
        self._two_phase(lambda ph: [r.gather(s, recv if r.rank == root else None, root, ph)
                                    for r, s in zip(self.ranks, sends)])
        return recv

    def allgather(self, sends, recvs):
        self._two_phase(lambda ph: [r.allgather(s, o, ph) for r, s, o in zip(self.ranks, sends, recvs)])
        return recvs

    def scatter(self, send, recvs, root=0):
        self._two_phase(lambda ph: [r.scatter(send if r.rank == root else None, o, root, ph)
                                    for r, o in zip(self.ranks, recvs)])
        return recvs

    def barrier(self):
        self._two_phase(lambda ph: [r.barrier(ph) for r in self.ranks])

    def check(self):
        for r in self.ranks:
            r.check()
