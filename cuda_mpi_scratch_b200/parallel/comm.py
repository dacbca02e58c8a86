"""MPI-like communicators (Python face of ``include/b200comm/comm.hpp``).

Capability parity with the MPI calls the reference exercises:

=====================================  ==========================================
reference                              here
=====================================  ==========================================
``MPI_Init`` ... ``MPI_Finalize``      :func:`init_process` / :func:`run_local`, ``Comm.finalize``
``MPI_Send/Recv/Probe/Get_count``      ``Comm.send/recv/probe``, ``Status.count`` (``mpi3.cpp``, ``mpi4.cpp``)
``MPI_Isend/Irecv/Wait/Waitall``       ``Comm.isend/irecv``, ``Request.wait``, :func:`waitall` (``mpi5.cpp``)
``MPI_Gather/Bcast/Reduce/Allreduce``  same names (``mpi6.cpp``, ``mpi9.cpp``, ``mpicuda2.cu:154,293``)
``MPI_Comm_group/Group_incl/create``   ``Comm.group/create/split/dup`` (``mpi9.cpp:26-54``)
``MPI_Cart_*``                         ``Comm.cart_create/coords/rank/shift`` (``mpi10.cpp``)
=====================================  ==========================================

Buffers may be torch tensors (CPU **or CUDA**), numpy arrays, ``bytearray``,
ctypes arrays or raw addresses.  CUDA tensors never travel through the host:
the receiver maps the sender's allocation and a kernel moves (and
packs/unpacks) the data; only a descriptor uses the host transport.
"""
from __future__ import annotations

import ctypes as C

import numpy as _np
import os
import threading
from typing import Callable, Sequence

from .. import _lib
from .datatype import Datatype

_L = _lib.lib
_vp, _i32, _i64 = C.c_void_p, C.c_int, C.c_int64
_pvp, _pi32 = C.POINTER(C.c_void_p), C.POINTER(C.c_int)

ANY_SOURCE, ANY_TAG, PROC_NULL, UNDEFINED = -1, -1, -2, -32766
SUM, PROD, MAX, MIN, LAND, LOR, BAND, BOR = range(8)


class CStatus(C.Structure):
    _fields_ = [("source", _i32), ("tag", _i32), ("error", _i32), ("bytes", _i64)]


_pst = C.POINTER(CStatus)
for _name, _args in {
    "b2c_init_local": [_i32, C.c_size_t, _pvp],
    "b2c_init_process": [C.c_char_p, _i32, _i32, C.c_size_t, C.c_double, _pvp],
    "b2c_init_env": [_pvp],
    "b2c_comm_finalize": [_vp], "b2c_comm_free": [_vp],
    "b2c_comm_rank": [_vp, _pi32], "b2c_comm_size": [_vp, _pi32],
    "b2c_comm_set_timeout": [_vp, C.c_double], "b2c_abort": [_vp, _i32],
    "b2c_processor_name": [C.c_char_p, _i32],
    "b2c_send": [_vp, _vp, _i64, _vp, _i32, _i32],
    "b2c_recv": [_vp, _vp, _i64, _vp, _i32, _i32, _pst],
    "b2c_isend": [_vp, _vp, _i64, _vp, _i32, _i32, _pvp],
    "b2c_irecv": [_vp, _vp, _i64, _vp, _i32, _i32, _pvp],
    "b2c_wait": [_vp, _pst], "b2c_test": [_vp, _pi32, _pst],
    "b2c_waitall": [_i32, _pvp, _pst],
    "b2c_probe": [_vp, _i32, _i32, _pst], "b2c_iprobe": [_vp, _i32, _i32, _pi32, _pst],
    "b2c_get_count": [_pst, _vp, C.POINTER(_i64)],
    "b2c_sendrecv": [_vp, _vp, _i64, _vp, _i32, _i32, _vp, _i64, _vp, _i32, _i32, _pst],
    "b2c_barrier": [_vp], "b2c_bcast": [_vp, _vp, _i64, _vp, _i32],
    "b2c_reduce": [_vp, _vp, _vp, _i64, _vp, _i32, _i32],
    "b2c_allreduce": [_vp, _vp, _vp, _i64, _vp, _i32],
    "b2c_gather": [_vp, _vp, _i64, _vp, _vp, _i64, _vp, _i32],
    "b2c_scatter": [_vp, _vp, _i64, _vp, _vp, _i64, _vp, _i32],
    "b2c_allgather": [_vp, _vp, _i64, _vp, _vp, _i64, _vp],
    "b2c_comm_group": [_vp, _pi32, _i32, _pi32],
    "b2c_comm_create": [_vp, _i32, _pi32, _pvp],
    "b2c_comm_split": [_vp, _i32, _i32, _pvp], "b2c_comm_dup": [_vp, _pvp],
    "b2c_cart_create": [_vp, _i32, _pi32, _pi32, _i32, _pvp],
    "b2c_cart_coords": [_vp, _i32, _i32, _pi32], "b2c_cart_rank": [_vp, _i32, _pi32, _pi32],
    "b2c_cart_shift": [_vp, _i32, _i32, _pi32, _pi32],
    "b2c_cart_get": [_vp, _i32, _pi32, _pi32, _pi32],
    "b2c_dims_create": [_i32, _i32, _pi32],
}.items():
    _lib._proto(_name, _args)
_L.b2c_wtime.restype = C.c_double


class Status:
    def __init__(self, cs: CStatus | None = None):
        self.source = cs.source if cs else ANY_SOURCE
        self.tag = cs.tag if cs else ANY_TAG
        self.error = cs.error if cs else 0
        self.bytes = cs.bytes if cs else 0

    def count(self, datatype: Datatype) -> int:
        """``MPI_Get_count`` (``mpi3.cpp:28-31``)."""
        s = datatype.size
        if s == 0:
            return 0
        return self.bytes // s if self.bytes % s == 0 else UNDEFINED

    def __repr__(self):
        return f"Status(source={self.source}, tag={self.tag}, bytes={self.bytes}, error={self.error})"


_dtype_cache: dict = {}   # numpy dtype / torch dtype / name -> committed basic Datatype (kept forever)


def _basic_for(key, name: str | None = None) -> Datatype:
    """Basic datatype for a numpy/torch dtype, created once: building a native type per call
    cost ~10 us, ten times the transport's own latency."""
    dt = _dtype_cache.get(key)
    if dt is None:
        dt = Datatype.basic(name if name is not None else str(key).replace("torch.", ""))
        _dtype_cache[key] = dt
    return dt


def _bufinfo(buf, count=None, datatype: Datatype | None = None):
    """-> (address, count, Datatype, keepalive)."""
    keep = buf
    dt = datatype
    nbytes = None
    if type(buf) is _np.ndarray:             # fast path: the common case of the host API
        try:                                 # 0.7 us; __array_interface__ builds a dict (2.3 us)
            addr = C.addressof(C.c_char.from_buffer(buf))
        except (TypeError, ValueError):      # read-only, empty or strided array
            addr = buf.__array_interface__["data"][0]
        if dt is None:
            dt = _basic_for(buf.dtype)
            return addr, (buf.nbytes // max(dt.size, 1)) if count is None else count, dt, buf
        return addr, 1 if count is None else count, dt, buf
    if buf is None:
        return None, 0 if count is None else count, dt or _basic_for("byte"), None
    if isinstance(buf, int):
        addr = buf
    elif hasattr(buf, "data_ptr"):  # torch tensor
        addr = buf.data_ptr()
        nbytes = buf.numel() * buf.element_size()
        if dt is None:
            if not buf.is_contiguous():
                raise ValueError("non-contiguous tensor needs an explicit datatype")
            dt = _basic_for(buf.dtype)
    elif hasattr(buf, "__array_interface__"):  # numpy
        addr = buf.__array_interface__["data"][0]
        nbytes = buf.nbytes
        if dt is None:
            dt = _basic_for(buf.dtype)
    elif isinstance(buf, (bytes,)):
        keep = C.create_string_buffer(buf, len(buf))
        addr = C.addressof(keep)
        nbytes = len(buf)
        dt = dt or _basic_for("char")
    elif isinstance(buf, (bytearray, memoryview)):
        keep = (C.c_char * len(buf)).from_buffer(buf)
        addr = C.addressof(keep)
        nbytes = len(buf)
        dt = dt or _basic_for("char")
    elif isinstance(buf, C.Array) or isinstance(buf, C.Structure):
        addr = C.addressof(buf)
        nbytes = C.sizeof(buf)
        dt = dt or _basic_for("byte")
    else:
        raise TypeError(f"unsupported buffer type {type(buf)}")
    if count is None:
        if nbytes is None:
            raise ValueError("count is required for raw addresses")
        count = nbytes // max(dt.size, 1) if datatype is None else 1
    return addr, count, dt, keep


def _sync_if_cuda(buf) -> None:
    """A CUDA tensor is handed to the transport by ADDRESS: whatever produces it on
    the current torch stream must be complete first (the rule of CUDA-aware MPI)."""
    if getattr(buf, "is_cuda", False):
        import torch
        torch.cuda.current_stream(buf.device).synchronize()


def device_eager_stats() -> dict:
    """Which path the device sends of this process took so far (``b2c_dev_eager_stats``): small
    contiguous device buffers go through the receiver's device mailbox (``eager``), the rest by
    rendezvous, for the reason counted."""
    out = (C.c_ulonglong * 5)()
    _lib.check(_lib.lib.b2c_dev_eager_stats(out))
    return dict(zip(("eager", "too_large", "no_mailbox", "no_slot", "other_device"), [int(v) for v in out]))


class Request:
    """``MPI_Request``; keeps the buffer alive until completion."""

    def __init__(self, handle: int, keep):
        self._h = C.c_void_p(handle)
        self._keep = keep
        self.status: Status | None = None

    def wait(self) -> Status:
        if self._h:
            cs = CStatus()
            h, self._h = self._h, C.c_void_p()
            rc = _L.b2c_wait(h, C.byref(cs))
            self.status = Status(cs)
            self._keep = None
            _lib.check(rc)
        return self.status or Status()

    def test(self) -> bool:
        if not self._h:
            return True
        cs, flag = CStatus(), C.c_int(0)
        _lib.check(_L.b2c_test(self._h, C.byref(flag), C.byref(cs)))
        if flag.value:
            self._h = C.c_void_p()
            self.status = Status(cs)
            self._keep = None
        return bool(flag.value)


def waitall(requests: Sequence[Request]) -> list[Status]:
    """``MPI_Waitall`` (``mpi5.cpp:72-75``, ``stencil2d/stencil2D.h:375-376``)."""
    return [r.wait() for r in requests]


class Comm:
    def __init__(self, handle: int):
        self._h = C.c_void_p(handle)
        r, s = C.c_int(), C.c_int()
        _lib.check(_L.b2c_comm_rank(self._h, C.byref(r)))
        _lib.check(_L.b2c_comm_size(self._h, C.byref(s)))
        self.rank, self.size = r.value, s.value

    # ---- point to point ------------------------------------------------------
    def send(self, buf, dest: int, tag: int = 0, count=None, datatype=None) -> None:
        a, n, dt, keep = _bufinfo(buf, count, datatype)
        _sync_if_cuda(buf)
        _lib.check(_L.b2c_send(self._h, a, n, dt._h, dest, tag))

    def recv(self, buf, source: int = ANY_SOURCE, tag: int = ANY_TAG, count=None,
             datatype=None) -> Status:
        a, n, dt, keep = _bufinfo(buf, count, datatype)
        cs = CStatus()
        rc = _L.b2c_recv(self._h, a, n, dt._h, source, tag, C.byref(cs))
        _lib.check(rc)
        return Status(cs)

    def isend(self, buf, dest: int, tag: int = 0, count=None, datatype=None) -> Request:
        a, n, dt, keep = _bufinfo(buf, count, datatype)
        _sync_if_cuda(buf)
        h = C.c_void_p()
        _lib.check(_L.b2c_isend(self._h, a, n, dt._h, dest, tag, C.byref(h)))
        return Request(h.value, (keep, dt))

    def irecv(self, buf, source: int = ANY_SOURCE, tag: int = ANY_TAG, count=None,
              datatype=None) -> Request:
        a, n, dt, keep = _bufinfo(buf, count, datatype)
        h = C.c_void_p()
        _lib.check(_L.b2c_irecv(self._h, a, n, dt._h, source, tag, C.byref(h)))
        return Request(h.value, (keep, dt))

    def probe(self, source: int = ANY_SOURCE, tag: int = ANY_TAG) -> Status:
        cs = CStatus()
        _lib.check(_L.b2c_probe(self._h, source, tag, C.byref(cs)))
        return Status(cs)

    def iprobe(self, source: int = ANY_SOURCE, tag: int = ANY_TAG) -> Status | None:
        cs, flag = CStatus(), C.c_int()
        _lib.check(_L.b2c_iprobe(self._h, source, tag, C.byref(flag), C.byref(cs)))
        return Status(cs) if flag.value else None

    def sendrecv(self, sendbuf, dest: int, recvbuf, source: int, sendtag: int = 0,
                 recvtag: int = ANY_TAG) -> Status:
        sa, sn, sdt, k1 = _bufinfo(sendbuf)
        ra, rn, rdt, k2 = _bufinfo(recvbuf)
        cs = CStatus()
        _lib.check(_L.b2c_sendrecv(self._h, sa, sn, sdt._h, dest, sendtag, ra, rn, rdt._h, source,
                                   recvtag, C.byref(cs)))
        return Status(cs)

    # ---- collectives ---------------------------------------------------------
    def barrier(self) -> None:
        _lib.check(_L.b2c_barrier(self._h))

    def bcast(self, buf, root: int = 0, count=None, datatype=None) -> None:
        a, n, dt, keep = _bufinfo(buf, count, datatype)
        _lib.check(_L.b2c_bcast(self._h, a, n, dt._h, root))

    def reduce(self, sendbuf, recvbuf, op: int = SUM, root: int = 0) -> None:
        sa, sn, sdt, k1 = _bufinfo(sendbuf)
        ra, _, _, k2 = _bufinfo(recvbuf) if recvbuf is not None else (None, 0, None, None)
        _lib.check(_L.b2c_reduce(self._h, sa, ra, sn, sdt._h, op, root))

    def allreduce(self, sendbuf, recvbuf, op: int = SUM) -> None:
        sa, sn, sdt, k1 = _bufinfo(sendbuf)
        ra, _, _, k2 = _bufinfo(recvbuf)
        _lib.check(_L.b2c_allreduce(self._h, sa, ra, sn, sdt._h, op))

    def gather(self, sendbuf, recvbuf, root: int = 0) -> None:
        """Non-root ranks may pass ``recvbuf=None`` (``mpi6.cpp:96-100``)."""
        sa, sn, sdt, k1 = _bufinfo(sendbuf)
        if recvbuf is None:
            _lib.check(_L.b2c_gather(self._h, sa, sn, sdt._h, None, 0, None, root))
        else:
            ra, _, rdt, k2 = _bufinfo(recvbuf)
            _lib.check(_L.b2c_gather(self._h, sa, sn, sdt._h, ra, sn, rdt._h, root))

    def scatter(self, sendbuf, recvbuf, root: int = 0) -> None:
        ra, rn, rdt, k2 = _bufinfo(recvbuf)
        if sendbuf is None:
            _lib.check(_L.b2c_scatter(self._h, None, 0, None, ra, rn, rdt._h, root))
        else:
            sa, _, sdt, k1 = _bufinfo(sendbuf)
            _lib.check(_L.b2c_scatter(self._h, sa, rn, sdt._h, ra, rn, rdt._h, root))

    def allgather(self, sendbuf, recvbuf) -> None:
        sa, sn, sdt, k1 = _bufinfo(sendbuf)
        ra, _, rdt, k2 = _bufinfo(recvbuf)
        _lib.check(_L.b2c_allgather(self._h, sa, sn, sdt._h, ra, sn, rdt._h))

    # ---- groups / sub-communicators --------------------------------------------
    def group(self) -> list[int]:
        """World ranks of the members (``MPI_Comm_group``)."""
        n = C.c_int()
        buf = (C.c_int * 256)()
        _lib.check(_L.b2c_comm_group(self._h, buf, 256, C.byref(n)))
        return list(buf[: n.value])

    def create(self, ranks: Sequence[int]) -> "Comm | None":
        """``MPI_Comm_group`` + ``MPI_Group_incl`` + ``MPI_Comm_create``
        (``mpi9.cpp:26-46``); collective over this communicator; ``None`` on
        non-members."""
        h = C.c_void_p()
        arr = (C.c_int * max(len(ranks), 1))(*ranks)
        _lib.check(_L.b2c_comm_create(self._h, len(ranks), arr, C.byref(h)))
        return Comm(h.value) if h.value else None

    def split(self, color: int, key: int = 0) -> "Comm | None":
        h = C.c_void_p()
        _lib.check(_L.b2c_comm_split(self._h, color, key, C.byref(h)))
        return Comm(h.value) if h.value else None

    def dup(self) -> "Comm":
        h = C.c_void_p()
        _lib.check(_L.b2c_comm_dup(self._h, C.byref(h)))
        return Comm(h.value)

    # ---- Cartesian topology ------------------------------------------------------
    def cart_create(self, dims: Sequence[int], periods: Sequence[bool] | bool = False,
                    reorder: bool = False) -> "Comm | None":
        if isinstance(periods, bool):
            periods = [periods] * len(dims)
        d = (C.c_int * len(dims))(*dims)
        p = (C.c_int * len(dims))(*[int(bool(x)) for x in periods])
        h = C.c_void_p()
        _lib.check(_L.b2c_cart_create(self._h, len(dims), d, p, int(reorder), C.byref(h)))
        return Comm(h.value) if h.value else None

    def cart_get(self) -> tuple[list[int], list[bool]]:
        n = C.c_int()
        d, p = (C.c_int * 8)(), (C.c_int * 8)()
        _lib.check(_L.b2c_cart_get(self._h, 8, C.byref(n), d, p))
        return list(d[: n.value]), [bool(x) for x in p[: n.value]]

    def cart_coords(self, rank: int | None = None) -> list[int]:
        nd = len(self.cart_get()[0])
        c = (C.c_int * max(nd, 1))()
        _lib.check(_L.b2c_cart_coords(self._h, self.rank if rank is None else rank, nd, c))
        return list(c[:nd])

    def cart_rank(self, coords: Sequence[int]) -> int:
        c = (C.c_int * len(coords))(*coords)
        r = C.c_int()
        _lib.check(_L.b2c_cart_rank(self._h, len(coords), c, C.byref(r)))
        return r.value

    def cart_shift(self, dim: int, disp: int = 1) -> tuple[int, int]:
        s, d = C.c_int(), C.c_int()
        _lib.check(_L.b2c_cart_shift(self._h, dim, disp, C.byref(s), C.byref(d)))
        return s.value, d.value

    # ---- life cycle ------------------------------------------------------------------
    def set_timeout(self, seconds: float) -> None:
        """Blocking calls raise ``CommError(ERR_TIMEOUT)`` after this long (0 = never)."""
        _lib.check(_L.b2c_comm_set_timeout(self._h, seconds))

    def abort(self, code: int = 1) -> None:
        _L.b2c_abort(self._h, code)

    def free(self) -> None:
        if self._h:
            _L.b2c_comm_free(self._h)
            self._h = C.c_void_p()

    def finalize(self) -> None:
        """Barrier + release (``MPI_Finalize`` for the world communicator)."""
        if self._h:
            h, self._h = self._h, C.c_void_p()
            _lib.check(_L.b2c_comm_finalize(h))


def dims_create(nranks: int, ndims: int = 2) -> list[int]:
    d = (C.c_int * ndims)()
    _lib.check(_L.b2c_dims_create(nranks, ndims, d))
    return list(d)


def wtime() -> float:
    return _L.b2c_wtime()


def processor_name() -> str:
    b = C.create_string_buffer(256)
    _lib.check(_L.b2c_processor_name(b, 256))
    return b.value.decode()


def init_local(nranks: int, ring_bytes: int = 4 << 20) -> list[Comm]:
    """World communicators for ``nranks`` ranks living in THIS process (hand one
    to each thread)."""
    hs = (C.c_void_p * nranks)()
    _lib.check(_L.b2c_init_local(nranks, ring_bytes, hs))
    return [Comm(h) for h in hs]


def init_process(job: str | None = None, rank: int | None = None, size: int | None = None,
                 ring_bytes: int = 8 << 20, timeout_s: float = 60.0) -> Comm:
    """World communicator for one-rank-per-process jobs.  Without arguments the
    rank/size/job come from the launcher's environment (torchrun's RANK /
    WORLD_SIZE / MASTER_PORT, Open MPI's OMPI_COMM_WORLD_*, PMI_*, SLURM_*, or
    B200_RANK / B200_WORLD_SIZE / B200_JOB set by ``python -m
    cuda_mpi_scratch_b200.launch``)."""
    h = C.c_void_p()
    if rank is None:
        _lib.check(_L.b2c_init_env(C.byref(h)))
    else:
        _lib.check(_L.b2c_init_process((job or "b200comm").encode(), rank, size, ring_bytes,
                                       timeout_s, C.byref(h)))
    return Comm(h.value)


def run_local(nranks: int, fn: Callable, *args, timeout_s: float = 60.0, **kwargs) -> list:
    """Run ``fn(comm, *args)`` on ``nranks`` thread-ranks and return the results in
    rank order (re-raises the first exception).  This is how multi-rank logic is
    tested on a box with neither MPI nor GPUs."""
    comms = init_local(nranks)
    for c in comms:
        c.set_timeout(timeout_s)
    results: list = [None] * nranks
    errors: list = [None] * nranks

    def body(r):
        try:
            results[r] = fn(comms[r], *args, **kwargs)
        except BaseException as e:  # noqa: BLE001
            errors[r] = e

    threads = [threading.Thread(target=body, args=(r,), daemon=True) for r in range(nranks)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout_s * 2 + 5)
    for r, t in enumerate(threads):
        if t.is_alive():
            raise TimeoutError(f"rank {r} did not finish")
    for e in errors:
        if e is not None:
            raise e
    for c in comms:
        c.free()
    return results
