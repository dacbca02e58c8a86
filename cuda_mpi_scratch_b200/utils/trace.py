"""Host-side call tracing (Chrome trace-event JSON, optional NVTX ranges).

Every communication call of the native library (send/recv/isend/irecv/wait/probe, the
collectives, halo exchanges) and every enqueue of a device operation (fused stencil steps,
dot, puts, device collectives) becomes a timed span with rank, peer, tag and byte count::

    from cuda_mpi_scratch_b200.utils import trace
    trace.enable()                      # or: B200COMM_TRACE=/tmp/run in the environment
    ... run ...
    with trace.region("my phase", rank=world.rank):
        ...
    trace.dump("run.json")              # open in chrome://tracing or ui.perfetto.dev

The reference times by hand: ``MPI_Wtime`` brackets (``test-benchmark/mpi-pingpong-gpu.cpp:51-57``)
and ``clock()`` deltas gathered to rank 0 (``mpicuda3.cu:176-179,315-325``).
"""
from __future__ import annotations

import contextlib
import ctypes as C

from .. import _lib

_L = _lib.lib
_L.b2c_trace_enable.argtypes = [C.c_int]
_L.b2c_trace_clear.argtypes = []
_L.b2c_trace_count.argtypes = [C.POINTER(C.c_int64)]
_L.b2c_trace_dump.argtypes = [C.c_char_p, C.POINTER(C.c_int64)]
_L.b2c_trace_begin.argtypes = [C.c_char_p, C.POINTER(C.c_uint64)]
_L.b2c_trace_end.argtypes = [C.c_char_p, C.c_uint64, C.c_int, C.c_int64]
for _f in ("b2c_trace_enable", "b2c_trace_clear", "b2c_trace_count", "b2c_trace_dump", "b2c_trace_begin",
           "b2c_trace_end"):
    getattr(_L, _f).restype = C.c_int


def enable(record: bool = True, nvtx: bool = False) -> None:
    """Start recording spans (and/or emitting NVTX ranges for nsys / ncu timelines)."""
    _lib.check(_L.b2c_trace_enable((1 if record else 0) | (2 if nvtx else 0)))


def disable() -> None:
    _lib.check(_L.b2c_trace_enable(0))


def clear() -> None:
    _lib.check(_L.b2c_trace_clear())


def count() -> int:
    n = C.c_int64()
    _lib.check(_L.b2c_trace_count(C.byref(n)))
    return n.value


def dump(path: str) -> int:
    """Write everything recorded so far as Chrome trace-event JSON; returns the span count."""
    n = C.c_int64()
    _lib.check(_L.b2c_trace_dump(str(path).encode(), C.byref(n)))
    return n.value


@contextlib.contextmanager
def region(name: str, rank: int = -1, nbytes: int = -1):
    """A user span around a block of Python code (category ``user``)."""
    t0 = C.c_uint64()
    bname = name.encode()
    _lib.check(_L.b2c_trace_begin(bname, C.byref(t0)))
    try:
        yield
    finally:
        _lib.check(_L.b2c_trace_end(bname, t0.value, rank, nbytes))
