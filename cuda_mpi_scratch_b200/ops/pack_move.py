"""Datatype-driven device moves: pack + move + unpack in one kernel.

``typed_move(src, stype, dst, dtype)`` copies ``scount`` elements laid out as
``stype`` at ``src`` into ``dcount`` elements laid out as ``dtype`` at ``dst``.
Either side may be a peer-mapped address (``SymmetricBuffer.peer_ptr``), so a
strided ghost column can be gathered locally and scattered straight into a
neighbour's memory.  Reference capability: derived datatypes on device pointers
handed to a CUDA-aware MPI (``stencil2d/stencil2D.h:369-373``), where the MPI
library packs, transfers and unpacks in separate internal steps.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _lib
from ..parallel.datatype import Datatype

_L = _lib.lib
_vp, _i64 = C.c_void_p, C.c_int64
_lib._proto("b2c_typed_move", [_vp, _vp, _i64, _vp, _vp, _i64, _vp, C.POINTER(_i64)])
_lib._proto("b2c_move_uniform", [_vp, _vp, _i64, _i64, _i64, _i64, _vp])


def _ptr(x) -> int:
    return x.data_ptr() if hasattr(x, "data_ptr") else int(x)


def typed_move(src, stype: Datatype, dst, dtype: Datatype, scount: int = 1, dcount: int = 1,
               device: int | None = None) -> int:
    """Enqueue the move on the current stream; returns the payload bytes."""
    if device is None:
        device = src.device.index if hasattr(src, "device") else torch.cuda.current_device()
    moved = _i64()
    with torch.cuda.device(device):
        _lib.check(_L.b2c_typed_move(C.c_void_p(_ptr(src)), stype._h, scount, C.c_void_p(_ptr(dst)),
                                     dtype._h, dcount,
                                     torch.cuda.current_stream(device).cuda_stream, C.byref(moved)))
    return moved.value


def move_uniform(src, dst, nruns: int, run_bytes: int, src_stride: int, dst_stride: int,
                 device: int | None = None) -> None:
    """Closed-form strided move: ``nruns`` runs of ``run_bytes``."""
    if device is None:
        device = torch.cuda.current_device()
    with torch.cuda.device(device):
        _lib.check(_L.b2c_move_uniform(C.c_void_p(_ptr(src)), C.c_void_p(_ptr(dst)), nruns, run_bytes,
                                       src_stride, dst_stride,
                                       torch.cuda.current_stream(device).cuda_stream))
