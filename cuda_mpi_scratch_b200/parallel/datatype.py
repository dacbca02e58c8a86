"""Derived datatypes (Python face of ``include/b200comm/datatype.hpp``).

Capability parity with the MPI datatype calls the reference uses:
``MPI_Type_indexed`` (``mpi7.cpp:32-41``), ``MPI_Type_create_struct`` +
``MPI_Type_extent`` (``mpi8.cpp:38-54``), ``MPI_Type_create_subarray``
(``mpi-complex-types.cpp:33-36``, ``stencil2d/stencil2D.h:210-228``),
``MPI_Type_create_hindexed`` over a subarray type with byte displacements between
separately allocated buffers (``mpi-complex-types.cpp:38-50``),
``MPI_Type_commit`` / ``MPI_Type_free``.

A committed type is a flat list of ``(byte displacement, byte length)`` runs; the
same list drives the host pack/unpack and the device pack/move kernels.
"""
from __future__ import annotations

import ctypes as C
from typing import Sequence

from .. import _lib

_L = _lib.lib
_vp, _i32, _i64 = C.c_void_p, C.c_int, C.c_int64
_pvp = C.POINTER(C.c_void_p)

for _name, _args in {
    "b2c_type_basic": [_i32, _pvp],
    "b2c_type_contiguous": [_i32, _vp, _pvp],
    "b2c_type_vector": [_i32, _i32, _i32, _vp, _pvp],
    "b2c_type_hvector": [_i32, _i32, _i64, _vp, _pvp],
    "b2c_type_indexed": [_i32, C.POINTER(_i32), C.POINTER(_i32), _vp, _pvp],
    "b2c_type_hindexed": [_i32, C.POINTER(_i32), C.POINTER(_i64), _vp, _pvp],
    "b2c_type_struct": [_i32, C.POINTER(_i32), C.POINTER(_i64), _pvp, _pvp],
    "b2c_type_subarray": [_i32, C.POINTER(_i32), C.POINTER(_i32), C.POINTER(_i32), _i32, _vp, _pvp],
    "b2c_type_resized": [_vp, _i64, _i64, _pvp],
    "b2c_type_commit": [_vp],
    "b2c_type_free": [_vp],
    "b2c_type_size": [_vp, C.POINTER(_i64)],
    "b2c_type_extent": [_vp, C.POINTER(_i64), C.POINTER(_i64)],
    "b2c_type_runs": [_vp, _i64, C.POINTER(_i64), _i64, C.POINTER(_i64)],
    "b2c_type_pack": [_vp, _vp, _i64, _vp],
    "b2c_type_unpack": [_vp, _vp, _i64, _vp, _i64],
    "b2c_type_describe": [_vp, C.c_char_p, _i32],
}.items():
    _lib._proto(_name, _args)

ORDER_C, ORDER_FORTRAN = 0, 1
_BASIC = {"byte": 0, "char": 1, "int8": 2, "uint8": 3, "int16": 4, "int32": 5, "int64": 6,
          "uint64": 7, "float32": 8, "float": 8, "float64": 9, "double": 9, "uint32": 10,
          "int": 5}


def _arr(ctype, seq):
    seq = list(seq)
    return (ctype * max(len(seq), 1))(*seq)


class Datatype:
    """Handle to a native datatype.  Build with the class methods, then ``commit()``."""

    def __init__(self, handle: int, keep=()):
        self._h = C.c_void_p(handle)
        self._keep = keep  # constituent types stay alive as long as this one

    # ---- constructors ----------------------------------------------------
    @classmethod
    def basic(cls, kind: str) -> "Datatype":
        h = C.c_void_p()
        _lib.check(_L.b2c_type_basic(_BASIC[kind], C.byref(h)))
        return cls(h.value)

    @classmethod
    def from_torch(cls, dtype) -> "Datatype":
        return cls.basic(str(dtype).replace("torch.", ""))

    def contiguous(self, count: int) -> "Datatype":
        h = C.c_void_p()
        _lib.check(_L.b2c_type_contiguous(count, self._h, C.byref(h)))
        return Datatype(h.value, (self,))

    def vector(self, count: int, blocklen: int, stride: int) -> "Datatype":
        h = C.c_void_p()
        _lib.check(_L.b2c_type_vector(count, blocklen, stride, self._h, C.byref(h)))
        return Datatype(h.value, (self,))

    def hvector(self, count: int, blocklen: int, stride_bytes: int) -> "Datatype":
        h = C.c_void_p()
        _lib.check(_L.b2c_type_hvector(count, blocklen, stride_bytes, self._h, C.byref(h)))
        return Datatype(h.value, (self,))

    def indexed(self, blocklens: Sequence[int], disps: Sequence[int]) -> "Datatype":
        """``MPI_Type_indexed``: displacements in multiples of this type's extent."""
        if len(blocklens) != len(disps):
            raise ValueError("blocklens and disps differ in length")
        h = C.c_void_p()
        _lib.check(_L.b2c_type_indexed(len(blocklens), _arr(_i32, blocklens), _arr(_i32, disps),
                                       self._h, C.byref(h)))
        return Datatype(h.value, (self,))

    def hindexed(self, blocklens: Sequence[int], byte_disps: Sequence[int]) -> "Datatype":
        """``MPI_Type_create_hindexed``: signed 64-bit byte displacements."""
        if len(blocklens) != len(byte_disps):
            raise ValueError("blocklens and byte_disps differ in length")
        h = C.c_void_p()
        _lib.check(_L.b2c_type_hindexed(len(blocklens), _arr(_i32, blocklens),
                                        _arr(_i64, byte_disps), self._h, C.byref(h)))
        return Datatype(h.value, (self,))

    @classmethod
    def struct(cls, blocklens: Sequence[int], byte_disps: Sequence[int],
               types: Sequence["Datatype"]) -> "Datatype":
        """``MPI_Type_create_struct`` (``mpi8.cpp:42-54``)."""
        if not (len(blocklens) == len(byte_disps) == len(types)):
            raise ValueError("struct argument lists differ in length")
        h = C.c_void_p()
        ts = (C.c_void_p * max(len(types), 1))(*[t._h.value for t in types])
        _lib.check(_L.b2c_type_struct(len(blocklens), _arr(_i32, blocklens), _arr(_i64, byte_disps),
                                      C.cast(ts, _pvp), C.byref(h)))
        return cls(h.value, tuple(types))

    def subarray(self, sizes: Sequence[int], subsizes: Sequence[int], starts: Sequence[int],
                 order: int = ORDER_C) -> "Datatype":
        """``MPI_Type_create_subarray``; with ``ORDER_C`` dimension 0 is the slow axis."""
        if not (len(sizes) == len(subsizes) == len(starts)):
            raise ValueError("subarray argument lists differ in length")
        h = C.c_void_p()
        _lib.check(_L.b2c_type_subarray(len(sizes), _arr(_i32, sizes), _arr(_i32, subsizes),
                                        _arr(_i32, starts), order, self._h, C.byref(h)))
        return Datatype(h.value, (self,))

    def resized(self, lb: int, extent: int) -> "Datatype":
        h = C.c_void_p()
        _lib.check(_L.b2c_type_resized(self._h, lb, extent, C.byref(h)))
        return Datatype(h.value, (self,))

    # ---- queries -----------------------------------------------------------
    def commit(self) -> "Datatype":
        _lib.check(_L.b2c_type_commit(self._h))
        return self

    @property
    def size(self) -> int:
        s = getattr(self, "_size", None)     # a datatype is immutable once built
        if s is None:
            v = _i64()
            _lib.check(_L.b2c_type_size(self._h, C.byref(v)))
            s = self._size = v.value
        return s

    @property
    def extent(self) -> int:
        lb, ex = _i64(), _i64()
        _lib.check(_L.b2c_type_extent(self._h, C.byref(lb), C.byref(ex)))
        return ex.value

    @property
    def lb(self) -> int:
        lb, ex = _i64(), _i64()
        _lib.check(_L.b2c_type_extent(self._h, C.byref(lb), C.byref(ex)))
        return lb.value

    def runs(self, count: int = 1) -> list[tuple[int, int]]:
        """Flattened ``(byte displacement, byte length)`` runs of ``count`` elements."""
        n = _i64()
        _lib.check(_L.b2c_type_runs(self._h, count, None, 0, C.byref(n)))
        buf = (_i64 * (2 * max(n.value, 1)))()
        _lib.check(_L.b2c_type_runs(self._h, count, buf, n.value, C.byref(n)))
        return [(buf[2 * i], buf[2 * i + 1]) for i in range(n.value)]

    def describe(self) -> str:
        b = C.create_string_buffer(512)
        _lib.check(_L.b2c_type_describe(self._h, b, 512))
        return b.value.decode()

    def pack(self, buf_ptr: int, count: int) -> bytes:
        out = C.create_string_buffer(max(self.size * count, 1))
        _lib.check(_L.b2c_type_pack(self._h, C.c_void_p(buf_ptr), count, out))
        return out.raw[: self.size * count]

    def unpack(self, packed: bytes, buf_ptr: int, count: int) -> None:
        _lib.check(_L.b2c_type_unpack(self._h, packed, len(packed), C.c_void_p(buf_ptr), count))

    def free(self) -> None:
        if self._h:
            _L.b2c_type_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):  # pragma: no cover
        try:
            self.free()
        except Exception:
            pass


# MPI-style names for the common basic types
BYTE = Datatype.basic("byte")
CHAR = Datatype.basic("char")
INT = Datatype.basic("int32")
INT64 = Datatype.basic("int64")
FLOAT = Datatype.basic("float32")
DOUBLE = Datatype.basic("float64")
