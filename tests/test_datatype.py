"""Datatype engine: flattening and host pack/unpack against the layouts the
reference builds (mpi7.cpp, mpi8.cpp, mpi-complex-types.cpp, stencil2D.h)."""
import ctypes as C

import numpy as np
import pytest

from cuda_mpi_scratch_b200 import _lib
from cuda_mpi_scratch_b200.parallel.datatype import DOUBLE, FLOAT, INT, ORDER_FORTRAN, Datatype


def test_indexed_mpi7_runs():
    # mpi7.cpp:32-41: blocks of 4 @5 and 2 @12 floats -> byte runs [20,16] and [48,8]
    t = FLOAT.indexed([4, 2], [5, 12]).commit()
    assert t.runs() == [(20, 16), (48, 8)]
    assert t.size == 24 and t.extent == 14 * 4 - 20
    a = np.arange(16, dtype=np.float32)
    packed = np.frombuffer(t.pack(a.ctypes.data, 1), dtype=np.float32)
    assert packed.tolist() == [5, 6, 7, 8, 12, 13]


def test_struct_mpi8_particle():
    # mpi8.cpp:13-17,42-54: 4 floats @0, 2 ints @16
    t = Datatype.struct([4, 2], [0, 16], [FLOAT, INT]).commit()
    assert t.runs() == [(0, 24)]          # dense: merged into one run
    assert t.size == 24 and t.extent == 24
    assert t.runs(3) == [(0, 72)]


def test_struct_with_padding_keeps_gaps():
    t = Datatype.struct([1, 1], [0, 8], [INT, DOUBLE]).commit()   # int, 4 pad, double
    assert t.runs() == [(0, 4), (8, 8)]
    assert t.size == 12 and t.extent == 16
    assert t.runs(2) == [(0, 4), (8, 12), (24, 8)]


def test_subarray_1d_and_extent():
    # mpi-complex-types.cpp:33-36: size 8, subsize 3, start 3
    t = INT.subarray([8], [3], [3]).commit()
    assert t.runs() == [(12, 12)]
    assert t.size == 12 and t.extent == 32     # extent of the FULL array


def test_subarray_2d_c_order_is_row_major():
    # 4 rows x 6 cols, take rows 1-2, cols 2-4
    t = DOUBLE.subarray([4, 6], [2, 3], [1, 2]).commit()
    assert t.runs() == [((1 * 6 + 2) * 8, 24), ((2 * 6 + 2) * 8, 24)]
    f = DOUBLE.subarray([6, 4], [3, 2], [2, 1], order=ORDER_FORTRAN).commit()
    assert f.runs() == t.runs()


def test_subarray_full_rows_merge():
    t = DOUBLE.subarray([4, 6], [2, 6], [1, 0]).commit()
    assert t.runs() == [(48, 96)]


def test_hindexed_of_subarray_three_buffers():
    # mpi-complex-types.cpp:38-50: three separately allocated buffers addressed by
    # SIGNED byte displacements from the first one
    b1 = (C.c_int * 8)(*range(8))
    b2 = (C.c_int * 8)(*[2 * i for i in range(8)])
    b3 = (C.c_int * 8)(*[2 * i + 1 for i in range(8)])
    d2 = C.addressof(b2) - C.addressof(b1)
    d3 = C.addressof(b3) - C.addressof(b1)
    sub = INT.subarray([8], [3], [3]).commit()
    t = sub.hindexed([1, 1, 1], [0, d2, d3]).commit()
    assert t.size == 36
    packed = np.frombuffer(t.pack(C.addressof(b1), 1), dtype=np.int32)
    assert packed.tolist() == [3, 4, 5, 6, 8, 10, 7, 9, 11]
    # receiver layout: start 0 instead of 3 (mpi-complex-types.cpp:72)
    r1, r2, r3 = ((C.c_int * 8)(*[-1] * 8) for _ in range(3))
    rsub = INT.subarray([8], [3], [0]).commit()
    rt = rsub.hindexed([1, 1, 1], [0, C.addressof(r2) - C.addressof(r1),
                                   C.addressof(r3) - C.addressof(r1)]).commit()
    rt.unpack(packed.tobytes(), C.addressof(r1), 1)
    assert list(r1) == [3, 4, 5, -1, -1, -1, -1, -1]
    assert list(r2) == [6, 8, 10, -1, -1, -1, -1, -1]
    assert list(r3) == [7, 9, 11, -1, -1, -1, -1, -1]


def test_vector_and_resized():
    col = DOUBLE.vector(4, 1, 6).commit()            # one column of a 4x6 matrix
    assert col.runs() == [(0, 8), (48, 8), (96, 8), (144, 8)]
    r = col.resized(0, 8).commit()                   # consecutive columns
    assert r.runs(2) == [(0, 8), (48, 8), (96, 8), (144, 8), (8, 8), (56, 8), (104, 8), (152, 8)]


def test_halo_column_type_for_big_tile():
    # "derived-datatype halo 4096^2 strided": one ghost column = 4096 runs of 8 B
    t = DOUBLE.subarray([4098, 4098], [4096, 1], [1, 1]).commit()
    runs = t.runs()
    assert len(runs) == 4096 and runs[0] == ((4098 + 1) * 8, 8)
    assert runs[1][0] - runs[0][0] == 4098 * 8


def test_errors():
    with pytest.raises(_lib.CommError):
        INT.subarray([8], [3], [6])                  # does not fit
    with pytest.raises(ValueError):
        INT.indexed([1, 2], [0])                     # length mismatch
    a = np.zeros(4, dtype=np.int32)
    with pytest.raises(_lib.CommError):
        INT.contiguous(2).pack(a.ctypes.data, 1)     # not committed
