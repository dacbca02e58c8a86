"""Cartesian topology vs the reference's expectations (mpi10.cpp, stencil2D.h)."""
import pytest

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.parallel.cart import PROC_NULL, CartTopology, dims_create


def test_dims_create():
    assert dims_create(1) == (1, 1)
    assert dims_create(2) == (1, 2)
    assert dims_create(4) == (2, 2)
    assert dims_create(8) == (2, 4)
    assert dims_create(9) == (3, 3)
    assert dims_create(12, 3) == (2, 2, 3)


def test_coords_rank_roundtrip():
    c = CartTopology.create((2, 4), True)
    for r in range(8):
        assert c.rank(c.coords(r)) == r
    assert c.coords(5) == (1, 1)


def test_periodic_wrap_matches_golden_neighbours():
    # stencil2d/sample-output/0_0 "Array after exchange": corners 8 / 7 / 5 / 4,
    # edges 6 (top) 2 (left) 1 (right) 3 (bottom)
    c = CartTopology.create((3, 3), True)
    assert c.neighbours8(0) == [8, 6, 7, 2, 1, 5, 3, 4]
    assert c.neighbours8(4) == [0, 1, 2, 3, 5, 6, 7, 8]


def test_non_periodic_proc_null():
    # mpi10.cpp:22-54: non-periodic grid, edges are MPI_PROC_NULL
    c = CartTopology.create((3, 3), False)
    assert c.neighbours8(0) == [PROC_NULL, PROC_NULL, PROC_NULL, PROC_NULL, 1, PROC_NULL, 3, 4]
    src, dst = c.shift(0, 0, 1)
    assert (src, dst) == (PROC_NULL, 3)
    src, dst = c.shift(4, 1, 1)
    assert (src, dst) == (3, 5)


def test_tiny_periodic_grids_alias_self():
    c = CartTopology.create((1, 1), True)
    assert c.neighbours8(0) == [0] * 8
    c = CartTopology.create((1, 2), True)
    assert c.neighbours8(0) == [1, 0, 1, 1, 1, 1, 0, 1]


def test_grid_string_format():
    c = CartTopology.create((3, 3), True)
    assert c.grid_string() == "0 1 2 \n3 4 5 \n6 7 8 \n"


def test_bad_args():
    with pytest.raises(ValueError):
        CartTopology.create((2, 2), True).coords(4)
    with pytest.raises(ValueError):
        dims_create(0)
    assert cms.PROC_NULL == PROC_NULL
