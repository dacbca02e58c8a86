"""Cartesian process topology (pure index arithmetic, no communication).

Capability parity with the reference's use of ``MPI_Cart_create`` /
``MPI_Cart_coords`` / ``MPI_Cart_rank`` / ``MPI_Cart_shift`` (``mpi10.cpp:22-54``
non-periodic, ``stencil2d/mpi-2d-stencil-subarray.cpp:43-52`` periodic) and of
``OffsetTaskId`` / ``MPIOffsetRegion`` (``stencil2d/stencil2D.h:232-299``).

One unambiguous convention is used everywhere (SURVEY.md 2.7 #1 explains the
reference's double transposition): **cart dimension 0 is the row (slow) axis,
dimension 1 the column (fast) axis; rank = row * ncols + col**.  The reference
requires a perfect-square rank count (``mpi-2d-stencil-subarray.cpp:44-47``);
:func:`dims_create` lifts that (1x2, 2x4, ...).

The native comm engine has the same logic in C++ (``csrc/comm/topology.cpp``);
this module is what the stencil planner and the CPU tests use.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Sequence

PROC_NULL = -2  # no-op peer (MPI_PROC_NULL, mpi10.cpp:45-51)

# Neighbour directions in the order of the reference's MPIGridCellID
# (stencil2d/stencil2D.h:86-88).  (drow, dcol) with row 0 at the top.
DIRECTIONS = ("NW", "N", "NE", "W", "E", "SW", "S", "SE")
DIR_OFFSETS = ((-1, -1), (-1, 0), (-1, 1), (0, -1), (0, 1), (1, -1), (1, 0), (1, 1))


def opposite(d: int) -> int:
    return 7 - d


def dims_create(nranks: int, ndims: int = 2) -> tuple[int, ...]:
    """Balanced factorisation of ``nranks`` into ``ndims`` factors, non-decreasing
    (``MPI_Dims_create``-like; 2 -> (1, 2), 8 -> (2, 4), 9 -> (3, 3))."""
    if nranks < 1 or ndims < 1:
        raise ValueError("nranks and ndims must be positive")
    dims = [1] * ndims
    n = nranks
    primes = []
    p = 2
    while p * p <= n:
        while n % p == 0:
            primes.append(p)
            n //= p
        p += 1
    if n > 1:
        primes.append(n)
    for f in sorted(primes, reverse=True):
        i = dims.index(min(dims))
        dims[i] *= f
    return tuple(sorted(dims))


@dataclass(frozen=True)
class CartTopology:
    dims: tuple[int, ...]
    periods: tuple[bool, ...]

    def __post_init__(self):
        if len(self.dims) != len(self.periods):
            raise ValueError("dims and periods must have the same length")
        if any(d < 1 for d in self.dims):
            raise ValueError("dims must be positive")

    @staticmethod
    def create(dims: Sequence[int], periods: Sequence[bool] | bool = True) -> "CartTopology":
        dims = tuple(int(d) for d in dims)
        if isinstance(periods, bool):
            periods = (periods,) * len(dims)
        return CartTopology(dims, tuple(bool(p) for p in periods))

    @property
    def size(self) -> int:
        n = 1
        for d in self.dims:
            n *= d
        return n

    @property
    def ndims(self) -> int:
        return len(self.dims)

    def coords(self, rank: int) -> tuple[int, ...]:
        """``MPI_Cart_coords``: row-major, dimension 0 slowest."""
        if not 0 <= rank < self.size:
            raise ValueError(f"rank {rank} outside communicator of size {self.size}")
        out = []
        for d in reversed(self.dims):
            out.append(rank % d)
            rank //= d
        return tuple(reversed(out))

    def rank(self, coords: Sequence[int]) -> int:
        """``MPI_Cart_rank``: periodic dimensions wrap; leaving a non-periodic
        dimension yields :data:`PROC_NULL`."""
        if len(coords) != self.ndims:
            raise ValueError("wrong number of coordinates")
        r = 0
        for c, d, p in zip(coords, self.dims, self.periods):
            if p:
                c %= d
            elif not 0 <= c < d:
                return PROC_NULL
            r = r * d + c
        return r

    def shift(self, rank: int, dim: int, disp: int = 1) -> tuple[int, int]:
        """``MPI_Cart_shift``: returns ``(source, dest)`` for a shift by ``disp``
        along ``dim`` (``mpi10.cpp:41-45``)."""
        c = list(self.coords(rank))
        up = list(c)
        up[dim] += disp
        dn = list(c)
        dn[dim] -= disp
        return self.rank(dn), self.rank(up)

    def neighbour(self, rank: int, drow: int, dcol: int) -> int:
        """Rank at offset (drow, dcol) on a 2-D grid (``OffsetTaskId``,
        ``stencil2d/stencil2D.h:232-244``)."""
        if self.ndims != 2:
            raise ValueError("neighbour() needs a 2-D topology")
        r, c = self.coords(rank)
        return self.rank((r + drow, c + dcol))

    def neighbours8(self, rank: int) -> list[int]:
        """The 8 neighbours in MPIGridCellID order (NW N NE W E SW S SE)."""
        return [self.neighbour(rank, dr, dc) for dr, dc in DIR_OFFSETS]

    def grid_string(self) -> str:
        """Rank layout, one row per line -- the format of ``PrintCartesianGrid``
        (``stencil2d/stencil2D.h:513-530``): each entry followed by a space."""
        if self.ndims != 2:
            raise ValueError("grid_string() needs a 2-D topology")
        rows, cols = self.dims
        return "".join("".join(f"{r * cols + c} " for c in range(cols)) + "\n" for r in range(rows))
