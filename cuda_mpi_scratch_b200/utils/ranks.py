"""Per-rank logging, node counting and cross-rank wall-time: the small utilities the
reference's dot-product drivers carry inline.

* ``rank_print``   -- every rank emits ONE string per call so concurrent output does not
  interleave (``mpi7.cpp:56-62``, ``mpi9.cpp:57-62``); ``B2C_NO_LOG=1`` silences the
  per-rank lines like the reference's ``-DNO_LOG`` (``mpicuda2.cu:139-141,183-188``).
* ``node_count``   -- number of distinct hosts in a communicator: names are gathered on the
  root, de-duplicated and the count is broadcast (``-DMPI_RROBIN_``, ``mpicuda2.cu:116-154``);
  feeds the ``round_robin`` device-binding policy.
* ``global_elapsed`` -- ``max(end) - min(begin)`` over all ranks (``mpicuda3.cu:176-179,
  315-325``), with wall-clock time instead of the reference's ``clock()`` CPU time.
* ``block_range``  -- 1-D block decomposition with the reference's divisibility rule
  (``mpicuda2.cu:158-165``).
"""
from __future__ import annotations

import os
import sys

import numpy as np

from ..parallel import comm as cm


def rank_print(comm, *args, file=None, force: bool = False) -> None:
    if os.environ.get("B2C_NO_LOG") == "1" and not force:
        return
    file = sys.stdout if file is None else file      # resolved at call time (stdout may be redirected)
    text = f"{cm.processor_name()} - rank {comm.rank}:\t" + " ".join(str(a) for a in args) + "\n"
    file.write(text)
    file.flush()


def node_count(comm) -> int:
    name = cm.processor_name().encode()[:255].ljust(256, b"\0")
    mine = np.frombuffer(name, dtype=np.uint8).copy()
    allnames = np.zeros(256 * comm.size, dtype=np.uint8) if comm.rank == 0 else None
    comm.gather(mine, allnames, root=0)
    n = np.zeros(1, dtype=np.int32)
    if comm.rank == 0:
        names = {bytes(allnames[256 * r:256 * (r + 1)]).rstrip(b"\0") for r in range(comm.size)}
        n[0] = len(names)
    comm.bcast(n, root=0)
    return int(n[0])


def global_elapsed(comm, begin: float, end: float) -> float:
    """Wall time of a region over ALL ranks: max(end) - min(begin); valid on every rank."""
    lo, hi = np.zeros(1), np.zeros(1)
    comm.allreduce(np.array([begin]), lo, cm.MIN)
    comm.allreduce(np.array([end]), hi, cm.MAX)
    return float(hi[0] - lo[0])


def block_range(n_total: int, rank: int, size: int) -> tuple[int, int]:
    if n_total % size:
        raise ValueError(f"array size {n_total} must be evenly divisible by the number of ranks {size}")
    n = n_total // size
    return rank * n, (rank + 1) * n
