#!/usr/bin/env python
"""Host transport micro-benchmarks (no GPU): BASELINE.json config 1, "plumbing".

    python -m cuda_mpi_scratch_b200.launch -n 4 bench/host_transport.py [--json out.json]

ranks = processes over the POSIX shared-memory fabric.  Measures
  * ping-pong round trip between rank 0 and rank 1, 8 B .. 16 MiB (blocking send/recv, the
    pattern of the reference's mpi4.cpp / test-benchmark/mpi-pingpong-gpu.cpp with host buffers);
  * the 8-neighbour halo exchange of the reference's CPU stencil driver
    (stencil2d/mpi-2d-stencil-subarray.cpp: 2 x 2 periodic grid, 5 x 5 stencil) on 16^2 and
    256^2 fp64 tiles, subarray datatypes on both sides;
  * barrier and 8-byte allreduce latency.
Wall clock (time.perf_counter), max over ranks, after warm-up.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np

from cuda_mpi_scratch_b200.parallel import comm as cm
from cuda_mpi_scratch_b200.parallel.cart import dims_create
from cuda_mpi_scratch_b200.parallel.datatype import DOUBLE
from cuda_mpi_scratch_b200.parallel.halo import Array2D, HaloPlan


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", default="")
    a = ap.parse_args()
    world = cm.init_process()
    world.set_timeout(120.0)
    rank, size = world.rank, world.size
    rows = []

    def tmax(dt):
        out = np.zeros(1)
        world.allreduce(np.array([dt]), out, cm.MAX)
        return float(out[0])

    # ---- ping-pong 0 <-> 1
    for nbytes in [8, 64, 1024, 16 << 10, 256 << 10, 4 << 20, 16 << 20]:
        buf = np.zeros(nbytes, dtype=np.uint8)
        iters = 2000 if nbytes <= (16 << 10) else (200 if nbytes <= (256 << 10) else 20)
        world.barrier()
        t = 0.0
        for it in range(iters + 10):
            if it == 10:
                world.barrier()
                t = time.perf_counter()
            if rank == 0:
                world.send(buf, 1, 1)
                world.recv(buf, 1, 1)
            elif rank == 1:
                world.recv(buf, 0, 1)
                world.send(buf, 0, 1)
        dt = (time.perf_counter() - t) / iters if rank < 2 else 0.0
        rtt = tmax(dt)
        rows.append({"op": "pingpong", "bytes": nbytes, "rtt_us": round(rtt * 1e6, 2),
                     "one_way_GBps": round(2 * nbytes / rtt / 1e9, 3)})

    # ---- halo exchange of the CPU stencil driver
    dims = dims_create(size, 2)
    cart = world.cart_create(dims, periods=True)
    for tile in (16, 256):
        g = 2
        tot = tile + 2 * g
        data = np.full((tot, tot), -1.0)
        plan = HaloPlan(cart, Array2D(tot, tot, tot), 5, 5, DOUBLE)
        data[g:-g, g:-g] = rank
        iters = 200
        for _ in range(10):
            plan.exchange(data)
        world.barrier()
        t = time.perf_counter()
        for _ in range(iters):
            plan.exchange(data)
        dt = tmax((time.perf_counter() - t) / iters)
        rows.append({"op": "halo_exchange_8_neighbours", "tile": tile, "grid": list(dims),
                     "us_per_exchange": round(dt * 1e6, 1),
                     "bytes_sent_per_rank": int(sum(e["bytes"] for e in plan.entries("send")))})

    # ---- small collectives
    x, y = np.ones(1), np.zeros(1)
    for name, fn in [("barrier", world.barrier), ("allreduce_8B", lambda: world.allreduce(x, y, cm.SUM))]:
        for _ in range(50):
            fn()
        world.barrier()
        t = time.perf_counter()
        for _ in range(2000):
            fn()
        rows.append({"op": name, "us": round(tmax((time.perf_counter() - t) / 2000) * 1e6, 2)})

    if rank == 0:
        for r in rows:
            print(r, flush=True)
        if a.json:
            json.dump({"ranks": size, "transport": "POSIX shm rings, ranks = processes",
                       "host": cm.processor_name(), "cpus": os.cpu_count(), "rows": rows},
                      open(a.json, "w"), indent=1)
    world.finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
