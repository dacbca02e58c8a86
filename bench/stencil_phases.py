#!/usr/bin/env python
"""Where a sweep of the temporal strip kernel spends its time, per phase, on N GPUs.

    B2C_STENCIL_DEBUG=32 python bench/stencil_phases.py                       # 1 GPU
    B2C_STENCIL_DEBUG=32 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 \
        --master-addr 127.0.0.1 --master-port 29640 bench/stencil_phases.py    # 2 GPUs

Debug bit 32 makes every CTA record `globaltimer` stamps (start, end of its boundary segments,
end) and its producer warp the time it spent waiting for halo flags.  Every CTA does its boundary
segments FIRST, so `boundary` is the part of a sweep in which peer stores and signals happen; the
difference between 1 and N GPUs in that phase, and in the flag waits, is what the NVLink exchange
costs.  One JSON line per rank."""
from __future__ import annotations

import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.models.stencil2d import Stencil2D
from cuda_mpi_scratch_b200.utils.timing import ClockSampler


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--tile", type=int, default=16384)
    ap.add_argument("--sweeps", type=int, default=40)
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--th", type=int, default=0, help="explicit tile height (with --tw)")
    ap.add_argument("--tw", type=int, default=0)
    a = ap.parse_args()
    assert os.environ.get("B2C_STENCIL_DEBUG") == "32", "run with B2C_STENCIL_DEBUG=32"
    world = cms.init()
    dims = cms.dims_create(world.size, 2)
    th, tw = (a.tile, a.tile) if a.scaling == "weak" else (a.tile // dims[0], a.tile // dims[1])
    if a.th and a.tw:
        th, tw = a.th, a.tw
    st = Stencil2D.create(world, (th, tw), radius=1, dims=dims, periods=True, temporal=3)
    dev = st.device
    st.fill(0.0, 0.0)
    st.core.copy_(torch.rand(th, tw, dtype=torch.float64, device=f"cuda:{dev}"))
    torch.cuda.synchronize(dev)
    st.exchange()
    sampler = ClockSampler(dev).start()
    sampler.wait_first(2.0)
    st.step(3 * 5)
    torch.cuda.synchronize(dev)
    world.barrier()
    acc = {}
    sampler.window_begin()
    for _ in range(a.sweeps):
        st.step(3, drain=False)
        for k, v in st.phase_stats().items():        # synchronises: stamps of the sweep just run
            acc.setdefault(k, []).append(v)
    sampler.window_end()
    st.drain()
    st.check()
    med = {k: sorted(v)[len(v) // 2] for k, v in acc.items()}
    print(json.dumps({"rank": world.rank, "n_gpus": world.size, "tile_per_gpu": [th, tw], "sweeps": a.sweeps,
                      "median_over_sweeps": med, "clocks": sampler.stop(),
                      "note": "host synchronises after every sweep here (stamps are read back), so ranks "
                              "start each sweep within a launch latency of each other"}), flush=True)
    world.barrier()
    st.close()
    world.finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
