#!/usr/bin/env python
"""Baseline for the headline: UNFUSED halo exchange + separate stencil kernel.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29630 bench/stencil_unfused_baseline.py [--tile 16384] [--steps 20] [--json out.json]

What the reference's loop would be on this box with a real Compute() (stencil2d/
mpi-2d-stencil-subarray-cuda.cu:169-172: do { ExchangeData; Compute } while (!Terminate)):
every time step
  1. pack: the 4 edge strips of the core are copied into contiguous send buffers (torch copies;
     the two column strips are strided gathers),
  2. exchange: host-driven NCCL send/recv of the 4 strips (torch.distributed.batch_isend_irecv),
  3. unpack: received strips are copied into the ghost cells of the current buffer,
  4. compute: ONE launch of this repo's own TMA stencil kernel on a plan WITHOUT neighbours
     (no peer stores, no flags, no waits inside the kernel) -- so the comparison isolates the
     fusion, not the quality of the stencil code.
5-point star, periodic Py x Px grid, same tile, same coefficients, same device timing (CUDA
events, max over ranks) as bench.py.  The result is checked against bench.py's fused path:
both must produce the same field bit for bit after K steps.

STATUS: written at the end of round 1 after the GPU budget was spent -- not yet run on a GPU.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.models.stencil2d import Stencil2D, jacobi_coeffs
from cuda_mpi_scratch_b200.ops.unfused_halo import UnfusedHalo


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--tile", type=int, default=16384)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--json", default="")
    a = ap.parse_args()
    world = cms.init()
    n, rank, dev = world.size, world.rank, world.device
    device = f"cuda:{dev}"
    dims = cms.dims_create(n, 2)
    cart = cms.CartTopology.create(dims, True)
    r, c = cart.coords(rank)
    th = tw = a.tile
    g = 1

    # the stencil kernel alone: a plan whose 8 neighbours are absent never pushes or waits
    grp = Stencil2D.create_group(cms.LocalGroup(1, devices=[dev]), (th, tw), 1, "star", dims=(1, 1),
                                 periods=False)
    loc = grp.ranks[0]
    coeffs = jacobi_coeffs(1)
    loc.set_coeffs(coeffs)
    ys = torch.arange(th, device=device, dtype=torch.float64) + r * th
    xs = torch.arange(tw, device=device, dtype=torch.float64) + c * tw
    init = torch.sin(ys * 1e-3)[:, None] * torch.cos(xs * 2e-3)[None, :]
    loc.fill(0.0, 0.0)
    loc.core.copy_(init)

    halo = UnfusedHalo(cart, rank, (th, tw), g, device)

    def exchange():
        halo.exchange(loc.core, loc.tile)         # views of the CURRENT buffer

    def step(k):
        for _ in range(k):
            exchange()
            loc.step(1, drain=False)

    stream = torch.cuda.current_stream(dev)
    step(max(a.warmup, 3))
    torch.cuda.synchronize(dev)
    world.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    step(a.steps)
    e1.record(stream)
    torch.cuda.synchronize(dev)
    ms = world.allreduce_max(e0.elapsed_time(e1))
    loc.check()

    # cross-check against the fused path (same start, same number of steps)
    fused = Stencil2D.create(world, (th, tw), radius=1, kind="star", dims=dims, periods=True)
    fused.fill(0.0, 0.0)
    fused.core.copy_(init)
    fused.set_coeffs(coeffs)
    torch.cuda.synchronize(dev)
    fused.exchange()
    fused.step(max(a.warmup, 3) + a.steps)
    torch.cuda.synchronize(dev)
    same = bool(torch.equal(fused.core, loc.core))
    bad = world.allreduce_sum(0.0 if same else 1.0)

    cells = float(th) * float(tw) * n
    if rank == 0:
        out = {"metric": "2D-stencil Gcell-updates/s, UNFUSED baseline (pack + NCCL send/recv + unpack + separate kernel)",
               "value": round(cells * a.steps / (ms * 1e-3) / 1e9, 3), "unit": "Gcell/s", "n_gpus": n,
               "steps": a.steps, "ms_per_step": round(ms / a.steps, 5), "tile_per_gpu": [th, tw],
               "process_grid": list(dims), "matches_fused_bitwise": bad == 0}
        print(json.dumps(out), flush=True)
        if a.json:
            json.dump(out, open(a.json, "w"), indent=1)
    world.barrier()
    world.finalize()
    return 0 if bad == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
