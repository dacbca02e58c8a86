#!/usr/bin/env python
"""Distributed dot product, 2^30 fp64 elements split over the ranks (BASELINE.json config 4).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node P --master-addr 127.0.0.1 \
        --master-port 29610 bench/dot.py [--log2n 30] [--dtype float64]

ours      one fused kernel per rank: local reduction + one-shot P2P allreduce, result on
          every GPU, nothing on the host (csrc/kernels/dot.cu)
baseline  the reference's path re-created on this box: reference-style kernel (variant 2,
          "last block finishes", mpicuda4.cu:157-185) + D2H of the scalar + allreduce of the
          host scalar over the host transport (the MPI_Reduce of mpicuda4.cu:368), and the
          same with ncclAllReduce on the device scalar.
All-ones inputs must give exactly N (mpicuda2.cu:296-298).  Device-timed, max over ranks.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200 import _lib
from cuda_mpi_scratch_b200.models.dot import _DT, DistributedDot
from cuda_mpi_scratch_b200.utils.timing import ClockSampler


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--log2n", type=int, default=30)
    ap.add_argument("--dtype", default="float64")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--json", default="")
    a = ap.parse_args()
    world = cms.init()
    P, rank, dev = world.size, world.rank, world.device
    dtype = getattr(torch, a.dtype)
    N = 1 << a.log2n
    assert N % P == 0, "ARRAY_SIZE must divide evenly (mpicuda2.cu:161-165)"
    n = N // P
    x = torch.ones(n, dtype=dtype, device=f"cuda:{dev}")
    y = torch.ones(n, dtype=dtype, device=f"cuda:{dev}")
    stream = torch.cuda.current_stream(dev)
    d = DistributedDot.create(world)
    out = []

    sampler = ClockSampler(dev).start()
    sampler.wait_first(2.0)
    clocks = []

    def timed(fn, sync_each=False):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(dev); world.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler.window_begin()
        t0 = time.perf_counter()
        e0.record(stream)
        for _ in range(a.iters):
            fn()
        e1.record(stream)
        torch.cuda.synchronize(dev)
        wall = (time.perf_counter() - t0) / a.iters * 1e3
        sampler.window_end()
        clocks.append(sampler.pop_window_summary())
        ms = e0.elapsed_time(e1) / a.iters
        return world.allreduce_max(ms), world.allreduce_max(wall)

    # read-only HBM peak of THIS box: the dot product only reads, so the copy peak of
    # MEASURED_PEAKS.json (read + write) is the wrong denominator for it
    big = torch.empty(1 << 28, dtype=torch.float64, device=f"cuda:{dev}").fill_(1.0)      # 2 GiB
    torch.cuda.synchronize(dev)
    ms_r, _ = timed(lambda: big.sum())
    read_peak = big.numel() * 8 / (ms_r * 1e-3) / 1e9
    clocks.clear()
    del big

    # ---- ours: fused, stream-ordered, no host involvement
    ms, wall = timed(lambda: d.launch(x, y))
    val = d.result()
    assert val == float(N), (val, N)
    bytes_per_gpu = 2 * n * x.element_size()
    out.append({"impl": "fused dot+allreduce (ours)", "ms": ms, "wall_ms": wall, "value": val})

    # ---- baseline: reference-style kernel + D2H + host allreduce
    blocks = _lib.lib.b2c_baseline_dot_blocks(n)
    part = torch.zeros(blocks, dtype=dtype, device=f"cuda:{dev}")
    res = {}

    def ref_host():
        _lib.check(_lib.lib.b2c_baseline_dot(2, x.data_ptr(), y.data_ptr(), n, _DT[dtype],
                                             part.data_ptr(), stream.cuda_stream))
        local = float(part[0].item())                       # D2H of one scalar (mpicuda4.cu:354)
        res["v"] = world.allreduce_sum(local)               # MPI_Reduce stand-in (host, gloo)
    ms, wall = timed(ref_host)
    assert res["v"] == float(N)
    out.append({"impl": "reference-style kernel + D2H + host allreduce", "ms": wall, "wall_ms": wall,
                "value": res["v"], "note": "host-synchronous: wall clock is the honest number"})

    def ref_nccl():
        _lib.check(_lib.lib.b2c_baseline_dot(2, x.data_ptr(), y.data_ptr(), n, _DT[dtype],
                                             part.data_ptr(), stream.cuda_stream))
        if P > 1:
            dist.all_reduce(part[:1])
    ms, wall = timed(ref_nccl)
    v = float(part[0].item())
    assert v == float(N), v
    out.append({"impl": "reference-style kernel + ncclAllReduce", "ms": ms, "wall_ms": wall, "value": v})

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm = peaks.get("hbm_gbs", 6650.0)
        for i, r in enumerate(out):
            r.update({"n_gpus": P, "N": N, "dtype": a.dtype, "clocks": clocks[i] if i < len(clocks) else None,
                      "read_only_hbm_gbs_measured_here": round(read_peak, 1),
                      "fraction_of_read_only_peak": round(bytes_per_gpu / (r["ms"] * 1e-3) / 1e9 / read_peak, 4),
                      "GBps_per_gpu": round(bytes_per_gpu / (r["ms"] * 1e-3) / 1e9, 1),
                      "fraction_of_measured_hbm": round(bytes_per_gpu / (r["ms"] * 1e-3) / 1e9 / hbm, 4),
                      "ms": round(r["ms"], 5), "wall_ms": round(r["wall_ms"], 5)})
            print(json.dumps(r), flush=True)
        if a.json:
            json.dump(out, open(a.json, "w"), indent=1)
    sampler.stop()
    world.barrier()
    d.close()
    world.finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
