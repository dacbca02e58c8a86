#!/usr/bin/env python
"""Device collectives over peer memory vs NCCL, one rank per GPU.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node P --master-addr 127.0.0.1 \
        --master-port 29620 bench/collectives.py [--json out.json]

ours      one kernel per collective on symmetric buffers (csrc/collectives/device_coll.cu):
          one-shot allreduce (every rank reads all inputs) for small payloads, two-shot
          (reduce-scatter + all-gather through peer memory) for large ones; bcast / gather /
          allgather / scatter as push kernels with epoch flags
baseline  the same collective through torch.distributed (NCCL)
Every result is first checked against the NCCL result (sums in rank order are bit-identical on
all ranks; against NCCL the comparison allows rounding), then timed on the device with CUDA
events, max over ranks.  The reference applies its collectives to host scalars only
(mpicuda2.cu:293, mpi9.cpp:51-54, mpi6.cpp:89-100); this is the device-resident form.
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
from cuda_mpi_scratch_b200.ops.collectives import DeviceCollectives
from cuda_mpi_scratch_b200.utils.timing import ClockSampler


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--max-log2", type=int, default=28, help="largest allreduce payload, log2 bytes")
    ap.add_argument("--min-log2", type=int, default=3, help="smallest allreduce payload, log2 bytes")
    ap.add_argument("--sweep-moves", action="store_true",
                    help="also sweep bcast and allgather over message sizes (8 B .. the workspace)")
    ap.add_argument("--json", default="")
    a = ap.parse_args()
    world = cms.init()
    P, rank, dev = world.size, world.rank, world.device
    device = f"cuda:{dev}"
    stream = torch.cuda.current_stream(dev)
    max_bytes = 1 << a.max_log2
    coll = DeviceCollectives.create(world, max_bytes=max_bytes)
    pg = None          # the default group routes CUDA tensors to NCCL (cms.init: "cpu:gloo,cuda:nccl")
    fails = 0
    rows = []

    sampler = ClockSampler(dev).start()      # clocks of every timed loop go into its row
    sampler.wait_first(2.0)
    clock_log = []

    def timed(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(dev)
        world.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler.window_begin()
        e0.record(stream)
        for _ in range(a.iters):
            fn()
        e1.record(stream)
        torch.cuda.synchronize(dev)
        sampler.window_end()
        c = sampler.pop_window_summary()
        clock_log.append({"sm_mhz": c.get("sm_mhz"), "reasons": c.get("reasons"), "samples": c.get("samples")})
        return world.allreduce_max(e0.elapsed_time(e1) / a.iters * 1e3)   # us

    def with_clocks(rec):
        rec["clocks"] = list(clock_log)
        clock_log.clear()
        return rec

    # operand of the zero-copy in-switch allreduce: ONE symmetric tensor, sliced per size
    symm = coll.symmetric_empty(max_bytes // 4, torch.float32) if coll.multicast else None

    # ---------------- allreduce: sizes x algorithms ----------------
    for lg in [3, 10, 12, 14, 16, 18, 20, 22, 24, 26, 28, 30]:
        if lg > a.max_log2 or lg < a.min_log2:
            continue
        n = (1 << lg) // 4
        if lg >= 30 and a.iters > 5:
            a.iters = 5                       # 1 GiB: a few iterations are enough and keep the run short
        g = torch.Generator(device=device).manual_seed(100 + rank)
        src = torch.rand(max(n, 1), dtype=torch.float32, device=device, generator=g)
        want = src.clone()
        dist.all_reduce(want, group=pg)
        algos = ["oneshot"] if lg < 14 else (["oneshot", "twoshot"] if lg <= 22 else ["twoshot"])
        if coll.multicast and lg >= 10:
            algos.append("nvls")              # reduction and fan-out inside the NVSwitch
        rec = {"op": "allreduce", "bytes": 1 << lg}
        for algo in algos:
            got = coll.allreduce(src.clone(), "sum", algo)
            coll.check()
            ok = torch.allclose(got, want, rtol=1e-5, atol=1e-6)
            # bit-identical on every rank: compare with rank 0's copy
            ref0 = got.clone()
            dist.broadcast(ref0, 0, group=pg)
            ok = ok and torch.equal(ref0, got)
            fails += 0 if ok else 1
            work = src.clone()
            rec[algo + "_us"] = round(timed(lambda: coll.allreduce(work, "sum", algo)), 2)
            rec[algo + "_ok"] = bool(ok)
        if symm is not None:
            sl = symm[:max(n, 1)]
            sl.copy_(src)
            torch.cuda.synchronize(dev); world.barrier()
            coll.allreduce_symmetric(sl)
            coll.check()
            ok = torch.allclose(sl, want, rtol=1e-5, atol=1e-6)
            ref0 = sl.clone()
            dist.broadcast(ref0, 0, group=pg)
            ok = ok and torch.equal(ref0, sl)
            fails += 0 if ok else 1
            sl.zero_()
            torch.cuda.synchronize(dev); world.barrier()
            rec["nvls_symmetric_us"] = round(timed(lambda: coll.allreduce_symmetric(sl)), 2)
            rec["nvls_symmetric_ok"] = bool(ok)
        work = src.clone()
        rec["nccl_us"] = round(timed(lambda: dist.all_reduce(work, group=pg)), 2)
        rows.append(with_clocks(rec))
        if rank == 0:
            print(rec, flush=True)

    # in-switch reduction of fp64 and int32 (scalar multimem operations)
    if coll.multicast:
        for dt in (torch.float64, torch.int32):
            v = (torch.arange(1 << 16, device=device) % 97 + rank).to(dt)
            want = v.clone()
            dist.all_reduce(want, group=pg)
            got = coll.allreduce(v.clone(), "sum", "nvls")
            coll.check()
            ok = bool(torch.equal(got, want))
            fails += 0 if ok else 1
            if rank == 0:
                print({"op": "allreduce nvls", "dtype": str(dt), "ok": ok}, flush=True)

    # ---------------- data movement ----------------
    blk = 1 << 20
    send = torch.full((blk // 4,), float(rank + 1), dtype=torch.float32, device=device)
    recv = torch.zeros(P * blk // 4, dtype=torch.float32, device=device)
    coll.allgather(send, recv)
    coll.check()
    want = torch.cat([torch.full((blk // 4,), float(r + 1), dtype=torch.float32) for r in range(P)])
    ok = torch.equal(recv.cpu(), want)
    fails += 0 if ok else 1
    rec = {"op": "allgather", "bytes_per_rank": blk, "ok": bool(ok),
           "ours_us": round(timed(lambda: coll.allgather(send, recv)), 2)}
    outs = list(recv.chunk(P))
    rec["nccl_us"] = round(timed(lambda: dist.all_gather(outs, send, group=pg)), 2)
    rows.append(with_clocks(rec))
    if rank == 0:
        print(rec, flush=True)

    b = torch.full((blk // 4,), float(rank), dtype=torch.float32, device=device)
    coll.bcast(b, root=P - 1)
    coll.check()
    ok = bool((b == float(P - 1)).all())
    fails += 0 if ok else 1
    rec = {"op": "bcast", "bytes": blk, "ok": ok, "ours_us": round(timed(lambda: coll.bcast(b, root=0)), 2),
           "nccl_us": round(timed(lambda: dist.broadcast(b, 0, group=pg)), 2),
           "multicast_store": bool(coll.multicast)}
    rows.append(with_clocks(rec))
    if rank == 0:
        print(rec, flush=True)

    if a.sweep_moves:
        # bcast (multimem.st fan-out through the switch when there is a multicast mapping) and
        # allgather (every rank stores its block into every peer) over message sizes, against
        # ncclBroadcast / ncclAllGather (all_gather_into_tensor)
        it0 = a.iters
        for lg in [3, 4, 10, 14, 18, 22, 24, 26, 28]:
            if (1 << lg) > max_bytes:
                continue
            a.iters = it0 if lg < 26 else min(it0, 10)
            nb = 1 << lg
            t = torch.full((max(nb // 4, 2),), float(rank), dtype=torch.float32, device=device)
            root = (lg // 2) % P
            coll.bcast(t, root=root)
            coll.check()
            ok = bool((t == float(root)).all())
            fails += 0 if ok else 1
            rec = {"op": "bcast", "bytes": t.numel() * 4, "ok": ok,
                   "ours_us": round(timed(lambda: coll.bcast(t, root=0)), 2),
                   "nccl_us": round(timed(lambda: dist.broadcast(t, 0, group=pg)), 2),
                   "multicast_store": bool(coll.multicast)}
            if symm is not None and (t.numel() * 4) % 16 == 0:
                # the same on a symmetric tensor: one multicast store pass, no staging
                sl = symm[:t.numel()]
                sl.fill_(float(rank))
                torch.cuda.synchronize(dev); world.barrier()
                coll.bcast_symmetric(sl, root=root)
                coll.check()
                ok = bool((sl == float(root)).all())
                fails += 0 if ok else 1
                rec["symmetric_us"] = round(timed(lambda: coll.bcast_symmetric(sl, root=0)), 2)
                rec["symmetric_ok"] = ok
            rows.append(with_clocks(rec))
            if rank == 0:
                print(rec, flush=True)
            del t
        for lg in [3, 4, 10, 14, 18, 22, 25]:
            nb = 1 << lg
            if nb * P > max_bytes:
                continue
            a.iters = it0 if lg < 22 else min(it0, 10)
            sblk = torch.full((max(nb // 4, 2),), float(rank + 1), dtype=torch.float32, device=device)
            rbuf = torch.zeros(P * sblk.numel(), dtype=torch.float32, device=device)
            coll.allgather(sblk, rbuf)
            coll.check()
            want_ag = torch.arange(1, P + 1, dtype=torch.float32, device=device).repeat_interleave(sblk.numel())
            ok = bool(torch.equal(rbuf, want_ag))
            fails += 0 if ok else 1
            rec = {"op": "allgather", "bytes_per_rank": sblk.numel() * 4, "ok": ok,
                   "ours_us": round(timed(lambda: coll.allgather(sblk, rbuf)), 2),
                   "nccl_us": round(timed(lambda: dist.all_gather_into_tensor(rbuf, sblk, group=pg)), 2)}
            if symm is not None and (sblk.numel() * 4) % 16 == 0:
                sl = symm[:P * sblk.numel()]
                sl.zero_()
                sl[rank * sblk.numel():(rank + 1) * sblk.numel()].copy_(sblk)
                torch.cuda.synchronize(dev); world.barrier()
                coll.allgather_symmetric(sl)
                coll.check()
                ok = bool(torch.equal(sl, want_ag))
                fails += 0 if ok else 1
                rec["symmetric_us"] = round(timed(lambda: coll.allgather_symmetric(sl)), 2)
                rec["symmetric_ok"] = ok
            rows.append(with_clocks(rec))
            if rank == 0:
                print(rec, flush=True)
            del sblk, rbuf, want_ag
        a.iters = it0

    g_recv = torch.zeros(P * 256, dtype=torch.float32, device=device) if rank == 0 else None
    g_send = torch.full((256,), float(rank + 1), dtype=torch.float32, device=device)
    coll.gather(g_send, g_recv, root=0)
    coll.check()
    if rank == 0:
        ok = torch.equal(g_recv.cpu(), torch.arange(1, P + 1, dtype=torch.float32).repeat_interleave(256))
        fails += 0 if ok else 1
    s_send = torch.arange(P * 256, dtype=torch.float32, device=device) if rank == 0 else None
    s_recv = torch.zeros(256, dtype=torch.float32, device=device)
    coll.scatter(s_send, s_recv, root=0)
    coll.check()
    ok = torch.equal(s_recv.cpu(), torch.arange(rank * 256, (rank + 1) * 256, dtype=torch.float32))
    fails += 0 if ok else 1
    red = torch.full((1024,), float(rank + 1), dtype=torch.float64, device=device)
    coll.reduce(red, root=0)
    coll.check()
    if rank == 0:
        ok = bool((red == P * (P + 1) / 2).all())
        fails += 0 if ok else 1
    rec = {"op": "barrier", "ours_us": round(timed(lambda: coll.barrier()), 2),
           "nccl_us": round(timed(lambda: dist.barrier(group=pg)), 2)}
    rows.append(with_clocks(rec))
    if rank == 0:
        print(rec, flush=True)

    tot = world.allreduce_sum(float(fails))
    if rank == 0:
        print("COLLECTIVES_CHECK PASSED" if tot == 0 else f"COLLECTIVES_CHECK FAILED ({int(tot)})", flush=True)
        if a.json:
            heap = None
            try:
                from cuda_mpi_scratch_b200.parallel.heap import SymmetricHeap
                hp = SymmetricHeap._by_world.get(id(world))
                heap = None if hp is None else dict(hp.stats, multicast=bool(coll.multicast))
            except Exception:
                pass
            json.dump({"n_gpus": P, "iters": a.iters, "timing": "CUDA events, max over ranks, us per call",
                       "symmetric_heap": heap,
                       "rows": rows, "check": "PASSED" if tot == 0 else "FAILED"}, open(a.json, "w"), indent=1)
    sampler.stop()
    world.barrier()
    coll.close()
    world.finalize()
    return 0 if tot == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
