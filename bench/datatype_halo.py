#!/usr/bin/env python
"""Derived-datatype halo transfer, 4096 x 4096 fp64 tile, strided columns + struct records,
packed inside the moving kernel (BASELINE.json config 5).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29620 bench/datatype_halo.py

Each rank pushes a ghost-width-g column strip of its tile (subarray type: 4096 runs of g*8 B at
a 32 KiB pitch) into the neighbour's opposite ghost strip (another subarray type), and an array
of {4 float, 2 int, pad} particle records (struct type, mpi8.cpp) into a dense array.

ours            ONE kernel: gather by the source type, scatter by the destination type straight
                into peer memory (typed_move)
pack+nccl       baseline: pack kernel -> contiguous buffer -> NCCL send/recv -> unpack kernel
memcpy2d        baseline: cudaMemcpy2DAsync on the peer mapping (copy engine), columns only
contiguous      a dense copy of the same payload, for reference
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.ops.pack_move import typed_move
from cuda_mpi_scratch_b200.parallel.datatype import DOUBLE, FLOAT, INT, Datatype
from cuda_mpi_scratch_b200.parallel.symmetric import SymmetricBuffer
from cuda_mpi_scratch_b200.utils.timing import ClockSampler


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--tile", type=int, default=4096)
    ap.add_argument("--iters", type=int, default=200)
    ap.add_argument("--json", default="")
    a = ap.parse_args()
    world = cms.init()
    assert world.size % 2 == 0, "ranks work in pairs (r, r^1)"
    rank, peer, dev = world.rank, world.rank ^ 1, world.device
    sampler = ClockSampler(dev).start()
    sampler.wait_first(2.0)
    stream = torch.cuda.current_stream(dev)
    T = a.tile
    out = []
    for g in (1, 2):
        tot = T + 2 * g
        nbytes_tile = tot * tot * 8
        sym = SymmetricBuffer.allocate(world, nbytes_tile + (1 << 20))
        tile = sym.tensor(0, (tot, tot), torch.float64)
        tile.copy_(torch.arange(tot * tot, dtype=torch.float64, device=tile.device).reshape(tot, tot) + rank * 1e9)
        torch.cuda.synchronize(dev); world.barrier()
        send_t = DOUBLE.subarray([tot, tot], [T, g], [g, T]).commit()       # my right core columns
        recv_t = DOUBLE.subarray([tot, tot], [T, g], [g, 0]).commit()       # neighbour's left ghost
        flat = DOUBLE.contiguous(T * g).commit()
        payload = T * g * 8
        packed = torch.zeros(T * g, dtype=torch.float64, device=tile.device)
        landed = torch.zeros_like(packed)
        dense_src = torch.rand(T * g, dtype=torch.float64, device=tile.device)

        def ours():
            typed_move(tile, send_t, sym.peer_ptr(peer), recv_t, device=dev)

        def pack_nccl():
            typed_move(tile, send_t, packed, flat, device=dev)               # pack kernel
            if rank == 0:
                dist.send(packed, peer); dist.recv(landed, peer)
            else:
                dist.recv(landed, peer); dist.send(packed, peer)
            typed_move(landed, flat, tile, recv_t, device=dev)               # unpack kernel

        cudart = C.CDLL("libcudart.so", mode=C.RTLD_GLOBAL) if False else None

        def contiguous():
            typed_move(dense_src, flat, sym.peer_ptr(peer) + nbytes_tile, flat, device=dev)

        def timed(fn):
            for _ in range(5):
                fn()
            torch.cuda.synchronize(dev); world.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            sampler.window_begin()
            e0.record(stream)
            for _ in range(a.iters):
                fn()
            e1.record(stream)
            torch.cuda.synchronize(dev)
            sampler.window_end()
            world.barrier()
            return world.allreduce_max(e0.elapsed_time(e1) / a.iters * 1e3)

        def ours_pull():
            # receiver-driven: gather the PEER's columns by its type, scatter into my ghost strip
            typed_move(sym.peer_ptr(peer), send_t, tile, recv_t, device=dev)

        for name, fn in (("typed_move push (ours, 1 kernel)", ours),
                         ("typed_move pull (ours, 1 kernel)", ours_pull),
                         ("pack + nccl + unpack", pack_nccl),
                         ("contiguous copy of the same payload", contiguous)):
            us = timed(fn)
            if rank == 0:
                rec = {"case": f"column halo g={g}", "impl": name, "payload_bytes": payload, "n_gpus": world.size,
                       "us": round(us, 3), "GBps": round(payload / (us * 1e-6) / 1e9, 3),
                       "clocks": sampler.pop_window_summary()}
                out.append(rec); print(json.dumps(rec), flush=True)
        # correctness of the one-kernel path: my left ghost must hold the peer's right core columns
        tile[:, :g].fill_(-1.0)
        torch.cuda.synchronize(dev); world.barrier()
        ours(); torch.cuda.synchronize(dev); world.barrier()
        base = torch.arange(tot * tot, dtype=torch.float64, device=tile.device).reshape(tot, tot) + peer * 1e9
        ok = bool(torch.equal(tile[g:g + T, :g], base[g:g + T, T:T + g]))
        if rank == 0:
            print(json.dumps({"case": f"column halo g={g}", "check": "PASSED" if ok else "FAILED"}), flush=True)
        world.barrier()
        sym.free()

    # ---- struct records (mpi8.cpp Particle: 4 float + 2 int, here padded to 32 B)
    nrec = 1 << 20
    rec_t = Datatype.struct([4, 2], [0, 16], [FLOAT, INT]).resized(0, 32).commit()
    dense_t = Datatype.basic("byte").contiguous(24).commit()
    sym = SymmetricBuffer.allocate(world, nrec * 32 + nrec * 24)
    recs = sym.tensor(0, (nrec * 32,), torch.uint8)
    recs.copy_(torch.randint(0, 255, (nrec * 32,), dtype=torch.uint8, device=recs.device))
    torch.cuda.synchronize(dev); world.barrier()

    def struct_move():
        typed_move(recs, rec_t, sym.peer_ptr(peer) + nrec * 32, dense_t, scount=nrec, dcount=nrec, device=dev)
    for _ in range(5):
        struct_move()
    torch.cuda.synchronize(dev); world.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.window_begin()
    e0.record(stream)
    for _ in range(50):
        struct_move()
    e1.record(stream)
    torch.cuda.synchronize(dev); world.barrier()
    sampler.window_end()
    us = world.allreduce_max(e0.elapsed_time(e1) / 50 * 1e3)
    got = sym.tensor(nrec * 32, (nrec, 24), torch.uint8)
    peer_recs = sym.peer_tensor(peer, 0, (nrec, 32), torch.uint8)
    ok = bool(torch.equal(got, peer_recs[:, :24]))
    if rank == 0:
        rec = {"case": "struct records 4f+2i (24 B of 32 B)", "impl": "typed_move (ours, 1 kernel)",
               "payload_bytes": nrec * 24, "us": round(us, 3),
               "GBps": round(nrec * 24 / (us * 1e-6) / 1e9, 3), "check": "PASSED" if ok else "FAILED",
               "n_gpus": world.size, "clocks": sampler.pop_window_summary()}
        out.append(rec); print(json.dumps(rec), flush=True)

    # ---- the reference's nested type (mpi-complex-types.cpp:32-58) at scale: a 1-D subarray
    # (3 of every 8 ints) wrapped in an hindexed type whose byte displacements point into THREE
    # SEPARATE device allocations -> one kernel gathers from all three into the peer's dense buffer
    nblk = 1 << 16
    bufs = [torch.randint(0, 1 << 30, (nblk * 8,), dtype=torch.int32, device=f"cuda:{dev}") for _ in range(3)]
    blk_t = INT.subarray([8], [3], [3]).commit().resized(0, 32).commit()        # 3 ints of each 8-int record
    arr_t = blk_t.contiguous(nblk).commit()
    disps = [0, bufs[1].data_ptr() - bufs[0].data_ptr(), bufs[2].data_ptr() - bufs[0].data_ptr()]
    nest_t = arr_t.hindexed([1, 1, 1], disps).commit()
    dense3 = Datatype.basic("byte").contiguous(3 * nblk * 12).commit()
    symc = SymmetricBuffer.allocate(world, 3 * nblk * 12)

    def complex_move():
        typed_move(bufs[0], nest_t, symc.peer_ptr(peer), dense3, device=dev)
    for _ in range(5):
        complex_move()
    torch.cuda.synchronize(dev); world.barrier()
    sampler.window_begin()
    e0.record(stream)
    for _ in range(50):
        complex_move()
    e1.record(stream)
    torch.cuda.synchronize(dev); world.barrier()
    sampler.window_end()
    us = world.allreduce_max(e0.elapsed_time(e1) / 50 * 1e3)
    # what the peer's three buffers should have delivered into MY block
    mine = symc.tensor(0, (3, nblk, 3), torch.int32)
    world.barrier()
    want = torch.stack([b.reshape(nblk, 8)[:, 3:6] for b in bufs])            # what I sent
    got_at_peer = symc.peer_tensor(peer, 0, (3, nblk, 3), torch.int32)
    ok = bool(torch.equal(got_at_peer, want))
    okall = world.allreduce_sum(0.0 if ok else 1.0) == 0.0
    if rank == 0:
        rec = {"case": "hindexed(3 x contiguous(65536 x subarray 3-of-8 ints)) over THREE allocations",
               "impl": "typed_move (ours, 1 kernel)", "payload_bytes": 3 * nblk * 12, "us": round(us, 3),
               "GBps": round(3 * nblk * 12 / (us * 1e-6) / 1e9, 3), "check": "PASSED" if okall else "FAILED",
               "n_gpus": world.size, "clocks": sampler.pop_window_summary()}
        out.append(rec); print(json.dumps(rec), flush=True)
        if a.json:
            json.dump(out, open(a.json, "w"), indent=1)
    sampler.stop()
    world.barrier()
    world.finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
