#!/usr/bin/env python
"""GPU ping-pong between 2 B200s, 8 B - 1 GiB, sync + async, several data paths.

BASELINE.json config: "GPU ping-pong sync+async (mpi-pingpong-gpu[-async]) 8 B-1 GB
between 2 B200s".  The reference does ONE un-warmed, host-timed round trip per process
launch and prints "Round-trip time(ms)" (test-benchmark/mpi-pingpong-gpu.cpp:51-68); here
every path is warmed up, repeated and timed on the device (rank 0, CUDA events or the
in-kernel globaltimer), and the echoed buffer is compared with the original (PASSED/FAILED,
:58-71).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29600 bench/pingpong.py [--max-bytes 1073741824] [--json out.json]

paths
    kernel_ldst / kernel_tma   product: put kernel -> peer buffer + release flag, acquire kernel
    device_loop                product: whole loop inside one kernel per GPU (latency floor)
    bidir_tma / bidir_ldst     both ranks put at the same time (GBps_one_way = per direction while BOTH run)
    copy_engine                baseline: cudaMemcpyAsync to the peer mapping + flag kernel
    nccl                       baseline: torch.distributed send/recv (NCCL)
    comm_sync / comm_async     MPI-like API on device tensors: <= 64 KiB eager through the receiver's
                               device mailbox (put kernel + flag, no host sync), larger messages by
                               rendezvous (descriptor over the host fabric, receiver pulls)
    host_staged[_pinned]       baseline: D2H + host transport + H2D (the reference's -DHOST_COPY)
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200 import _lib
from cuda_mpi_scratch_b200.ops.p2p import _OFF_DATA, _OFF_FLAGS, P2PChannel
from cuda_mpi_scratch_b200.parallel import comm as cm
from cuda_mpi_scratch_b200.utils.timing import ClockSampler


def iters_for(nbytes: int) -> int:
    return 200 if nbytes <= 1 << 16 else (50 if nbytes <= 1 << 24 else 10)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-bytes", type=int, default=1 << 30)
    ap.add_argument("--json", default="")
    ap.add_argument("--ctas", type=int, default=0, help="CTAs of the put kernel (0 = library default)")
    ap.add_argument("--paths", default="kernel_ldst,kernel_tma,device_loop,copy_engine,nccl,"
                                       "comm_sync,comm_async,host_staged,host_staged_pinned")
    a = ap.parse_args()
    world = cms.init()
    assert world.size == 2, "ping-pong needs exactly 2 ranks"
    rank, peer, dev = world.rank, 1 - world.rank, world.device
    stream = torch.cuda.current_stream(dev)
    host = cm.init_process()            # MPI-like world over the shm fabric (same launcher env)
    paths = a.paths.split(",")
    sizes = [s for s in (8, 64, 512, 4096, 32768, 1 << 18, 1 << 21, 1 << 24, 1 << 27, 1 << 30)
             if s <= a.max_bytes]
    ch = P2PChannel.create(world, max(sizes))
    results = []
    sampler = ClockSampler(dev).start()       # every record carries the clocks seen while it was timed
    sampler.wait_first(2.0)

    def ev_pair():
        return torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    for nbytes in sizes:
        n = nbytes // 8
        src = torch.arange(n, dtype=torch.float64, device=f"cuda:{dev}")
        back = torch.zeros_like(src)
        it = iters_for(nbytes)

        def roundtrip_kernel(mode):
            if rank == 0:
                ch.put(peer, src, mode=mode, max_ctas=a.ctas)
                ch.wait(peer)
            else:
                ch.wait(peer)
                ch.put(peer, ch.recv_view(torch.float64, n), mode=mode, max_ctas=a.ctas)

        def bidir_kernel(mode):
            # both directions at once: each rank puts to the other, then waits for the other's put
            ch.put(peer, src, mode=mode, max_ctas=a.ctas)
            ch.wait(peer)

        def roundtrip_copy_engine():
            pdata = ch.symbuf.peer_ptr(peer) + _OFF_DATA
            pflag = ch.symbuf.peer_ptr(peer) + _OFF_FLAGS + 8 * rank
            s = stream.cuda_stream

            def put(ptr):
                ch.sent[peer] += 1
                _lib.check(_lib.lib.b2c_memcpy_async(C.c_void_p(pdata), C.c_void_p(ptr), nbytes, 4, s))
                _lib.check(_lib.lib.b2c_flag_set(C.c_void_p(pflag), ch.sent[peer], s))
            if rank == 0:
                put(src.data_ptr()); ch.wait(peer)
            else:
                ch.wait(peer); put(ch.symbuf.local_ptr + _OFF_DATA)

        def roundtrip_nccl():
            if rank == 0:
                dist.send(src, peer); dist.recv(back, peer)
            else:
                dist.recv(back, peer); dist.send(back, peer)

        def roundtrip_comm(async_):
            if rank == 0:
                if async_:
                    r = host.irecv(back, peer, tag=0x10); s_ = host.isend(src, peer, tag=0x01)
                    r.wait(); s_.wait()
                else:
                    host.send(src, peer, tag=0x01); host.recv(back, peer, tag=0x10)
            else:
                host.recv(back, peer, tag=0x01); host.send(back, peer, tag=0x10)

        hbuf = torch.empty(n, dtype=torch.float64)
        hpin = torch.empty(n, dtype=torch.float64).pin_memory()

        def roundtrip_host(pinned):
            h = hpin if pinned else hbuf
            if rank == 0:
                h.copy_(src); host.send(h, peer, tag=1); host.recv(h, peer, tag=2); back.copy_(h)
            else:
                host.recv(h, peer, tag=1); back.copy_(h); h.copy_(back); host.send(h, peer, tag=2)
            torch.cuda.synchronize(dev)

        for path in paths:
            if path == "device_loop" and nbytes > 1 << 20:
                continue
            if path.startswith("host_staged") and nbytes > 1 << 27:
                continue
            if path in ("comm_sync", "comm_async") and nbytes > 1 << 27:
                pass
            fn = {"bidir_tma": lambda: bidir_kernel("tma"), "bidir_ldst": lambda: bidir_kernel("ldst"),
                  "kernel_ldst": lambda: roundtrip_kernel("ldst"),
                  "kernel_tma": lambda: roundtrip_kernel("tma"),
                  "copy_engine": roundtrip_copy_engine, "nccl": roundtrip_nccl,
                  "comm_sync": lambda: roundtrip_comm(False), "comm_async": lambda: roundtrip_comm(True),
                  "host_staged": lambda: roundtrip_host(False),
                  "host_staged_pinned": lambda: roundtrip_host(True)}.get(path)
            world.barrier()
            torch.cuda.synchronize(dev)
            ok = True
            if path == "device_loop":
                ch.pingpong_device(peer, rank, 20, nbytes, src)          # warm-up
                torch.cuda.synchronize(dev); world.barrier()
                sampler.window_begin()
                ch.pingpong_device(peer, rank, it, nbytes, src)
                torch.cuda.synchronize(dev)
                sampler.window_end()
                rtt_us = ch.pingpong_time_ns() / it / 1e3
                if rank == 0:
                    ok = bool(torch.equal(ch.recv_view(torch.float64, n), src))
                timer = "in-kernel globaltimer"
            else:
                host_timed = path.startswith("comm") or path.startswith("host")
                for _ in range(3):
                    fn()
                torch.cuda.synchronize(dev); world.barrier()
                sampler.window_begin()
                if host_timed:
                    t0 = time.perf_counter()
                    for _ in range(it):
                        fn()
                    torch.cuda.synchronize(dev)
                    rtt_us = (time.perf_counter() - t0) / it * 1e6
                    timer = "host wall clock (call is host-synchronous)"
                else:
                    e0, e1 = ev_pair()
                    e0.record(stream)
                    for _ in range(it):
                        fn()
                    e1.record(stream)
                    torch.cuda.synchronize(dev)
                    rtt_us = e0.elapsed_time(e1) / it * 1e3
                    timer = "CUDA events on rank 0"
                sampler.window_end()
                if rank == 0:
                    got = ch.recv_view(torch.float64, n) if path.startswith(("kernel", "copy", "bidir")) else back
                    ok = bool(torch.equal(got, src))
            ch.check()
            rtt_max = world.allreduce_max(rtt_us) if path == "device_loop" else rtt_us
            if rank == 0:
                # bidir_*: one "iteration" is ONE transfer per direction at the same time, not a round trip
                one_way_s = rtt_max * 1e-6 if path.startswith("bidir") else rtt_max * 1e-6 / 2
                rec = {"path": path, "bytes": nbytes, "iters": it, "rtt_us": round(rtt_max, 3),
                       "GBps_one_way": round(nbytes / one_way_s / 1e9, 3),
                       "check": "PASSED" if ok else "FAILED", "timer": timer,
                       "clocks": sampler.pop_window_summary()}
                results.append(rec)
                print(json.dumps(rec), flush=True)
            world.barrier()
        del src, back, hbuf, hpin
    sampler.stop()
    print(json.dumps({"rank": rank, "device_sends": cm.device_eager_stats()}), flush=True)
    if rank == 0 and a.json:
        with open(a.json, "w") as f:
            json.dump(results, f, indent=1)
    host.finalize()
    world.finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
