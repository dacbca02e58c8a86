"""Multi-process (one rank per GPU) checks of the fused dot, the P2P channel and device
buffers through the comm layer.  Run under torchrun with 2+ ranks."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import torch

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.models.dot import DistributedDot
from cuda_mpi_scratch_b200.ops.p2p import P2PChannel
from cuda_mpi_scratch_b200.parallel import comm as cm
from cuda_mpi_scratch_b200.parallel.datatype import DOUBLE, FLOAT


def main():
    world = cms.init()
    rank, P, dev = world.rank, world.size, world.device
    fails = 0
    # ---- fused dot + allreduce
    d = DistributedDot.create(world)
    n = 4_000_003
    g = torch.Generator(device=f"cuda:{dev}").manual_seed(100 + rank)
    x = torch.rand(n, dtype=torch.float64, device=f"cuda:{dev}", generator=g)
    y = torch.rand(n, dtype=torch.float64, device=f"cuda:{dev}", generator=g)
    want = world.allreduce_sum(float(torch.dot(x.cpu(), y.cpu())))
    vals = [d.dot(x, y) for _ in range(5)]
    same = world.allreduce_max(vals[0]) == vals[0] and -world.allreduce_max(-vals[0]) == vals[0]
    ok = same and all(v == vals[0] for v in vals) and abs(vals[0] - want) <= 1e-12 * abs(want)
    ones = torch.ones(1 << 22, dtype=torch.float32, device=f"cuda:{dev}")
    ok = ok and d.dot(ones, ones) == float(P * (1 << 22))
    r = d.dot(x, y, root=P - 1)
    ok = ok and (rank != P - 1 or r == vals[0])
    print(f"rank {rank}: dot {vals[0]:.12e} want {want:.12e} identical-on-all-ranks={same} {'OK' if ok else 'FAIL'}", flush=True)
    fails += 0 if ok else 1
    world.barrier()
    # ---- P2P channel ring: everybody puts to the right neighbour, waits for the left
    for mode in ("ldst", "tma"):
        nb = (1 << 20) + 16
        ch = P2PChannel.create(world, nb)
        src = torch.full((nb,), rank + 1, dtype=torch.uint8, device=f"cuda:{dev}")
        for it in range(3):
            ch.put((rank + 1) % P, src, mode=mode)
            ch.wait((rank - 1) % P)
        torch.cuda.synchronize(dev)
        ch.check()
        ok = bool((ch.recv_view() == ((rank - 1) % P) + 1).all())
        print(f"rank {rank}: p2p ring {mode} {'OK' if ok else 'FAIL'}", flush=True)
        fails += 0 if ok else 1
        world.barrier()
        ch.close()
    # ---- MPI-like send/recv of device tensors between PROCESSES (IPC mapping + pull kernel)
    host = cm.init_process()
    right, left = (rank + 1) % P, (rank - 1) % P
    a = torch.arange(1 << 16, dtype=torch.float64, device=f"cuda:{dev}") + rank
    b = torch.zeros_like(a)
    rr = host.irecv(b, left, tag=5)
    rs = host.isend(a, right, tag=5)
    cm.waitall([rr, rs])
    ok = bool(torch.equal(b, torch.arange(1 << 16, dtype=torch.float64, device=f"cuda:{dev}") + left))
    idx = FLOAT.indexed([4, 2], [5, 12]).commit()
    f = torch.arange(16, dtype=torch.float32, device=f"cuda:{dev}") + 100 * rank
    o = torch.zeros(6, dtype=torch.float32, device=f"cuda:{dev}")
    rr = host.irecv(o, left, tag=6)
    rs = host.isend(f, right, tag=6, count=1, datatype=idx)
    cm.waitall([rr, rs])
    ok = ok and o.tolist() == [100 * left + v for v in (5, 6, 7, 8, 12, 13)]
    print(f"rank {rank}: comm device send/recv across processes {'OK' if ok else 'FAIL'}", flush=True)
    fails += 0 if ok else 1
    host.finalize()
    tot = world.allreduce_sum(float(fails))
    if rank == 0:
        print("MP_DOT_P2P_CHECK", "PASSED" if tot == 0 else f"FAILED ({int(tot)})", flush=True)
    world.finalize()
    return 0 if tot == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
