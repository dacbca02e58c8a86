"""Rooted device collectives under skew (one rank per GPU, run under torchrun with 2+ ranks):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29541 tests/mp/coll_root_stress.py

reduce / bcast / gather / scatter / dot(root) are issued back to back, WITHOUT any host
synchronisation between operations, while the root is deliberately slow (it spins on the GPU
before every operation).  A non-root rank only sends in these operations, so nothing but the
library's own back-pressure (consumed-epoch flags in csrc/collectives/device_coll.cu, slot
acknowledgements in csrc/kernels/dot.cu) keeps it from running ahead and overwriting staging the
root has not read yet.  Every result is checked on the device and counted in a mismatch tensor
that is read once at the end.

The reference does these operations on host scalars after a D2H copy (mpicuda4.cu:368 MPI_Reduce,
mpicuda2.cu:154 MPI_Bcast, mpi6.cpp:89-100 MPI_Gather), one blocking call at a time."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import torch

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.models.dot import DistributedDot
from cuda_mpi_scratch_b200.ops.collectives import DeviceCollectives


def main():
    world = cms.init()
    rank, P, dev = world.rank, world.size, world.device
    device = f"cuda:{dev}"
    n = 1 << 14
    iters = int(os.environ.get("STRESS_ITERS", "60"))
    coll = DeviceCollectives.create(world, max_bytes=P * n * 8)
    dot = DistributedDot.create(world)
    bad = torch.zeros(1, dtype=torch.int64, device=device)
    ar = torch.arange(n, dtype=torch.float64, device=device)
    ones = torch.ones(n, dtype=torch.float64, device=device)

    def slow(root):
        if rank == root:
            torch.cuda._sleep(400_000)           # ~0.2 ms: the root lags every operation

    for it in range(iters):
        root = it % P
        # ---- reduce to root: sum over ranks of (rank + it) * arange
        slow(root)
        src = ar * float(rank + it)
        out = torch.empty_like(src)
        coll.reduce(src, root, out=out)
        if rank == root:
            want = ar * float(sum(r + it for r in range(P)))
            bad += (out != want).sum()
        # ---- bcast from root
        slow(root)
        t = ar + float(1000 * it + root) if rank == root else torch.zeros_like(ar)
        coll.bcast(t, root)
        bad += (t != ar + float(1000 * it + root)).sum()
        # ---- gather to root
        slow(root)
        blk = ones * float(100 * it + rank)
        recv = torch.empty(P * n, dtype=torch.float64, device=device) if rank == root else None
        coll.gather(blk, recv, root)
        if rank == root:
            want = (torch.arange(P, device=device, dtype=torch.float64) + 100.0 * it).repeat_interleave(n)
            bad += (recv != want).sum()
        # ---- scatter from root
        slow(root)
        send = (torch.arange(P, device=device, dtype=torch.float64) + 7.0 * it).repeat_interleave(n) \
            if rank == root else None
        got = torch.empty(n, dtype=torch.float64, device=device)
        coll.scatter(send, got, root)
        bad += (got != float(rank + 7 * it)).sum()
        # ---- fused dot, reduced to root: sum_r (r + it + 1) * n
        slow(root)
        dot.launch(ones * float(rank + it + 1), ones, root=root)
        if rank == root:
            want = float(sum(r + it + 1 for r in range(P)) * n)
            bad += (dot.result_tensor() != want).sum()
        # ---- small messages take the single-launch path (staging + reduction + consumed-epoch
        # exchange in one kernel): a rooted reduce and an allreduce of a few values
        slow(root)
        small = torch.arange(8, dtype=torch.float64, device=device) + float(rank * 10 + it)
        sout = torch.empty_like(small)
        coll.reduce(small, root, out=sout)
        if rank == root:
            want = torch.arange(8, dtype=torch.float64, device=device) * P + float(sum(r * 10 + it for r in range(P)))
            bad += (sout != want).sum()
        one = torch.full((1,), float(rank + it), dtype=torch.float32, device=device)
        coll.allreduce(one)
        bad += (one != float(sum(r + it for r in range(P)))).sum()
        # ---- every third round an allreduce in between (mixes the two families of operations)
        if it % 3 == 2:
            v = ones * float(rank + 1)
            coll.allreduce(v)
            bad += (v != float(P * (P + 1) // 2)).sum()
    torch.cuda.synchronize(dev)
    coll.check()
    nbad = int(bad.item())
    tot = world.allreduce_sum(float(nbad))
    print(f"rank {rank}: {iters} rounds of rooted collectives with a slow root: mismatches {nbad}", flush=True)
    world.barrier()
    if rank == 0:
        sys.stdout.write("COLL_ROOT_STRESS PASSED\n" if tot == 0 else f"COLL_ROOT_STRESS FAILED ({int(tot)})\n")
        sys.stdout.flush()
    coll.close()
    world.finalize()
    return 0 if tot == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
