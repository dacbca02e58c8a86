"""Symmetric heap between PROCESSES (one rank per GPU, run under torchrun with 2+ ranks):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29551 tests/mp/heap_check.py

Blocks come from VMM arenas that every rank has mapped (file descriptors over Unix sockets); the
same block has the same offset everywhere; a freed block is reused only once every rank has freed
it; blocks of a multicast arena can be written on all ranks with one store."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import torch

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.parallel.heap import SymmetricHeap
from cuda_mpi_scratch_b200.parallel.symmetric import SymmetricBuffer, as_tensor


def main():
    world = cms.init()
    rank, P, dev = world.rank, world.size, world.device
    fails = 0
    heap = SymmetricHeap.for_world(world)
    a = SymmetricBuffer.allocate(world, 1 << 20)
    b = SymmetricBuffer.allocate(world, 3 << 20)
    assert a._block is not None and b._block is not None, "allocations did not come from the heap"
    # same offsets on every rank
    offs = world.allgather_bytes(f"{a._block.offset},{b._block.offset}".encode())
    ok = len(set(offs)) == 1
    # zero filled; my writes are visible through every peer's mapping of my block
    mine = a.tensor(0, (1 << 17,), torch.float64)
    ok = ok and bool((mine == 0).all())
    mine.fill_(float(rank + 1))
    torch.cuda.synchronize(dev)
    world.barrier()
    for r in range(P):
        peer = a.peer_tensor(r, 0, (1 << 17,), torch.float64)
        ok = ok and bool((peer == float(r + 1)).all())
    world.barrier()
    # write into the right neighbour's block through the peer mapping
    nb = (rank + 1) % P
    b.peer_tensor(nb, 0, (1024,), torch.int64).fill_(100 + rank)
    torch.cuda.synchronize(dev)
    world.barrier()
    ok = ok and bool((b.tensor(0, (1024,), torch.int64) == 100 + (rank - 1) % P).all())
    print(f"rank {rank}: heap blocks symmetric, zeroed, peer-visible: {'OK' if ok else 'FAIL'}", flush=True)
    fails += 0 if ok else 1
    # free on one rank only: the block must NOT be reused yet; after all freed: reused
    off_a = a._block.offset
    if rank == 0:
        a.free()
    c = SymmetricBuffer.allocate(world, 1 << 20)
    ok = c._block.offset != off_a or c._block.arena is not a._block.arena
    if rank != 0:
        a.free()
    d = SymmetricBuffer.allocate(world, 1 << 20)
    ok = ok and d._block.offset == off_a and bool((d.tensor(0, (1 << 17,), torch.float64) == 0).all())
    print(f"rank {rank}: deferred free / reuse after everybody freed: {'OK' if ok else 'FAIL'}", flush=True)
    fails += 0 if ok else 1
    # multicast arena: one store through mc_ptr lands in every rank's block
    m = SymmetricBuffer.allocate(world, 1 << 16, multicast=True)
    if m.mc_ptr:
        if rank == P - 1:
            as_tensor(m.mc_ptr, (1 << 13,), torch.float64, dev).fill_(42.0)      # plain stores to a multicast address
        torch.cuda.synchronize(dev)
        world.barrier()
        ok = bool((m.tensor(0, (1 << 13,), torch.float64) == 42.0).all())
        print(f"rank {rank}: multicast store reached my block: {'OK' if ok else 'FAIL'}", flush=True)
        fails += 0 if ok else 1
    else:
        print(f"rank {rank}: no multicast support on this system (skipped)", flush=True)
    world.barrier()
    tot = world.allreduce_sum(float(fails))
    if rank == 0:
        print("heap stats", heap.stats, flush=True)
        sys.stdout.write("HEAP_CHECK PASSED\n" if tot == 0 else f"HEAP_CHECK FAILED ({int(tot)})\n")
        sys.stdout.flush()
    world.finalize()
    return 0 if tot == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
