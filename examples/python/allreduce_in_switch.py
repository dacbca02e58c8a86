#!/usr/bin/env python
"""Device collectives on the symmetric heap: allreduce of a CUDA tensor three ways.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        examples/python/allreduce_in_switch.py
    python -m cuda_mpi_scratch_b200.launch -n 8 examples/python/allreduce_in_switch.py   # same, no torchrun

What the reference does for this (mpi9.cpp:51-54, mpicuda4.cu:354-368): copy a scalar to the host
and call MPI_Allreduce / MPI_Reduce.  Here the operands stay on the GPUs:
  1. `allreduce(t)`            any tensor; small ones in ONE kernel launch (staging, fixed-order sum
                               over the peers' buffers, back-pressure exchange), large ones reduced
                               inside the NVSwitch (multimem.ld_reduce) on systems with NVLink multicast;
  2. `allreduce_symmetric(t)`  the tensor itself lives in a multicast arena of the symmetric heap:
                               no staging copies at all;
  3. `reduce(t, root)`         rooted; non-root ranks cannot overrun a slow root.
Every rank gets bit-identical results."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import torch

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.ops.collectives import DeviceCollectives


def main() -> int:
    world = cms.init()                       # TorchWorld under torchrun, CommWorld under the own launcher
    rank, P, dev = world.rank, world.size, world.device
    device = f"cuda:{dev}"
    coll = DeviceCollectives.create(world, max_bytes=64 << 20)
    x = torch.full((1 << 20,), float(rank + 1), dtype=torch.float32, device=device)
    coll.allreduce(x)                                         # in place, algorithm picked by size / rank count
    small = torch.tensor([float(rank)], dtype=torch.float64, device=device)
    coll.allreduce(small)                                     # one kernel launch
    ok = bool((x == P * (P + 1) / 2).all()) and float(small) == P * (P - 1) / 2
    if coll.multicast:
        s = coll.symmetric_empty(1 << 20, torch.float32)      # collective allocation in a multicast arena
        s.fill_(float(rank + 1))
        torch.cuda.synchronize(dev)
        world.barrier()
        coll.allreduce_symmetric(s)                           # zero-copy, reduced and fanned out by the switch
        ok = ok and bool((s == P * (P + 1) / 2).all())
        # the other two collectives on the same tensor, in place: one multicast store pass each
        blk = s.numel() // P
        s.zero_()
        s[rank * blk:(rank + 1) * blk] = float(rank + 1)
        torch.cuda.synchronize(dev)
        world.barrier()
        coll.allgather_symmetric(s)                           # block r now holds r + 1 on every rank
        want = torch.arange(1, P + 1, dtype=torch.float32, device=device).repeat_interleave(blk)
        ok = ok and bool(torch.equal(s, want))
        coll.barrier()                                        # everybody has checked before the tensor changes
        torch.cuda.synchronize(dev)
        world.barrier()
        if rank == P - 1:
            s.fill_(42.0)
        coll.bcast_symmetric(s, root=P - 1)
        ok = ok and bool((s == 42.0).all())
    r = torch.full((1024,), 1.0, dtype=torch.float64, device=device)
    coll.reduce(r, root=0)
    if rank == 0:
        ok = ok and bool((r == float(P)).all())
    coll.check()                                              # raises if a peer never arrived
    bad = world.allreduce_sum(0.0 if ok else 1.0)
    if rank == 0:
        print(f"allreduce on {P} GPUs, multicast={'yes' if coll.multicast else 'no'}: "
              f"{'PASSED' if bad == 0 else 'FAILED'}")
    coll.close()
    world.finalize()
    return 0 if bad == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
