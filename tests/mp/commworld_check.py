"""The library's own launcher and communicator as the bootstrap of the fused paths
(no torch.distributed):

    python -m cuda_mpi_scratch_b200.launch -n 2 tests/mp/commworld_check.py

cms.init() returns a CommWorld; the symmetric memory (heap or IPC), the fused stencil, the fused
dot and the device collectives are built on it.  With ONE visible GPU both ranks share it: the
script then only uses operations in which no kernel waits for a kernel of the other rank
(allocation, peer visibility, one halo exchange with a barrier between push and wait)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import torch

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.models.stencil2d import Stencil2D, jacobi_coeffs, reference_step
from cuda_mpi_scratch_b200.parallel.symmetric import SymmetricBuffer


def main():
    world = cms.init()
    assert isinstance(world, cms.CommWorld), type(world)
    rank, P, dev = world.rank, world.size, world.device
    shared_gpu = torch.cuda.device_count() < P
    fails = 0
    # symmetric memory over the CommWorld bootstrap
    sym = SymmetricBuffer.allocate(world, 1 << 20)
    sym.tensor(0, (1024,), torch.float64).fill_(float(rank + 1))
    torch.cuda.synchronize(dev)
    world.barrier()
    ok = all(bool((sym.peer_tensor(r, 0, (1024,), torch.float64) == float(r + 1)).all()) for r in range(P))
    lens = world.allgather_bytes(b"x" * (rank + 1))
    ok = ok and [len(b) for b in lens] == list(range(1, P + 1))
    ok = ok and world.allreduce_max(float(rank)) == float(P - 1) and world.allreduce_sum(1.0) == float(P)
    print(f"rank {rank}: CommWorld bootstrap + symmetric memory "
          f"({'heap' if sym._block is not None else 'ipc'}): {'OK' if ok else 'FAIL'}", flush=True)
    fails += 0 if ok else 1
    # stencil plan on it
    dims = cms.dims_create(P, 2)
    th, tw = 64, 256
    st = Stencil2D.create(world, (th, tw), 1, "star", dims=dims)
    g = torch.Generator().manual_seed(3)
    glob = torch.rand(dims[0] * th, dims[1] * tw, dtype=torch.float64, generator=g)
    r, c = st.cart.coords(rank)
    st.fill(0.0, 0.0)
    st.core.copy_(glob[r * th:(r + 1) * th, c * tw:(c + 1) * tw])
    torch.cuda.synchronize(dev)
    if shared_gpu:
        # push, host barrier, wait: the flag-wait kernel only starts when every push has completed
        world.barrier()
        st.exchange_push()
        torch.cuda.synchronize(dev)
        world.barrier()
        st.exchange_wait()
        torch.cuda.synchronize(dev)
        want = glob
        steps = 0
    else:
        st.exchange()
        steps = 5
        st.step(steps)
        torch.cuda.synchronize(dev)
        want = glob
        for _ in range(steps):
            want = reference_step(want, jacobi_coeffs(1), 1)
    st.check()
    ri = (torch.arange(-1, th + 1) + r * th) % want.shape[0]
    ci = (torch.arange(-1, tw + 1) + c * tw) % want.shape[1]
    tile = st.tile.cpu()
    exp = want[ri][:, ci]
    # ghost corners are only exchanged by the fused steps' pushes when corners are on; compare edges + core
    err = float((tile[1:-1, :] - exp[1:-1, :]).abs().max()) if steps == 0 else float((st.core.cpu() - exp[1:-1, 1:-1]).abs().max())
    ok = err < 1e-12
    print(f"rank {rank}: stencil over CommWorld ({steps} fused steps, shared_gpu={shared_gpu}): err {err:.2e} "
          f"{'OK' if ok else 'FAIL'}", flush=True)
    fails += 0 if ok else 1
    world.barrier()
    st.close()
    tot = world.allreduce_sum(float(fails))
    if rank == 0:
        sys.stdout.write("COMMWORLD_CHECK PASSED\n" if tot == 0 else f"COMMWORLD_CHECK FAILED ({int(tot)})\n")
        sys.stdout.flush()
    world.finalize()
    return 0 if tot == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
