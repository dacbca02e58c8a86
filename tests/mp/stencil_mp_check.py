"""Multi-process (one rank per GPU) check of the fused stencil, run under torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29533 tests/mp/stencil_mp_check.py

Every rank compares its tile with the PyTorch fp64 oracle of the GLOBAL periodic grid."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import torch

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.models.stencil2d import Stencil2D, jacobi_coeffs, reference_step


def main():
    world = cms.init()
    dims = cms.dims_create(world.size)
    fails = 0
    for (th, tw, R, kind, steps, tb) in [(64, 256, 1, "star", 7, 1), (96, 384, 2, "star", 5, 1),
                                          (50, 200, 1, "box", 4, 1), (1024, 2048, 1, "star", 20, 1),
                                          (64, 256, 1, "star", 7, 2), (1024, 2048, 1, "star", 20, 2),
                                          (96, 384, 1, "star", 11, 3), (1024, 2048, 1, "star", 12, 4)]:
        st = Stencil2D.create(world, (th, tw), R, kind, dims=dims, temporal=tb)
        g = torch.Generator().manual_seed(1234)
        glob = torch.rand(dims[0] * th, dims[1] * tw, dtype=torch.float64, generator=g)
        r, c = st.cart.coords(world.rank)
        st.fill(0.0, 0.0)
        st.core.copy_(glob[r * th:(r + 1) * th, c * tw:(c + 1) * tw])
        torch.cuda.synchronize()
        st.exchange()
        st.step(steps, residual=True)
        res = st.residual()
        st.check()
        exp, prev = glob, None
        co = jacobi_coeffs(R, kind)
        for _ in range(steps):
            prev = exp
            exp = reference_step(exp, co, R, kind)
        want = exp[r * th:(r + 1) * th, c * tw:(c + 1) * tw]
        err = float((st.core.cpu() - want).abs().max())
        want_res = float(((exp - prev) ** 2).sum().sqrt())
        ok = err < 1e-12 and abs(res - want_res) <= 1e-10 * want_res
        # ghosts of the current buffer must hold the neighbours' cores after the drain
        tile = st.tile.cpu()
        gw = R * tb
        ri = (torch.arange(-gw, th + gw) + r * th) % exp.shape[0]
        ci = (torch.arange(-gw, tw + gw) + c * tw) % exp.shape[1]
        gp = exp[ri][:, ci]
        if kind == "star":  # corners are exchanged too (push_corners=True by default)
            pass
        gerr = float((tile - gp).abs().max())
        ok = ok and gerr < 1e-12
        fails += 0 if ok else 1
        print(f"rank {world.rank} dims {dims} tile {th}x{tw} R{R} {kind} temporal {tb}: err {err:.2e} "
              f"ghost err {gerr:.2e} resid {res:.6e} (want {want_res:.6e}) {'OK' if ok else 'FAIL'}",
              flush=True)
        world.barrier()
        st.close()
    # ---- CUDA graph of 4 fused steps, replayed on every rank (real GPUs run concurrently)
    th, tw = 512, 1024
    g = torch.Generator().manual_seed(77)
    glob = torch.rand(dims[0] * th, dims[1] * tw, dtype=torch.float64, generator=g)
    st = Stencil2D.create(world, (th, tw), 1, "star", dims=dims)
    r, c = st.cart.coords(world.rank)
    st.fill(0.0, 0.0)
    st.core.copy_(glob[r * th:(r + 1) * th, c * tw:(c + 1) * tw])
    torch.cuda.synchronize()
    st.exchange()
    st.step(2)
    torch.cuda.synchronize()
    world.barrier()
    graph = st.capture(4)
    for _ in range(5):
        graph.replay()
    torch.cuda.synchronize()
    st.check()
    exp = glob
    for _ in range(22):
        exp = reference_step(exp, jacobi_coeffs(1), 1)
    err = float((st.core.cpu() - exp[r * th:(r + 1) * th, c * tw:(c + 1) * tw]).abs().max())
    ok = err < 1e-12 and st.info().steps_done == 22
    print(f"rank {world.rank}: graph replay x5 of 4 steps: err {err:.2e} {'OK' if ok else 'FAIL'}", flush=True)
    fails += 0 if ok else 1
    world.barrier()
    st.close()
    # ---- pipelined loop: new coefficients in and the residual out EVERY call (step_io / wait_io),
    # one call in flight, ranks free-running against each other
    for tb in (1, 3):
        th, tw = 512, 2048
        g = torch.Generator().manual_seed(99)
        glob = torch.rand(dims[0] * th, dims[1] * tw, dtype=torch.float64, generator=g)
        st = Stencil2D.create(world, (th, tw), 1, "star", dims=dims, temporal=tb)
        r, c = st.cart.coords(world.rank)
        st.fill(0.0, 0.0)
        st.core.copy_(glob[r * th:(r + 1) * th, c * tw:(c + 1) * tw])
        torch.cuda.synchronize()
        st.exchange()
        counts = [tb, tb, 2 * tb + 1, tb, 1, 2 * tb]
        coefs = [[0.1 + 0.01 * i, 0.2, 0.25 - 0.01 * i, 0.15, 0.3] for i in range(len(counts))]
        got, pend = [], None
        for i, (n, co) in enumerate(zip(counts, coefs)):
            t = st.step_io(n, co, residual=True, drain=(i == len(counts) - 1))
            if pend is not None:
                got.append(st.wait_io(pend))
            pend = t
        got.append(st.wait_io(pend))
        st.check()
        exp, ok = glob, True
        for n, co, loc in zip(counts, coefs, got):
            for _ in range(n):
                prev = exp
                exp = reference_step(exp, co, 1, "star")
            want = float(((exp - prev) ** 2).sum())
            tot_res = world.allreduce_sum(loc)
            ok = ok and abs(tot_res - want) <= 1e-10 * want
        err = float((st.core.cpu() - exp[r * th:(r + 1) * th, c * tw:(c + 1) * tw]).abs().max())
        ok = ok and err < 1e-12 and st.steps_done == sum(counts)
        print(f"rank {world.rank}: pipelined step_io loop, temporal {tb}: err {err:.2e} "
              f"{'OK' if ok else 'FAIL'}", flush=True)
        fails += 0 if ok else 1
        world.barrier()
        st.close()
    tot = world.allreduce_sum(float(fails))
    if world.rank == 0:
        sys.stdout.write("MP_STENCIL_CHECK PASSED\n" if tot == 0 else f"MP_STENCIL_CHECK FAILED ({int(tot)})\n")
        sys.stdout.flush()
    world.finalize()
    return 0 if tot == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
