"""Fused stencil on a periodic Py x Px grid, one rank per GPU (torchrun), run until the
residual drops below a tolerance -- the exchange/compute/terminate loop of
stencil2d/mpi-2d-stencil-subarray-cuda.cu:169-172 with a real Compute and TerminateCondition."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import cuda_mpi_scratch_b200 as cms

world = cms.init()
temporal = int(sys.argv[1]) if len(sys.argv) > 1 else 3       # time steps per kernel sweep (1..4)
st = cms.Stencil2D.create(world, (4096, 4096), radius=1, temporal=temporal)
r, c = st.cart.coords(world.rank)
st.fill(0.0, 0.0)
if (r, c) == (0, 0):
    st.core[1024:3072, 1024:3072] = 1.0            # a hot square on rank 0
torch.cuda.synchronize()
st.exchange()
steps, res = st.run(until=lambda x: x < 1e-3, check_every=100, max_steps=5000)
st.check()
if world.rank == 0:
    print(f"{world.size} GPUs, grid {st.cart.dims}, {temporal} step(s) per sweep: {steps} steps, "
          f"residual {res:.3e}", flush=True)
world.finalize()
