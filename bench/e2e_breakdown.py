"""Where the end-to-end loop of bench.py spends its time (one GPU, periodic 16384^2 tile, T = 3):

    python bench/e2e_breakdown.py [--size 16384] [--sweeps 300]

Per kernel sweep (3 time steps), wall clock over a few hundred sweeps, the same plan throughout:

  A  step(T)                                     kernels only, queued back to back
  B  step(T, residual=True)                      + residual accumulated in the sweep
  C  set_coeffs + step + local_residual_async    copies on the LAUNCHING stream (the round-1 loop)
     + event, host reads one sweep late
  D  step_io + wait_io(previous)                 copies on the plan's copy streams

The reference has no such loop: its driver exchanges once and exits
(stencil2d/mpi-2d-stencil-subarray-cuda.cu:169-172)."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.models.stencil2d import Stencil2D


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=16384)
    ap.add_argument("--temporal", type=int, default=3)
    ap.add_argument("--sweeps", type=int, default=300)
    a = ap.parse_args()
    world = cms.init()
    dev = world.device
    T, n = a.temporal, a.sweeps
    st = Stencil2D.create(world, (a.size, a.size), radius=1, kind="star", periods=True, temporal=T)
    st.fill(0.0, 0.0)
    st.core.copy_(torch.rand(a.size, a.size, dtype=torch.float64, device=f"cuda:{dev}"))
    coeffs = [0.0, 0.25, 0.25, 0.25, 0.25]
    st.set_coeffs(coeffs)
    torch.cuda.synchronize(dev)
    st.exchange()
    coef_host = torch.tensor(coeffs, dtype=torch.float64).pin_memory()
    stream = torch.cuda.current_stream(dev)
    evs = [torch.cuda.Event(), torch.cuda.Event()]

    def arm_a():
        for _ in range(n):
            st.step(T, drain=False)

    def arm_b():
        for _ in range(n):
            st.step(T, residual=True, drain=False)

    def arm_c():
        pend = None
        for i in range(n):
            s = i & 1
            st.set_coeffs(coef_host)
            st.step(T, residual=True, drain=False)
            buf = st.local_residual_async(s)
            evs[s].record(stream)
            if pend is not None:
                evs[pend[0]].synchronize()
                float(pend[1].item())
            pend = (s, buf)
        evs[pend[0]].synchronize()

    def arm_d():
        pend = None
        for _ in range(n):
            t = st.step_io(T, coef_host, residual=True, drain=False)
            if pend is not None:
                st.wait_io(pend)
            pend = t
        st.wait_io(pend)

    out = {}
    cells = a.size * a.size * T
    for name, fn in (("A_kernels_only", arm_a), ("B_with_residual", arm_b),
                     ("C_copies_on_launch_stream", arm_c), ("D_step_io", arm_d),
                     ("A_again", arm_a)):
        fn()                                   # warm-up (and heat: the GPU sits at its power cap)
        torch.cuda.synchronize(dev)
        l0 = st.kernel_launches
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize(dev)
        dt = time.perf_counter() - t0
        out[name] = {"us_per_sweep": round(dt / n * 1e6, 2), "gcell_per_s": round(cells * n / dt / 1e9, 1),
                     "kernels_per_sweep": round((st.kernel_launches - l0) / n, 2)}
    st.check()
    print(json.dumps({"size": a.size, "temporal": T, "sweeps": n, "arms": out}))
    st.close()
    world.finalize()


if __name__ == "__main__":
    main()
