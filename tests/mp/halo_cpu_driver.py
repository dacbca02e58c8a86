"""CPU stencil driver, one rank per process -- the plumbing configuration of
BASELINE.json ("stencil2d mpi-2d-stencil-subarray.cpp on CPU, world_size=4, 2x2
Cartesian grid").  Mirrors stencil2d/mpi-2d-stencil-subarray.cpp:35-100: periodic
grid, tile + ghosts filled with -1, core = rank id, ONE halo exchange, per-rank
file "<row>_<col>" in the reference's format.

    python -m cuda_mpi_scratch_b200.launch -n 4 tests/mp/halo_cpu_driver.py OUTDIR [TILE] [STENCIL]
"""
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import torch

from cuda_mpi_scratch_b200.parallel import comm as cm
from cuda_mpi_scratch_b200.parallel.datatype import DOUBLE
from cuda_mpi_scratch_b200.parallel.halo import (Array2D, HaloPlan, print_array, print_cart_grid)


def main():
    outdir = sys.argv[1]
    tile = int(sys.argv[2]) if len(sys.argv) > 2 else 16
    stencil = int(sys.argv[3]) if len(sys.argv) > 3 else 5
    world = cm.init_process()
    dim = int(math.isqrt(world.size))
    dims = [dim, dim] if dim * dim == world.size else cm.dims_create(world.size, 2)
    cart = world.cart_create(dims, [True, True])
    r, c = cart.cart_coords()
    ghost = stencil // 2
    tot = tile + 2 * ghost
    data = torch.full((tot, tot), -1.0, dtype=torch.float64)
    layout = Array2D(tot, tot, tot)
    plan = HaloPlan(cart, layout, stencil, stencil, DOUBLE)
    plan.core.view(data)[:] = float(cart.rank)
    with open(os.path.join(outdir, f"{r}_{c}"), "w") as f:
        f.write(f"Rank:  {cart.rank}\nCoord: {r}, {c}\n\nCompute grid\n{print_cart_grid(cart)}\n")
        f.write(f"{tile} x {tile} grid size\n{tot} x {tot} total(with ghost/halo regions) grid size\n")
        f.write(f"{stencil} x {stencil} stencil\n\nArray\n{print_array(data)}\n")
        plan.exchange(data)
        f.write(f"Array after exchange\n{print_array(data)}")
    world.finalize()


if __name__ == "__main__":
    main()
