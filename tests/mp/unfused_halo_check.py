"""CPU check (gloo) of the unfused baseline's halo exchange: every rank fills its core with its
rank id, exchanges once, and must find each neighbour's id in the matching ghost strip --
including grids where two directions are the same peer or the rank itself.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29710 tests/mp/unfused_halo_check.py PY PX
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import torch
import torch.distributed as dist

from cuda_mpi_scratch_b200.ops.unfused_halo import UnfusedHalo
from cuda_mpi_scratch_b200.parallel.cart import CartTopology

dist.init_process_group("gloo")
rank, size = dist.get_rank(), dist.get_world_size()
py, px = int(sys.argv[1]), int(sys.argv[2])
assert py * px == size
bad = 0
for g in (1, 2):
    cart = CartTopology.create((py, px), True)
    th, tw = 6, 10
    buf = torch.full((th + 2 * g, tw + 2 * g), -1.0, dtype=torch.float64)
    core = buf[g:-g, g:-g]
    # value = 1000 * rank + 100 * row + col, so a wrong EDGE (not just a wrong peer) is caught
    core.copy_(1000.0 * rank + 100.0 * torch.arange(th, dtype=torch.float64)[:, None]
               + torch.arange(tw, dtype=torch.float64)[None, :])
    UnfusedHalo(cart, rank, (th, tw), g, "cpu").exchange(core, buf)

    def cell(r, y, x):
        return 1000.0 * r + 100.0 * y + x
    n, s = cart.neighbour(rank, -1, 0), cart.neighbour(rank, 1, 0)
    w, e = cart.neighbour(rank, 0, -1), cart.neighbour(rank, 0, 1)
    for k in range(g):
        for x in range(tw):
            bad += buf[k, g + x].item() != cell(n, th - g + k, x)            # north ghost = north's last rows
            bad += buf[g + th + k, g + x].item() != cell(s, k, x)            # south ghost = south's first rows
        for y in range(th):
            bad += buf[g + y, k].item() != cell(w, y, tw - g + k)            # west ghost = west's last columns
            bad += buf[g + y, g + tw + k].item() != cell(e, y, k)            # east ghost = east's first columns
t = torch.tensor([float(bad)])
dist.all_reduce(t)
if rank == 0:
    sys.stdout.write("UNFUSED_HALO PASSED\n" if t.item() == 0 else f"UNFUSED_HALO FAILED ({int(t.item())})\n")
    sys.stdout.flush()
dist.destroy_process_group()
sys.exit(0 if t.item() == 0 else 1)
