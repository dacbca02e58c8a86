"""Single-GPU driver for ncu: ONLY the strided ghost-column move of a 4096^2 fp64 tile
(subarray type, 4096 runs of 8 bytes at a 32 KiB pitch, gathered and scattered by one kernel),
so that a capture of `uniform_move_kernel<int2>` cannot be mistaken for the struct-record case
(round 1's ncu row was).  g = 2 (16-byte runs, `uniform_move_kernel<int4>`) follows."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cuda_mpi_scratch_b200.ops.pack_move import typed_move
from cuda_mpi_scratch_b200.parallel.datatype import DOUBLE

torch.cuda.set_device(0)
T = 4096
for g in (1, 2):
    tot = T + 2 * g
    tile = torch.rand(tot, tot, dtype=torch.float64, device="cuda")
    dst = torch.zeros_like(tile)
    col = DOUBLE.subarray([tot, tot], [T, g], [g, T]).commit()       # right core columns
    ghost = DOUBLE.subarray([tot, tot], [T, g], [g, 0]).commit()     # left ghost columns
    for _ in range(4):
        typed_move(tile, col, dst, ghost)
    torch.cuda.synchronize()
    assert torch.equal(dst[g:g + T, :g], tile[g:g + T, T:T + g])
print("ok")
