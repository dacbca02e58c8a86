"""Tiny single-GPU driver for ncu: one launch each of the fused dot, the LD/ST and TMA put
kernels, the datatype move kernels and the device allreduce kernels (virtual ranks on one GPU)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.models.dot import DistributedDot
from cuda_mpi_scratch_b200.ops.p2p import P2PChannel
from cuda_mpi_scratch_b200.ops.pack_move import typed_move
from cuda_mpi_scratch_b200.ops.collectives import DeviceCollectives
from cuda_mpi_scratch_b200.parallel.datatype import DOUBLE, FLOAT, INT, Datatype

torch.cuda.set_device(0)
# fused dot, 2^28 fp64 (4 GiB of operands, >> L2)
n = 1 << 28
x = torch.ones(n, dtype=torch.float64, device="cuda")
d = DistributedDot.create(cms.SingleWorld(0))
for _ in range(3):
    v = d.dot(x, x)
assert v == float(n)
del x
# puts: 256 MiB between two virtual ranks
nb = 256 << 20
ch = P2PChannel.create_group(cms.LocalGroup(2), nb)
src = torch.randint(0, 255, (nb,), dtype=torch.uint8, device="cuda")
for mode in ("ldst", "tma"):
    for _ in range(2):
        ch[0].put(1, src, mode=mode)
        ch[1].wait(0)
torch.cuda.synchronize()
assert torch.equal(ch[1].recv_view(), src)
# datatype moves: ghost column of a 4096^2 tile, and 2^20 struct records
T = 4096
tile = torch.rand(T + 2, T + 2, dtype=torch.float64, device="cuda")
col = DOUBLE.subarray([T + 2, T + 2], [T, 1], [1, T]).commit()
ghost = DOUBLE.subarray([T + 2, T + 2], [T, 1], [1, 0]).commit()
dst = torch.zeros_like(tile)
for _ in range(3):
    typed_move(tile, col, dst, ghost)
rec = Datatype.struct([4, 2], [0, 16], [FLOAT, INT]).resized(0, 32).commit()
dense = Datatype.basic("byte").contiguous(24).commit()
recs = torch.randint(0, 255, ((1 << 20) * 32,), dtype=torch.uint8, device="cuda")
out = torch.zeros((1 << 20) * 24, dtype=torch.uint8, device="cuda")
for _ in range(3):
    typed_move(recs, rec, out, dense, scount=1 << 20, dcount=1 << 20)
# device allreduce, 4 virtual ranks, 16 MiB fp32, one-shot and two-shot
grp = DeviceCollectives.create_group(cms.LocalGroup(4), max_bytes=16 << 20)
ts = [torch.rand(4 << 20, dtype=torch.float32, device="cuda") for _ in range(4)]
for algo in ("oneshot", "twoshot"):
    grp.allreduce([t.clone() for t in ts], "sum", algo)
torch.cuda.synchronize()
print("ok")
