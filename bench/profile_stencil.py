"""Tiny driver for ncu: a few fused stencil steps on one 16384^2 tile (1 GPU).

    python bench/profile_stencil.py [n] [steps] [radius] [--temporal 2]
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.models.stencil2d import Stencil2D

argv = list(sys.argv[1:])
temporal = 1
if "--temporal" in argv:
    i = argv.index("--temporal")
    temporal = int(argv[i + 1])
    del argv[i:i + 2]
n = int(argv[0]) if len(argv) > 0 else 16384
steps = int(argv[1]) if len(argv) > 1 else 4
radius = int(argv[2]) if len(argv) > 2 else 1
w = cms.init()
st = Stencil2D.create(w, (n, n), radius=radius, temporal=temporal)
st.fill(1.0, 1.0)
st.core.uniform_(0, 1)
torch.cuda.synchronize()
st.exchange()
st.step(steps)
torch.cuda.synchronize()
st.check()
print("ok", float(st.core[0, 0]))
