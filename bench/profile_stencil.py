"""Tiny driver for ncu: a few fused stencil steps on one 16384^2 tile (1 GPU)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.models.stencil2d import Stencil2D

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
radius = int(sys.argv[3]) if len(sys.argv) > 3 else 1
w = cms.init()
st = Stencil2D.create(w, (n, n), radius=radius)
st.fill(1.0, 1.0)
st.core.uniform_(0, 1)
torch.cuda.synchronize()
st.exchange()
st.step(steps)
torch.cuda.synchronize()
st.check()
print("ok", float(st.core[0, 0]))
