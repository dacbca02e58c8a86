"""Single-GPU dot product checked against a CPU oracle -- the program of the reference's
``ref_parallel-dot-product-atomics.cu:55-105`` (no communication): fill two vectors on the GPU
(``init_vector``, ``:45-51``), reduce them with the synchronised kernel and with the deliberately
racy ``NO_SYNC`` variant (``:26-32``), print GPU vs CPU.

    python examples/python/dot_single_gpu.py [N]

Also shown: the framework's own kernel on one rank, and the entry point that allocates nothing.
"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200 import _lib
from cuda_mpi_scratch_b200.models.dot import DistributedDot, baseline_dot


def main() -> int:
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    x = torch.full((n,), 1.0, dtype=torch.float32, device="cuda")     # init_vector
    y = torch.full((n,), 1.0, dtype=torch.float32, device="cuda")
    cpu = float(torch.dot(x.cpu().double(), y.cpu().double()))        # dot() oracle on the host
    for variant, name in [(0, "atomicAdd per block"), (1, "per-block partials + host sum"),
                          (2, "last block finishes"), (3, "NO_SYNC (racy on purpose)")]:
        print(f"GPU ({name}): {baseline_dot(x, y, variant):.1f}   CPU: {cpu:.1f}")
    grp = DistributedDot.create_group(cms.LocalGroup(1))
    print(f"GPU (fused kernel, one rank): {grp.dot([x], [y])[0]:.1f}   CPU: {cpu:.1f}")
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    _lib.check(_lib.lib.b2c_dot_local_noalloc(C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()),
                                              C.c_int64(n), 8, 0, C.c_void_p(out.data_ptr()),
                                              C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    print(f"GPU (no-allocation entry point): {float(out.item()):.1f}   CPU: {cpu:.1f}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
