"""cuda_mpi_scratch_b200 -- Blackwell-native communication + stencil framework.

A from-scratch B200 (sm_100a) re-design of the capabilities of
``ugovaretto-accel/cuda-mpi-scratch``: MPI-style point-to-point and collective
communication of host and device buffers, derived datatypes, Cartesian
topologies, a 2-D decomposed stencil with halo exchange, a distributed dot
product and GPU ping-pong benchmarks.  The hot paths are hand-written CUDA
kernels that move data between GPUs themselves (peer stores over NVLink,
device-side flags) instead of calling a CUDA-aware MPI from the host.

Layout
    ``models/``    flagship workloads (Stencil2D, DistributedDot)
    ``ops/``       thin wrappers over single kernels (p2p copy, pack/move, dot)
    ``parallel/``  worlds, symmetric heap, Cartesian topology, comm, datatypes
    ``utils/``     timing, clock sampling, dumps
    ``lib/``       the native library (built in-tree by ``build.py``)
"""
from . import _lib
from ._lib import B200Error, CommError, CudaError
from .parallel.cart import PROC_NULL, CartTopology, dims_create
from .parallel.world import (CommWorld, LocalGroup, SingleWorld, TorchWorld, World, bind_device, init,
                             processor_name, wtime)
from .parallel.symmetric import SymmetricBuffer
from .models.stencil2d import Stencil2D, StencilGroup, jacobi_coeffs, reference_step
from .models.dot import DistributedDot
from . import models, ops, parallel, utils                       # noqa: F401  (cms.ops.p2p, cms.utils.trace, ...)
from .ops import collectives, p2p, pack_move                     # noqa: F401
from .parallel import comm, datatype, halo                       # noqa: F401
from .utils import ranks, timing, trace                          # noqa: F401

__version__ = "0.1.0"


def set_error_mode(mode: str) -> None:
    """``"throw"`` (default: Python exceptions) or ``"abort"`` (print + exit), the
    run-time form of the reference's ``MPI_ERR_USE_EXCEPTIONS`` switch
    (``mpierr.h:48-52``)."""
    _lib.lib.b2c_set_error_mode({"throw": 0, "abort": 1}[mode])
