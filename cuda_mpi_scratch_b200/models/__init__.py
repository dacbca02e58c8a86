"""Flagship workloads: the fused 2-D stencil and the distributed dot product."""
from . import dot, stencil2d  # noqa: F401
