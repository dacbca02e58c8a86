"""Thin wrappers over single kernels: P2P puts/flags, datatype-driven moves."""
from . import p2p, pack_move  # noqa: F401
