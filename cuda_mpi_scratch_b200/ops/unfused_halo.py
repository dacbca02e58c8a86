"""Host-driven (UNFUSED) 4-neighbour halo exchange over ``torch.distributed`` -- the baseline the
fused stencil kernel is compared against (``bench/stencil_unfused_baseline.py``).

This is the reference's ``ExchangeData`` (``stencil2d/stencil2D.h:363-377``: all receives, all
sends, one wait) with NCCL (or gloo, on the CPU) in the place of a CUDA-aware MPI and explicit
pack / unpack copies in the place of MPI subarray datatypes.  Edges only (a star stencil of
radius ``g`` needs no corners).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..parallel.cart import CartTopology

_OPPOSITE = {"n": "s", "s": "n", "w": "e", "e": "w"}


class UnfusedHalo:
    """Pack -> batch_isend_irecv -> unpack for one rank of a periodic (or open) 2-D grid."""

    def __init__(self, cart: CartTopology, rank: int, tile_hw: tuple[int, int], g: int, device, dtype=torch.float64,
                 group=None):
        self.cart, self.rank, self.g, self.group = cart, rank, g, group
        th, tw = tile_hw
        self.peers = {"n": cart.neighbour(rank, -1, 0), "s": cart.neighbour(rank, 1, 0),
                      "w": cart.neighbour(rank, 0, -1), "e": cart.neighbour(rank, 0, 1)}
        self.send = {k: torch.empty((g, tw) if k in "ns" else (th, g), dtype=dtype, device=device) for k in "nswe"}
        self.recv = {k: torch.empty_like(self.send[k]) for k in "nswe"}

    def exchange(self, core: torch.Tensor, tile: torch.Tensor) -> None:
        """``core``: H x W view, ``tile``: (H+2g) x (W+2g) view of the same buffer."""
        g, send, recv, me = self.g, self.send, self.recv, self.rank
        send["n"].copy_(core[:g, :]); send["s"].copy_(core[-g:, :])          # pack
        send["w"].copy_(core[:, :g]); send["e"].copy_(core[:, -g:])
        ops = []
        for key in "nswe":                            # my `key` edge fills the peer's OPPOSITE ghost
            peer = self.peers[key]
            if peer == me:
                recv[_OPPOSITE[key]].copy_(send[key])
            elif peer >= 0:
                ops.append(dist.P2POp(dist.isend, send[key], peer, self.group))
        # Receives in the opposite order: messages between one pair of ranks are matched in the
        # order they were posted, and when north and south are the SAME peer (2 ranks in a column)
        # that peer's first message to me is its north edge -- which is my SOUTH ghost.
        for key in "snew":
            peer = self.peers[key]
            if peer != me and peer >= 0:
                ops.append(dist.P2POp(dist.irecv, recv[key], peer, self.group))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        if self.peers["n"] >= 0: tile[:g, g:-g].copy_(recv["n"])              # unpack into the ghosts
        if self.peers["s"] >= 0: tile[-g:, g:-g].copy_(recv["s"])
        if self.peers["w"] >= 0: tile[g:-g, :g].copy_(recv["w"])
        if self.peers["e"] >= 0: tile[g:-g, -g:].copy_(recv["e"])
