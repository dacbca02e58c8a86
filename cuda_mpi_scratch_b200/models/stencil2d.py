"""2-D Cartesian-decomposed stencil with the halo exchange fused into the update.

Capability parity with the reference's stencil drivers
(``stencil2d/mpi-2d-stencil-subarray-cuda.cu:77-179`` and the CPU twin
``stencil2d/mpi-2d-stencil-subarray.cpp:35-100``): a periodic (or not)
``Py x Px`` process grid, one ``H x W`` tile per rank surrounded by ghost
cells of width ``stencil/2``, and the loop *exchange -> compute -> terminate?*.
In the reference ``Compute()`` is an empty stub and the loop runs once
(``:32-37``, ``:169-172``); here the update is a real radius-``R`` star (or
3x3/5x5 box) stencil and one CUDA kernel per step does update **and** halo
exchange (``csrc/stencil/stencil_kernel.cuh``).

Coefficient order
    star:  ``[centre, W1, E1, N1, S1, W2, E2, N2, S2, ...]`` (``4R+1`` values)
    box:   ``(2R+1) x (2R+1)`` row-major, row 0 = north
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, Sequence

import torch

from .. import _lib
from ..parallel.cart import DIR_OFFSETS, PROC_NULL, CartTopology, dims_create
from ..parallel.symmetric import SymmetricBuffer
from ..parallel.world import LocalGroup, SingleWorld, World

KIND = {"star": 0, "box": 1}


def jacobi_coeffs(radius: int = 1, kind: str = "star") -> list[float]:
    """Plain neighbour average (centre weight 0) -- the default."""
    if kind == "star":
        n = 4 * radius
        return [0.0] + [1.0 / n] * n
    side = 2 * radius + 1
    n = side * side - 1
    c = [1.0 / n] * (side * side)
    c[side * side // 2] = 0.0
    return c


def layout(W: int, H: int, radius: int) -> _lib.StencilLayout:
    lay = _lib.StencilLayout()
    _lib.check(_lib.lib.b2c_stencil_layout(W, H, radius, C.byref(lay)))
    return lay


def _check_temporal(radius: int, kind: str, periods, temporal: int) -> None:
    if temporal == 1:
        return
    if temporal not in (2, 3, 4) or kind != "star" or radius != 1:
        raise ValueError("temporal blocking supports temporal=2, 3 or 4 with a radius-1 star stencil")
    if temporal == 4 and not all(bool(p) for p in periods):
        raise ValueError("4 time steps per sweep need a fully periodic grid (2 and 3 also work with "
                         "fixed boundary values on a non-periodic one)")


def _stream_ptr(device: int) -> int:
    return torch.cuda.current_stream(device).cuda_stream


class Stencil2D:
    """One rank's tile + plan.  Build with :meth:`create` (one rank per process)
    or :meth:`create_group` (virtual ranks in one process)."""

    def __init__(self, rank: int, cart: CartTopology, tile_hw: tuple[int, int], radius: int,
                 kind: str, symbuf: SymmetricBuffer, push_corners: bool, world: World | None,
                 temporal: int = 1):
        self.rank, self.cart, self.world = rank, cart, world
        self.H, self.W = int(tile_hw[0]), int(tile_hw[1])
        self.radius, self.kind, self.temporal = radius, kind, int(temporal)
        _check_temporal(radius, kind, cart.periods, self.temporal)
        self.symbuf = symbuf
        self.device = symbuf.device
        self.layout = layout(self.W, self.H, radius * self.temporal)
        if symbuf.nbytes < self.layout.total_bytes:
            raise ValueError("symmetric block too small for this tile")
        self.neighbours = cart.neighbours8(rank)
        nbr = (C.c_void_p * 8)()
        for d, r in enumerate(self.neighbours):
            nbr[d] = None if r == PROC_NULL else symbuf.peer_ptr(r)
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib.b2c_stencil_create(C.byref(h), self.W, self.H, radius * self.temporal,
                                                   2 if self.temporal > 1 else KIND[kind],
                                                   C.c_void_p(symbuf.local_ptr), C.byref(nbr),
                                                   int(push_corners)))
        self._h = h
        # Host staging comes in PAIRS (coefficients in, residual out) so that a caller can keep
        # one sweep in flight while it reads the result of the previous one: slot i of sweep i,
        # slot i^1 still owned by the copy of sweep i-1 (see bench.py's end-to-end loop).
        pin = (lambda t: t.pin_memory()) if torch.cuda.is_available() else (lambda t: t)
        self._coef_pinned = pin(torch.empty(2, 128, dtype=torch.float64))
        self._resid_pinned = pin(torch.zeros(2, dtype=torch.float64))
        self._coef_slot = 0
        self._coeffs: list[float] | None = None     # last coefficients handed in (None: the plan's default)
        self._graph_launches = 0
        self._time_steps = 0

    # ------------------------------------------------------------ factories
    @classmethod
    def create(cls, world: World, tile_hw: tuple[int, int], radius: int = 1, kind: str = "star",
               dims: Sequence[int] | None = None, periods=True,
               push_corners: bool = True, temporal: int = 1) -> "Stencil2D":
        """Collective over ``world``: every rank gets an ``H x W`` tile.

        ``temporal=T`` with T = 2, 3 or 4 (radius-1 star): temporal blocking -- every sweep of the
        fused kernel advances T time steps from one read and one write of the grid (ghost width T,
        corners exchanged), bit-identical to T single steps.  Any step count is accepted: a
        remainder is done by one partial sweep.  T = 2 and 3 also run on NON-PERIODIC grids (the
        reference's topology, ``mpi10.cpp:24-27``): ghost cells beyond the edge of the global grid
        are boundary data -- set them in BOTH buffers (:meth:`fill`) -- and stay fixed through the
        intermediate time levels."""
        dims = tuple(dims) if dims is not None else dims_create(world.size, 2)
        cart = CartTopology.create(dims, periods)
        if cart.size != world.size:
            raise ValueError(f"dims {dims} do not match world size {world.size}")
        _check_temporal(radius, kind, cart.periods, temporal)
        lay = layout(tile_hw[1], tile_hw[0], radius * temporal)
        sym = SymmetricBuffer.allocate(world, lay.total_bytes)
        return cls(world.rank, cart, tile_hw, radius, kind, sym, push_corners, world, temporal)

    @classmethod
    def create_group(cls, group: LocalGroup, tile_hw: tuple[int, int], radius: int = 1,
                     kind: str = "star", dims: Sequence[int] | None = None, periods=True,
                     push_corners: bool = True, temporal: int = 1) -> "StencilGroup":
        dims = tuple(dims) if dims is not None else dims_create(group.size, 2)
        cart = CartTopology.create(dims, periods)
        if cart.size != group.size:
            raise ValueError(f"dims {dims} do not match group size {group.size}")
        _check_temporal(radius, kind, cart.periods, temporal)
        lay = layout(tile_hw[1], tile_hw[0], radius * temporal)
        syms = SymmetricBuffer.allocate_local(group, lay.total_bytes)
        return StencilGroup([cls(r, cart, tile_hw, radius, kind, syms[r], push_corners, None, temporal)
                             for r in range(group.size)])

    # ------------------------------------------------------------ views
    def info(self) -> _lib.StencilInfo:
        inf = _lib.StencilInfo()
        _lib.check(_lib.lib.b2c_stencil_info(self._h, C.byref(inf)))
        return inf

    def _buffer(self, which: int) -> torch.Tensor:
        L = self.layout
        off = L.off_buf0 if which == 0 else L.off_buf1
        return self.symbuf.tensor(off, (L.rows, L.pitch), torch.float64)

    @property
    def current(self) -> int:
        return self.info().cur

    @property
    def kernel_launches(self) -> int:
        """Kernels launched through this plan so far, counted by the native library
        (graph replays add the kernels of the captured batch)."""
        return int(self.info().launches) + self._graph_launches

    @property
    def padded(self) -> torch.Tensor:
        """Whole padded buffer holding the current state: ``rows x pitch``."""
        return self._buffer(self.current)

    @property
    def tile(self) -> torch.Tensor:
        """Core + ghost cells of the current state: ``(H+2g) x (W+2g)`` view."""
        L = self.layout
        return self.padded[:, L.lead - L.ghost: L.lead + L.W + L.ghost]

    @property
    def core(self) -> torch.Tensor:
        """The ``H x W`` core of the current state (writable view)."""
        L = self.layout
        return self.padded[L.ghost: L.ghost + L.H, L.lead: L.lead + L.W]

    # ------------------------------------------------------------ set-up
    def fill(self, core_value: float | None = None, ghost_value: float | None = None) -> None:
        """Initialise BOTH buffers (the reference fills the tile with -1 and the
        core with the rank id, ``mpi-2d-stencil-subarray-cuda.cu:148-159``)."""
        st = _stream_ptr(self.device)
        with torch.cuda.device(self.device):
            for b in (0, 1):
                if ghost_value is not None:
                    _lib.check(_lib.lib.b2c_stencil_fill(self._h, b, 0, float(ghost_value), st))
                if core_value is not None:
                    _lib.check(_lib.lib.b2c_stencil_fill(self._h, b, 1, float(core_value), st))

    def set_coeffs(self, coeffs: Sequence[float]) -> None:
        """Copy stencil coefficients host(pinned) -> device on the current stream.  The pinned
        staging area alternates between two slots, so a second call may follow before the copy
        of the first has run (not a third: synchronise at least every other call)."""
        n = len(coeffs)
        stage = self._coef_pinned[self._coef_slot]
        self._coef_slot ^= 1
        stage[:n] = torch.as_tensor(coeffs, dtype=torch.float64)
        self._coeffs = stage[:n].tolist()
        ptr = C.cast(stage.data_ptr(), C.POINTER(C.c_double))
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib.b2c_stencil_set_coef(self._h, ptr, n, _stream_ptr(self.device)))

    # ------------------------------------------------------------ exchange
    def exchange_push(self) -> None:
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib.b2c_stencil_exchange_push(self._h, _stream_ptr(self.device)))

    def exchange_wait(self) -> None:
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib.b2c_stencil_exchange_wait(self._h, _stream_ptr(self.device)))

    def exchange(self) -> None:
        """Unfused halo exchange of the current state (``ExchangeData``,
        ``stencil2d/stencil2D.h:363-377``).  One rank per process only; virtual
        ranks use :meth:`StencilGroup.exchange`."""
        if self.world is None:
            raise RuntimeError("use StencilGroup.exchange() for virtual ranks")
        self.world.barrier()  # peers are done touching the buffers
        self.exchange_push()
        self.exchange_wait()

    # ------------------------------------------------------------ stepping
    def step(self, nsteps: int = 1, residual: bool = False, drain: bool = True) -> None:
        """Enqueue ``nsteps`` fused update+halo time steps on the current stream.  With temporal
        blocking one kernel sweep advances ``temporal`` steps; a step count that is not a
        multiple ends with one partial sweep that computes only the remaining levels."""
        flags = (1 if residual else 0) | (2 if drain else 0)
        L, st = _lib.lib, _stream_ptr(self.device)
        with torch.cuda.device(self.device):
            if self.temporal == 1:
                _lib.check(L.b2c_stencil_step(self._h, nsteps, flags, st))
            else:
                full, rem = divmod(nsteps, self.temporal)
                if rem == 0:
                    _lib.check(L.b2c_stencil_step(self._h, full, flags, st))
                else:
                    if full:
                        _lib.check(L.b2c_stencil_step(self._h, full, 0, st))
                    _lib.check(L.b2c_stencil_step_levels(self._h, rem, flags, st))
        self._time_steps += nsteps

    def step_io(self, nsteps: int, coeffs: Sequence[float] | torch.Tensor | None = None,
                residual: bool = True, drain: bool = True) -> int:
        """Pipelined :meth:`step` for loops that hand in coefficients and read the residual
        EVERY call: the host->device copy of ``coeffs`` (staged through pinned memory of the
        plan) and the device->host copy of the residual run on the plan's own copy streams, next
        to the kernels of the previous and the next call, and the current stream carries the
        sweep kernels only.  Returns a ticket for :meth:`wait_io`; the results of the last two
        calls are kept, so a loop stays one call ahead of the result it reads."""
        flags = (1 if residual else 0) | (2 if drain else 0)
        ticket = C.c_ulonglong(0)
        if coeffs is None:
            ptr, n = None, 0
        else:
            ct = coeffs if isinstance(coeffs, torch.Tensor) else torch.as_tensor(coeffs, dtype=torch.float64)
            if ct.dtype != torch.float64 or ct.device.type != "cpu" or not ct.is_contiguous():
                ct = ct.to(device="cpu", dtype=torch.float64).contiguous()
            ptr, n = C.cast(ct.data_ptr(), C.POINTER(C.c_double)), ct.numel()
            self._coeffs = ct.tolist()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib.b2c_stencil_step_io(self._h, nsteps, flags, ptr, n,
                                                    _stream_ptr(self.device), C.byref(ticket)))
        self._time_steps += nsteps
        return int(ticket.value)

    def wait_io(self, ticket: int) -> float:
        """Block until call ``ticket`` of :meth:`step_io` has run on the device; returns its local
        sum of squared updates of the last time step (0.0 if no residual was asked for)."""
        r = C.c_double(0.0)
        _lib.check(_lib.lib.b2c_stencil_io_wait(self._h, int(ticket), C.byref(r)))
        return float(r.value)

    def _sweeps(self, nsteps: int) -> int:
        """Kernel sweeps that :meth:`step` launches for ``nsteps`` time steps."""
        return -(-nsteps // self.temporal)

    @property
    def steps_done(self) -> int:
        """Time steps completed since creation."""
        return self._time_steps

    def drain(self) -> None:
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib.b2c_stencil_drain(self._h, _stream_ptr(self.device)))

    def capture(self, nsteps: int, residual: bool = False, drain: bool = True) -> "StencilGraph":
        """Capture ``nsteps`` fused steps (an EVEN number: the two buffers must be back in
        place) into a CUDA graph.  Works because the epoch arithmetic of the kernels reads a
        step counter in device memory, so the captured launch parameters stay valid for every
        replay.  Run at least one plain :meth:`step` first (one-time kernel attributes are set
        outside the capture)."""
        if nsteps <= 0 or self._sweeps(nsteps) % 2:
            raise ValueError("capture a positive number of steps that makes an EVEN number of kernel "
                             f"sweeps ({self.temporal} step(s) per sweep): the two buffers must end "
                             "where they started")
        g = torch.cuda.CUDAGraph()
        launches0 = int(self.info().launches)
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side):
                self.step(nsteps, residual=residual, drain=drain)
        torch.cuda.current_stream(self.device).wait_stream(side)
        per_replay = int(self.info().launches) - launches0
        self._graph_launches -= per_replay      # captured, not executed
        # nothing ran during capture: take the host-side bookkeeping back
        _lib.check(_lib.lib.b2c_stencil_advance(self._h, -self._sweeps(nsteps)))
        self._time_steps -= nsteps
        return StencilGraph(self, g, nsteps, per_replay)

    def local_residual_async(self, slot: int = 0) -> torch.Tensor:
        """Start the D2H copy of the last step's local sum of squared updates into pinned memory
        (``slot`` 0 or 1); the returned 1-element tensor is valid once the stream (or an event
        recorded after this call) has been synchronised."""
        inf = self.info()
        dst = self._resid_pinned[slot:slot + 1]
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib.b2c_memcpy_async(C.c_void_p(dst.data_ptr()),
                                                 C.c_void_p(inf.resid), 8, 2,
                                                 _stream_ptr(self.device)))
        return dst

    def check(self) -> None:
        """Synchronise and raise if a device-side wait timed out."""
        st = C.c_uint(0)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib.b2c_stencil_status(self._h, _stream_ptr(self.device), C.byref(st)))
        if st.value & 1:
            raise _lib.CommError(_lib.ERR_TIMEOUT, "halo flag wait timed out on the device "
                                 f"(rank {self.rank}): a neighbour did not deliver")

    def phase_stats(self) -> dict:
        """Device-side timing of the LAST sweep, per phase (temporal strip kernel; create the plan
        with ``B2C_STENCIL_DEBUG=32`` in the environment).  Microseconds, mean / max over CTAs."""
        out = (C.c_double * 7)()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib.b2c_stencil_phase_stats(self._h, _stream_ptr(self.device), out))
        keys = ("boundary_mean_us", "boundary_max_us", "cta_mean_us", "cta_max_us",
                "flag_wait_mean_us", "flag_wait_max_us", "grid_span_us")
        return dict(zip(keys, [round(float(v), 2) for v in out]))

    def run(self, steps: int | None = None, until: Callable[[float], bool] | None = None,
            check_every: int = 10, max_steps: int = 1 << 30) -> tuple[int, float | None]:
        """``do { exchange; compute } while (!terminate)`` -- fused.

        ``steps``: run exactly that many.  ``until(residual_l2)``: evaluate the
        global residual every ``check_every`` steps and stop when it returns
        True (the reference's ``TerminateCondition``,
        ``mpi-2d-stencil-subarray-cuda.cu:36-37``, which always returns true).
        Returns ``(steps_run, last_residual_or_None)``.
        """
        if until is None:
            self.step(int(steps or 0))
            return int(steps or 0), None
        done, res = 0, None
        while done < max_steps:
            n = min(check_every, max_steps - done)
            self.step(n, residual=True)
            done += n
            res = self.residual()
            if until(res):
                break
        return done, res

    def residual(self) -> float:
        """Global L2 norm of the last update (synchronises)."""
        buf = self.local_residual_async()
        torch.cuda.current_stream(self.device).synchronize()
        s = float(buf.item())
        if self.world is not None and self.world.size > 1:
            s = self.world.allreduce_sum(s)
        return s ** 0.5

    # ------------------------------------------------------------ output
    def dump_str(self, fmt: str = "{:g}") -> str:
        """Tile incl. ghosts, one row per line, every value followed by a space
        -- the format of ``Print`` (``stencil2d/stencil2D.h:92-102``)."""
        t = self.tile.cpu()
        return "".join("".join(fmt.format(v) + " " for v in row.tolist()) + "\n" for row in t)

    def set_timeout_ms(self, ms: int) -> None:
        """Deadline of every device-side flag wait (default 20 s).  When a neighbour never
        delivers, the wait gives up, a status bit is set and :meth:`check` raises."""
        _lib.check(_lib.lib.b2c_stencil_set_timeout_ms(self._h, int(ms)))

    # ------------------------------------------------------------ checkpoint / resume
    def save_checkpoint(self, path: str) -> None:
        """Write this rank's state (core cells, ghost cells, coefficients, step count, geometry) to ``path``.
        The reference has no checkpointing, only the per-rank text dump
        (``mpi-2d-stencil-subarray-cuda.cu:114-119,164-176``); this is the binary form of it."""
        torch.cuda.current_stream(self.device).synchronize()
        torch.save({"tile": self.tile.cpu(), "steps_done": self.steps_done, "temporal": self.temporal,
                    "rank": self.rank, "dims": self.cart.dims, "periods": self.cart.periods,
                    "tile_hw": (self.H, self.W), "radius": self.radius, "kind": self.kind,
                    "coeffs": self._coeffs}, path)

    def load_checkpoint(self, path: str) -> int:
        """Restore the state saved by :meth:`save_checkpoint` into the CURRENT buffer (core and
        ghost cells) and return the step count of the checkpoint.  Call :meth:`exchange`
        afterwards (collective) before stepping again."""
        ck = torch.load(path)
        if tuple(ck["tile_hw"]) != (self.H, self.W) or ck["radius"] != self.radius \
                or ck["rank"] != self.rank or ck.get("temporal", 1) != self.temporal:
            raise ValueError("checkpoint does not match this plan (tile size, radius or rank)")
        self.tile.copy_(ck["tile"].to(self.tile.device))
        if ck.get("coeffs") is not None:          # coefficients are part of the state
            self.set_coeffs(ck["coeffs"])
        torch.cuda.current_stream(self.device).synchronize()
        return int(ck["steps_done"])

    def close(self) -> None:
        if self._h:
            _lib.lib.b2c_stencil_destroy(self._h)
            self._h = None
            self.symbuf.release_if_heap()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass


class StencilGraph:
    """A captured batch of fused steps (see :meth:`Stencil2D.capture`)."""

    def __init__(self, stencil: "Stencil2D", graph, nsteps: int, launches: int):
        self.stencil, self.graph, self.nsteps, self.launches = stencil, graph, nsteps, launches

    def replay(self) -> None:
        self.graph.replay()
        _lib.check(_lib.lib.b2c_stencil_advance(self.stencil._h, self.stencil._sweeps(self.nsteps)))
        self.stencil._time_steps += self.nsteps
        self.stencil._graph_launches += self.launches


class StencilGroup:
    """All virtual ranks of a :class:`LocalGroup`, stepped in lock-step.

    Within one process every launch goes into a stream in program order, so
    rank ``r``'s step ``s+1`` is always enqueued after every rank's step
    ``s``: no kernel ever spins on a flag whose producer has not been launched
    (the constraint spelled out in the profiling guide for single-GPU runs).
    """

    def __init__(self, ranks: list[Stencil2D]):
        self.ranks = ranks

    def __len__(self):
        return len(self.ranks)

    def __getitem__(self, i):
        return self.ranks[i]

    def sync(self) -> None:
        for d in {s.device for s in self.ranks}:
            torch.cuda.synchronize(d)

    def exchange(self) -> None:
        self.sync()
        for s in self.ranks:
            s.exchange_push()
        self.sync()
        for s in self.ranks:
            s.exchange_wait()
        self.sync()

    def step(self, nsteps: int = 1, residual: bool = False) -> None:
        multi_dev = len({s.device for s in self.ranks}) > 1
        t = self.ranks[0].temporal
        chunks = [t] * (nsteps // t) + ([nsteps % t] if nsteps % t else [])
        for i, n in enumerate(chunks):
            last = i == len(chunks) - 1
            for s in self.ranks:
                s.step(n, residual=residual and last, drain=False)
        # virtual ranks on several devices run concurrently; the epoch flags keep them within one
        # sweep of each other exactly as they keep separate processes, so no host sync per sweep
        if multi_dev:
            self.sync()
        for s in self.ranks:
            s.drain()

    def residual(self) -> float:
        bufs = [s.local_residual_async() for s in self.ranks]
        self.sync()
        return sum(float(b.item()) for b in bufs) ** 0.5

    def check(self) -> None:
        for s in self.ranks:
            s.check()

    def gather_global(self) -> torch.Tensor:
        """Assemble the global grid from the cores (CPU tensor)."""
        cart = self.ranks[0].cart
        rows, cols = cart.dims
        H, W = self.ranks[0].H, self.ranks[0].W
        out = torch.empty(rows * H, cols * W, dtype=torch.float64)
        for s in self.ranks:
            r, c = cart.coords(s.rank)
            out[r * H:(r + 1) * H, c * W:(c + 1) * W] = s.core.cpu()
        return out

    def scatter_global(self, g: torch.Tensor) -> None:
        """Set every rank's core (both buffers are NOT touched: only current)."""
        cart = self.ranks[0].cart
        H, W = self.ranks[0].H, self.ranks[0].W
        for s in self.ranks:
            r, c = cart.coords(s.rank)
            s.core.copy_(g[r * H:(r + 1) * H, c * W:(c + 1) * W].to(s.core.device))


# ------------------------------------------------------------------ oracle
def reference_step(grid: torch.Tensor, coeffs: Sequence[float], radius: int = 1,
                   kind: str = "star") -> torch.Tensor:
    """Plain PyTorch fp64 periodic stencil on the GLOBAL grid (test oracle)."""
    g = grid.to(torch.float64)
    out = torch.zeros_like(g)
    if kind == "star":
        out += coeffs[0] * g
        for k in range(1, radius + 1):
            b = 1 + 4 * (k - 1)
            out += coeffs[b + 0] * torch.roll(g, k, dims=1)    # west neighbour
            out += coeffs[b + 1] * torch.roll(g, -k, dims=1)   # east
            out += coeffs[b + 2] * torch.roll(g, k, dims=0)    # north
            out += coeffs[b + 3] * torch.roll(g, -k, dims=0)   # south
    else:
        side = 2 * radius + 1
        for iy in range(side):
            for ix in range(side):
                w = coeffs[iy * side + ix]
                if w != 0.0:
                    out += w * torch.roll(g, (radius - iy, radius - ix), dims=(0, 1))
    return out
