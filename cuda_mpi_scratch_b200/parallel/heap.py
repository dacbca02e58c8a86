"""Symmetric heap: ONE peer-mapped (and, on NVSwitch systems, multicast-bound) allocation per
rank, grown in arenas, from which stencil plans, dot plans, P2P channels and collective
workspaces are sub-allocated.

Round 1 gave every plan its own ``cudaMalloc`` + ``cudaIpcOpenMemHandle`` per peer
(``SymmetricBuffer.allocate``).  Here (``csrc/core/vmm.cpp`` does the driver calls):

* every rank creates a physical allocation with ``cuMemCreate`` that is exportable as a POSIX file
  descriptor; the descriptors are passed between the rank processes over Unix sockets
  (``SCM_RIGHTS``); every rank imports and maps every peer's allocation once per arena;
* one NVLink **multicast object** per arena: rank 0 creates it, everybody imports it, adds its
  device and binds its arena.  ``mc_ptr + offset`` then addresses the same offset in EVERY rank's
  arena: ``multimem.st`` stores to all of them, ``multimem.ld_reduce`` returns their sum, both
  executed by the switch (``csrc/collectives/device_coll.cu``);
* a first-fit free list hands out blocks; allocation is collective and deterministic, so a block
  has the same offset on every rank.

Reference counterpart: none -- the reference passes device pointers to a CUDA-aware MPI and lets
the library map them per message (``stencil2d/stencil2D.h:369-373``).  Capability backed:
``MPI_Bcast`` ``mpicuda2.cu:154``, ``MPI_Allreduce`` ``mpi9.cpp:51-54``, ``MPI_Reduce``
``mpicuda4.cu:368`` on device memory.
"""
from __future__ import annotations

import ctypes as C
import os
import socket
import threading

from .. import _lib
from .world import World

_L = _lib.lib


def _send_fd_to_peers(path: str, fd: int, npeers: int, ready: threading.Event) -> threading.Thread:
    """Serve ``fd`` to ``npeers`` connections on the abstract Unix socket ``path``."""
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    srv.bind(path)
    srv.listen(npeers + 1)
    ready.set()

    def serve():
        try:
            for _ in range(npeers):
                conn, _ = srv.accept()
                with conn:
                    socket.send_fds(conn, [b"f"], [fd])
        finally:
            srv.close()

    t = threading.Thread(target=serve, daemon=True)
    t.start()
    return t


def _recv_fd(path: str) -> int:
    with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as c:
        c.connect(path)
        _, fds, _, _ = socket.recv_fds(c, 16, 1)
        if not fds:
            raise _lib.CommError(_lib.ERR_PEER, f"no file descriptor received from {path!r}")
        return fds[0]


def _exchange_fds(world: World, my_fd: int | None, tag: str, only_from: int | None = None) -> dict[int, int]:
    """Collective: every rank that has ``my_fd`` serves it; returns {rank: received fd} for the
    ranks asked (all others, or just ``only_from``)."""
    path = f"\0b200heap-{os.getpid()}-{tag}"
    t = None
    if my_fd is not None:
        n = world.size - 1
        ready = threading.Event()
        t = _send_fd_to_peers(path, my_fd, n, ready)
        ready.wait()
    paths = world.allgather_bytes(path.encode() if my_fd is not None else b"")
    out = {}
    for r, p in enumerate(paths):
        if r == world.rank or not p or (only_from is not None and r != only_from):
            continue
        out[r] = _recv_fd(p.decode())
    if t is not None:
        if only_from is not None and world.rank == only_from:
            pass
        t.join(timeout=60)
    world.barrier()
    return out


class _Arena:
    def __init__(self):
        self.bytes = 0
        self.handle = 0                 # my physical allocation
        self.peer_handles: dict[int, int] = {}
        self.ptrs: list[int] = []       # base address of every rank's arena as mapped here
        self.mc_handle = 0
        self.mc_ptr = 0                 # 0: no multicast
        self.free: list[list[int]] = [] # [offset, size], sorted by offset


class HeapBlock:
    """A block of the heap: the same offset in every rank's arena."""

    def __init__(self, heap: "SymmetricHeap", arena: _Arena, offset: int, nbytes: int, serial: int):
        self.heap, self.arena, self.offset, self.nbytes, self.serial = heap, arena, offset, nbytes, serial

    @property
    def local_ptr(self) -> int:
        return self.arena.ptrs[self.heap.world.rank] + self.offset

    @property
    def peer_ptrs(self) -> list[int]:
        return [p + self.offset for p in self.arena.ptrs]

    @property
    def mc_ptr(self) -> int:
        """Multicast address of this block (0 when the arena has no multicast binding)."""
        return self.arena.mc_ptr + self.offset if self.arena.mc_ptr else 0


def _round_up(v: int, a: int) -> int:
    return (v + a - 1) // a * a


class SymmetricHeap:
    """One per world; get it with :meth:`for_world`.  Every method that changes the heap
    (``alloc``, arena growth) is COLLECTIVE: all ranks call it in the same order with the same
    arguments.  ``free`` may be called at any time (also from a finaliser): the block returns to the
    free list at the next ``alloc`` once every rank has freed it."""

    _by_world: dict[int, "SymmetricHeap"] = {}

    def __init__(self, world: World):
        self.world = world
        self.device = world.device
        vm, mc = C.c_int(0), C.c_int(0)
        _lib.check(_L.b2c_vmm_support(self.device, C.byref(vm), C.byref(mc)))
        if not vm.value:
            raise _lib.CommError(_lib.ERR_STATE, "this GPU/driver cannot export allocations as file descriptors")
        ag, mg = C.c_size_t(0), C.c_size_t(0)
        _lib.check(_L.b2c_vmm_granularity(self.device, world.size, C.byref(ag), C.byref(mg)))
        self.granularity = max(ag.value, mg.value, 2 << 20)
        want_mc = bool(mc.value) and mg.value > 0 and os.environ.get("B200COMM_MULTICAST", "1") != "0"
        # every rank must see the same answer
        self.multicast_possible = world.allreduce_max(0.0 if want_mc else 1.0) == 0.0
        self.arena_bytes = int(os.environ.get("B200COMM_HEAP_ARENA_MB", "256")) << 20
        self.arenas: list[_Arena] = []
        self._serial = 0
        self._live: dict[int, HeapBlock] = {}
        self._pending_free: set[int] = set()
        self._seq = 0
        self.stats = {"arenas": 0, "multicast_arenas": 0, "allocs": 0, "bytes_mapped": 0}

    @classmethod
    def for_world(cls, world: World) -> "SymmetricHeap":
        h = cls._by_world.get(id(world))
        if h is None:
            h = cls._by_world[id(world)] = cls(world)
        return h

    @staticmethod
    def enabled(world: World) -> bool:
        """Heap allocations for multi-process worlds unless ``B200COMM_HEAP=ipc``."""
        return world.size > 1 and os.environ.get("B200COMM_HEAP", "vmm") != "ipc"

    # ------------------------------------------------------------------ arenas
    def _new_arena(self, min_bytes: int, multicast: bool) -> _Arena:
        w, dev = self.world, self.device
        a = _Arena()
        a.bytes = _round_up(max(min_bytes, self.arena_bytes), self.granularity)
        self._seq += 1
        h, fd = C.c_ulonglong(0), C.c_int(-1)
        _lib.check(_L.b2c_vmm_create(a.bytes, dev, C.byref(h), C.byref(fd)))
        a.handle = h.value
        fds = _exchange_fds(w, fd.value, f"a{self._seq}")
        _L.b2c_close_fd(fd.value)
        a.ptrs = [0] * w.size
        for r in range(w.size):
            hr = a.handle
            if r != w.rank:
                hh = C.c_ulonglong(0)
                _lib.check(_L.b2c_vmm_import(fds[r], C.byref(hh)))
                _L.b2c_close_fd(fds[r])
                hr = a.peer_handles[r] = hh.value
            p = C.c_void_p()
            _lib.check(_L.b2c_vmm_map(hr, a.bytes, self.granularity, dev, C.byref(p)))
            a.ptrs[r] = p.value
        _lib.check(_L.b2c_memset_async(C.c_void_p(a.ptrs[w.rank]), 0, a.bytes, None))
        _lib.check(_L.b2c_device_synchronize())
        w.barrier()
        if multicast and self.multicast_possible:
            self._bind_multicast(a)
        a.free = [[0, a.bytes]]
        self.arenas.append(a)
        self.stats["arenas"] += 1
        self.stats["multicast_arenas"] += 1 if a.mc_ptr else 0
        self.stats["bytes_mapped"] += a.bytes * w.size
        return a

    def _bind_multicast(self, a: _Arena) -> None:
        """rank 0 creates the multicast object, all import it, add their device, then bind."""
        w, dev = self.world, self.device
        ok, mc = True, 0
        fd0 = None
        try:
            if w.rank == 0:
                h, fd = C.c_ulonglong(0), C.c_int(-1)
                _lib.check(_L.b2c_mc_create(w.size, a.bytes, C.byref(h), C.byref(fd)))
                mc, fd0 = h.value, fd.value
        except _lib.CommError:
            ok = False
        ok = w.allreduce_max(0.0 if ok else 1.0) == 0.0
        if not ok:
            self.multicast_possible = False
            return
        self._seq += 1
        fds = _exchange_fds(w, fd0, f"m{self._seq}", only_from=0)
        try:
            if w.rank == 0:
                _L.b2c_close_fd(fd0)
            else:
                h = C.c_ulonglong(0)
                _lib.check(_L.b2c_vmm_import(fds[0], C.byref(h)))
                _L.b2c_close_fd(fds[0])
                mc = h.value
            _lib.check(_L.b2c_mc_add_device(mc, dev))
        except _lib.CommError:
            ok = False
        ok = w.allreduce_max(0.0 if ok else 1.0) == 0.0      # also: every device added before any bind
        if ok:
            try:
                _lib.check(_L.b2c_mc_bind(mc, a.handle, a.bytes))
            except _lib.CommError:
                ok = False
            ok = w.allreduce_max(0.0 if ok else 1.0) == 0.0
        if ok:
            try:
                p = C.c_void_p()
                _lib.check(_L.b2c_vmm_map(mc, a.bytes, self.granularity, dev, C.byref(p)))
                a.mc_handle, a.mc_ptr = mc, p.value
            except _lib.CommError:
                ok = False
            ok = w.allreduce_max(0.0 if ok else 1.0) == 0.0
            if not ok:
                a.mc_ptr = 0
        if not ok:
            self.multicast_possible = False

    # ------------------------------------------------------------------ blocks
    def _apply_frees(self) -> None:
        """Blocks freed by EVERY rank go back to the free lists (same order everywhere)."""
        mine = sorted(self._pending_free)
        lists = self.world.allgather_bytes(",".join(map(str, mine)).encode())
        common = None
        for b in lists:
            s = set(int(x) for x in b.decode().split(",") if x)
            common = s if common is None else (common & s)
        for serial in sorted(common or ()):
            blk = self._live.pop(serial, None)
            self._pending_free.discard(serial)
            if blk is None:
                continue
            fl = blk.arena.free
            fl.append([blk.offset, blk.nbytes])
            fl.sort()
            merged = [fl[0]]
            for off, n in fl[1:]:
                if merged[-1][0] + merged[-1][1] == off:
                    merged[-1][1] += n
                else:
                    merged.append([off, n])
            blk.arena.free = merged

    def alloc(self, nbytes: int, align: int = 4096, multicast: bool = False) -> HeapBlock:
        """Collective.  A zero-filled block of ``nbytes`` at the same offset on every rank;
        ``multicast=True`` asks for an arena with a multicast binding (``block.mc_ptr`` is 0 when
        the system has none)."""
        nbytes = _round_up(max(int(nbytes), 1), 4096)
        align = max(int(align), 4096)
        self._apply_frees()
        want_mc = multicast and self.multicast_possible
        blk = None
        for a in self.arenas:
            if bool(a.mc_ptr) != want_mc:
                continue
            for i, (off, n) in enumerate(a.free):
                start = _round_up(off, align)
                if start + nbytes <= off + n:
                    pieces = []
                    if start > off:
                        pieces.append([off, start - off])
                    if start + nbytes < off + n:
                        pieces.append([start + nbytes, off + n - start - nbytes])
                    a.free[i:i + 1] = pieces
                    blk = (a, start)
                    break
            if blk:
                break
        if blk is None:
            a = self._new_arena(nbytes, want_mc)
            if want_mc and not a.mc_ptr:          # multicast turned out to be unavailable
                want_mc = False
            a.free = [[nbytes, a.bytes - nbytes]] if a.bytes > nbytes else []
            blk = (a, 0)
            fresh = True
        else:
            fresh = False
        self._serial += 1
        b = HeapBlock(self, blk[0], blk[1], nbytes, self._serial)
        self._live[b.serial] = b
        self.stats["allocs"] += 1
        if not fresh:                             # a fresh arena is already zero and fenced by a barrier
            _lib.check(_L.b2c_memset_async(C.c_void_p(b.local_ptr), 0, nbytes, None))
            _lib.check(_L.b2c_device_synchronize())
            self.world.barrier()
        return b

    def free(self, block: HeapBlock) -> None:
        self._pending_free.add(block.serial)
