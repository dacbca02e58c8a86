"""Kernel-initiated point-to-point transfers between GPUs.

Capability parity with the reference's GPU ping-pong programs
(``test-benchmark/mpi-pingpong-gpu.cpp``: blocking ``MPI_Send``/``MPI_Recv`` of
device pointers; ``mpi-pingpong-gpu-async.cpp``: ``Isend``/``Irecv``/``Wait``).
A :class:`P2PChannel` gives every rank a receive buffer and one epoch flag per
source inside a symmetric, peer-mapped block.  ``put`` launches a kernel that
writes the payload into the destination rank's buffer over NVLink and then
publishes an epoch with a release store; ``wait`` enqueues a one-thread acquire
kernel on the receiver's stream.  Nothing touches the host.

``put``/``wait`` are stream-ordered (async); ``send``/``recv`` add the host
synchronisation of the blocking variant.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _lib
from ..parallel.symmetric import SymmetricBuffer, as_tensor
from ..parallel.world import LocalGroup, World

_L = _lib.lib
_vp, _i32 = C.c_void_p, C.c_int
_u64 = C.c_ulonglong
_lib._proto("b2c_p2p_put", [_vp, _vp, C.c_size_t, _vp, _u64, _vp, _i32, _i32, _vp])
_lib._proto("b2c_flag_wait", [_vp, _u64, _vp, _u64, _vp])
_lib._proto("b2c_flag_set", [_vp, _u64, _vp])
_lib._proto("b2c_pingpong_device", [_i32, _i32, C.c_size_t, _vp, _vp, _vp, _vp, _u64, _vp, _vp,
                                    _u64, _vp])

_OFF_FLAGS, _OFF_STATUS, _OFF_DATA = 0, 512, 1024
MODES = {"ldst": 0, "tma": 1}


def _stream(device: int) -> int:
    return torch.cuda.current_stream(device).cuda_stream


class P2PChannel:
    def __init__(self, rank: int, nranks: int, symbuf: SymmetricBuffer, nbytes: int,
                 world: World | None):
        self.rank, self.nranks, self.symbuf, self.nbytes, self.world = rank, nranks, symbuf, nbytes, world
        self.device = symbuf.device
        self.sent = [0] * nranks       # epochs published to each destination
        self.seen = [0] * nranks       # epochs consumed from each source
        p = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(_L.b2c_malloc(C.byref(p), 256))
        self._scratch = p              # [0,4) CTA arrival counter, [64,72) ping-pong time
        self.kernel_launches = 0

    # ---- factories ---------------------------------------------------------
    @staticmethod
    def block_bytes(nbytes: int) -> int:
        return _OFF_DATA + ((nbytes + 255) & ~255)

    @classmethod
    def create(cls, world: World, nbytes: int) -> "P2PChannel":
        sym = SymmetricBuffer.allocate(world, cls.block_bytes(nbytes))
        return cls(world.rank, world.size, sym, nbytes, world)

    @classmethod
    def create_group(cls, group: LocalGroup, nbytes: int) -> list["P2PChannel"]:
        syms = SymmetricBuffer.allocate_local(group, cls.block_bytes(nbytes))
        return [cls(r, group.size, syms[r], nbytes, None) for r in range(group.size)]

    # ---- views ---------------------------------------------------------------
    def recv_view(self, dtype=torch.uint8, numel: int | None = None) -> torch.Tensor:
        es = torch.empty(0, dtype=dtype).element_size()
        n = self.nbytes // es if numel is None else numel
        return self.symbuf.tensor(_OFF_DATA, (n,), dtype)

    def flags(self) -> torch.Tensor:
        return self.symbuf.tensor(_OFF_FLAGS, (self.nranks,), torch.int64)

    # ---- stream-ordered primitives -----------------------------------------------
    def put(self, dest: int, src: torch.Tensor, nbytes: int | None = None, mode: str = "ldst",
            max_ctas: int = 64, signal: bool = True) -> None:
        """Write ``src`` into ``dest``'s receive buffer and publish the next epoch."""
        n = src.numel() * src.element_size() if nbytes is None else nbytes
        if n > self.nbytes:
            raise ValueError(f"message of {n} B exceeds the channel's {self.nbytes} B buffer")
        peer = self.symbuf.peer_ptr(dest)
        flag = None
        if signal:
            self.sent[dest] += 1
            flag = C.c_void_p(peer + _OFF_FLAGS + 8 * self.rank)
        with torch.cuda.device(self.device):
            _lib.check(_L.b2c_p2p_put(C.c_void_p(peer + _OFF_DATA), C.c_void_p(src.data_ptr()), n,
                                      flag, self.sent[dest], self._scratch, MODES[mode], max_ctas,
                                      _stream(self.device)))
        self.kernel_launches += 1

    def wait(self, source: int, timeout_ms: int = 20000) -> None:
        """Block THE STREAM until the next message of ``source`` has landed."""
        self.seen[source] += 1
        with torch.cuda.device(self.device):
            _lib.check(_L.b2c_flag_wait(C.c_void_p(self.symbuf.local_ptr + _OFF_FLAGS + 8 * source),
                                        self.seen[source],
                                        C.c_void_p(self.symbuf.local_ptr + _OFF_STATUS), timeout_ms,
                                        _stream(self.device)))
        self.kernel_launches += 1

    # ---- blocking variants (mpi-pingpong-gpu.cpp) -------------------------------------
    def send(self, dest: int, src: torch.Tensor, **kw) -> None:
        self.put(dest, src, **kw)
        torch.cuda.current_stream(self.device).synchronize()

    def recv(self, source: int, out: torch.Tensor | None = None) -> torch.Tensor:
        self.wait(source)
        torch.cuda.current_stream(self.device).synchronize()
        self.check()
        if out is None:
            return self.recv_view()
        out.view(torch.uint8).view(-1).copy_(self.recv_view()[: out.numel() * out.element_size()])
        return out

    def check(self) -> None:
        st = self.symbuf.tensor(_OFF_STATUS, (1,), torch.int32)
        if int(st.item()) & 4:
            raise _lib.CommError(_lib.ERR_TIMEOUT, "P2P wait timed out on the device")

    # ---- latency floor: the whole loop inside one kernel per side -------------------------
    def pingpong_device(self, peer: int, role: int, iters: int, nbytes: int,
                        payload: torch.Tensor | None = None, timeout_ms: int = 20000) -> None:
        """Enqueue the device-resident ping-pong kernel.  Role 0 sends ``payload`` and waits
        for the echo; role 1 echoes what lands in its receive buffer.  BOTH ranks must call
        this, on different GPUs (the kernels wait for each other)."""
        if nbytes > self.nbytes:
            raise ValueError("payload larger than the channel buffer")
        src = payload.data_ptr() if role == 0 else self.symbuf.local_ptr + _OFF_DATA
        base = self.sent[peer]
        assert self.sent[peer] == self.seen[peer], "channel must be idle"
        with torch.cuda.device(self.device):
            _lib.check(_L.b2c_pingpong_device(
                role, iters, nbytes, C.c_void_p(src),
                C.c_void_p(self.symbuf.peer_ptr(peer) + _OFF_DATA),
                C.c_void_p(self.symbuf.local_ptr + _OFF_FLAGS + 8 * peer),
                C.c_void_p(self.symbuf.peer_ptr(peer) + _OFF_FLAGS + 8 * self.rank), base,
                C.c_void_p(self._scratch.value + 64),
                C.c_void_p(self.symbuf.local_ptr + _OFF_STATUS), timeout_ms, _stream(self.device)))
        self.sent[peer] += iters
        self.seen[peer] += iters
        self.kernel_launches += 1

    def pingpong_time_ns(self) -> float:
        """Device time of the last ping-pong loop on this rank (after a stream sync)."""
        return float(as_tensor(self._scratch.value + 64, (1,), torch.float64, self.device).item())

    def close(self) -> None:
        if self._scratch:
            _L.b2c_free(self._scratch)
            self._scratch = None
