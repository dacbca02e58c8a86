"""ctypes binding of ``libb200comm.so`` (the C ABI in ``include/b200comm.h``).

The library is built in-tree by :mod:`cuda_mpi_scratch_b200.build`.  If it is
missing we try to build it once; if that fails the import raises -- there is
no eager/PyTorch fallback for the kernels, by design: a GPU box without the
native library must fail loudly rather than silently run something else.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_PKG = Path(__file__).resolve().parent
LIB_PATH = _PKG / "lib" / "libb200comm.so"


class B200Error(RuntimeError):
    """Raised when a native call returns a non-zero error class."""

    def __init__(self, code: int, message: str):
        super().__init__(message)
        self.code = code


class CommError(B200Error):
    pass


class CudaError(B200Error):
    pass


ERR_CUDA = 10
ERR_TIMEOUT = 11
ERR_TRUNCATE = 8


def _build_module():
    # import build.py WITHOUT going through the package (its import is in progress here)
    import importlib.util

    spec = importlib.util.spec_from_file_location("_b200comm_build", _PKG / "build.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _load() -> C.CDLL:
    """Load the library; (re)build it first if it is missing or was built from different
    sources (content digest, not mtimes: a snapshot copy must not trigger rebuilds)."""
    mod = _build_module()
    digest_file = LIB_PATH.with_name(LIB_PATH.name + ".digest")
    want = mod.source_digest() if (mod.CSRC.is_dir() and os.environ.get("B200COMM_NO_AUTOBUILD") != "1") else None
    have = digest_file.read_text().strip() if digest_file.exists() else None
    if os.environ.get("B200COMM_REBUILD") == "1" or not LIB_PATH.exists() or (want and want != have):
        import fcntl

        LIB_PATH.parent.mkdir(parents=True, exist_ok=True)
        with open(LIB_PATH.parent / ".build.lock", "w") as lock:     # one builder at a time
            fcntl.flock(lock, fcntl.LOCK_EX)
            have = digest_file.read_text().strip() if digest_file.exists() else None
            if os.environ.get("B200COMM_REBUILD") == "1" or not LIB_PATH.exists() or (want and want != have):
                mod.build()
    return C.CDLL(str(LIB_PATH), mode=C.RTLD_GLOBAL)


lib = _load()
lib.b2c_last_error.restype = C.c_char_p
lib.b2c_error_class_string.restype = C.c_char_p
lib.b2c_error_class_string.argtypes = [C.c_int]


def check(rc: int) -> None:
    if rc == 0:
        return
    msg = lib.b2c_last_error().decode(errors="replace")
    if not msg:
        msg = lib.b2c_error_class_string(rc).decode()
    if rc == ERR_CUDA:
        raise CudaError(rc, msg)
    raise CommError(rc, msg)


def _proto(name: str, argtypes, restype=C.c_int):
    fn = getattr(lib, name)
    fn.argtypes = argtypes
    fn.restype = restype
    return fn


vp = C.c_void_p
sz = C.c_size_t
i32 = C.c_int
u64 = C.c_ulonglong
pvp = C.POINTER(C.c_void_p)
pi32 = C.POINTER(C.c_int)


class StencilLayout(C.Structure):
    _fields_ = [("W", i32), ("H", i32), ("radius", i32), ("ghost", i32), ("lead", i32),
                ("pitch", i32), ("rows", i32), ("pad_bytes", sz), ("tile_bytes", sz),
                ("off_buf0", sz), ("off_buf1", sz), ("off_red_slots", sz), ("total_bytes", sz)]


class StencilInfo(C.Structure):
    _fields_ = [("cur", i32), ("steps_done", u64), ("grid", i32), ("threads", i32),
                ("nstages", i32), ("tiles_x", i32), ("tiles_y", i32), ("ncoef", i32),
                ("launches_per_step", i32), ("smem_bytes", sz), ("buf", vp * 2),
                ("resid", vp), ("launches", u64)]


# ---- core
_proto("b2c_set_error_mode", [i32])
_proto("b2c_version", [])
_proto("b2c_device_count", [pi32])
_proto("b2c_bind_device", [i32, i32, i32, pi32])
_proto("b2c_set_device", [i32])
_proto("b2c_get_device", [pi32])
_proto("b2c_device_synchronize", [])
_proto("b2c_stream_synchronize", [vp])
_proto("b2c_device_sm_count", [i32, pi32])
_proto("b2c_malloc", [pvp, sz])
_proto("b2c_free", [vp])
_proto("b2c_memset_async", [vp, i32, sz, vp])
_proto("b2c_memcpy_async", [vp, vp, sz, i32, vp])
_proto("b2c_pinned_alloc", [pvp, sz])
_proto("b2c_pinned_free", [vp])
_proto("b2c_ipc_export", [vp, C.c_char_p])
_proto("b2c_ipc_open", [C.c_char_p, pvp])
_proto("b2c_ipc_close", [vp])
_proto("b2c_enable_peer_access", [i32, i32])
_proto("b2c_can_access_peer", [i32, i32, pi32])
pu64 = C.POINTER(C.c_ulonglong)
psz = C.POINTER(C.c_size_t)
_proto("b2c_vmm_support", [i32, pi32, pi32])
_proto("b2c_vmm_granularity", [i32, i32, psz, psz])
_proto("b2c_vmm_create", [sz, i32, pu64, pi32])
_proto("b2c_vmm_import", [i32, pu64])
_proto("b2c_vmm_map", [u64, sz, sz, i32, pvp])
_proto("b2c_vmm_unmap", [vp, sz])
_proto("b2c_vmm_release", [u64])
_proto("b2c_mc_create", [i32, sz, pu64, pi32])
_proto("b2c_mc_add_device", [u64, i32])
_proto("b2c_mc_bind", [u64, u64, sz])
_proto("b2c_close_fd", [i32])
# ---- stencil
_proto("b2c_stencil_layout", [i32, i32, i32, C.POINTER(StencilLayout)])
_proto("b2c_stencil_create", [pvp, i32, i32, i32, i32, vp, C.POINTER(vp * 8), i32])
_proto("b2c_stencil_destroy", [vp])
_proto("b2c_stencil_set_coef", [vp, C.POINTER(C.c_double), i32, vp])
_proto("b2c_stencil_set_timeout_ms", [vp, u64])
_proto("b2c_stencil_info", [vp, C.POINTER(StencilInfo)])
_proto("b2c_stencil_step", [vp, i32, i32, vp])
_proto("b2c_stencil_step_levels", [vp, i32, i32, vp])
_proto("b2c_stencil_step_io", [vp, i32, i32, C.POINTER(C.c_double), i32, vp, C.POINTER(C.c_ulonglong)])
_proto("b2c_stencil_io_wait", [vp, C.c_ulonglong, C.POINTER(C.c_double)])
_proto("b2c_stencil_drain", [vp, vp])
_proto("b2c_stencil_advance", [vp, C.c_longlong])
_proto("b2c_stencil_exchange_push", [vp, vp])
_proto("b2c_stencil_exchange_wait", [vp, vp])
_proto("b2c_stencil_fill", [vp, i32, i32, C.c_double, vp])
_proto("b2c_stencil_status", [vp, vp, C.POINTER(C.c_uint)])
_proto("b2c_stencil_phase_stats", [vp, vp, C.POINTER(C.c_double)])


def have(name: str) -> bool:
    return hasattr(lib, name)
