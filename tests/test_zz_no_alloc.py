"""``b2c_dot_local_noalloc``: the dot product for callers that may not allocate.

Who needs it: code that has to produce a dot product at a point where device
allocation is forbidden or can fail -- inside another runtime's callback, in a
latency-critical loop (``cudaMalloc`` synchronises the whole device), or in a
process whose GPU memory is already owned by a framework allocator or a fixed
arena.  The plan-based entry point (``b2c_dot_create``) allocates its scratch;
this one keeps the scratch in a fixed array inside the library image.

The test drives the new entry point through the C ABI only (ctypes on the shared
object, raw device pointers and a stream) and checks it against the existing
plan-based entry point, twice in a row because the built-in scratch keeps the
previous call's contents.  The file name sorts last on purpose: a run that stops
at the first failure has then already executed every other test.
"""
import ctypes as C

import pytest
import torch

from cuda_mpi_scratch_b200 import _lib

pytestmark = pytest.mark.gpu

L = _lib.lib
L.b2c_dot_local_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int,
                                    C.c_void_p, C.c_void_p]
L.b2c_dot_local_noalloc.restype = C.c_int
L.b2c_dot_local_noalloc_limits.argtypes = [C.POINTER(C.c_int), C.POINTER(C.c_size_t)]
L.b2c_dot_local_noalloc_limits.restype = C.c_int
L.b2c_dot_symmetric_bytes.argtypes = [C.c_int, C.POINTER(C.c_size_t)]
L.b2c_dot_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.POINTER(C.c_void_p)]
L.b2c_dot_run.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int,
                          C.c_void_p]
L.b2c_dot_result.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]
L.b2c_dot_destroy.argtypes = [C.c_void_p]

F64 = 9
F32 = 8


def _existing_entry_point(x, y, dtype_code, stream):
    """Reference value from the plan-based API (b2c_dot_create / run / result), 1 rank."""
    nbytes = C.c_size_t()
    assert L.b2c_dot_symmetric_bytes(1, C.byref(nbytes)) == 0
    slots = torch.zeros(nbytes.value, dtype=torch.uint8, device=x.device)
    tables = (C.c_void_p * 1)(slots.data_ptr())
    plan = C.c_void_p()
    assert L.b2c_dot_create(C.byref(plan), 0, 1, tables) == 0, L.b2c_last_error()
    vals = []
    for _ in range(2):
        assert L.b2c_dot_run(plan, x.data_ptr(), y.data_ptr(), x.numel(), dtype_code, -1, 3, stream) == 0
        v = C.c_double()
        assert L.b2c_dot_result(plan, stream, C.byref(v)) == 0, L.b2c_last_error()
        vals.append(v.value)
    L.b2c_dot_destroy(plan)
    return vals


@pytest.mark.parametrize("dtype,code,n", [(torch.float64, F64, 3_000_017), (torch.float32, F32, 1_000_003)])
def test_no_alloc_matches_existing_entry_point_two_calls_in_a_row(dtype, code, n):
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    g = torch.Generator(device="cuda").manual_seed(11)
    x = torch.rand(n, dtype=dtype, device="cuda", generator=g) - 0.5
    y = torch.rand(n, dtype=dtype, device="cuda", generator=g) - 0.5
    out = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    want = _existing_entry_point(x, y, code, stream)

    got = []
    for call in range(2):
        if call == 1:          # second call on different data: stale scratch must not leak in
            x, y = y * 2.0, x + 0.25
            torch.cuda.synchronize()
            want = _existing_entry_point(x, y, code, stream)
        out.fill_(float("nan"))
        rc = L.b2c_dot_local_noalloc(x.data_ptr(), y.data_ptr(), n, code, 0, out.data_ptr(), stream)
        assert rc == 0, L.b2c_last_error()
        torch.cuda.synchronize()
        got.append(float(out.item()))
        # same kernel, same grid, same summation order: tolerance of the other dot tests
        assert got[-1] == pytest.approx(want[0], rel=1e-12, abs=1e-9)
        assert want[0] == want[1]
    oracle = float(torch.dot(x.double().cpu(), y.double().cpu()))
    assert got[-1] == pytest.approx(oracle, rel=1e-10 if dtype == torch.float64 else 1e-6, abs=1e-6)


def test_no_alloc_explicit_grid_and_ones():
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    n = (1 << 22) + 5
    x = torch.ones(n, dtype=torch.float64, device="cuda")
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    for ctas in (1, 37, 1024):
        assert L.b2c_dot_local_noalloc(x.data_ptr(), x.data_ptr(), n, F64, ctas, out.data_ptr(), stream) == 0
        torch.cuda.synchronize()
        assert float(out.item()) == float(n)


def test_no_alloc_rejects_what_the_scratch_cannot_hold():
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    max_ctas, nbytes = C.c_int(), C.c_size_t()
    assert L.b2c_dot_local_noalloc_limits(C.byref(max_ctas), C.byref(nbytes)) == 0
    assert max_ctas.value == 1024 and nbytes.value == 8448
    x = torch.ones(1024, dtype=torch.float64, device="cuda")
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    rc = L.b2c_dot_local_noalloc(x.data_ptr(), x.data_ptr(), 1024, F64, max_ctas.value + 1,
                                 out.data_ptr(), stream)
    assert rc != 0
    msg = L.b2c_last_error().decode()
    assert "at most 1024" in msg and "1025" in msg
    assert L.b2c_dot_local_noalloc(x.data_ptr(), x.data_ptr(), 1024, F64, 0, None, stream) != 0
    assert L.b2c_dot_local_noalloc(x.data_ptr(), x.data_ptr(), 1024, 5, 0, out.data_ptr(), stream) != 0
