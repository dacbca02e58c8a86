"""Host-side checks of the stencil layout / error ABI (no GPU needed)."""
import ctypes as C

import pytest

from cuda_mpi_scratch_b200 import _lib
from cuda_mpi_scratch_b200.models.stencil2d import jacobi_coeffs, layout, reference_step


def test_layout_alignment():
    L = layout(16384, 16384, 1)
    assert L.lead * 8 % 128 == 0            # core rows start on a 128 B line
    assert L.pitch % 16 == 0 and L.pitch >= L.lead + L.W + 2
    assert L.rows == 16384 + 2
    assert L.off_buf0 % 1024 == 0 and L.off_buf1 % 1024 == 0
    assert L.total_bytes >= L.pad_bytes + 2 * L.tile_bytes   # + 4 column landing buffers


def test_layout_rejects_bad_shapes():
    lay = _lib.StencilLayout()
    assert _lib.lib.b2c_stencil_layout(15, 16, 1, C.byref(lay)) != 0   # odd width
    assert b"even" in _lib.lib.b2c_last_error()
    assert _lib.lib.b2c_stencil_layout(16, 16, 5, C.byref(lay)) != 0   # radius too large
    with pytest.raises(_lib.CommError):
        layout(16, 16, 0)


def test_error_class_strings():
    assert _lib.lib.b2c_error_class_string(8) == b"message truncated"
    assert _lib.lib.b2c_error_class_string(0) == b"no error"


def test_reference_step_is_average():
    import torch
    g = torch.arange(36, dtype=torch.float64).reshape(6, 6)
    out = reference_step(g, jacobi_coeffs(1), 1)
    exp = (torch.roll(g, 1, 0) + torch.roll(g, -1, 0) + torch.roll(g, 1, 1) + torch.roll(g, -1, 1)) / 4
    assert torch.allclose(out, exp)
    ones = torch.ones(8, 8, dtype=torch.float64)
    assert torch.allclose(reference_step(ones, jacobi_coeffs(1, "box"), 1, "box"), ones)
