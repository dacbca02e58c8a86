"""Halo plan + unfused exchange on HOST tiles: region algebra, the reference's
nine golden files (3x3 grid) and the 2x2 / 1x2 / 1x1 plumbing configurations."""
import os
import subprocess
import sys

import pytest
import torch

from cuda_mpi_scratch_b200.parallel import comm as cm
from cuda_mpi_scratch_b200.parallel.cart import CartTopology
from cuda_mpi_scratch_b200.parallel.datatype import DOUBLE
from cuda_mpi_scratch_b200.parallel import halo as H

REF = "/root/reference/stencil2d/sample-output"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def expected_tile(cart: CartTopology, rank: int, tile: int, ghost: int) -> torch.Tensor:
    """What every tile must look like after one exchange when core == rank id:
    each ghost block holds the id of the neighbour on that side."""
    tot = tile + 2 * ghost
    t = torch.full((tot, tot), float(rank), dtype=torch.float64)
    nb = cart.neighbours8(rank)   # NW N NE W E SW S SE
    g, T = ghost, tile
    rows = [slice(0, g), slice(g, g + T), slice(g + T, tot)]
    cols = [slice(0, g), slice(g, g + T), slice(g + T, tot)]
    order = [(0, 0), (0, 1), (0, 2), (1, 0), (1, 2), (2, 0), (2, 1), (2, 2)]
    for (ri, ci), n in zip(order, nb):
        t[rows[ri], cols[ci]] = float(n) if n >= 0 else -1.0
    return t


def test_region_algebra_matches_reference_values():
    # TestSubRegionExtraction (stencil2D.h:441-510) for a 32x32 core, 5x5 stencil
    g = H.Array2D(36, 36, 36)
    assert H.sub_array_region(g, 5, 5, H.TOP_LEFT) == H.Array2D(2, 2, 36, 0, 0)
    assert H.sub_array_region(g, 5, 5, H.TOP_CENTER) == H.Array2D(32, 2, 36, 2, 0)
    assert H.sub_array_region(g, 5, 5, H.CENTER_RIGHT) == H.Array2D(2, 32, 36, 34, 2)
    assert H.sub_array_region(g, 5, 5, H.BOTTOM_RIGHT) == H.Array2D(2, 2, 36, 34, 34)
    core = H.sub_array_region(g, 5, 5, H.CENTER)
    assert core == H.Array2D(32, 32, 36, 2, 2)
    # regions of the core are expressed in the coordinates of the same allocation
    assert H.sub_array_region(core, 5, 5, H.TOP) == H.Array2D(32, 2, 36, 2, 2)
    assert H.sub_array_region(core, 5, 5, H.LEFT) == H.Array2D(2, 32, 36, 2, 2)
    assert H.sub_array_region(core, 5, 5, H.RIGHT) == H.Array2D(2, 32, 36, 32, 2)
    assert H.sub_array_region(core, 5, 5, H.BOTTOM) == H.Array2D(32, 2, 36, 2, 32)
    assert H.sub_array_region(core, 5, 5, H.BOTTOM_LEFT) == H.Array2D(2, 2, 36, 2, 32)
    rep = H.region_report(32, 32, 5, 5)
    assert "Width: 36, Height: 36" in rep and "center: width:  32, height: 32" in rep


def test_rectangular_stencil_and_tile():
    g = H.Array2D(20 + 2 * 3, 10 + 2 * 1, 32)          # 7x3 stencil, padded row stride
    assert H.sub_array_region(g, 7, 3, H.CENTER) == H.Array2D(20, 10, 32, 3, 1)
    assert H.sub_array_region(g, 7, 3, H.RIGHT) == H.Array2D(3, 12, 32, 23, 0)


def _exchange_body(c, dims, tile, stencil, periodic=True):
    cart = c.cart_create(dims, periodic)
    ghost = stencil // 2
    tot = tile + 2 * ghost
    data = torch.full((tot, tot), -1.0, dtype=torch.float64)
    plan = H.HaloPlan(cart, H.Array2D(tot, tot, tot), stencil, stencil, DOUBLE)
    plan.core.view(data)[:] = float(cart.rank)
    before = H.print_array(data)
    sends = plan.entries("send")
    plan.exchange(data)
    return cart.cart_coords(), before, data, H.print_cart_grid(cart), sends


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference sample output not mounted")
def test_nine_golden_files_bit_for_bit():
    out = cm.run_local(9, _exchange_body, [3, 3], 16, 5)
    for coords, before, data, grid, _ in out:
        text = open(f"{REF}/{coords[0]}_{coords[1]}").read()
        assert f"Compute grid\n{grid}\n" in text
        assert f"Array\n{before}\n" in text
        assert text.endswith("Array after exchange\n" + H.print_array(data))


@pytest.mark.parametrize("dims,n", [([3, 3], 9), ([2, 2], 4), ([1, 2], 2), ([2, 4], 8), ([1, 1], 1)])
def test_exchange_matches_neighbour_ids(dims, n):
    out = cm.run_local(n, _exchange_body, dims, 16, 5)
    cart = CartTopology.create(dims, True)
    for rank, (coords, _, data, _, sends) in enumerate(out):
        assert torch.equal(data, expected_tile(cart, rank, 16, 2))
        # tag = id of the send region, unique per direction even when peers coincide
        assert [s["tag"] for s in sends] == [H.TOP_LEFT, H.TOP, H.TOP_RIGHT, H.LEFT, H.RIGHT,
                                             H.BOTTOM_LEFT, H.BOTTOM, H.BOTTOM_RIGHT]


def test_exchange_non_periodic_leaves_outer_ghosts():
    out = cm.run_local(4, _exchange_body, [2, 2], 8, 3, False)
    cart = CartTopology.create([2, 2], False)
    for rank, (_, _, data, _, _) in enumerate(out):
        assert torch.equal(data, expected_tile(cart, rank, 8, 1))


def test_halo_message_sizes_of_the_sample_run():
    # 16x16 tile, ghost 2 -> edges 256 B, corners 32 B
    out = cm.run_local(1, _exchange_body, [1, 1], 16, 5)
    sizes = sorted(s["bytes"] for s in out[0][4])
    assert sizes == [32, 32, 32, 32, 256, 256, 256, 256]


def test_cpu_driver_2x2_processes_plumbing_config(tmp_path):
    """BASELINE.json config 1: 4 ranks as PROCESSES over the shared-memory
    fabric, 2x2 periodic grid (left/right and up/down neighbours coincide, the
    four diagonals are one rank -- tags keep the 8 messages apart)."""
    cmd = [sys.executable, "-m", "cuda_mpi_scratch_b200.launch", "-n", "4", "--timeout", "120",
           os.path.join(ROOT, "tests/mp/halo_cpu_driver.py"), str(tmp_path), "256", "3"]
    subprocess.run(cmd, check=True, cwd=ROOT, timeout=180)
    cart = CartTopology.create([2, 2], True)
    for rank in range(4):
        r, c = cart.coords(rank)
        text = (tmp_path / f"{r}_{c}").read_text()
        assert text.startswith(f"Rank:  {rank}\nCoord: {r}, {c}\n\nCompute grid\n0 1 \n2 3 \n")
        got = text.split("Array after exchange\n")[1]
        assert got == H.print_array(expected_tile(cart, rank, 256, 1))
