"""Fused stencil + halo kernel vs a plain PyTorch fp64 oracle (needs a GPU).

Multi-rank cases run as virtual ranks in one process, stepped in lock-step
(``StencilGroup``): the same kernels, peer pointers and epoch flags as the
multi-process path, but never two kernels that wait on each other.
"""
import os


import pytest
import torch

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.models.stencil2d import Stencil2D, jacobi_coeffs, reference_step

pytestmark = pytest.mark.gpu

REF = "/root/reference/stencil2d/sample-output"


def _rand_global(h, w, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.rand(h, w, dtype=torch.float64, generator=g)


def _run_group(dims, tile_hw, radius, kind, steps, coeffs=None, periods=True, seed=0, temporal=1):
    n = dims[0] * dims[1]
    grp = Stencil2D.create_group(cms.LocalGroup(n), tile_hw, radius, kind, dims=dims, periods=periods,
                                 temporal=temporal)
    glob = _rand_global(dims[0] * tile_hw[0], dims[1] * tile_hw[1], seed)
    for s in grp.ranks:
        s.fill(0.0, 0.0)
        if coeffs is not None:
            s.set_coeffs(coeffs)
    grp.scatter_global(glob)
    grp.exchange()
    grp.step(steps, residual=True)
    grp.sync()
    grp.check()
    return grp, glob


@pytest.mark.parametrize("radius,kind", [(1, "star"), (2, "star"), (3, "star"), (4, "star"),
                                         (1, "box"), (2, "box")])
def test_single_rank_periodic_matches_oracle(radius, kind):
    H, W = 96, 384   # 3 x 3 tiles
    coeffs = [0.1 * (i + 1) / 7 for i in range(4 * radius + 1 if kind == "star" else (2 * radius + 1) ** 2)]
    grp, glob = _run_group((1, 1), (H, W), radius, kind, steps=4, coeffs=coeffs)
    exp = glob.clone()
    for _ in range(4):
        exp = reference_step(exp, coeffs, radius, kind)
    got = grp.gather_global()
    assert torch.allclose(got, exp, rtol=1e-12, atol=1e-13)


@pytest.mark.parametrize("dims", [(1, 2), (2, 1), (2, 2), (2, 4), (3, 3)])
def test_virtual_ranks_bitwise_equal_single_rank(dims):
    th, tw = 64, 256
    steps = 5
    grp, glob = _run_group(dims, (th, tw), 1, "star", steps)
    one, _ = _run_group((1, 1), (dims[0] * th, dims[1] * tw), 1, "star", steps)
    a, b = grp.gather_global(), one.gather_global()
    assert torch.equal(a, b), "decomposed run must be bit-identical to the single-tile run"
    exp = glob.clone()
    for _ in range(steps):
        exp = reference_step(exp, jacobi_coeffs(1), 1)
    assert torch.allclose(a, exp, rtol=1e-12, atol=1e-13)


def test_box_stencil_needs_corners_2x2():
    grp, glob = _run_group((2, 2), (32, 128), 1, "box", 3)
    exp = glob.clone()
    for _ in range(3):
        exp = reference_step(exp, jacobi_coeffs(1, "box"), 1, "box")
    assert torch.allclose(grp.gather_global(), exp, rtol=1e-12, atol=1e-13)


def test_partial_tiles_and_radius2_2x2():
    grp, glob = _run_group((2, 2), (50, 200), 2, "star", 3)
    exp = glob.clone()
    for _ in range(3):
        exp = reference_step(exp, jacobi_coeffs(2), 2)
    assert torch.allclose(grp.gather_global(), exp, rtol=1e-12, atol=1e-13)


def test_non_periodic_keeps_dirichlet_ghosts():
    dims, th, tw = (2, 2), 32, 128
    grp = Stencil2D.create_group(cms.LocalGroup(4), (th, tw), 1, "star", dims=dims, periods=False)
    for s in grp.ranks:
        s.fill(1.0, 5.0)   # core 1, ghosts (the fixed boundary) 5
    grp.exchange()
    grp.step(2)
    grp.sync()
    grp.check()
    # oracle: pad the global grid with the boundary value and iterate
    g = torch.full((2 * th + 2, 2 * tw + 2), 5.0, dtype=torch.float64)
    g[1:-1, 1:-1] = 1.0
    for _ in range(2):
        n = g.clone()
        n[1:-1, 1:-1] = 0.25 * (g[:-2, 1:-1] + g[2:, 1:-1] + g[1:-1, :-2] + g[1:-1, 2:])
        g = n
    assert torch.allclose(grp.gather_global(), g[1:-1, 1:-1], rtol=1e-13, atol=1e-14)


def test_residual_matches_oracle():
    grp, glob = _run_group((2, 2), (64, 128), 1, "star", 3)
    exp = glob.clone()
    for _ in range(2):
        exp = reference_step(exp, jacobi_coeffs(1), 1)
    last = reference_step(exp, jacobi_coeffs(1), 1)
    want = float(((last - exp) ** 2).sum().sqrt())
    assert grp.residual() == pytest.approx(want, rel=1e-11)


@pytest.mark.parametrize("algo", ["strip", "tile"])
@pytest.mark.parametrize("dims,tile_hw,radius", [((1, 1), (96, 600), 1), ((2, 2), (64, 256), 1),
                                                   ((2, 4), (40, 504), 2), ((3, 3), (32, 260), 1),
                                                   ((1, 2), (50, 200), 2)])
def test_both_kernels_agree_bitwise_and_match_oracle(monkeypatch, algo, dims, tile_hw, radius):
    """The plan picks the strip kernel only for very large tiles; force each kernel on
    small multi-rank cases (several strips, partial last strip, partial last chunk)."""
    monkeypatch.setenv("B2C_STENCIL_ALGO", algo)
    steps = 4
    grp, glob = _run_group(dims, tile_hw, radius, "star", steps)
    exp = glob.clone()
    for _ in range(steps):
        exp = reference_step(exp, jacobi_coeffs(radius), radius)
    got = grp.gather_global()
    assert torch.allclose(got, exp, rtol=1e-12, atol=1e-13)
    monkeypatch.setenv("B2C_STENCIL_ALGO", "tile" if algo == "strip" else "strip")
    other, _ = _run_group(dims, tile_hw, radius, "star", steps)
    assert torch.equal(got, other.gather_global()), "strip and tile kernels must agree bit for bit"
    assert grp.residual() == pytest.approx(other.residual(), rel=1e-12)
    # ghost cells (incl. landing-buffer unpack) complete after the drain
    s = grp.ranks[0]
    R = radius
    ri = (torch.arange(-R, tile_hw[0] + R)) % exp.shape[0]
    ci = (torch.arange(-R, tile_hw[1] + R)) % exp.shape[1]
    assert torch.allclose(s.tile.cpu(), exp[ri][:, ci], rtol=1e-12, atol=1e-13)


@pytest.mark.parametrize("T", [2, 3, 4])
@pytest.mark.parametrize("dims,tile_hw", [((1, 1), (96, 384)), ((2, 2), (64, 256)), ((2, 4), (40, 504)),
                                          ((3, 3), (32, 260)), ((1, 2), (50, 200)), ((1, 1), (256, 1024))])
def test_temporal_blocking_is_bit_identical_to_single_steps(dims, tile_hw, T):
    """temporal=T: one sweep = T time steps (ghost width T, corners exchanged).  Same bits as
    the single-step kernel, same residual (of the last update), complete ghost ring."""
    steps = 12 if dims != (2, 2) else 2 * T + (T - 1)      # (2,2): ends with a partial sweep of T-1 levels
    coeffs = [0.1, 0.2, 0.25, 0.15, 0.3]
    tb, glob = _run_group(dims, tile_hw, 1, "star", steps, coeffs, temporal=T)
    one, _ = _run_group(dims, tile_hw, 1, "star", steps, coeffs)
    got = tb.gather_global()
    assert torch.equal(got, one.gather_global())
    exp = glob.clone()
    for _ in range(steps):
        exp = reference_step(exp, coeffs, 1)
    assert torch.allclose(got, exp, rtol=1e-12, atol=1e-13)
    assert tb.residual() == pytest.approx(one.residual(), rel=1e-12)
    s = tb.ranks[0]
    assert s.steps_done == steps and s.layout.ghost == T
    ri = (torch.arange(-T, tile_hw[0] + T)) % exp.shape[0]
    ci = (torch.arange(-T, tile_hw[1] + T)) % exp.shape[1]
    assert torch.allclose(s.tile.cpu(), exp[ri][:, ci], rtol=1e-12, atol=1e-13)


@pytest.mark.parametrize("T", [2, 3, 4])
def test_temporal_blocking_any_step_count(T):
    """Step counts that are not multiples of T end with a partial sweep (fewer levels, same
    exchange); stepping 1, then T+1, then 1 again must equal 3+T single steps, residual included."""
    dims, hw = (2, 2), (64, 256)
    total = T + 3
    one, glob = _run_group(dims, hw, 1, "star", total)
    grp = Stencil2D.create_group(cms.LocalGroup(4), hw, 1, "star", dims=dims, temporal=T)
    for s in grp.ranks:
        s.fill(0.0, 0.0)
    grp.scatter_global(glob)
    grp.exchange()
    grp.step(1)
    grp.step(T + 1)
    grp.step(1, residual=True)
    grp.sync()
    grp.check()
    assert grp.ranks[0].steps_done == total
    assert torch.equal(grp.gather_global(), one.gather_global())
    assert grp.residual() == pytest.approx(one.residual(), rel=1e-12)


@pytest.mark.parametrize("T", [2, 3])
@pytest.mark.parametrize("dims,tile_hw", [((1, 1), (96, 384)), ((2, 2), (64, 256)), ((2, 4), (40, 504)),
                                          ((3, 3), (32, 260)), ((1, 2), (50, 200)), ((1, 1), (64, 3200)),
                                          ((2, 2), (36, 1600)), ((1, 1), (600, 1536))])
def test_temporal_strip_kernel_is_bit_identical_to_single_steps(monkeypatch, dims, tile_hw, T):
    """The temporal STRIP kernel (levels in registers, strips marched top to bottom; the default
    for large tiles) forced on small multi-rank cases: one strip, several strips with a partial
    last one, partial last chunk, every direction a self- or aliased neighbour.  Same bits as
    the single-step kernel and as the temporal TILE kernel, a partial sweep at the end."""
    if tile_hw[0] % 4 and tile_hw[0] % 4 < T:
        pytest.skip("the south edge would straddle two 4-row chunks: the plan keeps the tile kernel")
    steps = 4 * T + (T - 1)
    coeffs = [0.1, 0.2, 0.25, 0.15, 0.3]
    monkeypatch.setenv("B2C_TB_ALGO", "strip")
    tb, glob = _run_group(dims, tile_hw, 1, "star", steps, coeffs, temporal=T)
    assert tb.ranks[0].info().threads == 384          # the strip kernel really ran
    monkeypatch.setenv("B2C_TB_ALGO", "tile")
    tile, _ = _run_group(dims, tile_hw, 1, "star", steps, coeffs, temporal=T)
    assert tile.ranks[0].info().threads == 288
    one, _ = _run_group(dims, tile_hw, 1, "star", steps, coeffs)
    got = tb.gather_global()
    assert torch.equal(got, one.gather_global())
    assert torch.equal(got, tile.gather_global())
    exp = glob.clone()
    for _ in range(steps):
        exp = reference_step(exp, coeffs, 1)
    assert torch.allclose(got, exp, rtol=1e-12, atol=1e-13)
    assert tb.residual() == pytest.approx(one.residual(), rel=1e-12)
    s = tb.ranks[0]
    ri = (torch.arange(-T, tile_hw[0] + T)) % exp.shape[0]
    ci = (torch.arange(-T, tile_hw[1] + T)) % exp.shape[1]
    assert torch.allclose(s.tile.cpu(), exp[ri][:, ci], rtol=1e-12, atol=1e-13)


@pytest.mark.parametrize("T,algo,hw", [(1, "strip", (96, 600)), (3, "strip", (96, 1900)), (3, "tile", (96, 384)),
                                       (2, "strip", (64, 3200))])
def test_pipelined_step_io_equals_set_coeffs_step_residual(monkeypatch, T, algo, hw):
    """``step_io`` (coefficients in and residual out on the plan's copy streams, one call kept in
    flight) against the plain sequence set_coeffs / step(residual) / residual on a twin plan:
    same field bits and the same residual for every call, with coefficients that change every
    call and step counts that are not multiples of the levels per sweep."""
    monkeypatch.setenv("B2C_TB_ALGO", algo)
    glob = _rand_global(hw[0], hw[1], 3)

    def make():
        grp = Stencil2D.create_group(cms.LocalGroup(1), hw, 1, "star", dims=(1, 1), temporal=T)
        s = grp.ranks[0]
        s.fill(0.0, 0.0)
        grp.scatter_global(glob)
        grp.exchange()
        return grp, s

    ga, a = make()
    gb, b = make()
    counts = [T, T, 2 * T + 1, 1, T, 3 * T, T + 1 if T > 1 else 2]
    coefs = [[0.1 + 0.01 * i, 0.2, 0.25 - 0.01 * i, 0.15, 0.3] for i in range(len(counts))]
    want = []
    for n, c in zip(counts, coefs):
        b.set_coeffs(c)
        b.step(n, residual=True)
        buf = b.local_residual_async()
        torch.cuda.synchronize()
        want.append(float(buf.item()))
    got, pend = [], None
    for i, (n, c) in enumerate(zip(counts, coefs)):
        cc = c if i % 2 else torch.tensor(c, dtype=torch.float64).pin_memory()
        t = a.step_io(n, cc, residual=True, drain=(i == len(counts) - 1))
        if pend is not None:
            got.append(a.wait_io(pend))
        pend = t
    got.append(a.wait_io(pend))
    a.check()
    b.check()
    assert got == want
    assert a.steps_done == b.steps_done == sum(counts)
    assert torch.equal(ga.gather_global(), gb.gather_global())
    with pytest.raises(cms.CommError):
        a.wait_io(0)                       # more than two calls back
    # plain steps afterwards keep the coefficients of the last pipelined call
    a.step(T)
    b.step(T)
    torch.cuda.synchronize()
    assert torch.equal(ga.gather_global(), gb.gather_global())


@pytest.mark.parametrize("T", [2, 3])
def test_temporal_strip_kernel_graph_replay_and_odd_step_counts(monkeypatch, T):
    monkeypatch.setenv("B2C_TB_ALGO", "strip")
    hw = (96, 1900)
    plain, glob = _run_group((1, 1), hw, 1, "star", 7 * T + 2)
    grp = Stencil2D.create_group(cms.LocalGroup(1), hw, 1, "star", dims=(1, 1), temporal=T)
    s = grp.ranks[0]
    s.fill(0.0, 0.0)
    grp.scatter_global(glob)
    grp.exchange()
    s.step(T)
    torch.cuda.synchronize()
    graph = s.capture(2 * T)
    for _ in range(3):
        graph.replay()
    s.step(1)
    s.step(1)
    torch.cuda.synchronize()
    s.check()
    assert s.steps_done == 7 * T + 2
    assert torch.equal(grp.gather_global(), plain.gather_global())


@pytest.mark.parametrize("T", [2, 3])
@pytest.mark.parametrize("dims,tile_hw,periods", [((2, 2), (32, 128), False), ((2, 4), (40, 504), False),
                                                  ((1, 1), (64, 3200), False), ((3, 1), (48, 1600), False),
                                                  ((2, 2), (36, 260), (True, False)),
                                                  ((1, 3), (32, 384), (False, True))])
def test_temporal_blocking_on_non_periodic_grids_keeps_the_boundary_fixed(dims, tile_hw, periods, T):
    """The reference's own topology is NON-periodic (mpi10.cpp:24-27).  Ghost cells beyond the
    edge of the global grid are boundary data; with T time steps per sweep the intermediate
    levels must stay pinned there.  Same bits as single steps, and the oracle of a padded grid."""
    n = dims[0] * dims[1]
    steps = 3 * T + 1                                        # ends with a partial sweep
    coeffs = [0.1, 0.2, 0.25, 0.15, 0.3]
    H, W = dims[0] * tile_hw[0], dims[1] * tile_hw[1]
    glob = _rand_global(H, W, seed=5)
    runs = {}
    for temporal in (T, 1):
        grp = Stencil2D.create_group(cms.LocalGroup(n), tile_hw, 1, "star", dims=dims, periods=periods,
                                     temporal=temporal)
        for s in grp.ranks:
            s.fill(0.0, 7.5)                                 # boundary value 7.5 in both buffers
            s.set_coeffs(coeffs)
        grp.scatter_global(glob)
        grp.exchange()
        grp.step(steps, residual=True)
        grp.sync()
        grp.check()
        runs[temporal] = grp
    got = runs[T].gather_global()
    assert torch.equal(got, runs[1].gather_global())
    assert runs[T].residual() == pytest.approx(runs[1].residual(), rel=1e-12)
    # oracle: pad with the boundary value where the grid is not periodic, wrap where it is
    per = (periods, periods) if isinstance(periods, bool) else periods
    g = glob.clone()
    for _ in range(steps):
        p = torch.nn.functional.pad(g[None, None], (1, 1, 1, 1), value=7.5)[0, 0]
        if per[0]:
            p[0, 1:-1], p[-1, 1:-1] = g[-1], g[0]
        if per[1]:
            p[1:-1, 0], p[1:-1, -1] = g[:, -1], g[:, 0]
        g = (coeffs[0] * p[1:-1, 1:-1] + coeffs[1] * p[1:-1, :-2] + coeffs[2] * p[1:-1, 2:]
             + coeffs[3] * p[:-2, 1:-1] + coeffs[4] * p[2:, 1:-1])
    assert torch.allclose(got, g, rtol=1e-12, atol=1e-13)


def test_temporal_blocking_rejects_what_it_cannot_do():
    with pytest.raises(ValueError):
        Stencil2D.create_group(cms.LocalGroup(1), (64, 256), 2, "star", dims=(1, 1), temporal=2)
    with pytest.raises(ValueError):
        Stencil2D.create_group(cms.LocalGroup(1), (64, 256), 1, "star", dims=(1, 1), periods=False, temporal=4)
    with pytest.raises(ValueError):
        Stencil2D.create_group(cms.LocalGroup(1), (64, 256), 1, "box", dims=(1, 1), temporal=2)


@pytest.mark.parametrize("T", [2, 3])
def test_temporal_blocking_graph_replay(T):
    hw = (96, 512)
    plain, glob = _run_group((1, 1), hw, 1, "star", 7 * T)
    grp = Stencil2D.create_group(cms.LocalGroup(1), hw, 1, "star", dims=(1, 1), temporal=T)
    s = grp.ranks[0]
    s.fill(0.0, 0.0)
    grp.scatter_global(glob)
    grp.exchange()
    s.step(T)
    torch.cuda.synchronize()
    graph = s.capture(2 * T)
    for _ in range(3):
        graph.replay()
    torch.cuda.synchronize()
    s.check()
    assert s.steps_done == 7 * T
    assert torch.equal(grp.gather_global(), plain.gather_global())
    with pytest.raises(ValueError):
        s.capture(T)          # one sweep: the buffers would not end where they started


def test_missing_neighbour_times_out_instead_of_hanging():
    """Failure detection: if a neighbour never delivers its halo, the device-side wait gives
    up after its deadline, sets the status word, and check() raises ERR_TIMEOUT."""
    grp = Stencil2D.create_group(cms.LocalGroup(2), (32, 128), 1, "star", dims=(1, 2))
    for s in grp.ranks:
        s.fill(1.0, 1.0)
        s.set_timeout_ms(30)
    grp.exchange()
    a = grp.ranks[0]
    a.step(1, drain=False)        # fine: step 0 needs no epochs
    a.step(1, drain=False)        # needs rank 1's step-0 halo, which is never produced
    with pytest.raises(cms.CommError) as ei:
        a.check()
    assert ei.value.code == 11    # ERR_TIMEOUT


def test_checkpoint_resume_continues_bit_identically(tmp_path):
    dims, hw = (2, 2), (64, 256)
    # non-default coefficients: they are part of the saved state
    grp, glob = _run_group(dims, hw, 1, "star", 3, coeffs=[0.1, 0.2, 0.25, 0.15, 0.3])
    for s in grp.ranks:
        s.save_checkpoint(str(tmp_path / f"ck{s.rank}.pt"))
    grp.step(4)
    grp.sync()
    want = grp.gather_global()
    # fresh plans, restored from the files, must reproduce the same continuation
    grp2 = Stencil2D.create_group(cms.LocalGroup(4), hw, 1, "star", dims=dims)
    for s in grp2.ranks:
        s.fill(0.0, 0.0)
        assert s.load_checkpoint(str(tmp_path / f"ck{s.rank}.pt")) == 3
    grp2.exchange()
    grp2.step(4)
    grp2.sync()
    grp2.check()
    assert torch.equal(grp2.gather_global(), want)


@pytest.mark.parametrize("algo", ["tile", "strip"])
def test_cuda_graph_replay_equals_plain_stepping(monkeypatch, algo):
    """A captured batch of steps can be replayed because the kernels derive their epochs
    from a device-side step counter; 2 plain + 3 x replay(4) must equal 14 plain steps."""
    monkeypatch.setenv("B2C_STENCIL_ALGO", algo)
    hw = (96, 512)
    plain, glob = _run_group((1, 1), hw, 1, "star", 14)
    want = plain.gather_global()
    grp = Stencil2D.create_group(cms.LocalGroup(1), hw, 1, "star", dims=(1, 1))
    s = grp.ranks[0]
    s.fill(0.0, 0.0)
    grp.scatter_global(glob)
    grp.exchange()
    s.step(2)                                    # also sets the one-time kernel attributes
    torch.cuda.synchronize()
    before = s.kernel_launches
    graph = s.capture(4)
    assert s.info().steps_done == 2 and s.kernel_launches == before
    for _ in range(3):
        graph.replay()
    torch.cuda.synchronize()
    s.check()
    assert s.info().steps_done == 14 and s.kernel_launches == before + 3 * graph.launches
    assert torch.equal(grp.gather_global(), want)
    with pytest.raises(ValueError):
        s.capture(3)


def test_exchange_reproduces_reference_golden_layout():
    """3x3 periodic grid, 16x16 tile, 5x5 stencil (ghost 2), core = rank id, ghosts = -1:
    after ONE device-side exchange every ghost block holds the id of the neighbour on
    that side -- the content of stencil2d/sample-output/<r>_<c> (compared with the
    files themselves when the reference tree is mounted)."""
    from test_halo_host import expected_tile
    grp = Stencil2D.create_group(cms.LocalGroup(9), (16, 16), 2, "star", dims=(3, 3), periods=True)
    for s in grp.ranks:
        s.fill(float(s.rank), -1.0)
    grp.exchange()
    for s in grp.ranks:
        r, c = s.cart.coords(s.rank)
        assert torch.equal(s.tile.cpu(), expected_tile(s.cart, s.rank, 16, 2))
        if os.path.isdir(REF):
            golden = open(f"{REF}/{r}_{c}").read().split("Array after exchange\n")[1]
            assert s.dump_str() == golden, f"tile {r}_{c} differs from the reference dump"
