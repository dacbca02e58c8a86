"""Fused dot + P2P allreduce vs torch fp64 (needs a GPU)."""
import pytest
import torch

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.models.dot import DistributedDot, baseline_dot

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_ones_give_the_element_count_exactly(dtype):
    # mpicuda2.cu:169-172,296-298: all-ones inputs, result must equal ARRAY_SIZE
    n = (1 << 24) + 3
    d = DistributedDot.create(cms.SingleWorld(0))
    x = torch.ones(n, dtype=dtype, device="cuda")
    assert d.dot(x, x) == float(n)
    assert d.dot(x, x) == float(n)          # second call: next epoch / other slot parity
    d.close()


def test_random_matches_fp64_oracle_and_is_deterministic():
    n = 5_000_011
    d = DistributedDot.create(cms.SingleWorld(0))
    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.rand(n, dtype=torch.float64, device="cuda", generator=g) - 0.5
    y = torch.rand(n, dtype=torch.float64, device="cuda", generator=g) - 0.5
    want = float(torch.dot(x.cpu(), y.cpu()))
    a, b = d.dot(x, y), d.dot(x, y)
    assert a == b
    assert a == pytest.approx(want, rel=1e-12, abs=1e-9)
    d.close()


@pytest.mark.parametrize("nranks", [2, 4, 8])
def test_virtual_ranks_allreduce_and_reduce_root(nranks):
    n_local = 1_000_003
    grp = DistributedDot.create_group(cms.LocalGroup(nranks))
    g = torch.Generator(device="cuda").manual_seed(nranks)
    xs = [torch.rand(n_local, dtype=torch.float64, device="cuda", generator=g) for _ in range(nranks)]
    ys = [torch.rand(n_local, dtype=torch.float64, device="cuda", generator=g) for _ in range(nranks)]
    want = sum(float(torch.dot(x.cpu(), y.cpu())) for x, y in zip(xs, ys))
    got = grp.dot(xs, ys)
    assert all(v == got[0] for v in got), "every rank must hold bit-identical results"
    assert got[0] == pytest.approx(want, rel=1e-12)
    got2 = grp.dot(xs, ys, root=1)              # MPI_Reduce to a root (mpicuda2.cu:293)
    assert got2[1] == got[0]
    ones = [torch.ones(n_local, dtype=torch.float32, device="cuda") for _ in range(nranks)]
    assert grp.dot(ones, ones)[0] == float(nranks * n_local)


def test_reference_style_baselines_agree():
    n = 1 << 22
    x = torch.ones(n, dtype=torch.float64, device="cuda")
    for variant in (0, 1, 2):
        assert baseline_dot(x, x, variant) == float(n)
        assert baseline_dot(x, x, variant) == float(n)


def test_bad_arguments():
    d = DistributedDot.create(cms.SingleWorld(0))
    x = torch.ones(8, dtype=torch.float64, device="cuda")
    with pytest.raises(TypeError):
        d.dot(x, x.float())
    with pytest.raises(ValueError):
        d.dot(x, x[:4])
    with pytest.raises(cms.CommError):
        d.launch(x, x, root=3)
    d.close()
