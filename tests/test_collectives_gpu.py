"""Device collectives over peer-mapped staging (virtual ranks on one GPU)."""
import pytest
import torch

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.ops.collectives import DeviceCollectives

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("algo", ["oneshot", "twoshot"])
@pytest.mark.parametrize("P,n,dtype", [(2, 1000, torch.float64), (4, 65537, torch.float32),
                                        (8, 300001, torch.float64), (3, 17, torch.int32)])
def test_allreduce_sum_matches_torch_and_is_identical_everywhere(algo, P, n, dtype):
    grp = DeviceCollectives.create_group(cms.LocalGroup(P), max_bytes=4 << 20)
    g = torch.Generator(device="cuda").manual_seed(P * 1000 + n)
    if dtype.is_floating_point:
        xs = [torch.rand(n, dtype=dtype, device="cuda", generator=g) - 0.5 for _ in range(P)]
    else:
        xs = [torch.randint(-1000, 1000, (n,), dtype=dtype, device="cuda", generator=g) for _ in range(P)]
    want = torch.stack([x.double() for x in xs]).sum(0)
    for rep in range(2):                       # twice: both staging parities
        ts = [x.clone() for x in xs]
        grp.allreduce(ts, "sum", algo)
        grp.check()
        for t in ts[1:]:
            assert torch.equal(t, ts[0])
        tol = 1e-12 if dtype == torch.float64 else (1e-5 if dtype == torch.float32 else 0)
        assert torch.allclose(ts[0].double(), want, rtol=tol, atol=tol)


@pytest.mark.parametrize("algo", ["oneshot", "twoshot"])
def test_allreduce_unaligned_tensors_and_odd_lengths(algo):
    """The kernels work on 16-byte vectors of the staging buffers; user tensors that start at
    an odd element (not 16-byte aligned) and odd lengths take the scalar copy-out path."""
    P = 3
    grp = DeviceCollectives.create_group(cms.LocalGroup(P), max_bytes=1 << 20)
    for n in (1, 3, 1023, 4099):
        bases = [torch.arange(n + 1, dtype=torch.float32, device="cuda") * (r + 1) for r in range(P)]
        ts = [b[1:] for b in bases]                  # data_ptr() % 16 == 4
        want = sum((r + 1) for r in range(P)) * torch.arange(1, n + 1, dtype=torch.float32, device="cuda")
        grp.allreduce(ts, "sum", algo)
        grp.check()
        for r in range(P):
            assert torch.equal(ts[r], want)
            assert float(bases[r][0]) == 0.0         # the element before the view is untouched


def test_allreduce_max_min_and_reduce_to_root():
    P, n = 4, 5000
    grp = DeviceCollectives.create_group(cms.LocalGroup(P), max_bytes=1 << 20)
    xs = [torch.arange(n, dtype=torch.int64, device="cuda") * (r + 1) % 977 for r in range(P)]
    ts = [x.clone() for x in xs]
    grp.allreduce(ts, "max")
    assert torch.equal(ts[2], torch.stack(xs).max(0).values)
    ts = [x.clone() for x in xs]
    grp.allreduce(ts, "min")
    assert torch.equal(ts[0], torch.stack(xs).min(0).values)
    ts = [x.clone() for x in xs]
    grp.reduce(ts, root=3)
    assert torch.equal(ts[3], torch.stack(xs).sum(0))
    assert torch.equal(ts[0], xs[0])            # non-roots keep their input
    grp.check()


def test_bcast_gather_allgather_scatter_barrier():
    P, n = 4, 3001
    grp = DeviceCollectives.create_group(cms.LocalGroup(P), max_bytes=1 << 20)
    ts = [torch.full((n,), float(r), dtype=torch.float64, device="cuda") for r in range(P)]
    grp.bcast(ts, root=2)
    assert all(torch.equal(t, torch.full((n,), 2.0, dtype=torch.float64, device="cuda")) for t in ts)
    sends = [torch.arange(n, dtype=torch.float32, device="cuda") + 100 * r for r in range(P)]
    recv = torch.zeros(P * n, dtype=torch.float32, device="cuda")
    grp.gather(sends, recv, root=1)
    assert torch.equal(recv, torch.cat(sends))
    recvs = [torch.zeros(P * n, dtype=torch.float32, device="cuda") for _ in range(P)]
    grp.allgather(sends, recvs)
    assert all(torch.equal(o, torch.cat(sends)) for o in recvs)
    src = torch.arange(P * n, dtype=torch.int32, device="cuda")
    outs = [torch.zeros(n, dtype=torch.int32, device="cuda") for _ in range(P)]
    grp.scatter(src, outs, root=0)
    for r in range(P):
        assert torch.equal(outs[r], src[r * n:(r + 1) * n])
    grp.barrier()
    grp.barrier()
    grp.check()


def test_message_too_large_is_rejected():
    grp = DeviceCollectives.create_group(cms.LocalGroup(2), max_bytes=1024)
    t = torch.zeros(1000, dtype=torch.float64, device="cuda")
    with pytest.raises(cms.CommError):
        grp.ranks[0].allreduce(t)
    with pytest.raises(TypeError):
        grp.ranks[0].allreduce(t.half())
