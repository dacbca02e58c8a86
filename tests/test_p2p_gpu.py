"""Kernel-initiated puts, flags, typed moves and device buffers through the comm
layer, on ONE GPU (virtual ranks / thread ranks; peers are plain pointers).
The 2-GPU versions run from tests/mp/ under torchrun."""
import numpy as np
import pytest
import torch

import cuda_mpi_scratch_b200 as cms
from cuda_mpi_scratch_b200.ops.p2p import P2PChannel
from cuda_mpi_scratch_b200.ops.pack_move import move_uniform, typed_move
from cuda_mpi_scratch_b200.parallel import comm as cm
from cuda_mpi_scratch_b200.parallel.datatype import DOUBLE, FLOAT, INT, Datatype

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", ["ldst", "tma"])
@pytest.mark.parametrize("nbytes", [8, 24, 4096, 1 << 20, (1 << 22) + 16])
def test_put_echo_equality(mode, nbytes):
    # echo equality, the PASSED/FAILED check of mpi-pingpong-gpu.cpp:58-71
    ch = P2PChannel.create_group(cms.LocalGroup(2), nbytes)
    src = torch.randint(0, 255, (nbytes,), dtype=torch.uint8, device="cuda")
    ch[0].put(1, src, mode=mode)
    ch[1].wait(0)
    ch[1].put(0, ch[1].recv_view(), mode=mode)     # echo what landed
    ch[0].wait(1)
    torch.cuda.synchronize()
    ch[0].check(); ch[1].check()
    assert torch.equal(ch[0].recv_view(), src)
    assert ch[0].flags().tolist() == [0, 1] and ch[1].flags().tolist() == [1, 0]


def test_blocking_send_recv_and_message_too_large():
    ch = P2PChannel.create_group(cms.LocalGroup(2), 1024)
    x = torch.arange(128, dtype=torch.float64, device="cuda")
    ch[0].send(1, x)
    out = torch.empty_like(x)
    ch[1].recv(0, out)
    assert torch.equal(out, x)
    with pytest.raises(ValueError):
        ch[0].put(1, torch.zeros(2048, dtype=torch.uint8, device="cuda"))


def test_typed_move_strided_column_to_contiguous_and_back():
    # the "derived-datatype halo" pattern: one ghost column of a row-major tile
    H, W = 512, 520
    tile = torch.arange(H * W, dtype=torch.float64, device="cuda").reshape(H, W)
    col = DOUBLE.subarray([H, W], [H - 2, 1], [1, 3]).commit()
    flat = DOUBLE.contiguous(H - 2).commit()
    out = torch.zeros(H - 2, dtype=torch.float64, device="cuda")
    assert typed_move(tile, col, out, flat) == (H - 2) * 8
    torch.cuda.synchronize()
    assert torch.equal(out, tile[1:H - 1, 3])
    dst = torch.zeros_like(tile)
    col2 = DOUBLE.subarray([H, W], [H - 2, 1], [1, W - 1]).commit()
    typed_move(out, flat, dst, col2)
    torch.cuda.synchronize()
    assert torch.equal(dst[1:H - 1, W - 1], tile[1:H - 1, 3]) and float(dst.sum()) == float(out.sum())


def test_typed_move_general_segments_struct_and_indexed():
    # irregular layouts go through the segment table (mpi7 / mpi8 style types)
    a = torch.arange(16, dtype=torch.float32, device="cuda")
    idx = FLOAT.indexed([4, 2], [5, 12]).commit()
    out = torch.zeros(6, dtype=torch.float32, device="cuda")
    assert typed_move(a, idx, out, FLOAT.contiguous(6).commit()) == 24
    torch.cuda.synchronize()
    assert out.tolist() == [5, 6, 7, 8, 12, 13]
    padded = Datatype.struct([1, 1], [0, 8], [INT, DOUBLE]).commit()   # int, pad, double
    raw = torch.zeros(4 * 16, dtype=torch.uint8, device="cuda")
    host = np.zeros(4, dtype=[("i", "<i4"), ("pad", "<i4"), ("d", "<f8")])
    host["i"], host["d"], host["pad"] = [1, 2, 3, 4], [0.5, 1.5, 2.5, 3.5], -1
    raw.copy_(torch.from_numpy(host.view(np.uint8)))
    packed = torch.zeros(4 * 12, dtype=torch.uint8, device="cuda")
    assert typed_move(raw, padded, packed, Datatype.basic("byte").contiguous(48).commit(), scount=4) == 48
    torch.cuda.synchronize()
    got = packed.cpu().numpy().view([("i", "<i4"), ("d", "<f8")])
    assert got["i"].tolist() == [1, 2, 3, 4] and got["d"].tolist() == [0.5, 1.5, 2.5, 3.5]


def test_move_uniform_2d_subarray_closed_form():
    src = torch.rand(64, 100, dtype=torch.float64, device="cuda")
    dst = torch.zeros(64, 40, dtype=torch.float64, device="cuda")
    move_uniform(src[:, 10:], dst, 64, 40 * 8, 100 * 8, 40 * 8)
    torch.cuda.synchronize()
    assert torch.equal(dst, src[:, 10:50])


def test_comm_send_recv_device_tensors_between_thread_ranks():
    """MPI-style send/recv where BOTH buffers live on the GPU (the reference's ping-pong):
    the payload is pulled by a kernel, only the descriptor uses the host fabric."""
    n = 1 << 18
    def body(c):
        torch.cuda.set_device(0)
        peer = 1 - c.rank
        if c.rank == 0:
            x = torch.arange(n, dtype=torch.float64, device="cuda")
            back = torch.zeros_like(x)
            c.send(x, peer, tag=1)
            st = c.recv(back, peer, tag=2)
            return bool(torch.equal(back, x)), st.count(DOUBLE), st.source
        buf = torch.zeros(n, dtype=torch.float64, device="cuda")
        st = c.probe(peer, 1)
        assert st.count(DOUBLE) == n
        c.recv(buf, peer, tag=1)
        c.send(buf, peer, tag=2)
        return None
    assert cm.run_local(2, body)[0] == (True, n, 1)


def test_comm_device_derived_types_and_mixed_host_device():
    def body(c):
        torch.cuda.set_device(0)
        idx = FLOAT.indexed([4, 2], [5, 12]).commit()
        if c.rank == 0:
            a = torch.arange(16, dtype=torch.float32, device="cuda")
            c.send(a, 1, tag=1, count=1, datatype=idx)      # device, derived -> device, contiguous
            c.send(a, 1, tag=2, count=1, datatype=idx)      # device, derived -> HOST buffer
            h = torch.arange(16, dtype=torch.float32)
            c.send(h, 1, tag=3, count=1, datatype=idx)      # host, derived -> DEVICE buffer
            return None
        d = torch.zeros(6, dtype=torch.float32, device="cuda")
        c.recv(d, 0, tag=1)
        h = torch.zeros(6, dtype=torch.float32)
        c.recv(h, 0, tag=2)
        d2 = torch.zeros(6, dtype=torch.float32, device="cuda")
        c.recv(d2, 0, tag=3)
        return d.tolist(), h.tolist(), d2.tolist()
    out = cm.run_local(2, body)[1]
    assert out == ([5, 6, 7, 8, 12, 13],) * 3


def test_comm_eager_and_rendezvous_paths_and_send_before_receive():
    """Small device messages travel through the receiver's device mailbox (no IPC, no FIN): a rank
    may therefore SEND FIRST and receive later on both sides, which a pure rendezvous would deadlock
    on.  Larger ones, and bursts that exceed the four slots per peer, fall back to the rendezvous;
    order per (source, tag) is kept across the two paths."""
    def body(c):
        torch.cuda.set_device(0)
        peer = 1 - c.rank
        before = cm.device_eager_stats()
        small = torch.full((512,), float(c.rank + 1), dtype=torch.float64, device="cuda")
        got = torch.zeros_like(small)
        # publish both mailboxes first (a rank's mailbox appears with its first device operation)
        r0 = c.irecv(got, peer, tag=0); s0 = c.isend(small, peer, tag=0); r0.wait(); s0.wait()
        c.barrier()
        c.send(small, peer, tag=1)                    # both ranks send first ...
        c.recv(got, peer, tag=1)                      # ... and receive afterwards
        ok = bool((got == float(peer + 1)).all())
        # a burst of 12 messages with distinct contents, received in order
        msgs = [torch.full((64,), float(100 * c.rank + i), dtype=torch.float64, device="cuda") for i in range(12)]
        big = torch.arange(1 << 16, dtype=torch.float64, device="cuda") + c.rank     # 512 KiB: rendezvous
        if c.rank == 0:
            reqs = [c.isend(m, peer, tag=7) for m in msgs] + [c.isend(big, peer, tag=7)]
            cm.waitall(reqs)
        else:
            outs = [torch.zeros(64, dtype=torch.float64, device="cuda") for _ in range(12)]
            bout = torch.zeros(1 << 16, dtype=torch.float64, device="cuda")
            for o in outs:
                c.recv(o, peer, tag=7)
            c.recv(bout, peer, tag=7)
            ok = ok and all(bool((o == float(i)).all()) for i, o in enumerate(outs))
            ok = ok and bool(torch.equal(bout, torch.arange(1 << 16, dtype=torch.float64, device="cuda")))
        c.barrier()
        after = cm.device_eager_stats()
        return ok, after["eager"] - before["eager"], after["too_large"] - before["too_large"]
    res = cm.run_local(2, body)
    assert res[0][0] and res[1][0]
    assert res[0][1] + res[1][1] >= 4          # counters are per process (both thread ranks)
    assert res[0][2] + res[1][2] >= 1          # the 512 KiB message took the rendezvous


def test_comm_device_strided_column_and_multi_buffer_hindexed_types():
    """Derived types on DEVICE memory through send/recv: (a) a ghost column of a 2-D tile
    (subarray, 8-byte runs at row pitch) gathered by the sender and scattered into another column
    by the receiver; (b) the reference's nested type of mpi-complex-types.cpp:32-58 -- a 1-D
    subarray (3 of 8 ints) wrapped in an hindexed type whose byte displacements point into THREE
    SEPARATE allocations (signed, arbitrary distances) -- gathered from three device buffers by
    one send and scattered into three device buffers by one receive."""
    H = W = 256
    def body(c):
        torch.cuda.set_device(0)
        peer = 1 - c.rank
        col_send = DOUBLE.subarray([H, W], [H, 1], [0, W - 2]).commit()     # my last core column
        col_recv = DOUBLE.subarray([H, W], [H, 1], [0, 0]).commit()         # my left ghost column
        tile = torch.arange(H * W, dtype=torch.float64, device="cuda").reshape(H, W) + 1e6 * c.rank
        r = c.irecv(tile, peer, tag=3, count=1, datatype=col_recv)
        s = c.isend(tile, peer, tag=3, count=1, datatype=col_send)
        r.wait(); s.wait()
        want = torch.arange(H * W, dtype=torch.float64, device="cuda").reshape(H, W)[:, W - 2] + 1e6 * peer
        ok = bool(torch.equal(tile[:, 0], want))
        # (b) three separate allocations
        sub_s = INT.subarray([8], [3], [3]).commit()
        sub_r = INT.subarray([8], [3], [0]).commit()
        bufs = [torch.full((8,), -1, dtype=torch.int32, device="cuda") for _ in range(3)]
        pad = torch.empty(1 << 20, dtype=torch.uint8, device="cuda")          # keeps them apart
        disps = [0, bufs[1].data_ptr() - bufs[0].data_ptr(), bufs[2].data_ptr() - bufs[0].data_ptr()]
        if c.rank == 0:
            for k, b in enumerate(bufs):
                b.copy_(torch.arange(8, dtype=torch.int32, device="cuda") * (k + 1))
            t = sub_s.hindexed([1, 1, 1], disps).commit()
            c.send(bufs[0], 1, tag=4, count=1, datatype=t)
            del pad
            return ok, None
        t = sub_r.hindexed([1, 1, 1], disps).commit()
        c.recv(bufs[0], 0, tag=4, count=1, datatype=t)
        del pad
        return ok, [b.tolist() for b in bufs]
    res = cm.run_local(2, body)
    assert res[0][0] and res[1][0]
    assert res[1][1] == [[3, 4, 5, -1, -1, -1, -1, -1], [6, 8, 10, -1, -1, -1, -1, -1],
                         [9, 12, 15, -1, -1, -1, -1, -1]]


def test_halo_plan_exchange_on_device_tiles():
    """The reference's CUDA stencil driver (mpi-2d-stencil-subarray-cuda.cu): the tile
    lives in device memory and the subarray types are applied on the GPU."""
    from cuda_mpi_scratch_b200.parallel import halo as H
    from cuda_mpi_scratch_b200.parallel.cart import CartTopology
    from test_halo_host import expected_tile

    def body(c):
        torch.cuda.set_device(0)
        cart = c.cart_create([2, 2], True)
        tot = 16 + 4
        data = torch.full((tot, tot), -1.0, dtype=torch.float64, device="cuda")
        plan = H.HaloPlan(cart, H.Array2D(tot, tot, tot), 5, 5, DOUBLE)
        plan.core.view(data)[:] = float(cart.rank)
        torch.cuda.synchronize()
        plan.exchange(data)
        return data.cpu()
    out = cm.run_local(4, body)
    cart = CartTopology.create([2, 2], True)
    for r in range(4):
        assert torch.equal(out[r], expected_tile(cart, r, 16, 2))
