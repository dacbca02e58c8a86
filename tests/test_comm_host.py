"""MPI-like semantics over the built-in host transport, ranks = threads.

Each test re-creates the observable behaviour of one of the reference's
tutorial programs (mpi3.cpp ... mpi10.cpp, mpi-complex-types.cpp) and checks the
values those programs print.  No MPI and no GPU is needed.
"""
import ctypes as C

import numpy as np
import pytest

from cuda_mpi_scratch_b200 import _lib
from cuda_mpi_scratch_b200.parallel import comm as cm
from cuda_mpi_scratch_b200.parallel.datatype import CHAR, DOUBLE, FLOAT, INT, Datatype


def test_mpi1_rank_size_name():
    out = cm.run_local(4, lambda c: (c.rank, c.size, cm.processor_name()))
    assert [o[0] for o in out] == [0, 1, 2, 3] and all(o[1] == 4 for o in out)
    assert all(len(o[2]) > 0 for o in out)
    assert cm.wtime() > 0


def test_mpi3_blocking_send_probe_get_count():
    # mpi3.cpp:26-46: receiver sizes its buffer from MPI_Probe + MPI_Get_count
    def body(c):
        msg = b"Hello from rank %d" % c.rank
        peer = 1 - c.rank
        if c.rank == 0:
            c.send(msg, peer, tag=7)
            st = c.probe(peer, 7)
            buf = bytearray(st.count(CHAR))
            c.recv(buf, peer, 7)
        else:
            st = c.probe(peer, 7)
            assert st.source == 0 and st.tag == 7
            buf = bytearray(st.count(CHAR))
            c.recv(buf, peer, 7)
            c.send(msg, peer, tag=7)
        return bytes(buf)
    out = cm.run_local(2, body)
    assert out == [b"Hello from rank 1", b"Hello from rank 0"]


def test_mpi4_pingpong_counter_reaches_10():
    # mpi4.cpp:24-44
    def body(c):
        v = np.zeros(1, dtype=np.int32)
        peer = 1 - c.rank
        turn = 0
        while v[0] < 10:
            if turn % 2 == c.rank:
                v[0] += 1
                c.send(v, peer)
            else:
                c.recv(v, peer)
            turn += 1
        return int(v[0])
    assert cm.run_local(2, body) == [10, 10]


def test_mpi5_mpi6_chain_neighbours_and_gather():
    # mpi5.cpp:31-75 (open chain, Isend/Irecv/Waitall), mpi6.cpp:87-101 (Gather of triples)
    def body(c):
        me = np.array([c.rank], dtype=np.int32)
        prev, nxt = np.array([-1], dtype=np.int32), np.array([-1], dtype=np.int32)
        left = c.rank - 1 if c.rank > 0 else cm.PROC_NULL
        right = c.rank + 1 if c.rank < c.size - 1 else cm.PROC_NULL
        reqs = [c.irecv(prev, left, tag=1), c.irecv(nxt, right, tag=2),
                c.isend(me, left, tag=2), c.isend(me, right, tag=1)]
        cm.waitall(reqs)
        triple = np.array([c.rank, prev[0], nxt[0]], dtype=np.int32)
        allt = np.zeros(3 * c.size, dtype=np.int32) if c.rank == 0 else None
        c.gather(triple, allt, root=0)
        return allt.reshape(-1, 3).tolist() if c.rank == 0 else None
    out = cm.run_local(5, body)
    assert out[0] == [[0, -1, 1], [1, 0, 2], [2, 1, 3], [3, 2, 4], [4, 3, -1]]


def test_mpi7_indexed_to_contiguous_with_self_send():
    # mpi7.cpp:32-54: root Isends 1 x indexed{4@5, 2@12} to EVERY rank incl. itself,
    # everybody Recvs 6 plain floats -> 5,6,7,8,12,13
    def body(c):
        idx = FLOAT.indexed([4, 2], [5, 12]).commit()
        reqs = []
        a = np.arange(16, dtype=np.float32)
        if c.rank == 0:
            reqs = [c.isend(a, i, tag=1, count=1, datatype=idx) for i in range(c.size)]
        b = np.zeros(6, dtype=np.float32)
        st = c.recv(b, 0, 1)
        cm.waitall(reqs)      # the reference drops these requests
        assert st.count(FLOAT) == 6
        return b.tolist()
    for row in cm.run_local(4, body):
        assert row == [5, 6, 7, 8, 12, 13]


def test_mpi8_struct_particles():
    # mpi8.cpp:13-17,38-74: root sends particle i to rank i
    class Particle(C.Structure):
        _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float), ("velocity", C.c_float),
                    ("id", C.c_int), ("type", C.c_int)]

    def body(c):
        pt = Datatype.struct([4, 2], [0, Particle.id.offset], [FLOAT, INT]).commit()
        assert pt.extent == C.sizeof(Particle)
        reqs, ps = [], (Particle * c.size)()
        if c.rank == 0:
            for i in range(c.size):
                ps[i] = Particle(1.0 * i, -1.0 * i, 2.0 * i, 0.25, i, i % 2)
                reqs.append(c.isend(C.addressof(ps) + i * C.sizeof(Particle), i, tag=1, count=1,
                                    datatype=pt))
        p = Particle()
        c.recv(p, 0, 1, count=1, datatype=pt)
        cm.waitall(reqs)
        return (p.x, p.y, p.z, p.velocity, p.id, p.type)
    out = cm.run_local(4, body)
    assert out == [(1.0 * i, -1.0 * i, 2.0 * i, 0.25, i, i % 2) for i in range(4)]


def test_mpi9_groups_and_allreduce_per_half():
    # mpi9.cpp:26-54: split world in halves, Allreduce(SUM) of ranks per half and world
    def body(c):
        half = c.size // 2
        lo, hi = list(range(half)), list(range(half, c.size))
        sub_lo = c.create(lo)     # collective on the parent, None on non-members
        sub_hi = c.create(hi)
        sub = sub_lo if c.rank < half else sub_hi
        assert (sub_lo is None) == (c.rank >= half)
        me = np.array([c.rank], dtype=np.int32)
        s_half, s_world = np.zeros(1, dtype=np.int32), np.zeros(1, dtype=np.int32)
        sub.allreduce(me, s_half, cm.SUM)
        c.allreduce(me, s_world, cm.SUM)
        return sub.rank, sub.size, int(s_half[0]), int(s_world[0])
    out = cm.run_local(8, body)
    assert out[:4] == [(i, 4, 6, 28) for i in range(4)]
    assert out[4:] == [(i, 4, 22, 28) for i in range(4)]


def test_mpi10_cartesian_nonperiodic_neighbour_exchange():
    # mpi10.cpp:22-54: sqrt(N) x sqrt(N) non-periodic grid, 4-neighbour Isend/Irecv,
    # edges are PROC_NULL and leave the receive buffer untouched
    def body(c):
        cart = c.cart_create([3, 3], [False, False])
        coords = cart.cart_coords()
        up, down = cart.cart_shift(0, 1)
        left, right = cart.cart_shift(1, 1)
        nbrs = [up, down, left, right]
        me = np.array([cart.rank], dtype=np.int32)
        got = [np.array([-1], dtype=np.int32) for _ in nbrs]
        reqs = []
        for n, g in zip(nbrs, got):
            reqs.append(cart.irecv(g, n, tag=1))
            reqs.append(cart.isend(me, n, tag=1))
        cm.waitall(reqs)
        return coords, nbrs, [int(g[0]) for g in got]
    out = cm.run_local(9, body)
    assert out[4] == ([1, 1], [1, 7, 3, 5], [1, 7, 3, 5])
    assert out[0] == ([0, 0], [cm.PROC_NULL, 3, cm.PROC_NULL, 1], [-1, 3, -1, 1])
    assert out[8] == ([2, 2], [5, cm.PROC_NULL, 7, cm.PROC_NULL], [5, -1, 7, -1])


def test_cart_periodic_wrap_and_rank_lookup():
    def body(c):
        cart = c.cart_create([2, 2], True)
        return (cart.cart_rank([-1, 0]), cart.cart_rank([0, 2]), cart.cart_shift(1, 1),
                cart.cart_get())
    out = cm.run_local(4, body)
    assert out[0] == (2, 0, (1, 1), ([2, 2], [True, True]))


def test_complex_types_three_buffers_one_message():
    # mpi-complex-types.cpp: 3 sub-blocks from 3 separately allocated arrays in ONE send
    def body(c):
        def bufs(fill):
            return [(C.c_int * 8)(*fill(i)) for i in range(3)]
        if c.rank == 0:
            b = bufs(lambda k: [[i, 2 * i, 2 * i + 1][k] for i in range(8)])
            start = 3
        else:
            b = bufs(lambda k: [-1] * 8)
            start = 0
        sub = INT.subarray([8], [3], [start]).commit()
        t = sub.hindexed([1, 1, 1], [0, C.addressof(b[1]) - C.addressof(b[0]),
                                     C.addressof(b[2]) - C.addressof(b[0])]).commit()
        if c.rank == 0:
            c.send(C.addressof(b[0]), 1, tag=123, count=1, datatype=t)
        else:
            st = c.recv(C.addressof(b[0]), 0, 123, count=1, datatype=t)
            assert st.bytes == 36
        return [list(x) for x in b]
    out = cm.run_local(2, body)
    assert out[1] == [[3, 4, 5, -1, -1, -1, -1, -1], [6, 8, 10, -1, -1, -1, -1, -1],
                      [7, 9, 11, -1, -1, -1, -1, -1]]


def test_large_message_is_chunked_through_the_ring():
    n = 3_000_000   # 24 MB of doubles through a 4 MB ring
    def body(c):
        if c.rank == 0:
            a = np.arange(n, dtype=np.float64)
            c.send(a, 1, tag=3)
            return None
        b = np.zeros(n, dtype=np.float64)
        st = c.recv(b, 0, 3)
        return st.count(DOUBLE), float(b[0]), float(b[-1]), float(b.sum())
    out = cm.run_local(2, body)
    assert out[1] == (n, 0.0, n - 1.0, n * (n - 1) / 2)


def test_bidirectional_large_isend_irecv_does_not_deadlock():
    n = 2_000_000
    def body(c):
        peer = 1 - c.rank
        a = np.full(n, c.rank + 1, dtype=np.float64)
        b = np.zeros(n, dtype=np.float64)
        reqs = [c.irecv(b, peer, 5), c.isend(a, peer, 5)]
        cm.waitall(reqs)
        return float(b[0]), float(b[-1])
    assert cm.run_local(2, body) == [(2.0, 2.0), (1.0, 1.0)]


def test_any_source_any_tag_and_ordering():
    def body(c):
        if c.rank == 0:
            seen = []
            for _ in range(2 * (c.size - 1)):
                v = np.zeros(1, dtype=np.int32)
                st = c.recv(v, cm.ANY_SOURCE, cm.ANY_TAG)
                seen.append((st.source, st.tag, int(v[0])))
            return seen
        for k in range(2):   # two messages per sender: must arrive in order
            c.send(np.array([10 * c.rank + k], dtype=np.int32), 0, tag=k)
        return None
    seen = cm.run_local(4, body)[0]
    for r in (1, 2, 3):
        mine = [s for s in seen if s[0] == r]
        assert mine == [(r, 0, 10 * r), (r, 1, 10 * r + 1)]


def test_tag_matching_out_of_order_receives():
    def body(c):
        if c.rank == 0:
            c.send(np.array([1], dtype=np.int32), 1, tag=1)
            c.send(np.array([2], dtype=np.int32), 1, tag=2)
            return None
        a, b = np.zeros(1, dtype=np.int32), np.zeros(1, dtype=np.int32)
        c.recv(b, 0, tag=2)     # skips over the tag-1 message
        c.recv(a, 0, tag=1)
        return int(a[0]), int(b[0])
    assert cm.run_local(2, body)[1] == (1, 2)


def test_truncation_is_an_error():
    def body(c):
        if c.rank == 0:
            c.send(np.arange(8, dtype=np.int32), 1)
            return None
        small = np.zeros(4, dtype=np.int32)
        with pytest.raises(_lib.CommError) as ei:
            c.recv(small, 0)
        return ei.value.code
    assert cm.run_local(2, body)[1] == _lib.ERR_TRUNCATE


def test_collectives_bcast_reduce_scatter_allgather_split():
    def body(c):
        x = np.array([3.5, -1.0]) if c.rank == 2 else np.zeros(2)
        c.bcast(x, root=2)
        r = np.zeros(2) if c.rank == 1 else None
        c.reduce(np.array([c.rank, 1.0]), r, cm.SUM, root=1)
        mx = np.zeros(1, dtype=np.int64)
        c.allreduce(np.array([c.rank * 7], dtype=np.int64), mx, cm.MAX)
        part = np.zeros(2, dtype=np.int32)
        src = np.arange(2 * c.size, dtype=np.int32) if c.rank == 0 else None
        c.scatter(src, part, root=0)
        ag = np.zeros(c.size, dtype=np.int32)
        c.allgather(np.array([c.rank * c.rank], dtype=np.int32), ag)
        sub = c.split(c.rank % 2, key=-c.rank)      # reversed order inside each colour
        c.barrier()
        return (x.tolist(), None if r is None else r.tolist(), int(mx[0]), part.tolist(),
                ag.tolist(), sub.rank, sub.size, sub.group())
    out = cm.run_local(6, body)
    for rk, o in enumerate(out):
        assert o[0] == [3.5, -1.0]
        assert o[1] == ([15.0, 6.0] if rk == 1 else None)
        assert o[2] == 35 and o[3] == [2 * rk, 2 * rk + 1] and o[4] == [i * i for i in range(6)]
    assert out[0][5:] == (2, 3, [4, 2, 0]) and out[5][5:] == (0, 3, [5, 3, 1])


def test_deadlock_turns_into_timeout_error():
    def body(c):
        c.set_timeout(0.5)
        v = np.zeros(1, dtype=np.int32)
        with pytest.raises(_lib.CommError) as ei:
            c.recv(v, 1 - c.rank, tag=99)    # nobody sends
        return ei.value.code
    assert cm.run_local(2, body, timeout_s=20) == [_lib.ERR_TIMEOUT, _lib.ERR_TIMEOUT]


def test_invalid_arguments():
    def body(c):
        v = np.zeros(1, dtype=np.int32)
        codes = []
        for fn in (lambda: c.send(v, 5), lambda: c.send(v, 0, tag=-3),
                   lambda: c.cart_create([4, 4], True), lambda: c.cart_coords()):
            with pytest.raises(_lib.CommError) as ei:
                fn()
            codes.append(ei.value.code)
        return codes
    assert cm.run_local(1, body)[0] == [7, 5, 9, 9]   # RANK, TAG, TOPOLOGY, TOPOLOGY


def test_rank_utilities_node_count_elapsed_block_range(capsys):
    from cuda_mpi_scratch_b200.utils import ranks as ru

    def body(c):
        t0 = cm.wtime()
        n = ru.node_count(c)
        el = ru.global_elapsed(c, t0 - 0.001 * c.rank, cm.wtime())
        ru.rank_print(c, "nodes", n)
        return n, el > 0.0, ru.block_range(1 << 10, c.rank, c.size)
    out = cm.run_local(4, body)
    assert all(o[0] == 1 and o[1] for o in out)
    assert [o[2] for o in out] == [(256 * r, 256 * (r + 1)) for r in range(4)]
    lines = [l for l in capsys.readouterr().out.splitlines() if "nodes 1" in l]
    assert len(lines) == 4 and all(" - rank " in l for l in lines)
    with pytest.raises(ValueError):
        ru.block_range(10, 0, 3)
