"""Randomised traffic over the host transport (ranks = threads): many messages of random size
(empty .. several ring chunks), random tags, random posting order, derived datatypes on some of
them, blocking and non-blocking calls mixed, sub-communicator traffic interleaved with world
traffic.  Every payload is a function of (source, destination, index), so any mismatch,
reordering across different tags, truncation or lost message is detected; deadlocks become
time-outs.  The reference has no such test (its programs are its tests)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from cuda_mpi_scratch_b200.parallel import comm as cm
from cuda_mpi_scratch_b200.parallel.datatype import INT


def _payload(src, dst, idx, n):
    return (np.arange(n, dtype=np.int64) * 7 + src * 1_000_003 + dst * 10_007 + idx * 101).astype(np.int32)


def _schedule(nranks, per_rank, seed):
    """Global list of messages, identical on every rank: (src, dst, idx, nelems, tag)."""
    rng = np.random.default_rng(seed)
    msgs = []
    for src in range(nranks):
        for idx in range(per_rank):
            dst = int(rng.integers(0, nranks))            # includes sends to self
            kind = rng.integers(0, 10)
            n = 0 if kind == 0 else int(rng.integers(1, 64)) if kind < 6 else int(rng.integers(1000, 700_000))
            msgs.append((src, dst, idx, n, src * per_rank + idx))   # unique tag per message
    return msgs


@pytest.mark.parametrize("nranks,per_rank,seed", [(3, 30, 1), (6, 24, 2), (8, 12, 3)])
def test_random_traffic_every_message_arrives_intact(nranks, per_rank, seed):
    msgs = _schedule(nranks, per_rank, seed)

    def body(c):
        rng = np.random.default_rng(seed * 100 + c.rank)
        mine_in = [m for m in msgs if m[1] == c.rank]
        mine_out = [m for m in msgs if m[0] == c.rank]
        order = rng.permutation(len(mine_in))
        bufs, reqs = {}, []
        for k in order:                                    # receives posted in random order
            src, dst, idx, n, tag = mine_in[k]
            bufs[tag] = np.full(n, -1, dtype=np.int32)
            wild = rng.integers(0, 4) == 0                 # some by ANY_SOURCE (the tag is unique)
            reqs.append((tag, c.irecv(bufs[tag], cm.ANY_SOURCE if wild else src, tag)))
        sends = []
        for (src, dst, idx, n, tag) in mine_out:
            data = _payload(src, dst, idx, n)
            if rng.integers(0, 3) == 0 and n < 200_000 and dst != c.rank:
                c.send(data, dst, tag)                     # blocking; the receive is already posted or buffered
            else:
                sends.append((data, c.isend(data, dst, tag)))
        for tag, r in reqs:
            st = r.wait()
            src, dst, idx, n, _ = msgs[tag]
            assert st.source == src and st.tag == tag and st.count(INT) == n
            assert np.array_equal(bufs[tag], _payload(src, dst, idx, n)), (src, dst, idx, n)
        for _, r in sends:
            r.wait()
        c.barrier()
        return len(mine_in)

    got = cm.run_local(nranks, body, timeout_s=120.0)
    assert sum(got) == len(msgs)


def test_same_tag_messages_keep_their_order_per_sender():
    """Non-overtaking: messages from one sender with the SAME tag are matched in send order,
    however the receiver interleaves its receives with other traffic."""
    n, rounds = 4, 50

    def body(c):
        if c.rank == 0:
            seen = {s: 0 for s in range(1, n)}
            buf = np.zeros(2, dtype=np.int32)
            for _ in range(rounds * (n - 1)):
                st = c.recv(buf, cm.ANY_SOURCE, 5)
                assert buf[0] == st.source and buf[1] == seen[st.source]
                seen[st.source] += 1
            return seen
        for i in range(rounds):
            c.send(np.array([c.rank, i], dtype=np.int32), 0, 5)
        return None

    out = cm.run_local(n, body, timeout_s=60.0)
    assert out[0] == {s: rounds for s in range(1, n)}


def test_sub_communicator_traffic_does_not_mix_with_world_traffic():
    """Same (source, tag) on two communicators: the context keeps them apart."""
    n = 6

    def body(c):
        half = c.split(c.rank % 2, c.rank)
        peer_w = (c.rank + 2) % n                          # same parity: also a member of `half`
        peer_h = (half.rank + 1) % half.size
        a = np.full(1000, c.rank, dtype=np.int32)
        b = np.full(1000, 100 + c.rank, dtype=np.int32)
        ra, rb = np.zeros(1000, dtype=np.int32), np.zeros(1000, dtype=np.int32)
        r1 = half.irecv(rb, cm.ANY_SOURCE, 9)              # posted FIRST, must not catch the world message
        r2 = c.irecv(ra, cm.ANY_SOURCE, 9)
        s1 = c.isend(a, peer_w, 9)
        s2 = half.isend(b, peer_h, 9)
        for r in (r1, r2, s1, s2):
            r.wait()
        src_w = (c.rank - 2) % n
        assert (ra == src_w).all() and (rb == 100 + src_w).all()
        c.barrier()
        return True

    assert all(cm.run_local(n, body, timeout_s=60.0))


@pytest.mark.parametrize("nranks,seed", [(4, 5), (7, 6)])
def test_random_traffic_between_process_ranks(nranks, seed):
    """The same schedule with ranks as PROCESSES over the POSIX shared-memory fabric."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    p = subprocess.run([sys.executable, "-m", "cuda_mpi_scratch_b200.launch", "-n", str(nranks), "--timeout", "150",
                        os.path.join(root, "tests/mp/comm_stress_proc.py"), str(seed)],
                       cwd=root, capture_output=True, text=True, timeout=240)
    assert p.returncode == 0 and "COMM_STRESS PASSED" in p.stdout, p.stdout + p.stderr


def test_launcher_stops_the_job_when_one_rank_dies(tmp_path):
    """Failure detection at job level: rank 1 dies while the others wait for its message; the
    launcher terminates them and reports the failing exit code instead of hanging."""
    import time
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    prog = tmp_path / "dies.py"
    prog.write_text(
        "import os, sys\n"
        f"sys.path.insert(0, {root!r})\n"
        "import numpy as np\n"
        "from cuda_mpi_scratch_b200.parallel import comm as cm\n"
        "w = cm.init_process()\n"
        "w.set_timeout(60.0)\n"
        "if w.rank == 1:\n"
        "    os._exit(7)\n"
        "buf = np.zeros(4)\n"
        "w.recv(buf, 1, 0)\n")
    t0 = time.time()
    p = subprocess.run([sys.executable, "-m", "cuda_mpi_scratch_b200.launch", "-n", "3", "--timeout", "90", str(prog)],
                       cwd=root, capture_output=True, text=True, timeout=120)
    assert p.returncode == 7
    assert time.time() - t0 < 45, "the surviving ranks must be stopped, not left to time out"


@pytest.mark.parametrize("py,px", [(1, 2), (2, 1), (2, 2), (2, 4)])
def test_unfused_baseline_halo_exchange_orderings(py, px):
    """The baseline the fused stencil is compared against (ops/unfused_halo.py) on the CPU with
    gloo: ghost strips hold the right neighbour's right edge on every grid shape, including the
    ones where north == south, west == east, or a direction wraps onto the rank itself."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    port = 29710 + py * 10 + px
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={py * px}",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(root, "tests/mp/unfused_halo_check.py"), str(py), str(px)],
                       cwd=root, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0 and "UNFUSED_HALO PASSED" in p.stdout, p.stdout[-2000:] + p.stderr[-2000:]
