"""Call tracing: spans of the communication calls as Chrome trace-event JSON."""
import json
import os
import subprocess
import sys

import numpy as np

from cuda_mpi_scratch_b200.parallel import comm as cm
from cuda_mpi_scratch_b200.parallel.datatype import DOUBLE
from cuda_mpi_scratch_b200.parallel.halo import Array2D, HaloPlan
from cuda_mpi_scratch_b200.utils import trace

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _body(c):
    peer = 1 - c.rank
    a, b = np.full(1000, c.rank, dtype=np.float64), np.zeros(1000)
    with trace.region("exchange phase", rank=c.rank, nbytes=8000):
        r = c.irecv(b, peer, 7)
        c.send(a, peer, 7)
        r.wait()
    out = np.zeros(1)
    c.allreduce(np.ones(1), out, cm.SUM)
    c.barrier()
    cart = c.cart_create([1, 2], periods=True)
    tile = np.zeros((20, 20))
    HaloPlan(cart, Array2D(20, 20, 20), 5, 5, DOUBLE).exchange(tile)
    return float(b[0])


def test_spans_are_recorded_with_rank_peer_tag_bytes(tmp_path):
    trace.clear()
    trace.enable()
    try:
        assert cm.run_local(2, _body) == [1.0, 0.0]
    finally:
        trace.disable()
    path = tmp_path / "t.json"
    n = trace.dump(str(path))
    ev = json.load(open(path))["traceEvents"]
    assert n == len(ev) == trace.count() and n > 20
    names = {e["name"] for e in ev}
    assert {"send", "isend", "irecv", "wait", "allreduce", "barrier", "halo_exchange", "exchange phase"} <= names
    assert all(e["ph"] == "X" and e["dur"] >= 0 for e in ev)
    sends = [e for e in ev if e["name"] == "send" and e["args"]["tag"] == 7]
    assert sorted(e["pid"] for e in sends) == [0, 1]
    assert all(e["args"]["bytes"] == 8000 and e["args"]["peer"] == 1 - e["pid"] for e in sends)
    user = [e for e in ev if e["cat"] == "user"]
    assert sorted(e["pid"] for e in user) == [0, 1] and all(e["args"]["bytes"] == 8000 for e in user)
    # the user span encloses the p2p calls made inside it (same thread)
    for u in user:
        inner = [e for e in sends if e["pid"] == u["pid"]][0]
        assert u["ts"] <= inner["ts"] and inner["ts"] + inner["dur"] <= u["ts"] + u["dur"] + 1e-3
    halo = [e for e in ev if e["name"] == "halo_exchange"]
    assert len(halo) == 2 and all(e["args"]["bytes"] > 0 for e in halo)
    # switched off: nothing more is recorded
    before = trace.count()
    cm.run_local(2, _body)
    assert trace.count() == before
    trace.clear()
    assert trace.count() == 0


def test_environment_variable_writes_a_file_per_process(tmp_path):
    prefix = str(tmp_path / "run")
    env = dict(os.environ, B200COMM_TRACE=prefix)
    p = subprocess.run([sys.executable, "-m", "cuda_mpi_scratch_b200.launch", "-n", "3", "--timeout", "60",
                        os.path.join(ROOT, "examples/python/ring_processes.py")], cwd=ROOT, env=env,
                       capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stdout + p.stderr
    files = [f for f in os.listdir(tmp_path) if f.startswith("run.") and f.endswith(".json")]
    assert len(files) >= 3          # one per rank process (the launcher itself may add one)
    pids = set()
    for f in files:
        for e in json.load(open(tmp_path / f))["traceEvents"]:
            if e["name"] in ("isend", "irecv", "gather"):
                pids.add(e["pid"])
    assert pids == {0, 1, 2}
