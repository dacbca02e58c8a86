"""Random traffic with ranks as PROCESSES over the POSIX shared-memory fabric
(start with ``python -m cuda_mpi_scratch_b200.launch -n N tests/mp/comm_stress_proc.py SEED``).
Same schedule and payloads as tests/test_comm_stress.py; prints one line per rank."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np

from cuda_mpi_scratch_b200.parallel import comm as cm
from cuda_mpi_scratch_b200.parallel.datatype import INT
from test_comm_stress import _payload, _schedule

seed = int(sys.argv[1]) if len(sys.argv) > 1 else 5
world = cm.init_process()
world.set_timeout(120.0)
msgs = _schedule(world.size, 20, seed)
rng = np.random.default_rng(seed * 100 + world.rank)
mine_in = [m for m in msgs if m[1] == world.rank]
mine_out = [m for m in msgs if m[0] == world.rank]
bufs, reqs = {}, []
for k in rng.permutation(len(mine_in)):
    src, dst, idx, n, tag = mine_in[k]
    bufs[tag] = np.full(n, -1, dtype=np.int32)
    reqs.append((tag, world.irecv(bufs[tag], src, tag)))
sends = [(d, world.isend(d, dst, tag)) for (src, dst, idx, n, tag) in mine_out
         for d in [_payload(src, dst, idx, n)]]
bad = 0
for tag, r in reqs:
    st = r.wait()
    src, dst, idx, n, _ = msgs[tag]
    ok = st.source == src and st.count(INT) == n and np.array_equal(bufs[tag], _payload(src, dst, idx, n))
    bad += 0 if ok else 1
for _, r in sends:
    r.wait()
tot = np.zeros(1, dtype=np.int32)
world.allreduce(np.array([bad], dtype=np.int32), tot, cm.SUM)
# one write() per line: several processes share the pipe, and print() writes text and newline separately
sys.stdout.write(f"rank {world.rank}: {len(mine_in)} received, {len(mine_out)} sent, bad={bad}\n")
if world.rank == 0:
    sys.stdout.write("COMM_STRESS PASSED\n" if tot[0] == 0 else f"COMM_STRESS FAILED ({int(tot[0])})\n")
sys.stdout.flush()
world.finalize()
sys.exit(0 if tot[0] == 0 else 1)
