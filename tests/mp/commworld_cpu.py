"""CommWorld (the library's own communicator as bootstrap) without a GPU:
python -m cuda_mpi_scratch_b200.launch -n 3 tests/mp/commworld_cpu.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import cuda_mpi_scratch_b200 as cms

w = cms.init()
assert isinstance(w, cms.CommWorld), type(w)
parts = w.allgather_bytes(bytes([65 + w.rank]) * (3 * w.rank + 1))
ok = [len(p) for p in parts] == [3 * r + 1 for r in range(w.size)] and all(p == bytes([65 + r]) * (3 * r + 1) for r, p in enumerate(parts))
ok = ok and w.allgather_bytes(b"") == [b""] * w.size
ok = ok and w.allreduce_max(float(w.rank)) == float(w.size - 1) and w.allreduce_sum(float(w.rank)) == float(sum(range(w.size)))
ok = ok and w.allreduce_max_vec([w.rank, -w.rank, 2.5]) == [float(w.size - 1), 0.0, 2.5]
w.barrier()
if w.rank == 0:
    print("COMMWORLD_CPU " + ("PASSED" if ok else "FAILED"), flush=True)
ok_all = w.allreduce_sum(0.0 if ok else 1.0) == 0.0
w.finalize()
sys.exit(0 if ok_all else 1)
