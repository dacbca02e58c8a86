"""mpi5/mpi6 in spirit with ranks as PROCESSES (start with cuda_mpi_scratch_b200.launch):
every rank sends its id round a periodic ring, then rank 0 gathers (self, prev, next)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from cuda_mpi_scratch_b200.parallel import comm as cm

world = cm.init_process()
me = np.array([world.rank], dtype=np.int32)
prev, nxt = np.zeros(1, dtype=np.int32), np.zeros(1, dtype=np.int32)
left, right = (world.rank - 1) % world.size, (world.rank + 1) % world.size
cm.waitall([world.irecv(prev, left, 1), world.irecv(nxt, right, 2),
            world.isend(me, right, 1), world.isend(me, left, 2)])
allt = np.zeros(3 * world.size, dtype=np.int32) if world.rank == 0 else None
world.gather(np.array([world.rank, prev[0], nxt[0]], dtype=np.int32), allt, root=0)
if world.rank == 0:
    print(cm.processor_name(), allt.reshape(-1, 3).tolist(), flush=True)
world.finalize()
