"""``python -m cuda_mpi_scratch_b200.launch -n N prog.py [args]`` -- start N ranks
of one node, one process each, attached to a shared-memory fabric.

The reference is started by ``mpiexec.hydra -rmk {pbs,slurm} ./a.out`` from a
batch script (``mpi_pbs_sample.sh:18``, ``mpi_cuda_pbs_ref.sh:18``,
``stencil2d/sample-output/job_9_1_1_cuda-2d-stencil-subarray.slurm:15``).  Here
the launcher sets ``B200_RANK`` / ``B200_WORLD_SIZE`` / ``B200_JOB`` /
``LOCAL_RANK`` for every child; ``torchrun`` works too (``init_process`` reads
its variables).  If one rank fails, the others are terminated.
"""
from __future__ import annotations

import argparse
import os
import subprocess
import sys
import time


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="cuda_mpi_scratch_b200.launch")
    ap.add_argument("-n", "--nranks", type=int, required=True)
    ap.add_argument("--timeout", type=float, default=0.0, help="kill the job after this many seconds")
    ap.add_argument("prog")
    ap.add_argument("args", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    job = f"b200job-{os.getpid()}-{int(time.time())}"
    procs = []
    for r in range(a.nranks):
        # LOCAL_RANK for this library; the MVAPICH2 / Open MPI names because unmodified MPI + CUDA
        # programs pick their GPU from them BEFORE MPI_Init (the reference's BindDevice,
        # stencil2d/mpi-2d-stencil-subarray-cuda.cu:40-73, reads MV2_COMM_WORLD_LOCAL_RANK)
        env = dict(os.environ, B200_RANK=str(r), B200_WORLD_SIZE=str(a.nranks), B200_JOB=job,
                   LOCAL_RANK=str(r), MV2_COMM_WORLD_LOCAL_RANK=str(r),
                   OMPI_COMM_WORLD_LOCAL_RANK=str(r), MPISPAWN_LOCAL_NPROCS=str(a.nranks))
        cmd = [sys.executable, a.prog] + a.args if a.prog.endswith(".py") else [a.prog] + a.args
        procs.append(subprocess.Popen(cmd, env=env))
    t0 = time.time()
    rc = 0
    alive = list(procs)
    while alive:
        for p in list(alive):
            r = p.poll()
            if r is None:
                continue
            alive.remove(p)
            if r != 0 and rc == 0:
                rc = r
                for q in alive:
                    q.terminate()
        if a.timeout and time.time() - t0 > a.timeout:
            for q in alive:
                q.kill()
            return 124
        time.sleep(0.01)
    return rc


if __name__ == "__main__":
    sys.exit(main())
