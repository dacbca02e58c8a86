"""mpi4.cpp in spirit: two ranks bounce an int, each increments it, until it is 10."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from cuda_mpi_scratch_b200.parallel import comm as cm


def body(c):
    v, peer, turn = np.zeros(1, dtype=np.int32), 1 - c.rank, 0
    while v[0] < 10:
        if turn % 2 == c.rank:
            v[0] += 1
            c.send(v, peer)
            print(f"rank {c.rank} sent {v[0]}", flush=True)
        else:
            c.recv(v, peer)
        turn += 1
    return int(v[0])


if __name__ == "__main__":
    print(cm.run_local(2, body))
