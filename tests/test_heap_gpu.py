"""Symmetric heap (VMM arenas, file descriptors over Unix sockets, deferred collective frees) between
two PROCESSES that share the one visible GPU: tests/mp/heap_check.py under torchrun.  The multicast
part needs two distinct GPUs and is skipped by the script itself when the binding fails (it runs on
2 and 8 GPUs from scripts/gpu_r2_heap.sh / gpu_r2_scale.sh)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def test_heap_between_two_processes_on_one_gpu():
    env = dict(os.environ, PYTHONPATH=ROOT, B200COMM_MULTICAST="0")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29571", os.path.join(ROOT, "tests", "mp", "heap_check.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT, env=env)
    assert p.returncode == 0 and "HEAP_CHECK PASSED" in p.stdout, p.stdout[-3000:] + p.stderr[-3000:]


def test_commworld_bootstrap_of_the_fused_paths_two_processes_one_gpu():
    """The library's own launcher + communicator (CommWorld) instead of torchrun: symmetric
    memory and a stencil plan between two processes (tests/mp/commworld_check.py; with two GPUs
    the same script also runs fused steps, scripts/gpu_r2_final.sh)."""
    env = {k: v for k, v in os.environ.items() if k not in ("WORLD_SIZE", "RANK", "LOCAL_RANK")}
    env.update(PYTHONPATH=ROOT, B200COMM_MULTICAST="0")
    cmd = [sys.executable, "-m", "cuda_mpi_scratch_b200.launch", "-n", "2", "--timeout", "200",
           os.path.join(ROOT, "tests", "mp", "commworld_check.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT, env=env)
    assert p.returncode == 0 and "COMMWORLD_CHECK PASSED" in p.stdout, p.stdout[-3000:] + p.stderr[-3000:]
