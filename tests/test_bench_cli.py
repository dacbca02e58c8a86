"""bench.py contract pieces that can be checked without a GPU."""
import json
import os
import subprocess
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_reports_unavailable_and_exits_zero():
    p = subprocess.run([sys.executable, "bench.py", "--impl", "reference", "--gpus", "1", "--steps", "2",
                        "--warmup", "1"], cwd=ROOT, capture_output=True, text=True, timeout=120)
    assert p.returncode == 0
    d = json.loads(p.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and "unavailable" in d and len(d["unavailable"]) > 10


def test_default_arguments_match_the_contract():
    sys.path.insert(0, ROOT)
    import bench
    a = bench.parse_args([])
    assert a.gpus == 1 and a.warmup >= 3 and a.steps >= 1 and a.impl == "ours"
    assert a.tile == 16384 and a.radius == 1 and a.scaling == "weak"


def test_without_a_gpu_the_bench_fails_loudly():
    if torch.cuda.is_available():
        return
    p = subprocess.run([sys.executable, "bench.py", "--steps", "2"], cwd=ROOT, capture_output=True, text=True,
                       timeout=300)
    assert p.returncode != 0
    assert "no CUDA device" in p.stdout


def test_stdout_carries_only_the_json_line():
    """bench.py points file descriptor 1 at stderr for the run (NCCL prints its version line to
    stdout on some boxes) and writes the result to the original stdout."""
    code = ("import os, sys; sys.path.insert(0, %r); import bench\n"
            "out = bench._claim_stdout()\n"
            "print('python noise')\n"
            "os.write(1, b'library noise\\n')\n"
            "print('{\"ok\": 1}', file=out, flush=True)\n" % ROOT)
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0
    assert p.stdout == '{"ok": 1}\n'
    assert "python noise" in p.stderr and "library noise" in p.stderr
