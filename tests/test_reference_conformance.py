"""Conformance: the reference's OWN tutorial sources, compiled unchanged from the mounted
reference tree against ``include/b200comm/shim/mpi.h`` and run over this library's host
transport (ranks = processes started by ``cuda_mpi_scratch_b200.launch``).

Nothing of the reference is copied into this repository: the test reads
``/root/reference/*.cpp`` at run time and is skipped where that tree is not mounted (the
GPU boxes).  What is checked is what the reference's programs print.
"""
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
LIBDIR = os.path.join(ROOT, "cuda_mpi_scratch_b200", "lib")

pytestmark = pytest.mark.skipif(not os.path.isfile(f"{REF}/mpi1.cpp"), reason="reference tree not mounted")


def _build(src, tmp_path, extra=()):
    import cuda_mpi_scratch_b200  # noqa: F401
    exe = str(tmp_path / (os.path.basename(src).replace(".", "_")))
    cmd = ["g++", "-std=c++17", "-O1", "-w", f"-I{ROOT}/include/b200comm/shim", f"-I{ROOT}/include",
           f"-I{REF}", *extra, src, "-o", exe, f"-L{LIBDIR}", "-lb200comm", f"-Wl,-rpath,{LIBDIR}", "-pthread"]
    subprocess.run(cmd, check=True, timeout=300)
    return exe


def _run(exe, n, cwd=None, timeout=120):
    cmd = [sys.executable, "-m", "cuda_mpi_scratch_b200.launch", "-n", str(n), "--timeout", str(timeout), exe]
    env = dict(os.environ, PYTHONPATH=ROOT)
    return subprocess.run(cmd, check=True, capture_output=True, text=True, timeout=timeout + 60,
                          cwd=cwd or ROOT, env=env).stdout


def test_mpi1_mpi2_rank_and_size(tmp_path):
    for src in ("mpi1.cpp", "mpi2.cpp"):
        out = _run(_build(f"{REF}/{src}", tmp_path), 4)
        ranks = sorted(int(x) for x in re.findall(r"rank\s+(\d+)", out))
        assert ranks == [0, 1, 2, 3] or len(out.strip().splitlines()) == 4, out


def test_mpi3_strings_and_mpi4_counter(tmp_path):
    out = _run(_build(f"{REF}/mpi3.cpp", tmp_path), 2)
    assert len(out.strip().splitlines()) >= 2, out
    out = _run(_build(f"{REF}/mpi4.cpp", tmp_path), 2)
    assert "10" in out, out


def test_mpi5_mpi6_neighbours_and_gather(tmp_path):
    out = _run(_build(f"{REF}/mpi5.cpp", tmp_path), 4)
    assert len(out.strip().splitlines()) >= 4, out
    out = _run(_build(f"{REF}/mpi6.cpp", tmp_path), 4)
    assert out.strip(), out


def test_mpi7_indexed_type(tmp_path):
    out = _run(_build(f"{REF}/mpi7.cpp", tmp_path), 4)
    lines = [l for l in out.splitlines() if "rank" in l]
    assert len(lines) == 4 and all(l.rstrip().endswith("5,6,7,8,12,13,") for l in lines), out


def test_mpi8_struct_type(tmp_path):
    out = _run(_build(f"{REF}/mpi8.cpp", tmp_path), 4)
    assert len([l for l in out.splitlines() if l.strip()]) >= 4, out


def test_mpi9_groups(tmp_path):
    out = _run(_build(f"{REF}/mpi9.cpp", tmp_path), 8)
    # ranks 0-3 sum to 6, ranks 4-7 to 22, the world to 28 (printed once by the root)
    assert out.count("received: 6") == 4 and out.count("received: 22") == 4, out
    assert "Allreduce total: 28" in out, out


def test_mpi10_cartesian(tmp_path):
    out = _run(_build(f"{REF}/mpi10.cpp", tmp_path), 9)
    assert len([l for l in out.splitlines() if l.strip()]) >= 9, out


def test_mpi_complex_types(tmp_path):
    out = _run(_build(f"{REF}/mpi-complex-types.cpp", tmp_path), 2)
    got = {m.group(1): [int(v) for v in re.findall(rf"{m.group(1)}\[\d\] = (-?\d+)", out)]
           for m in re.finditer(r"(B[123])\[0\]", out)}
    assert got["B1"][:4] == [3, 4, 5, -1] and got["B2"][:4] == [6, 8, 10, -1] and got["B3"][:4] == [7, 9, 11, -1]


def test_reference_cpu_stencil_driver_writes_its_golden_files(tmp_path):
    """stencil2d/mpi-2d-stencil-subarray.cpp (the reference's driver AND its stencil2D.h)
    over the shim: 9 ranks must reproduce stencil2d/sample-output/* apart from the
    'CUDA device id' line that only the CUDA driver prints."""
    exe = _build(f"{REF}/stencil2d/mpi-2d-stencil-subarray.cpp", tmp_path, extra=(f"-I{REF}/stencil2d",))
    _run(exe, 9, cwd=str(tmp_path))
    for r in range(3):
        for c in range(3):
            got = (tmp_path / f"{r}_{c}").read_text()
            ref = re.sub(r"\nCUDA device id: \d+\n", "", open(f"{REF}/stencil2d/sample-output/{r}_{c}").read())
            assert got == ref, f"{r}_{c} differs"
