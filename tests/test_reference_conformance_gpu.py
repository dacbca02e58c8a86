"""GPU conformance: the reference's CUDA + MPI programs, compiled UNCHANGED against
``include/b200comm/shim/mpi.h`` and run over this library (ranks = processes started by
``cuda_mpi_scratch_b200.launch``, device pointers handed straight to MPI_Send / MPI_Recv /
MPI_Isend / MPI_Irecv as the reference does with a CUDA-aware MPI).

  test-benchmark/mpi-pingpong-gpu.cpp            blocking Send/Recv of a device buffer, echo equality
  test-benchmark/mpi-pingpong-gpu-async.cpp      Isend/Irecv/Wait; also -DHOST_COPY -DPAGE_LOCKED
  mpicuda4.cu -DGPU -DREDUCE_GPU                 block-decomposed dot product + MPI_Reduce
  stencil2d/mpi-2d-stencil-subarray-cuda.cu      halo exchange of subarray types on DEVICE tiles,
                                                 nine golden files

Nothing of the reference is copied into this repository: the sources are read from
``/root/reference`` or, on a GPU box where that tree is not mounted, from the git-ignored
``baseline/_ref/src`` that travels with the snapshot.  Numbers these binaries print are "reference
source over shim transport", never "CUDA-aware MPI".
"""
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "cuda_mpi_scratch_b200", "lib")
REF = next((d for d in ("/root/reference", os.path.join(ROOT, "baseline", "_ref", "src"))
            if os.path.isfile(os.path.join(d, "mpicuda4.cu"))), None)
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(REF is None, reason="reference tree neither mounted nor staged")]


def _build(src, tmp_path, name, defs=(), inc=()):
    import cuda_mpi_scratch_b200  # noqa: F401  (builds the library if needed)
    exe = str(tmp_path / name)
    common = [f"-I{ROOT}/include/b200comm/shim", f"-I{ROOT}/include", f"-I{REF}", *[f"-I{i}" for i in inc],
              *[f"-D{d}" for d in defs], f"{REF}/{src}", "-o", exe, f"-L{LIBDIR}", "-lb200comm"]
    if src.endswith(".cu"):
        cmd = [f"{CUDA}/bin/nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O2", "-w",
               *common, "-Xlinker", f"-rpath,{LIBDIR}"]
    else:
        cmd = ["g++", "-std=c++17", "-O1", "-w", f"-I{CUDA}/include", *common, f"-Wl,-rpath,{LIBDIR}",
               f"-L{CUDA}/lib64", "-lcudart", "-pthread"]
    subprocess.run(cmd, check=True, timeout=600)
    return exe


def _run(exe, n, args=(), cwd=None, timeout=180):
    cmd = [sys.executable, "-m", "cuda_mpi_scratch_b200.launch", "-n", str(n), "--timeout", str(timeout),
           exe, *[str(a) for a in args]]
    env = dict(os.environ, PYTHONPATH=ROOT)
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout + 60, cwd=cwd or ROOT, env=env)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    return p.stdout


@pytest.mark.parametrize("n", [1, 1000, 1 << 20])
def test_reference_pingpong_blocking(tmp_path, n):
    out = _run(_build("test-benchmark/mpi-pingpong-gpu.cpp", tmp_path, "pp"), 2, [n])
    assert "PASSED" in out and "FAILED" not in out, out


@pytest.mark.parametrize("defs", [(), ("HOST_COPY",), ("HOST_COPY", "PAGE_LOCKED")])
def test_reference_pingpong_async(tmp_path, defs):
    exe = _build("test-benchmark/mpi-pingpong-gpu-async.cpp", tmp_path, "ppa", defs=defs)
    for n in (1, 4096, 1 << 19):
        out = _run(exe, 2, [n])
        assert "PASSED" in out and "FAILED" not in out, out


def test_reference_dot_product_mpicuda4(tmp_path):
    """All-ones vectors: the dot product must equal the element count (mpicuda2.cu:169-172)."""
    exe = _build("mpicuda4.cu", tmp_path, "dot4", defs=("GPU", "REDUCE_GPU", "NO_LOG"))
    out = _run(exe, 2, timeout=300)
    m = re.search(r"dot product result:\s*([0-9.e+]+)", out)
    # printed with six significant digits ("2.68435e+08")
    assert m and abs(float(m.group(1)) - 2.0 ** 28) <= 1e-5 * 2.0 ** 28, out


def test_reference_cuda_stencil_driver_writes_its_golden_files(tmp_path):
    """9 ranks (processes sharing the visible GPUs), tiles in device memory, 2-D subarray types on
    device pointers: the per-rank dumps must equal stencil2d/sample-output/* byte for byte apart
    from the 'CUDA device id' line."""
    exe = _build("stencil2d/mpi-2d-stencil-subarray-cuda.cu", tmp_path, "st", inc=(f"{REF}/stencil2d",))
    _run(exe, 9, cwd=str(tmp_path), timeout=300)
    strip = lambda s: re.sub(r"\nCUDA device id: \d+\n", "\n", s)
    for r in range(3):
        for c in range(3):
            got = strip((tmp_path / f"{r}_{c}").read_text())
            ref = strip(open(f"{REF}/stencil2d/sample-output/{r}_{c}").read())
            assert got == ref, f"{r}_{c} differs"
