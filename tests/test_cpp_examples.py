"""The C++ API is usable from plain C++ programs: build the tutorials in examples/cpp
against include/ + the shared library and check what they print against the values
the reference's tutorials print."""
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "cuda_mpi_scratch_b200", "lib")
REF = "/root/reference/stencil2d/sample-output"


def _build(name, tmp_path):
    import cuda_mpi_scratch_b200  # noqa: F401  (makes sure the library exists)
    exe = str(tmp_path / name)
    if os.path.exists(f"{ROOT}/examples/cpp/{name}.cu"):    # examples with their own kernels
        cmd = ["/usr/local/cuda/bin/nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O1",
               f"-I{ROOT}/include", f"{ROOT}/examples/cpp/{name}.cu", "-o", exe,
               f"-L{LIBDIR}", "-lb200comm", "-Xlinker", f"-rpath,{LIBDIR}"]
    else:
        cmd = ["g++", "-std=c++17", "-O1", f"-I{ROOT}/include", f"{ROOT}/examples/cpp/{name}.cpp", "-o", exe,
               f"-L{LIBDIR}", "-lb200comm", f"-Wl,-rpath,{LIBDIR}", "-pthread"]
    subprocess.run(cmd, check=True, timeout=300)
    return exe


def _launch(n, exe, *args, timeout=240):
    return subprocess.run([sys.executable, "-m", "cuda_mpi_scratch_b200.launch", "-n", str(n), "--timeout",
                           str(timeout), exe, *map(str, args)], cwd=ROOT, timeout=timeout + 60,
                          capture_output=True, text=True)


@pytest.mark.parametrize("name", ["halo_exchange_gpu", "pingpong_gpu", "dot_gpu", "stencil_fused_gpu",
                                  "collectives_gpu"])
def test_gpu_drivers_compile_without_a_gpu(name, tmp_path):
    """The GPU drivers (the reference's CUDA main() programs) build against include/ + the library;
    without arguments the ping-pong prints its usage and exits 1 instead of reading a null argv."""
    exe = _build(name, tmp_path)
    if name == "pingpong_gpu":
        p = subprocess.run([exe], capture_output=True, text=True, timeout=60)
        assert p.returncode == 1 and "usage" in p.stderr


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.isdir(REF), reason="reference sample output not mounted")
def test_gpu_halo_driver_reproduces_golden_files(tmp_path):
    """9 process ranks sharing the GPU, tiles in device memory, subarray types applied by the
    pack/move kernel through CUDA-IPC mappings: the dumps equal stencil2d/sample-output."""
    exe = _build("halo_exchange_gpu", tmp_path)
    outdir = tmp_path / "out"
    outdir.mkdir()
    p = _launch(9, exe, 16, 5, outdir)
    assert p.returncode == 0, p.stdout + p.stderr
    dev = re.compile(r"CUDA device id: \d+")
    for r in range(3):
        for c in range(3):
            got = dev.sub("CUDA device id: X", (outdir / f"{r}_{c}").read_text())
            ref = dev.sub("CUDA device id: X", open(f"{REF}/{r}_{c}").read())
            assert got == ref, f"{r}_{c} differs"


@pytest.mark.gpu
@pytest.mark.parametrize("opts", [(), ("--async",), ("--host-copy",), ("--async", "--host-copy", "--pinned")])
def test_gpu_pingpong_driver(tmp_path, opts):
    exe = _build("pingpong_gpu", tmp_path)
    for n in (1, 1 << 20):
        p = _launch(2, exe, n, *opts, "--iters", 3)
        assert p.returncode == 0 and "PASSED" in p.stdout, p.stdout + p.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("nranks,opts", [(2, ("--variant", 0)), (4, ("--variant", 1)), (4, ("--variant", 2, "--rrobin")),
                                         (2, ("--float",))])
def test_gpu_dot_driver(tmp_path, nranks, opts):
    exe = _build("dot_gpu", tmp_path)
    p = _launch(nranks, exe, 22, *opts, "--iters", 2)
    assert p.returncode == 0 and "PASSED" in p.stdout, p.stdout + p.stderr
    assert f"(fused kernel + P2P allreduce): {float(1 << 22):g}" in p.stdout or "4.1943e+06" in p.stdout


def test_tutorial_p2p(tmp_path):
    exe = _build("tutorial_p2p", tmp_path)
    out = subprocess.run([exe, "4"], check=True, capture_output=True, text=True, timeout=120).stdout
    lines = out.strip().split("\n")
    assert len(lines) == 4
    assert "[Hello from rank 1] counter=10 prev=-1 next=1" in lines[0]
    assert "gathered: (0,-1,1) (1,0,2) (2,1,3) (3,2,-1)" in lines[0]
    assert "[Hello from rank 0] counter=10 prev=0 next=2" in lines[1]
    assert all("caught error class 7: invalid rank" in l for l in lines)


def test_tutorial_types_topology(tmp_path):
    exe = _build("tutorial_types_topology", tmp_path)
    out = subprocess.run([exe, "4"], check=True, capture_output=True, text=True, timeout=120).stdout
    lines = out.strip().split("\n")
    for r, l in enumerate(lines):
        assert "5,6,7,8,12,13," in l                                        # mpi7
        assert f"particle({r},{-r},{2 * r},0.25,{r},{r % 2})" in l.replace("-0", "0")   # mpi8
        assert f"half-sum={1 if r < 2 else 5} world-sum=6" in l             # mpi9
    assert "cart(0,0) nbrs -1 2 -1 1" in lines[0]                           # mpi10 (2x2, non periodic)
    assert "cart(1,1) nbrs 1 -1 2 -1" in lines[3]
    assert "complex: [3 4 5 -1] [6 8 10 -1] [7 9 11 -1]" in lines[1]        # mpi-complex-types


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference sample output not mounted")
def test_cpp_halo_driver_reproduces_golden_files(tmp_path):
    exe = _build("halo_exchange_cpu", tmp_path)
    outdir = tmp_path / "out"
    outdir.mkdir()
    subprocess.run([sys.executable, "-m", "cuda_mpi_scratch_b200.launch", "-n", "9", "--timeout", "120",
                    exe, str(outdir)], check=True, cwd=ROOT, timeout=180)
    for r in range(3):
        for c in range(3):
            got = (outdir / f"{r}_{c}").read_text()
            ref = open(f"{REF}/{r}_{c}").read()
            ref = re.sub(r"\nCUDA device id: \d+\n", "", ref)     # the CPU driver has no such line
            assert got == ref, f"{r}_{c} differs"


@pytest.mark.gpu
@pytest.mark.parametrize("temporal,steps", [(1, 9), (3, 10), (2, 7), (4, 12)])
def test_gpu_fused_stencil_driver_from_cpp(tmp_path, temporal, steps):
    """The flagship through the C++ classes of b200comm/device.hpp (no Python on the path): one
    rank, periodic 1x1 grid (the rank is its own neighbour in all 8 directions), every cell
    compared bit for bit with a host computation, partial sweeps included."""
    import torch
    exe = _build("stencil_fused_gpu", tmp_path)
    p = _launch(1, exe, 256, steps, temporal)
    assert p.returncode == 0 and "PASSED" in p.stdout, p.stdout + p.stderr
    assert "resumed run bit-identical" in p.stdout, p.stdout     # FusedStencil::save/load_checkpoint
    if torch.cuda.device_count() >= 2:      # one rank per GPU only
        p = _launch(2, exe, 256, steps, temporal)
        assert p.returncode == 0 and "PASSED" in p.stdout, p.stdout + p.stderr
        assert "resumed run bit-identical" in p.stdout, p.stdout


def test_mpi_shim_close_relatives(tmp_path):
    """tests/mp/shim_extra.cpp: an MPI-style program that uses Sendrecv, Test, Waitany, Iprobe,
    Allgather, logical reductions, Comm_split/dup, Dims_create, Cart_get, hvector + resized types
    and Get_address through include/b200comm/shim/mpi.h, 4 process ranks."""
    import cuda_mpi_scratch_b200  # noqa: F401
    exe = str(tmp_path / "shim_extra")
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", f"-I{ROOT}/include/b200comm/shim", f"-I{ROOT}/include",
                    f"{ROOT}/tests/mp/shim_extra.cpp", "-o", exe, f"-L{LIBDIR}", "-lb200comm",
                    f"-Wl,-rpath,{LIBDIR}", "-pthread"], check=True, timeout=300)
    p = _launch(4, exe, timeout=120)
    assert p.returncode == 0 and "SHIM_EXTRA PASSED" in p.stdout, p.stdout + p.stderr


@pytest.mark.gpu
def test_cpp_device_collectives_driver(tmp_path):
    """examples/cpp/collectives_gpu.cpp (b200::device::DeviceCollectives): with one visible GPU the
    two ranks share it and the driver says so and skips (kernels of two ranks on one GPU must not
    wait for each other); with two or more GPUs it must print PASSED."""
    exe = _build("collectives_gpu", tmp_path)
    p = _launch(2, exe)
    assert p.returncode == 0 and ("PASSED" in p.stdout or "skipped" in p.stdout), p.stdout + p.stderr
