"""In-tree native build of ``libb200comm.so`` (sm_100a only).

Everything native in this framework -- the runtime (device binding, symmetric
heap, host transport, MPI-like comm engine, datatype engine, halo planner) and
every CUDA kernel -- is compiled into ONE shared object that exports the C ABI
declared in ``include/b200comm.h``.  Python talks to it through ``ctypes``
(``cuda_mpi_scratch_b200/_lib.py``); C++ programs link it directly.

The reference builds each tutorial as its own executable with hand-written
``nvcc``/``mpicxx`` lines (``mpicuda2.cu:25-28``) or a 13-line legacy CMake file
(``stencil2d/CMakeLists.txt:1-12``).  Here there is one library, built for
Blackwell only: ``-gencode arch=compute_100a,code=sm_100a -lineinfo``.

Usage::

    python -m cuda_mpi_scratch_b200.build            # incremental
    python -m cuda_mpi_scratch_b200.build --force    # rebuild everything
    python -m cuda_mpi_scratch_b200.build --verbose  # show ptxas -v output
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
REPO = PKG_DIR.parent
CSRC = REPO / "csrc"
INCLUDE = REPO / "include"
OBJ_DIR = REPO / "build" / "obj"
LIB_DIR = PKG_DIR / "lib"
LIB_PATH = LIB_DIR / "libb200comm.so"

CUDA_HOME = Path(os.environ.get("CUDA_HOME", "/usr/local/cuda"))
NVCC = str(CUDA_HOME / "bin" / "nvcc")

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ARCH_FLAGS + [
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC,-fvisibility=hidden,-Wall,-Wno-unknown-pragmas",
    "-Xptxas", "-v",
    "--expt-relaxed-constexpr",
    f"-I{INCLUDE}", f"-I{CSRC}",
]
NVCC_FLAGS += os.environ.get("B200COMM_NVCC_EXTRA", "").split()   # e.g. -DB2C_TILE_H=64 (experiments)
CXX_FLAGS = [
    "-O2", "-std=c++17", "-fPIC", "-fvisibility=hidden", "-Wall", "-pthread",
    f"-I{INCLUDE}", f"-I{CSRC}", f"-I{CUDA_HOME / 'include'}",
]
LINK_FLAGS = ["-shared", "-cudart", "static", "-Xlinker", "--no-undefined",
              "-lpthread", "-lrt", "-ldl"]


def _sources() -> list[Path]:
    srcs = sorted(p for p in CSRC.rglob("*") if p.suffix in (".cu", ".cpp")
                  and "shim" not in p.parts and "programs" not in p.parts)
    return srcs


def _header_digest() -> str:
    h = hashlib.sha256()
    for p in sorted(list(INCLUDE.rglob("*.h")) + list(INCLUDE.rglob("*.hpp"))
                    + list(CSRC.rglob("*.h")) + list(CSRC.rglob("*.hpp"))
                    + list(CSRC.rglob("*.cuh"))):
        h.update(str(p).encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def source_digest() -> str:
    """Hash of every native source and public header (+ the compile flags): identifies
    what a built library was made from, independent of file times."""
    h = hashlib.sha256(" ".join(NVCC_FLAGS + CXX_FLAGS).replace(str(REPO), "").encode())
    files = [p for root in (CSRC, INCLUDE) for p in root.rglob("*")
             if p.suffix in (".cu", ".cuh", ".cpp", ".hpp", ".h")]
    for p in sorted(files):
        h.update(str(p.relative_to(REPO)).encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def _obj_for(src: Path) -> Path:
    rel = src.relative_to(CSRC)
    return OBJ_DIR / ("_".join(rel.parts) + ".o")


def _compile_one(src: Path, hdr_digest: str, force: bool, verbose: bool) -> tuple[Path, str]:
    obj = _obj_for(src)
    stamp = obj.with_suffix(".stamp")
    if src.suffix == ".cu":
        cmd = [NVCC] + NVCC_FLAGS + ["-c", str(src), "-o", str(obj)]
    else:
        cmd = [os.environ.get("CXX", "g++")] + CXX_FLAGS + ["-c", str(src), "-o", str(obj)]
    key = hashlib.sha256((" ".join(cmd) + hdr_digest).encode() + src.read_bytes()).hexdigest()
    if not force and obj.exists() and stamp.exists() and stamp.read_text() == key:
        return obj, ""
    proc = subprocess.run(cmd, capture_output=True, text=True)
    log = proc.stdout + proc.stderr
    if proc.returncode != 0:
        raise RuntimeError(f"compile failed: {' '.join(cmd)}\n{log}")
    stamp.write_text(key)
    (obj.with_suffix(".log")).write_text(log)
    return obj, log


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every source under ``csrc/`` for sm_100a and link the library."""
    if shutil.which(NVCC) is None and not Path(NVCC).exists():
        raise RuntimeError(f"nvcc not found at {NVCC}")
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    srcs = _sources()
    digest = _header_digest()
    objs: list[Path] = []
    rebuilt = False
    with cf.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        futs = {ex.submit(_compile_one, s, digest, force, verbose): s for s in srcs}
        for f in cf.as_completed(futs):
            obj, log = f.result()
            objs.append(obj)
            if log != "" or not LIB_PATH.exists():
                rebuilt = True
            if verbose and log:
                print(f"--- {futs[f].relative_to(REPO)}\n{log}")
    newest_obj = max(o.stat().st_mtime for o in objs)
    if rebuilt or not LIB_PATH.exists() or LIB_PATH.stat().st_mtime < newest_obj:
        cmd = [NVCC] + ARCH_FLAGS + LINK_FLAGS + ["-o", str(LIB_PATH)] + [str(o) for o in sorted(objs)]
        proc = subprocess.run(cmd, capture_output=True, text=True)
        if proc.returncode != 0:
            raise RuntimeError(f"link failed: {' '.join(cmd)}\n{proc.stdout}{proc.stderr}")
    LIB_PATH.with_name(LIB_PATH.name + ".digest").write_text(source_digest())
    return LIB_PATH


def ptxas_report() -> str:
    """Concatenated ``ptxas -v`` output of the last build (registers, spills, smem)."""
    out = []
    for log in sorted(OBJ_DIR.glob("*.log")):
        txt = log.read_text()
        if "ptxas info" in txt:
            out.append(f"=== {log.stem}\n{txt}")
    return "\n".join(out)


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args(argv)
    path = build(force=a.force, verbose=a.verbose)
    print(path)
    return 0


if __name__ == "__main__":
    sys.exit(main())
