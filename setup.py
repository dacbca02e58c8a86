"""Packaging for cuda_mpi_scratch_b200.

The native library is NOT a Python extension module: it is one plain shared object with a C ABI
(`cuda_mpi_scratch_b200/lib/libb200comm.so`) that the package loads with ctypes, built for
sm_100a only by `cuda_mpi_scratch_b200/build.py`.  `python setup.py build_ext --inplace` (or
`pip install -e . --no-build-isolation`) runs that build in-tree; the package also rebuilds the
library itself when its sources change (see `_lib.py`).
"""
import importlib.util
import os

from setuptools import Command, find_packages, setup
from setuptools.command.build_py import build_py

ROOT = os.path.dirname(os.path.abspath(__file__))


def _native_build(force=False):
    spec = importlib.util.spec_from_file_location("b200_build", os.path.join(ROOT, "cuda_mpi_scratch_b200", "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.build(force=force)


class BuildNative(Command):
    description = "compile csrc/ for sm_100a into cuda_mpi_scratch_b200/lib/libb200comm.so"
    user_options = [("inplace", "i", "accepted for compatibility; the build is always in-tree"),
                    ("force", "f", "rebuild everything")]
    boolean_options = ["inplace", "force"]

    def initialize_options(self):
        self.inplace = False
        self.force = False

    def finalize_options(self):
        pass

    def run(self):
        print("built", _native_build(force=bool(self.force)))


class BuildPyWithNative(build_py):
    def run(self):
        _native_build()
        super().run()


setup(
    name="cuda_mpi_scratch_b200",
    version="0.1.0",
    description="Blackwell-native (sm_100a) GPU communication + stencil framework: MPI-like API, "
                "fused stencil/halo, fused dot/allreduce, datatype-driven device moves",
    packages=find_packages(include=["cuda_mpi_scratch_b200", "cuda_mpi_scratch_b200.*"]),
    package_data={"cuda_mpi_scratch_b200": ["lib/*.so", "lib/*.digest"]},
    python_requires=">=3.10",
    install_requires=["numpy"],
    cmdclass={"build_ext": BuildNative, "build_py": BuildPyWithNative},
)
