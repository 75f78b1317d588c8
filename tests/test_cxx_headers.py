"""The public C++ front ends (include/dlaf/**: cholesky_factorization, triangular_solver, generalized_to_standard,
Matrix, Distribution, CommunicatorGrid) must compile and instantiate for the four element types with the host
compiler alone (they are header-only wrappers over the C ABI). Syntax check only: nothing is run."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_front_ends_instantiate_for_all_types():
    cuda_inc = "/usr/local/cuda/include"
    if shutil.which("g++") is None or not os.path.exists(os.path.join(cuda_inc, "cuda_runtime.h")):
        pytest.skip("needs g++ and the CUDA headers")
    cmd = ["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "include", "dlafb_mpi_shim"), "-I" + cuda_inc,
           os.path.join(ROOT, "tests", "cxx", "api_compile.cpp")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_c_headers_are_plain_c_and_every_symbol_links(pkg, tmp_path):
    """The drop-in boundary is a C ABI: every public C header must compile as C99 (no C++-isms) and the functions it
    declares must link from a C program against the shared library (C linkage, no mangling). Not run."""
    if shutil.which("gcc") is None:
        pytest.skip("needs gcc")
    src = os.path.join(ROOT, "tests", "cxx", "c_headers.c")
    inc = ["-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "include", "dlafb_mpi_shim")]
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only"] + inc + [src],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    if not os.path.exists(pkg.LIB_PATH):
        pytest.skip("library not built")
    libdir = os.path.dirname(pkg.LIB_PATH)
    r = subprocess.run(["gcc", "-std=c99"] + inc + [src, "-o", str(tmp_path / "c_link"), "-L" + libdir, "-ldlaf_b200",
                        "-Wl,-rpath," + libdir, "-Wl,--allow-shlib-undefined"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_real_mpi_bootstrap_compiles(tmp_path):
    """-DDLAFB_WITH_MPI: dlaf_create_grid(MPI_Comm, ...) takes a REAL communicator as in the reference
    (include/dlaf_c/grid.h:31): rank / size from MPI, the NCCL id over MPI_Bcast, grid_ordering over MPI_Allreduce. No MPI
    on this image: src/c_api.cu is compiled (not linked, not run) against tests/cxx/mpi_stub/mpi.h, and the object must
    reference the MPI functions and no longer define the shim's."""
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("needs nvcc")
    obj = str(tmp_path / "c_api_mpi.o")
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O1", "-std=c++17", "-DDLAFB_WITH_MPI",
           "-I" + os.path.join(ROOT, "tests", "cxx", "mpi_stub"), "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "dla-future_b200", "csrc"), "-I" + os.path.join(ROOT, "dla-future_b200", "src"),
           "-Xcompiler", "-fPIC,-Wall", "-c", os.path.join(ROOT, "dla-future_b200", "src", "c_api.cu"), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    syms = subprocess.run(["nm", obj], capture_output=True, text=True).stdout
    for name in ("MPI_Bcast", "MPI_Allreduce", "MPI_Comm_rank", "MPI_Comm_size", "MPI_Barrier"):
        assert f" U {name}" in syms, name
    assert "dlafb_mpi_init" not in syms and " T MPI_Comm_rank" not in syms
    assert " T dlaf_create_grid" in syms and " T dlaf_pdpotrf" in syms
