"""-m gpu: ScaLAPACK sub-matrix arguments of the C API (dlaf_p?potrf with ia, ja > 1; DLAF_descriptor.i, .j != 0) on
one GPU, with host and with device pointers, and on several GPUs (tests/dist_submatrix_check.py under torchrun).
(The file name sorts last on purpose: written after the round's GPU budget was spent.)"""
import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest

import dlaf_oracle as orc

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CTX = 43


@pytest.fixture(scope="module")
def capi(pkg, L):
    L.dlaf_initialize(0, None, 0, None)
    pkg.check(L.dlafb_mpi_init(0, 1, None), "mpi_init")
    L.dlaf_create_grid_from_blacs_info(CTX, L.dlafb_mpi_comm_world(), 1, 1, b"R")
    yield L
    L.dlaf_free_grid(CTX)
    L.dlaf_finalize()


def build(x, uplo, m, nb, i0, j0, n):
    from gpu_util import NP
    dt = NP[x]
    parent = np.full((m, m), dt(-9.9))
    a = orc.gen_hpd(n, nb, dt)
    keep = (np.tril if uplo == "L" else np.triu)(np.ones((n, n))) > 0
    parent[i0 * nb:i0 * nb + n, j0 * nb:j0 * nb + n] = np.where(keep, a, dt(-9.9))
    ref = a.copy(order="F")
    assert orc.cholesky_local(uplo, ref, nb) == 0
    expect = parent.copy()
    expect[i0 * nb:i0 * nb + n, j0 * nb:j0 * nb + n] = np.where(keep, ref, dt(-9.9))
    return parent, expect, 50 * n * float(np.finfo(dt).eps) * np.sqrt(2.0)


CASES = [("d", 1536, 128, 3, 3, 1152), ("d", 1536, 128, 2, 5, 700), ("z", 640, 64, 1, 4, 384), ("s", 100, 8, 5, 2, 37),
         ("c", 96, 16, 2, 2, 64), ("d", 64, 16, 0, 3, 16)]


@pytest.mark.parametrize("x,m,nb,i0,j0,n", CASES)
@pytest.mark.parametrize("uplo", "LU")
def test_pxpotrf_submatrix_host_pointer(pkg, capi, x, uplo, m, nb, i0, j0, n):
    L = capi
    parent, expect, tol = build(x, uplo, m, nb, i0, j0, n)
    ld = m + 3
    h = np.full((m, ld), parent.dtype.type(-7.7))
    h[:, :m] = parent.T
    desca = pkg.DESC9(1, CTX, m, m, nb, nb, 0, 0, ld)
    info = ctypes.c_int(-1)
    getattr(L, f"dlaf_p{x}potrf")(uplo.encode(), n, h.ctypes.data, i0 * nb + 1, j0 * nb + 1, desca, ctypes.byref(info))
    assert info.value == 0
    assert orc.check_near(np.ascontiguousarray(h[:, :m].T), expect, tol, tol) == 0
    assert np.all(h[:, m:] == parent.dtype.type(-7.7))


@pytest.mark.parametrize("x,m,nb,i0,j0,n", CASES[:3])
def test_descriptor_offsets_device_pointer(pkg, capi, x, m, nb, i0, j0, n):
    from gpu_util import DevMat, sync
    L = capi
    parent, expect, tol = build(x, "L", m, nb, i0, j0, n)
    A = DevMat(parent, ld=m + 2)
    desc = pkg.DLAFDescriptor(n, n, nb, nb, 0, 0, i0 * nb, j0 * nb, A.ld)
    r = getattr(L, f"dlaf_cholesky_factorization_{x}")(CTX, b"L", ctypes.c_void_p(A.ptr), desc)
    sync()
    assert r == 0
    assert orc.check_near(A.get(), expect, tol, tol) == 0
    assert np.all(A.pad() == parent.dtype.type(-7.7))


@pytest.mark.parametrize("world", [2, 4])
def test_distributed_submatrix(world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29700 + world), os.path.join(ROOT, "tests",
                                                                                          "dist_submatrix_check.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=1500)
    tail = (r.stdout + r.stderr)[-3000:]
    assert r.returncode == 0, tail
    assert "total_failures=0" in r.stdout, tail
