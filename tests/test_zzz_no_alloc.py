"""-m gpu: dlafb_potrf_tile_noalloc_d (include/dlafb_kernels.h) - the diagonal-tile Cholesky with its workspace in a
file-scope device array of the library, for callers that cannot allocate (no counterpart in the reference).
Called through the C ABI and compared with the existing entry dlafb_potrf_tile_d
(caller-supplied workspace) on two calls in a row - the array keeps the previous call's contents, so the second call
follows a different size / uplo - at the tolerance of the tile tests, 4(n+1)*err (test_lapack_tile.cpp:142-205).
Collected last on purpose."""
import numpy as np
import pytest
import torch

import dlaf_oracle as orc
from gpu_util import DevMat, stream_ptr, sync

pytestmark = pytest.mark.gpu


def hpd(n, seed):
    rng = np.random.default_rng(seed)
    m = rng.uniform(-1, 1, (n, n))
    return (m + m.T) / 2 + 2 * n * np.eye(n)


def run_existing(pkg, L, a, uplo, ld):
    n = a.shape[0]
    A = DevMat(a, ld=ld)
    ib = L.dlafb_inner_block(b"d", A.ptr, A.ld, n)
    work = torch.full((L.dlafb_potrf_work_elems(b"d", n, ib) + 64,), float("nan"), dtype=torch.float64, device="cuda")
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    pkg.check(L.dlafb_potrf_tile_d(stream_ptr(), uplo.encode(), n, A.ptr, A.ld, work.data_ptr(), ib, info.data_ptr(), 0),
              "potrf_tile")
    sync()
    return A, int(info.item())


def run_noalloc(pkg, L, a, uplo, ld):
    n = a.shape[0]
    A = DevMat(a, ld=ld)
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    before = torch.cuda.memory_allocated()
    pkg.check(L.dlafb_potrf_tile_noalloc_d(stream_ptr(), uplo.encode(), n, A.ptr, A.ld, info.data_ptr(), 0),
              "potrf_tile_noalloc")
    sync()
    assert torch.cuda.memory_allocated() == before
    return A, int(info.item())


@pytest.mark.parametrize("first,second", [((512, "L"), (300, "U")), ((300, "U"), (512, "L")), ((128, "L"), (17, "L")),
                                          ((448, "U"), (448, "L"))])
def test_two_calls_in_a_row_match_the_existing_entry(pkg, L, first, second):
    for k, (n, uplo) in enumerate((first, second)):
        a = hpd(n, 11 * n + k)
        ref = np.tril(np.ones((n, n), bool)) if uplo == "L" else np.triu(np.ones((n, n), bool))
        a_in = np.where(ref, a, -9.9)  # the other triangle must come back untouched
        ld = n + (0 if n % 2 == 0 else 3)
        want, info_w = run_existing(pkg, L, a_in, uplo, ld)
        got, info_g = run_noalloc(pkg, L, a_in, uplo, ld)
        assert info_w == 0 and info_g == 0
        tol = 4 * (n + 1) * orc.type_error(np.float64)
        w, g = want.get(), got.get()
        assert orc.check_near(np.where(ref, g, 0), np.where(ref, w, 0), tol, tol) == 0
        assert np.all(g[~ref] == -9.9) and np.all(got.pad() == -7.7)
        # and against LAPACK, so that both being wrong the same way cannot pass
        lref = np.linalg.cholesky(a)
        fac = np.tril(g) if uplo == "L" else np.triu(g).T
        assert np.abs(fac - lref).max() <= 50 * n * np.finfo(np.float64).eps * np.abs(lref).max()


def test_info_and_size_limit(pkg, L):
    n = 256
    a = hpd(n, 5)
    a[200, 200] = -1.0
    _, info_w = run_existing(pkg, L, a, "L", n)
    A = DevMat(a)
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    assert L.dlafb_potrf_tile_noalloc_d(stream_ptr(), b"L", n, A.ptr, A.ld, info.data_ptr(), 7) == 0
    sync()
    assert info_w == 201 and int(info.item()) == 201 + 7
    # beyond the built-in scratch: rejected, nothing launched, the tile is left alone
    n = 513
    B = DevMat(hpd(n, 6))
    before = B.get().copy()
    info.zero_()
    assert L.dlafb_potrf_tile_noalloc_d(stream_ptr(), b"L", n, B.ptr, B.ld, info.data_ptr(), 0) != 0
    sync()
    assert np.array_equal(B.get(), before) and int(info.item()) == 0
    assert L.dlafb_potrf_tile_noalloc_d(stream_ptr(), b"L", 0, None, 1, info.data_ptr(), 0) == 0
