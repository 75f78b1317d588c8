"""The independent residual checker of bench.py's dist_parity leg (N > 1), exercised over a REAL process group on the
CPU: world size 2 (grids 1 x 2 and 2 x 1) and 4 (2 x 2), torch.distributed with gloo. Every rank holds its 2D
block-cyclic block of A and of a factor computed with LAPACK; the checker must return a residual at rounding level,
must see a perturbed factor, and must agree with numpy on the global matrix. Also pins bench.py's restatement of the
reference's closed-form test matrices against the oracle's."""
import os
import socket

import numpy as np
import pytest

import dlaf_oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, grid, port, q, cplx):
    import sys
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import bench
    import dlaf_oracle as orc_
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    P, Q = grid
    pr, pc = rank // Q, rank % Q
    n, nb = 200, 32   # ragged last tile
    dt = np.complex128 if cplx else np.float64
    a = orc_.gen_hpd(n, nb, dt)
    lfac = np.linalg.cholesky(a)
    lfull = np.tril(lfac) + np.triu(np.full((n, n), 123.0, dtype=dt), 1)   # garbage above the diagonal, as in place

    def idx(Pn, r):
        tiles = range(r, -(-n // nb), Pn)
        return np.concatenate([np.arange(g * nb, min(n, (g + 1) * nb)) for g in tiles] + [np.arange(0)]).astype(int)
    rows, cols = idx(P, pr), idx(Q, pc)
    ld = len(rows) + 5

    def local(m):
        h = np.full((max(1, len(cols)), ld), -7.7, dtype=dt)
        h[:len(cols), :len(rows)] = m[np.ix_(rows, cols)].T
        return torch.from_numpy(h.reshape(-1).copy())
    out = {}
    for name, fac in (("good", lfull), ("bad", lfull + np.where(np.arange(n)[:, None] == 70, 1e-3, 0) *
                                        (np.arange(n)[None, :] == 10))):
        amax, dmax = bench.independent_residual(local(a), local(fac), ld, len(rows), len(cols), n, nb, P, Q, pr, pc,
                                                world, cplx)
        lo = np.tril(fac)
        want = np.abs(np.tril(a - lo @ lo.conj().T)).max()
        out[name] = (amax, dmax, float(np.abs(np.tril(a)).max()), float(want))
    q.put((rank, out))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("grid,cplx", [((1, 2), False), ((2, 1), True), ((2, 2), False)])
def test_independent_residual_over_gloo(grid, cplx):
    import torch.multiprocessing as mp
    world = grid[0] * grid[1]
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, grid, port, q, cplx)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    n = 200
    for rank, out in res:
        amax, dmax, want_amax, want_d = out["good"]
        assert abs(amax - want_amax) <= 1e-12 * want_amax
        assert dmax <= n * np.finfo(np.float64).eps * amax
        assert abs(dmax - want_d) <= 1e-9 * amax + 1e-3 * want_d
        amax, dmax, _, want_d = out["bad"]
        assert dmax > 1e-4 and abs(dmax - want_d) <= 1e-6 * want_d   # sees the perturbed factor, same number as numpy


@pytest.mark.parametrize("dt", [np.float32, np.float64, np.complex64, np.complex128])
@pytest.mark.parametrize("uplo", "LU")
def test_closed_form_restatement_matches_the_oracle(dt, uplo):
    import sys
    sys.path.insert(0, ROOT)
    import bench
    for n in (0, 5, 34, 128):
        a, t = orc.cholesky_setters(n, uplo, dt)
        a2, t2 = bench.analytic_pair(n, uplo, dt, np.arange(n), np.arange(n))
        assert a.shape == a2.shape
        if n:
            assert np.allclose(a, a2, rtol=4 * np.finfo(dt).eps, atol=0) and np.array_equal(t, t2)
    rows, cols = np.array([3, 4, 9]), np.array([0, 9, 11, 12])
    a, t = orc.cholesky_setters(16, uplo, dt, rows, cols)
    a2, t2 = bench.analytic_pair(16, uplo, dt, rows, cols)
    assert np.allclose(a, a2, rtol=4 * np.finfo(dt).eps, atol=0) and np.array_equal(t, t2)
