"""N > 1 path of the distributed rows over a REAL process group on the CPU: two processes (grids 1 x 2, 2 x 1) and four
(2 x 2, where the row and the column communicators are both non-trivial), torch.distributed with the gloo backend. Each process runs the product's algorithm template for ITS rank (tests/emu: the per-rank entry points of
solver_emu.cu), and every broadcast / reduction / all-to-all of the template is a callback into this file that performs
it with torch.distributed on the row / column / world group - the same call sequence a rank issues to NCCL on the
device (src/dist_device_ops.h). A rank that skipped or reordered a collective would hang or fail here as it would there.

  distributed triangular solver   dla-future_b200/src/triangular_solver_dist.h   (both patterns: broadcast and reduce)
  distributed reduction to standard form   src/gen_to_std_dist.h                 (panels, grouped column panel, trapezoid)
  matrix redistribution           src/redistribute.h                             (pack - all-to-all - unpack)
"""
import ctypes
import os
import socket
import subprocess

import numpy as np
import pytest

import dlaf_oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

BCAST = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_longlong)
REDUCE = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int)
A2A = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.POINTER(ctypes.c_longlong), ctypes.POINTER(ctypes.c_longlong),
                       ctypes.c_void_p, ctypes.POINTER(ctypes.c_longlong), ctypes.POINTER(ctypes.c_longlong), ctypes.c_int,
                       ctypes.c_int)


def colmajor(a, ld, fill):
    m, n = a.shape
    h = np.full((max(n, 1), ld), fill, dtype=a.dtype)
    h[:n, :m] = a.T
    return h


def _worker(rank, world, grid, port, q):
    import sys
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import scipy.linalg as sla
    lib = ctypes.CDLL(os.path.join(ROOT, "tests", "emu", "_build", "libsolver_emu.so"))
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    P, Q = grid
    pr, pc = rank // Q, rank % Q
    row_group = col_group = None
    for r_ in range(P):                                  # every rank must create every group, in the same order
        g = dist.new_group([r_ * Q + c for c in range(Q)])
        if r_ == pr:
            row_group = g
    for c_ in range(Q):
        g = dist.new_group([r * Q + c_ for r in range(P)])
        if c_ == pc:
            col_group = g
    ncalls = {"bcast": 0, "reduce": 0, "a2a": 0}

    def view(ptr, nbytes, dtype=np.uint8):
        buf = (ctypes.c_char * nbytes).from_address(ptr)
        return torch.from_numpy(np.frombuffer(buf, dtype=dtype))

    def bcast(axis, root, ptr, nbytes):
        ncalls["bcast"] += 1
        src = pr * Q + root if axis == 0 else root * Q + pc
        dist.broadcast(view(ptr, nbytes), src=src, group=row_group if axis == 0 else col_group)
        return 0

    def reduce(axis, root, ptr, count, is_double):
        ncalls["reduce"] += 1
        dst = pr * Q + root if axis == 0 else root * Q + pc
        dt = np.float64 if is_double else np.float32
        dist.reduce(view(ptr, count * np.dtype(dt).itemsize, dt), dst=dst, op=dist.ReduceOp.SUM,
                    group=row_group if axis == 0 else col_group)
        return 0

    def a2a(sendbuf, send_off, send_cnt, recvbuf, recv_off, recv_cnt, elem, nranks):
        ncalls["a2a"] += 1
        reqs = []
        for p in range(nranks):
            if p == rank:
                continue
            if recv_cnt[p] > 0:
                reqs.append(dist.irecv(view(recvbuf + recv_off[p] * elem, recv_cnt[p] * elem), src=p))
            if send_cnt[p] > 0:
                reqs.append(dist.isend(view(sendbuf + send_off[p] * elem, send_cnt[p] * elem), dst=p))
        for r_ in reqs:
            r_.wait()
        return 0

    cb_b, cb_r, cb_a = BCAST(bcast), REDUCE(reduce), A2A(a2a)
    failures = []
    src = (max(0, P - 1), min(1, Q - 1))
    rng = np.random.default_rng(77)

    # ---- distributed triangular solver: all side / uplo / op cases, double and complex double
    for x, dt in (("s", np.float32), ("d", np.float64), ("c", np.complex64), ("z", np.complex128)):
        f = getattr(lib, f"emu_triangular_solver_dist_rank_{x}")
        f.argtypes = [ctypes.c_int] * 6 + [ctypes.c_char] * 4 + [ctypes.c_int] * 4 + [
            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_longlong, BCAST, REDUCE]
        f.restype = ctypes.c_int
        m, n, mb, nb = 23, 17, 4, 3
        for side in "LR":
            for uplo in "LU":
                for op in "NTC":
                    na = m if side == "L" else n
                    cplx = x in "cz"
                    a = rng.standard_normal((na, na)) + (1j * rng.standard_normal((na, na)) if cplx else 0)
                    a = (a / np.sqrt(na) + 2 * np.eye(na)).astype(dt)
                    b = (rng.standard_normal((m, n)) + (1j * rng.standard_normal((m, n)) if cplx else 0)).astype(dt)
                    tri = np.tril(a) if uplo == "L" else np.triu(a)
                    alpha = orc.triangular_test_alpha(dt)
                    opa = tri if op == "N" else tri.T if op == "T" else tri.conj().T
                    want = sla.solve(opa, alpha * b) if side == "L" else sla.solve(opa.T, (alpha * b).T).T
                    ta = mb if side == "L" else nb
                    la, _, _ = orc.local_block(a, ta, ta, grid, (pr, pc), src)
                    lb, ri, ci = orc.local_block(b, mb, nb, grid, (pr, pc), src)
                    ha = colmajor(la, max(1, la.shape[0]) + 1, dt(-7.7))
                    hb = colmajor(lb, max(1, lb.shape[0]) + 2, dt(-5.5))
                    al = np.array([alpha], dtype=dt)
                    r = f(P, Q, pr, pc, src[0], src[1], side.encode(), uplo.encode(), op.encode(), b"N", m, n, mb, nb,
                          al.ctypes.data, ha.ctypes.data, ha.shape[1], hb.ctypes.data, hb.shape[1], cb_b, cb_r)
                    got = hb[:lb.shape[1], :lb.shape[0]].T
                    tol = 5e4 * float(np.finfo(dt).eps) * max(m, n)   # 1e-11 max(m, n) in double
                    if r != 0 or (got.size and not np.abs(got - want[np.ix_(ri, ci)]).max() <= tol):
                        failures.append(f"trsolve {x} {side}{uplo}{op} rc={r}")

    # ---- distributed reduction to standard form, both uplo
    for x, dt in (("s", np.float32), ("d", np.float64), ("c", np.complex64), ("z", np.complex128)):
        f = getattr(lib, f"emu_gen_to_std_dist_rank_{x}")
        f.argtypes = [ctypes.c_int] * 6 + [ctypes.c_char, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_longlong,
                                           ctypes.c_void_p, ctypes.c_longlong, BCAST, REDUCE]
        f.restype = ctypes.c_int
        for uplo in "LU":
            for m, mb in [(34, 13), (32, 5), (16, 10)]:          # from test_gen_to_std.cpp:53-57
                t, a, res = orc.gen_to_std_setters(m, uplo, dt)
                la, ri, ci = orc.local_block(a, mb, mb, grid, (pr, pc), src)
                lt, _, _ = orc.local_block(t, mb, mb, grid, (pr, pc), src)
                ha = colmajor(la, max(1, la.shape[0]) + 1, dt(-7.7))
                ht = colmajor(lt, max(1, lt.shape[0]), dt(-5.5))
                r = f(P, Q, pr, pc, src[0], src[1], uplo.encode(), m, mb, ha.ctypes.data, ha.shape[1], ht.ctypes.data,
                      ht.shape[1], cb_b, cb_r)
                got = np.ascontiguousarray(ha[:la.shape[1], :la.shape[0]].T)
                tol = 10 * (m + 1) * orc.type_error(dt)
                if r != 0 or (got.size and orc.check_near(got, res[np.ix_(ri, ci)], tol, 0.0) != 0):
                    failures.append(f"gen_to_std {x} {uplo} m={m} mb={mb} rc={r}")

    # ---- redistribution, both directions
    f = lib.emu_redistribute_rank_d
    f.argtypes = [ctypes.c_int] * 4 + [ctypes.c_char, ctypes.c_longlong, ctypes.c_longlong] + [ctypes.c_int] * 8 + [
        ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_longlong, A2A]
    f.restype = ctypes.c_int
    a = rng.standard_normal((41, 29))
    for st, dtile, sr, dr in [((4, 3), (8, 6), (0, 0), src), ((8, 6), (4, 3), src, (0, 0)), ((5, 29), (1, 1), (0, 0), (0, 0))]:
        ls, _, _ = orc.local_block(a, st[0], st[1], grid, (pr, pc), sr)
        want, _, _ = orc.local_block(a, dtile[0], dtile[1], grid, (pr, pc), dr)
        hs = colmajor(ls, max(1, ls.shape[0]) + 1, -7.7)
        hd = np.full((max(1, want.shape[1]), max(1, want.shape[0]) + 2), -5.5)
        r = f(P, Q, pr, pc, b"R", 41, 29, st[0], st[1], sr[0], sr[1], dtile[0], dtile[1], dr[0], dr[1], hs.ctypes.data,
              hs.shape[1], hd.ctypes.data, hd.shape[1], cb_a)
        got = hd[:want.shape[1], :want.shape[0]].T
        if r != 0 or not np.array_equal(got, want):
            failures.append(f"redistribute {st}->{dtile} rc={r}")

    q.put((rank, failures, dict(ncalls)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("grid", [(1, 2), (2, 1), (2, 2)])
def test_process_group_gloo(grid):
    import torch.multiprocessing as mp
    so = os.path.join(ROOT, "tests", "emu", "_build", "libsolver_emu.so")
    if not os.path.exists(so):
        subprocess.check_call(["make", "emu"], cwd=ROOT)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    world = grid[0] * grid[1]
    procs = [ctx.Process(target=_worker, args=(r, world, grid, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    for rank, failures, ncalls in res:
        assert not failures, (rank, failures[:5])
        # the collectives really went through the process group
        assert ncalls["bcast"] > 100 and ncalls["reduce"] > 20 and ncalls["a2a"] == 3, ncalls
