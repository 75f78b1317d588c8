"""Developer tool: runs the multi-GPU parity scripts (tests/dist_trsolve_check.py, dist_g2s_check.py,
dist_redistribute_check.py) on a box WITHOUT GPUs, as real multi-process jobs: torch.distributed with gloo instead of
NCCL, CPU tensors instead of device tensors, and the product's per-rank algorithm templates executed by the CPU
emulation (tests/emu) with every collective performed on gloo groups - so the scripts' own logic (scatter, local sizes,
tolerances, comparisons, collective bookkeeping) is exercised before it costs a multi-GPU session.

    DLAFB_CHECK_SMALL=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29555 tests/dryrun_dist_scripts.py dist_trsolve_check
"""
import ctypes
import os
import sys
import types

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
NP = {"s": np.float32, "d": np.float64, "c": np.complex64, "z": np.complex128}
EMU = ctypes.CDLL(os.path.join(ROOT, "tests", "emu", "_build", "libsolver_emu.so"))
BCAST = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_longlong)
REDUCE = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int)
A2A = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.POINTER(ctypes.c_longlong), ctypes.POINTER(ctypes.c_longlong),
                       ctypes.c_void_p, ctypes.POINTER(ctypes.c_longlong), ctypes.POINTER(ctypes.c_longlong), ctypes.c_int,
                       ctypes.c_int)


def view(ptr, nbytes, dtype=np.uint8):
    return torch.from_numpy(np.frombuffer((ctypes.c_char * nbytes).from_address(ptr), dtype=dtype))


class Grid:
    """Stand-in for pkg.Grid: coordinates as grid.cu computes them, gloo groups for the row / column communicators."""

    def __init__(self, L, P, Q, rank, world, uid, ordering="R"):
        self.P, self.Q, self.rank, self.ordering = P, Q, rank, ordering
        self.pr, self.pc = (rank // Q, rank % Q) if ordering == "R" else (rank % P, rank // P)
        lin = (lambda r, c: r * Q + c) if ordering == "R" else (lambda r, c: c * P + r)
        self.lin = lin
        self.row_group = self.col_group = None
        for r in range(P):                      # every rank creates every group, in the same order
            g = dist.new_group([lin(r, c) for c in range(Q)])
            if r == self.pr:
                self.row_group = g
        for c in range(Q):
            g = dist.new_group([lin(r, c) for r in range(P)])
            if c == self.pc:
                self.col_group = g

        def bcast(axis, root, ptr, nbytes):
            src = lin(self.pr, root) if axis == 0 else lin(root, self.pc)
            dist.broadcast(view(ptr, nbytes), src=src, group=self.row_group if axis == 0 else self.col_group)
            return 0

        def reduce(axis, root, ptr, count, is_double):
            dst = lin(self.pr, root) if axis == 0 else lin(root, self.pc)
            dt = np.float64 if is_double else np.float32
            dist.reduce(view(ptr, count * np.dtype(dt).itemsize, dt), dst=dst, op=dist.ReduceOp.SUM,
                        group=self.row_group if axis == 0 else self.col_group)
            return 0

        def a2a(sendbuf, send_off, send_cnt, recvbuf, recv_off, recv_cnt, elem, nranks):
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

        self.cb_b, self.cb_r, self.cb_a = BCAST(bcast), REDUCE(reduce), A2A(a2a)

    def close(self):
        pass


def local_size(n, nb, nprocs, rank, src):
    import dlaf_oracle as orc
    return len(orc.local_indices(n, nb, nprocs, rank, src))


class TriangularSolverDistPlan:
    def __init__(self, L, grid, x, side, uplo, op, diag, m, n, mb, nb, src=(0, 0)):
        self.g, self.args, self.src = grid, (x, side, uplo, op, diag, m, n, mb, nb), src
        ta, na = (mb, m) if side == "L" else (nb, n)
        self.a_rows, self.a_cols = local_size(na, ta, grid.P, grid.pr, src[0]), local_size(na, ta, grid.Q, grid.pc, src[1])
        self.b_rows, self.b_cols = local_size(m, mb, grid.P, grid.pr, src[0]), local_size(n, nb, grid.Q, grid.pc, src[1])

    def run(self, alpha, a, lda, b, ldb, stream):
        x, side, uplo, op, diag, m, n, mb, nb = self.args
        f = getattr(EMU, f"emu_triangular_solver_dist_rank_{x}")
        f.argtypes = [ctypes.c_int] * 6 + [ctypes.c_char] * 4 + [ctypes.c_int] * 4 + [
            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_longlong, BCAST, REDUCE]
        al = np.array([alpha], dtype=NP[x])
        g = self.g
        r = f(g.P, g.Q, g.pr, g.pc, self.src[0], self.src[1], side.encode(), uplo.encode(), op.encode(), diag.encode(), m,
              n, mb, nb, al.ctypes.data, a, lda, b, ldb, g.cb_b, g.cb_r)
        assert r == 0, r

    def close(self):
        pass


class GenToStdDistPlan:
    def __init__(self, L, grid, x, uplo, n, nb, src=(0, 0)):
        self.g, self.args, self.src = grid, (x, uplo, n, nb), src
        self.local_rows, self.local_cols = local_size(n, nb, grid.P, grid.pr, src[0]), local_size(n, nb, grid.Q, grid.pc, src[1])

    def run(self, a, lda, f_, ldf, stream):
        x, uplo, n, nb = self.args
        f = getattr(EMU, f"emu_gen_to_std_dist_rank_{x}")
        f.argtypes = [ctypes.c_int] * 6 + [ctypes.c_char, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_longlong,
                                           ctypes.c_void_p, ctypes.c_longlong, BCAST, REDUCE]
        g = self.g
        r = f(g.P, g.Q, g.pr, g.pc, self.src[0], self.src[1], uplo.encode(), n, nb, a, lda, f_, ldf, g.cb_b, g.cb_r)
        assert r == 0, r

    def close(self):
        pass


class RedistributePlan:
    def __init__(self, L, grid, x, m, n, st, sr, dtile, dr):
        self.g, self.args = grid, (x, m, n, st, sr, dtile, dr)
        self.src_rows, self.src_cols = local_size(m, st[0], grid.P, grid.pr, sr[0]), local_size(n, st[1], grid.Q, grid.pc, sr[1])
        self.dst_rows, self.dst_cols = local_size(m, dtile[0], grid.P, grid.pr, dr[0]), local_size(n, dtile[1], grid.Q, grid.pc, dr[1])

    def run(self, s, lds, d, ldd, stream):
        x, m, n, st, sr, dtile, dr = self.args
        f = getattr(EMU, f"emu_redistribute_rank_{x}")
        f.argtypes = [ctypes.c_int] * 4 + [ctypes.c_char, ctypes.c_longlong, ctypes.c_longlong] + [ctypes.c_int] * 8 + [
            ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_longlong, A2A]
        g = self.g
        r = f(g.P, g.Q, g.pr, g.pc, g.ordering.encode(), m, n, st[0], st[1], sr[0], sr[1], dtile[0], dtile[1], dr[0], dr[1],
              s, lds, d, ldd, g.cb_a)
        assert r == 0, r

    def close(self):
        pass


def max_norm(L, grid, x, uplo, m, n, mb, nb, src, ptr, lda, stream):
    """Local maximum over the rank's part of the (global) uplo region with numpy, max over the ranks with gloo."""
    import dlaf_oracle as orc
    ri, ci = orc.local_indices(m, mb, grid.P, grid.pr, src[0]), orc.local_indices(n, nb, grid.Q, grid.pc, src[1])
    val = 0.0
    if ri.size and ci.size:
        a = np.frombuffer((ctypes.c_char * (len(ci) * lda * np.dtype(NP[x]).itemsize)).from_address(ptr), dtype=NP[x])
        a = a.reshape(len(ci), lda)[:, :len(ri)].T
        mask = (ri[:, None] >= ci[None, :]) if uplo == "L" else np.ones(a.shape, bool)
        if mask.any():
            val = float(np.abs(a[mask]).max())
    t = torch.tensor([val], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


class OracleCApi:
    """Stand-in for the C API in tests/dist_submatrix_check.py: dlaf_p?potrf gathers the ranks' local blocks over gloo,
    factorizes sub(A) of the global matrix with the oracle and writes every rank's part back."""

    def __init__(self):
        self.grids = {}

    def dlaf_initialize(self, *a):
        return None

    def dlaf_finalize(self):
        return None

    def dlafb_mpi_init(self, rank, world, buf):
        self.rank, self.world = rank, world
        return 0

    def dlafb_mpi_comm_world(self):
        return None

    def dlaf_create_grid_from_blacs_info(self, ctx, comm, P, Q, order):
        self.grids[ctx] = (P, Q)

    def dlaf_free_grid(self, ctx):
        self.grids.pop(ctx, None)

    def __getattr__(self, name):
        if not (name.startswith("dlaf_p") and name.endswith("potrf")):
            raise AttributeError(name)
        x = name[6]
        import dlaf_oracle as orc

        def potrf(uplo, n, ptr, ia, ja, desca, info_ref):
            dt = np.dtype(NP[x])
            ctx, m, nb, isrc, jsrc, ld = desca[1], desca[2], desca[4], desca[6], desca[7], desca[8]
            P, Q = self.grids[ctx]
            me = (self.rank // Q, self.rank % Q)
            ri = orc.local_indices(m, nb, P, me[0], isrc)
            ci = orc.local_indices(m, nb, Q, me[1], jsrc)
            mine = np.frombuffer((ctypes.c_char * (max(1, len(ci)) * ld * dt.itemsize)).from_address(ptr), dtype=dt)
            mine = mine.reshape(max(1, len(ci)), ld)
            parts = [None] * self.world
            dist.all_gather_object(parts, (me, mine[:len(ci), :len(ri)].T.copy()))
            glob = np.zeros((m, m), dtype=dt)
            for (pr, pc), block in parts:
                r2, c2 = orc.local_indices(m, nb, P, pr, isrc), orc.local_indices(m, nb, Q, pc, jsrc)
                if r2.size and c2.size:
                    glob[np.ix_(r2, c2)] = block
            i0, j0 = ia - 1, ja - 1
            assert i0 % nb == 0 and j0 % nb == 0
            up = uplo.decode()
            keep = (np.tril if up == "L" else np.triu)(np.ones((n, n))) > 0
            sub = glob[i0:i0 + n, j0:j0 + n]
            work = np.asfortranarray(np.where(keep, sub, 0))
            info_ref._obj.value = orc.cholesky_local(up, work, nb)
            sub[keep] = work[keep]
            if ri.size and ci.size:
                mine[:len(ci), :len(ri)] = glob[np.ix_(ri, ci)].T
        return potrf


def main():
    script = sys.argv[1]
    # ---- CPU stand-ins for what the scripts take from torch.cuda / NCCL / the device library
    torch.cuda.set_device = lambda *a, **k: None
    torch.cuda.synchronize = lambda *a, **k: None
    torch.cuda.current_stream = lambda *a, **k: types.SimpleNamespace(cuda_stream=None)
    torch.Tensor.cuda = lambda self, *a, **k: self
    real_tensor = torch.tensor
    torch.tensor = lambda *a, **k: real_tensor(*a, **{kk: vv for kk, vv in k.items() if kk != "device"})
    real_init = dist.init_process_group
    dist.init_process_group = lambda backend, **k: real_init("gloo")
    capi = OracleCApi()
    pkg = types.SimpleNamespace(lib=lambda: capi, DESC9=ctypes.c_int * 9, Grid=Grid, TriangularSolverDistPlan=TriangularSolverDistPlan,
                                GenToStdDistPlan=GenToStdDistPlan, RedistributePlan=RedistributePlan, max_norm=max_norm)
    import conftest
    conftest.load_package = lambda: pkg
    import dist_gpu_check
    dist_gpu_check.new_uid = lambda L, rank: b"\0" * 128
    mod = __import__(script)
    if hasattr(mod, "new_uid"):
        mod.new_uid = dist_gpu_check.new_uid
    mod.load_package = conftest.load_package
    mod.main()


if __name__ == "__main__":
    main()
