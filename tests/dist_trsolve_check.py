"""Multi-GPU parity check of the distributed triangular solver, one process per GPU (run under torchrun; started by
tests/test_zz_triangular_dist_gpu.py or by hand):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29611 \
      tests/dist_trsolve_check.py

Cases follow test/unit/solver/test_triangular.cpp:106-175 (CorrectnessDistributed): the analytical system for side x
uplo x op x diag at the reference's sizes with the non-zero source rank (max(0, P-1), min(1, Q-1)), tolerance
20 (m+1) error, on every P x Q factorisation of the world size; plus larger random systems on the TMA + DMMA path
against LAPACK. Every rank builds the same global matrices on the host, solves with its own blocks on its GPU and
compares its block. The same algorithm template runs on the CPU for simulated grids in tests/test_triangular_dist_emu.py.
"""
import os
import sys

import numpy as np
import scipy.linalg as sla
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
from conftest import load_package  # noqa: E402
from dist_gpu_check import new_uid  # noqa: E402
import dlaf_oracle as orc  # noqa: E402

NP = {"s": np.float32, "d": np.float64, "c": np.complex64, "z": np.complex128}
QUICK = os.environ.get("DLAFB_DIST_QUICK") == "1"
SMALL = os.environ.get("DLAFB_CHECK_SMALL") == "1"   # CPU dry run (tests/dryrun_dist_scripts.py): same code, small sizes
TYPES = "dz" if QUICK else "sdcz"


def to_dev(loc, pad, fill):
    """Column-major local block with leading dimension rows + pad on the device; returns (tensor, ld, host image)."""
    rows, cols = loc.shape
    ld = max(1, rows) + pad
    host = np.full((max(1, cols), ld), fill, dtype=loc.dtype)
    host[:cols, :rows] = loc.T
    return torch.from_numpy(host.reshape(-1).copy()).cuda(), ld, host


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = load_package()
    L = pkg.lib()
    stream = torch.cuda.current_stream().cuda_stream
    failures, ncases = [], 0
    grids = [(p, world // p) for p in range(1, world + 1) if world % p == 0]

    for P, Q in grids:
        grid = pkg.Grid(L, P, Q, rank, world, new_uid(L, rank), "R")
        me = (grid.pr, grid.pc)
        src = (max(0, P - 1), min(1, Q - 1))

        def run_case(x, side, uplo, op, diag, alpha, a, b, want, mb, nb, tol_abs, tol_rel, what):
            m, n = b.shape
            ta = mb if side == "L" else nb
            la, _, _ = orc.local_block(a, ta, ta, (P, Q), me, src)
            lb, ri, ci = orc.local_block(b, mb, nb, (P, Q), me, src)
            da, lda, ha = to_dev(la, 1, NP[x](-7.7))
            db, ldb, _ = to_dev(lb, 3, NP[x](-5.5))
            plan = pkg.TriangularSolverDistPlan(L, grid, x, side, uplo, op, diag, m, n, mb, nb, src)
            ok_sizes = (plan.a_rows, plan.a_cols, plan.b_rows, plan.b_cols) == (la.shape[0], la.shape[1], lb.shape[0],
                                                                                lb.shape[1])
            plan.run(alpha, da.data_ptr(), lda, db.data_ptr(), ldb, stream)
            torch.cuda.synchronize()
            plan.close()
            out = db.cpu().numpy().reshape(max(1, lb.shape[1]), ldb)
            got = np.ascontiguousarray(out[:lb.shape[1], :lb.shape[0]].T)
            bad = orc.check_near(got, want[np.ix_(ri, ci)], tol_abs, tol_rel) if got.size else 0
            pad_ok = bool(np.all(out[:lb.shape[1], lb.shape[0]:] == NP[x](-5.5)))
            a_ok = np.array_equal(da.cpu().numpy().reshape(ha.shape), ha, equal_nan=True)
            return (bad > 0 or not pad_ok or not a_ok or not ok_sizes,
                    f"{what} bad={bad} pad_ok={pad_ok} a_untouched={a_ok} sizes_ok={ok_sizes}")

        for x in TYPES:
            alpha = orc.triangular_test_alpha(NP[x])
            for side in "LR":
                for uplo in "LU":
                    for op in "NTC":
                        for diag in ("N",) if QUICK else "NU":
                            for m, n, mb, nb in orc.TRIANGULAR_TEST_SIZES:
                                a, b, want = orc.triangular_system(side, uplo, op, diag, alpha, m, n, NP[x])
                                tol = 20 * (m + 1) * orc.type_error(NP[x])
                                fail, msg = run_case(x, side, uplo, op, diag, alpha, a, b, want, mb, nb, tol, tol,
                                                     f"analytic {x} {side}{uplo}{op}{diag} m={m} n={n} mb={mb} nb={nb} "
                                                     f"grid={P}x{Q}")
                                ncases += 1
                                if fail:
                                    failures.append(msg)
        # larger random systems: aligned tiles (TMA + DMMA contractions for d and z), more tiles than ranks, ragged end
        rng = np.random.default_rng(23)
        big = [("d", 1536, 1280, 256, 128), ("d", 1100, 900, 128, 128), ("z", 640, 512, 64, 128)]
        if SMALL:
            big = [("d", 96, 80, 16, 8), ("d", 69, 57, 8, 8), ("z", 40, 32, 4, 8)]
        for x, m, n, mb, nb in big:
            for side in "LR":
                na = m if side == "L" else n
                cplx = x in "cz"
                a = rng.standard_normal((na, na)) + (1j * rng.standard_normal((na, na)) if cplx else 0)
                a = (a / np.sqrt(na) + 2 * np.eye(na) * np.sign(rng.standard_normal(na))).astype(NP[x])
                b = (rng.standard_normal((m, n)) + (1j * rng.standard_normal((m, n)) if cplx else 0)).astype(NP[x])
                for uplo in "LU":
                    tri = np.tril(a) if uplo == "L" else np.triu(a)
                    ap = a.copy()
                    ap[np.triu(np.ones(a.shape, bool), 1) if uplo == "L" else np.tril(np.ones(a.shape, bool), -1)] = np.nan
                    for op in "NC" if QUICK else "NTC":
                        alpha = orc.triangular_test_alpha(NP[x])
                        opa = tri if op == "N" else tri.T if op == "T" else tri.conj().T
                        want = sla.solve(opa, alpha * b) if side == "L" else sla.solve(opa.T, (alpha * b).T).T
                        tol = 1e-11 * np.abs(want).max() * max(m, n)
                        fail, msg = run_case(x, side, uplo, op, "N", alpha, ap, b, want, mb, nb, tol, 0.0,
                                             f"random {x} {side}{uplo}{op} m={m} n={n} mb={mb} nb={nb} grid={P}x{Q}")
                        ncases += 1
                        if fail:
                            failures.append(msg)
        grid.close()

    flag = torch.tensor([len(failures)], device="cuda")
    dist.all_reduce(flag)
    if failures:
        print(f"[rank {rank}] {len(failures)} of {ncases} FAILED:\n  " + "\n  ".join(failures[:20]), flush=True)
    if rank == 0:
        print(f"dist_trsolve_check world={world} grids={grids} cases={ncases} total_failures={int(flag.item())}",
              flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(1 if int(flag.item()) else 0)


if __name__ == "__main__":
    main()
