"""Multi-GPU parity check of the distributed generalized-to-standard reduction, one process per GPU (run under torchrun;
started by tests/test_zz_gen_to_std_gpu.py or by hand):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29811 \
      tests/dist_g2s_check.py

Cases follow test/unit/eigensolver/test_gen_to_std.cpp:85-125 (CorrectnessDistributed): the analytical case for uplo
L/U at the reference's sizes with the non-zero source rank, tolerance 10 (m+1) error on the WHOLE local block (the
unreferenced triangle must keep its -9.9), factor unchanged, on every P x Q factorisation of the world size; plus larger
random problems on the TMA + DMMA path against the oracle's restatement of the reference tile loops.
The same algorithm template runs on the CPU for simulated grids in tests/test_gen_to_std_dist_emu.py.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
from conftest import load_package  # noqa: E402
from dist_gpu_check import new_uid  # noqa: E402
from dist_trsolve_check import to_dev  # noqa: E402
import dlaf_oracle as orc  # noqa: E402

NP = {"s": np.float32, "d": np.float64, "c": np.complex64, "z": np.complex128}
QUICK = os.environ.get("DLAFB_DIST_QUICK") == "1"
SMALL = os.environ.get("DLAFB_CHECK_SMALL") == "1"   # CPU dry run (tests/dryrun_dist_scripts.py): same code, small sizes
TYPES = "dz" if QUICK else "sdcz"


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

        def run_case(x, uplo, a, f, want, nb, tol, nan_outside, what):
            n = a.shape[0]
            la, ri, ci = orc.local_block(a, nb, nb, (P, Q), me, src)
            lf, _, _ = orc.local_block(f, nb, nb, (P, Q), me, src)
            da, lda, _ = to_dev(la, 1, NP[x](-7.7))
            df, ldf, hf = to_dev(lf, 2, NP[x](-5.5))
            plan = pkg.GenToStdDistPlan(L, grid, x, uplo, n, nb, src)
            ok_sizes = (plan.local_rows, plan.local_cols) == la.shape
            plan.run(da.data_ptr(), lda, df.data_ptr(), ldf, stream)
            torch.cuda.synchronize()
            plan.close()
            out = da.cpu().numpy().reshape(max(1, la.shape[1]), lda)
            got = np.ascontiguousarray(out[:la.shape[1], :la.shape[0]].T)
            w = want[np.ix_(ri, ci)]
            if nan_outside:
                inside = ~np.isnan(w)
                bad = int(np.count_nonzero(~np.isnan(got[~inside]))) + \
                    int(np.count_nonzero(~(np.abs(got[inside] - w[inside]) <= tol)))
            else:
                bad = orc.check_near(got, w, tol, 0.0) if got.size else 0
            pad_ok = bool(np.all(out[:la.shape[1], la.shape[0]:] == NP[x](-7.7)))
            f_ok = np.array_equal(df.cpu().numpy().reshape(hf.shape), hf, equal_nan=True)
            return (bad > 0 or not pad_ok or not f_ok or not ok_sizes,
                    f"{what} bad={bad} pad_ok={pad_ok} factor_untouched={f_ok} sizes_ok={ok_sizes}")

        for x in TYPES:
            for uplo in "LU":
                for m, mb in orc.GEN_TO_STD_TEST_SIZES:
                    t, a, res = orc.gen_to_std_setters(m, uplo, NP[x])
                    tol = 10 * (m + 1) * orc.type_error(NP[x])
                    fail, msg = run_case(x, uplo, a, t, res, mb, tol, False,
                                         f"analytic {x} {uplo} m={m} mb={mb} grid={P}x{Q}")
                    ncases += 1
                    if fail:
                        failures.append(msg)
        rng = np.random.default_rng(29)
        for x, n, nb in ([("d", 96, 8), ("d", 69, 8), ("z", 40, 4)] if SMALL else [("d", 1536, 128), ("d", 1100, 128), ("z", 640, 64)]):
            dt = NP[x]
            cplx = x in "cz"
            g = rng.standard_normal((n, n)) + (1j * rng.standard_normal((n, n)) if cplx else 0)
            a_full = (g + g.conj().T).astype(dt)
            a_full[np.diag_indices(n)] = a_full.diagonal().real
            h = rng.standard_normal((n, n)) + (1j * rng.standard_normal((n, n)) if cplx else 0)
            lfac = np.linalg.cholesky((h @ h.conj().T / n + 2 * np.eye(n)).astype(dt))
            for uplo in "LU":
                keep = (np.tril if uplo == "L" else np.triu)(np.ones((n, n))) > 0
                fac = lfac if uplo == "L" else lfac.conj().T
                want = orc.gen_to_std_local(uplo, np.asfortranarray(np.where(keep, a_full, 0).astype(dt)),
                                            np.asfortranarray(np.where(keep, fac, 0).astype(dt)), nb)
                want = np.where(keep, want, np.nan)
                tol = 1e-11 * n * np.nanmax(np.abs(want))
                fail, msg = run_case(x, uplo, np.where(keep, a_full, np.nan).astype(dt),
                                     np.where(keep, fac, np.nan).astype(dt), want, nb, tol, True,
                                     f"random {x} {uplo} n={n} nb={nb} grid={P}x{Q}")
                ncases += 1
                if fail:
                    failures.append(msg)
        grid.close()

    flag = torch.tensor([len(failures)], device="cuda")
    dist.all_reduce(flag)
    if failures:
        print(f"[rank {rank}] {len(failures)} of {ncases} FAILED:\n  " + "\n  ".join(failures[:20]), flush=True)
    if rank == 0:
        print(f"dist_g2s_check world={world} grids={grids} cases={ncases} total_failures={int(flag.item())}", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(1 if int(flag.item()) else 0)


if __name__ == "__main__":
    main()
