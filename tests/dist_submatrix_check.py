"""Multi-GPU check of the ScaLAPACK sub-matrix arguments of dlaf_p?potrf (ia, ja > 1, tile aligned), one process per
GPU under torchrun (started by tests/test_zz_submatrix_gpu.py). Every rank builds the same global parent matrix on the
host, with a Hermitian positive definite block at tiles (i0, j0) and poison around it, calls dlaf_p?potrf on its local
block of the WHOLE matrix with ia = i0*nb + 1, ja = j0*nb + 1 and compares its block: the factor inside sub(A)'s
referenced triangle (against the oracle), everything else untouched. Index algebra on the CPU: tests/test_submatrix_algebra.py."""
import ctypes
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
import dlaf_oracle as orc  # noqa: E402

NP = {"s": np.float32, "d": np.float64, "c": np.complex64, "z": np.complex128}
SMALL = os.environ.get("DLAFB_CHECK_SMALL") == "1"   # CPU dry run (tests/dryrun_dist_scripts.py): same code, small sizes


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = load_package()
    L = pkg.lib()
    failures, ncases = [], 0
    grids = [(p, world // p) for p in range(1, world + 1) if world % p == 0]
    L.dlaf_initialize(0, None, 0, None)
    assert L.dlafb_mpi_init(rank, world, ctypes.create_string_buffer(new_uid(L, rank), 128)) == 0
    for gi, (P, Q) in enumerate(grids):
        ctx = 300 + gi
        L.dlaf_create_grid_from_blacs_info(ctx, L.dlafb_mpi_comm_world(), P, Q, b"R")
        me = (rank // Q, rank % Q)
        src = (max(0, P - 1), min(1, Q - 1))
        cases = [("d", 1536, 128, 3, 3, 1152), ("d", 1536, 128, 2, 5, 700), ("z", 640, 64, 1, 4, 384),
                 ("s", 100, 8, 5, 2, 37), ("d", 64, 16, 0, 3, 16)]
        if SMALL:
            cases = [("d", 192, 16, 3, 3, 144), ("d", 192, 16, 2, 5, 87), ("z", 80, 8, 1, 4, 48), ("s", 100, 8, 5, 2, 37),
                     ("d", 64, 16, 0, 3, 16)]
        for x, m, nb, i0, j0, n in cases:
            for uplo in "LU":
                dt = NP[x]
                parent = np.full((m, m), dt(-9.9))
                a = orc.gen_hpd(n, nb, dt)
                keep = (np.tril if uplo == "L" else np.triu)(np.ones((n, n))) > 0
                parent[i0 * nb:i0 * nb + n, j0 * nb:j0 * nb + n] = np.where(keep, a, dt(-9.9))
                ref = a.copy(order="F")
                assert orc.cholesky_local(uplo, ref, nb) == 0
                expect = parent.copy()
                expect[i0 * nb:i0 * nb + n, j0 * nb:j0 * nb + n] = np.where(keep, ref, dt(-9.9))
                loc, ri, ci = orc.local_block(parent, nb, nb, (P, Q), me, src)
                ld = max(1, loc.shape[0]) + 2
                host = np.full((max(1, loc.shape[1]), ld), dt(-7.7))
                host[:loc.shape[1], :loc.shape[0]] = loc.T
                desca = pkg.DESC9(1, ctx, m, m, nb, nb, src[0], src[1], ld)
                info = ctypes.c_int(-1)
                getattr(L, f"dlaf_p{x}potrf")(uplo.encode(), n, host.ctypes.data, i0 * nb + 1, j0 * nb + 1, desca,
                                              ctypes.byref(info))
                got = np.ascontiguousarray(host[:loc.shape[1], :loc.shape[0]].T)
                want = expect[np.ix_(ri, ci)]
                tol = 50 * n * float(np.finfo(dt).eps) * np.sqrt(2.0)
                bad = orc.check_near(got, want, tol, tol) if got.size else 0
                pad_ok = bool(np.all(host[:loc.shape[1], loc.shape[0]:] == dt(-7.7)))
                ncases += 1
                if info.value != 0 or bad or not pad_ok:
                    failures.append(f"{x} {uplo} m={m} nb={nb} sub=({i0},{j0}) n={n} grid={P}x{Q} info={info.value} "
                                    f"bad={bad} pad_ok={pad_ok}")
        # ScaLAPACK-sized blocks through dlaf_pdpotrf with the internal redistribution (DLAFB_POTRF_REDISTRIBUTE=1, read
        # at every call): redistribute -> Cholesky at 512 -> redistribute back, distributed
        for flag in ("1", "0"):
            os.environ["DLAFB_POTRF_REDISTRIBUTE"] = flag
            n, nb = (2304, 64) if not SMALL else (288, 8)
            for uplo in "LU":
                a = orc.gen_hpd(n, nb, np.float64)
                ref = a.copy(order="F")
                assert orc.cholesky_local(uplo, ref, nb) == 0
                keep = (np.tril if uplo == "L" else np.triu)(np.ones((n, n))) > 0
                full_in = np.where(keep, a, -9.9)
                expect = np.where(keep, ref, -9.9)
                loc, ri, ci = orc.local_block(full_in, nb, nb, (P, Q), me, src)
                ld = max(1, loc.shape[0]) + 1
                host = np.full((max(1, loc.shape[1]), ld), -7.7)
                host[:loc.shape[1], :loc.shape[0]] = loc.T
                desca = pkg.DESC9(1, ctx, n, n, nb, nb, src[0], src[1], ld)
                info = ctypes.c_int(-1)
                L.dlaf_pdpotrf(uplo.encode(), n, host.ctypes.data, 1, 1, desca, ctypes.byref(info))
                got = np.ascontiguousarray(host[:loc.shape[1], :loc.shape[0]].T)
                tol = 50 * n * float(np.finfo(np.float64).eps) * np.sqrt(2.0) * float(np.abs(ref).max())
                bad = int(np.count_nonzero(~(np.abs(got - expect[np.ix_(ri, ci)]) <= tol))) if got.size else 0
                pad_ok = bool(np.all(host[:loc.shape[1], loc.shape[0]:] == -7.7))
                ncases += 1
                if info.value != 0 or bad or not pad_ok:
                    failures.append(f"pdpotrf nb=64 redistribute={flag} {uplo} grid={P}x{Q} info={info.value} bad={bad} "
                                    f"pad_ok={pad_ok}")
        os.environ["DLAFB_POTRF_REDISTRIBUTE"] = "0"
        L.dlaf_free_grid(ctx)
    L.dlaf_finalize()
    flag = torch.tensor([len(failures)], device="cuda")
    dist.all_reduce(flag)
    if failures:
        print(f"[rank {rank}] {len(failures)} of {ncases} FAILED:\n  " + "\n  ".join(failures[:20]), flush=True)
    if rank == 0:
        print(f"dist_submatrix_check world={world} grids={grids} cases={ncases} total_failures={int(flag.item())}",
              flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(1 if int(flag.item()) else 0)


if __name__ == "__main__":
    main()
