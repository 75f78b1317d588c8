"""Multi-GPU check of the matrix redistribution (include/dlafb_matrix.h), one process per GPU under torchrun (started by
tests/test_zz_redistribute_gpu.py): every rank builds the same global matrix on the host, uploads its block in the
source layout, runs the collective and compares its block in the destination layout with the oracle's owner lists -
bit exact - for every P x Q factorisation of the world size, both rank orderings, four element types, block sizes in
both directions (small -> large and large -> small), different source ranks; then factorizes a matrix that arrives
with ScaLAPACK-sized blocks (64) by redistribute -> Cholesky at 512 -> redistribute back and checks it against the oracle.
The same list builder and kernel body run on the CPU in tests/test_redistribute_emu.py."""
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
SMALL = os.environ.get("DLAFB_CHECK_SMALL") == "1"   # CPU dry run (tests/dryrun_dist_scripts.py): same code, small sizes
CASES = [(10, 10, (3, 3), (3, 3)), (10, 5, (5, 10), (10, 2)), (26, 13, (10, 3), (5, 6)),  # test_matrix.cpp:1676-1680
         (37, 29, (4, 3), (8, 6)), (37, 29, (8, 6), (4, 3)), (512, 384, (32, 32), (128, 128)), (1000 if not SMALL else 100, 1000 if not SMALL else 100, (64, 64), (256, 256)),
         (50, 41, (5, 41), (10, 1)), (33, 70, (3, 10), (3, 10)), (1, 1, (2, 2), (4, 4)), (0, 5, (2, 2), (4, 4))]


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = load_package()
    L = pkg.lib()
    stream = torch.cuda.current_stream().cuda_stream
    failures, ncases = [], 0
    grids = [(p, world // p) for p in range(1, world + 1) if world % p == 0]
    rng = np.random.default_rng(41)

    def redistribute(grid, me, x, a, st, sr, dtile, dr, P, Q):
        ls, _, _ = orc.local_block(a, st[0], st[1], (P, Q), me, sr)
        want, _, _ = orc.local_block(a, dtile[0], dtile[1], (P, Q), me, dr)
        ds, lds, _ = to_dev(ls, 1, NP[x](-7.7))
        dd, ldd, _ = to_dev(np.full(want.shape, NP[x](-5.5)), 2, NP[x](-5.5))
        plan = pkg.RedistributePlan(L, grid, x, a.shape[0], a.shape[1], st, sr, dtile, dr)
        ok_sizes = (plan.src_rows, plan.src_cols, plan.dst_rows, plan.dst_cols) == ls.shape + want.shape
        plan.run(ds.data_ptr(), lds, dd.data_ptr(), ldd, stream)
        torch.cuda.synchronize()
        plan.close()
        out = dd.cpu().numpy().reshape(max(1, want.shape[1]), ldd)
        got = np.ascontiguousarray(out[:want.shape[1], :want.shape[0]].T)
        pad_ok = bool(np.all(out[:want.shape[1], want.shape[0]:] == NP[x](-5.5)))
        return got, want, pad_ok and ok_sizes

    for P, Q in grids:
        for order in "RC":
            grid = pkg.Grid(L, P, Q, rank, world, new_uid(L, rank), order)
            me = (grid.pr, grid.pc)
            for x in "sdcz":
                for m, n, st, dtile in CASES:
                    a = rng.standard_normal((m, n)) + (1j * rng.standard_normal((m, n)) if x in "cz" else 0)
                    a = a.astype(NP[x])
                    for sr, dr in [((0, 0), (0, 0)), ((P - 1, min(1, Q - 1)), (P // 2, Q - 1))]:
                        got, want, ok = redistribute(grid, me, x, a, st, sr, dtile, dr, P, Q)
                        ncases += 1
                        if not ok or not np.array_equal(got, want):
                            failures.append(f"{x} {m}x{n} {st}->{dtile} src={sr} dst={dr} grid={P}x{Q}{order} ok={ok}")
            # distributed max norm (dlafb_max_norm): Lower ignores what lies above the global diagonal
            for x, m, nb in [("d", 301 if not SMALL else 45, 16), ("z", 77, 8), ("s", 5, 8)]:
                a = rng.uniform(-1, 1, (m, m)) + (1j * rng.uniform(-1, 1, (m, m)) if x in "cz" else 0)
                a = a.astype(NP[x])
                a[1, m - 1] = 40.0
                a[m - 1, 0] = -30.0
                src = (P - 1, min(1, Q - 1))
                la, _, _ = orc.local_block(a, nb, nb, (P, Q), me, src)
                da, lda, _ = to_dev(la, 1, NP[x](-7.7))
                gl = pkg.max_norm(L, grid, x, "L", m, m, nb, nb, src, da.data_ptr(), lda, stream)
                gg = pkg.max_norm(L, grid, x, "G", m, m, nb, nb, src, da.data_ptr(), lda, stream)
                ncases += 1
                if abs(gl - 30.0) > 1e-5 or abs(gg - 40.0) > 1e-5:
                    failures.append(f"max_norm {x} m={m} nb={nb} grid={P}x{Q}{order}: L={gl} G={gg}")
            if order == "R" and not SMALL:   # (the dry run has no factorization engine to stand in)
                # ScaLAPACK-sized blocks in, factorization at 512, ScaLAPACK-sized blocks out
                n, nb_in, nb = 2048, 64, 512
                a = orc.gen_hpd(n, nb, np.float64)
                ref = a.copy(order="F")
                assert orc.cholesky_local("L", ref, nb) == 0
                ls, ri, ci = orc.local_block(a, nb_in, nb_in, (P, Q), me, (0, 0))
                d_in, ld_in, _ = to_dev(ls, 0, -7.7)
                fwd = pkg.RedistributePlan(L, grid, "d", n, n, (nb_in, nb_in), (0, 0), (nb, nb), (0, 0))
                bwd = pkg.RedistributePlan(L, grid, "d", n, n, (nb, nb), (0, 0), (nb_in, nb_in), (0, 0))
                chol = pkg.CholeskyPlan(L, grid, "d", "L", n, nb)
                ldw = max(1, chol.local_rows)
                work = torch.zeros(ldw * max(1, chol.local_cols) + 64, dtype=torch.float64, device="cuda")
                fwd.run(d_in.data_ptr(), ld_in, work.data_ptr(), ldw, stream)
                chol.run(work.data_ptr(), ldw, stream)
                info = chol.info()
                bwd.run(work.data_ptr(), ldw, d_in.data_ptr(), ld_in, stream)
                torch.cuda.synchronize()
                for pl in (fwd, bwd, chol):
                    pl.close()
                out = d_in.cpu().numpy()[:ld_in * max(1, ls.shape[1])].reshape(max(1, ls.shape[1]), ld_in)
                got = np.ascontiguousarray(out[:ls.shape[1], :ls.shape[0]].T)
                low = (ri[:, None] >= ci[None, :]) if ri.size and ci.size else np.zeros(got.shape, bool)
                err = float(np.abs(got - ref[np.ix_(ri, ci)])[low].max()) if low.any() else 0.0
                ncases += 1
                if info != 0 or not (err <= 50 * n * np.finfo(np.float64).eps * np.sqrt(2.0)):
                    failures.append(f"redistribute-potrf-redistribute grid={P}x{Q} info={info} err={err}")
            grid.close()

    flag = torch.tensor([len(failures)], device="cuda")
    dist.all_reduce(flag)
    if failures:
        print(f"[rank {rank}] {len(failures)} of {ncases} FAILED:\n  " + "\n  ".join(failures[:20]), flush=True)
    if rank == 0:
        print(f"dist_redistribute_check world={world} grids={grids} cases={ncases} total_failures={int(flag.item())}",
              flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(1 if int(flag.item()) else 0)


if __name__ == "__main__":
    main()
