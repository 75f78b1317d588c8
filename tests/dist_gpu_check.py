"""Multi-GPU parity check of the distributed factorization, one process per GPU (run under torchrun; started by
tests/test_cholesky_dist_gpu.py or by hand):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
      tests/dist_gpu_check.py

Cases follow test/unit/factorization/test_cholesky.cpp:54-139 (distributed part): the reference sizes with the
analytical known answer and a non-zero source rank, on every P x Q factorisation of the world size, both uplo,
four element types, through both the engine ABI (device pointers) and the C API (dlaf_p?potrf, host pointers);
plus larger seeded random HPD matrices on the TMA + DMMA path checked against the oracle factor.
Every rank builds the same global matrix on the host, factorizes its own block on its GPU and compares its block.
"""
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
import dlaf_oracle as orc  # noqa: E402

NP = {"s": np.float32, "d": np.float64, "c": np.complex64, "z": np.complex128}
REF_SIZES = [(0, 2), (5, 8), (34, 34), (4, 3), (16, 10), (34, 13), (32, 5)]
QUICK = os.environ.get("DLAFB_DIST_QUICK") == "1"   # fewer types / sizes: for large world sizes
TYPES = "dz" if QUICK else "sdcz"
SIZES = [(0, 2), (34, 13), (32, 5)] if QUICK else REF_SIZES


def new_uid(L, rank):
    buf = ctypes.create_string_buffer(128)
    if rank == 0:
        assert L.dlafb_nccl_unique_id(buf) == 0
    t = torch.tensor(list(buf.raw), dtype=torch.uint8, device="cuda")
    dist.broadcast(t, 0)
    return bytes(t.cpu().tolist())


def local_index(L, n, nb, P, rank, src):
    tiles = [L.dlafb_dist_global_tile_from_local(l, P, rank, src)
             for l in range(L.dlafb_dist_local_nr_tiles(n, nb, P, rank, src))]
    idx = [np.arange(g * nb, min(n, (g + 1) * nb)) for g in tiles] + [np.arange(0)]
    return np.concatenate(idx).astype(int)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = load_package()
    L = pkg.lib()
    stream = torch.cuda.current_stream().cuda_stream
    failures, ncases = [], 0
    grids = [(p, world // p) for p in range(1, world + 1) if world % p == 0]
    L.dlaf_initialize(0, None, 0, None)
    assert L.dlafb_mpi_init(rank, world, ctypes.create_string_buffer(new_uid(L, rank), 128)) == 0

    for gi, (P, Q) in enumerate(grids):
        for order in ("R", "C"):
            grid = pkg.Grid(L, P, Q, rank, world, new_uid(L, rank), order)
            ctx = 100 + 2 * gi + (order == "C")
            L.dlaf_create_grid_from_blacs_info(ctx, L.dlafb_mpi_comm_world(), P, Q, order.encode())
            src = (max(0, P - 1), min(1, Q - 1))

            def run_case(x, uplo, n, nb, a_glob, expect, tol, what, use_c_api):
                rows = local_index(L, n, nb, P, grid.pr, src[0])
                cols = local_index(L, n, nb, Q, grid.pc, src[1])
                loc = np.asfortranarray(a_glob[np.ix_(rows, cols)])
                ld = max(1, len(rows)) + 3
                host = np.full((max(1, len(cols)), ld), -7.7, dtype=a_glob.dtype)
                host[:len(cols), :len(rows)] = loc.T
                if use_c_api:
                    desca = pkg.DESC9(1, ctx, n, n, nb, nb, src[0], src[1], ld)
                    info = ctypes.c_int(-1)
                    getattr(L, f"dlaf_p{x}potrf")(uplo.encode(), n, host.ctypes.data, 1, 1, desca, ctypes.byref(info))
                    out, inf = host, info.value
                else:
                    plan = pkg.CholeskyPlan(L, grid, x, uplo, n, nb, src)
                    assert (plan.local_rows, plan.local_cols) == (len(rows), len(cols))
                    dev = torch.from_numpy(host.reshape(-1).copy()).cuda()
                    plan.run(dev.data_ptr(), ld, stream)
                    inf = plan.info()
                    torch.cuda.synchronize()
                    out = dev.cpu().numpy().reshape(host.shape)
                    res_msg = ""
                    if uplo == "L" and n > 0:
                        # the distributed device-side residual check against numpy on the global matrix
                        dev0 = torch.from_numpy(host.reshape(-1).copy()).cuda()
                        amax, dmax = plan.residual(dev0.data_ptr(), ld, dev.data_ptr(), ld, stream)
                        low = np.tril(np.ones((n, n), bool))
                        want_amax = float(np.abs(a_glob[low]).max())
                        eps_ = float(np.finfo(a_glob.dtype).eps)
                        if abs(amax - want_amax) > 1e-6 * want_amax or not (dmax <= 8 * n * eps_ * want_amax):
                            res_msg = f" residual amax={amax} (want {want_amax}) dmax={dmax}"
                        # ... and it must see a wrong factor: perturb one local element of L on the owner of (0, 0)
                        if len(rows) and len(cols) and rows[0] == 0 and cols[0] == 0:
                            bad_l = dev.clone()
                            bad_l[0] += 1.0
                        else:
                            bad_l = dev
                        dev0 = torch.from_numpy(host.reshape(-1).copy()).cuda()
                        _, dbad = plan.residual(dev0.data_ptr(), ld, bad_l.data_ptr(), ld, stream)
                        if not (dbad > 0.5):
                            res_msg += f" residual blind to a perturbed factor (dmax={dbad})"
                    plan.close()
                got = np.ascontiguousarray(out[:len(cols), :len(rows)].T)
                bad = orc.check_near(got, expect[np.ix_(rows, cols)], tol, tol) if got.size else 0
                pad_ok = bool(np.all(out[:len(cols), len(rows):] == a_glob.dtype.type(-7.7)))
                res_msg = locals().get("res_msg", "")
                return ((inf != 0) or bad > 0 or not pad_ok or bool(res_msg),
                        f"{what} info={inf} bad={bad} pad_ok={pad_ok}{res_msg}")

            for x in TYPES:
                for uplo in "LU":
                    for (m, mb) in SIZES:
                        a, t = orc.cholesky_setters(m, uplo, NP[x])
                        tol = 4 * (m + 1) * orc.type_error(NP[x])
                        for capi in (False, True):
                            if capi and (order == "C" and x in "sc"):
                                continue
                            fail, msg = run_case(x, uplo, m, mb, a, t, tol,
                                                 f"analytic {x} {uplo} m={m} mb={mb} grid={P}x{Q}{order} capi={capi}", capi)
                            ncases += 1
                            if fail:
                                failures.append(msg)
            # ScaLAPACK convention for a matrix that is not positive definite: EVERY rank returns the 1-based global column
            # of the first non-positive pivot, although only the owner of that diagonal tile sees it
            for capi in (False, True):
                n, nb = 6 * 32 + 5, 32
                a = orc.gen_hpd(n, nb, np.float64)
                bad_col = 5 * 32 + 3          # in the last full tile: owned by one rank of the grid
                a[bad_col, bad_col] = -1.0
                rows = local_index(L, n, nb, P, grid.pr, src[0])
                cols = local_index(L, n, nb, Q, grid.pc, src[1])
                ld = max(1, len(rows))
                host = np.zeros((max(1, len(cols)), ld))
                host[:len(cols), :len(rows)] = a[np.ix_(rows, cols)].T
                if capi:
                    desca = pkg.DESC9(1, ctx, n, n, nb, nb, src[0], src[1], ld)
                    info = ctypes.c_int(-1)
                    L.dlaf_pdpotrf(b"L", n, host.ctypes.data, 1, 1, desca, ctypes.byref(info))
                    inf = info.value
                else:
                    plan = pkg.CholeskyPlan(L, grid, "d", "L", n, nb, src)
                    dev = torch.from_numpy(host.reshape(-1).copy()).cuda()
                    plan.run(dev.data_ptr(), ld, stream)
                    inf = plan.info()
                    torch.cuda.synchronize()
                    plan.close()
                ncases += 1
                if inf != bad_col + 1:
                    failures.append(f"not-HPD info={inf} (want {bad_col + 1}) grid={P}x{Q}{order} capi={capi}")
            if order == "R":
                big = [("d", 4096, 128), ("d", 1100, 128)] if QUICK else [("d", 2048, 128), ("d", 1664, 256), ("d", 1100, 128), ("z", 640, 64), ("s", 768, 128)]
                for x, n, nb in big:
                    for uplo in "LU":
                        a = orc.gen_hpd(n, nb, NP[x])
                        ref = a.copy(order="F")
                        assert orc.cholesky_local(uplo, ref, nb) == 0
                        eps = np.finfo(NP[x]).eps
                        tol = 50 * n * eps * np.sqrt(2.0)
                        fail, msg = run_case(x, uplo, n, nb, a, ref, tol,
                                             f"random {x} {uplo} n={n} nb={nb} grid={P}x{Q}", False)
                        ncases += 1
                        if fail:
                            failures.append(msg)
                # host pointers through dlaf_pdpotrf with >= 16 tile columns: the upload of the local block is
                # overlapped with the first steps (column groups on their own stream, split bulk updates)
                for n, nb in ([(4608, 128)] if QUICK else [(2304, 128), (4608, 128), (2200, 64)]):
                    a = orc.gen_hpd(n, nb, np.float64)
                    ref = a.copy(order="F")
                    assert orc.cholesky_local("L", ref, nb) == 0
                    tol = 50 * n * np.finfo(np.float64).eps * np.sqrt(2.0)
                    fail, msg = run_case("d", "L", n, nb, a, ref, tol,
                                         f"random host-pointer d L n={n} nb={nb} grid={P}x{Q}", True)
                    ncases += 1
                    if fail:
                        failures.append(msg)
            L.dlaf_free_grid(ctx)
            grid.close()

    L.dlaf_finalize()
    flag = torch.tensor([len(failures)], device="cuda")
    dist.all_reduce(flag)
    if failures:
        print(f"[rank {rank}] {len(failures)} of {ncases} FAILED:\n  " + "\n  ".join(failures[:20]), flush=True)
    if rank == 0:
        print(f"dist_gpu_check world={world} grids={grids} cases={ncases} total_failures={int(flag.item())}", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(1 if int(flag.item()) else 0)


if __name__ == "__main__":
    main()
