"""Hunts the order-dependent failure of tests/test_zz_redistribute_gpu.py::test_pxpotrf_with_redistribution_to_large_blocks:
the same calls with freshly NaN-poisoned device memory in front of every library allocation, and a per-tile map of where
the result is wrong.   python profiles/tools/flaky_probe.py [reps]"""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import bench  # noqa: E402
import dlaf_oracle as orc  # noqa: E402

NP = {"s": np.float32, "d": np.float64, "c": np.complex64, "z": np.complex128}


def poison(gb=24):
    bufs = []
    for _ in range(gb):
        t = torch.empty(1 << 27, dtype=torch.float64, device="cuda")
        t.fill_(float("nan"))
        bufs.append(t)
    torch.cuda.synchronize()
    del bufs
    torch.cuda.empty_cache()


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    pkg = bench.load_package()
    L = pkg.lib()
    torch.cuda.set_device(0)
    ctx = 47
    cases = [("z", 2048, 32, "L"), ("d", 2100, 48, "U"), ("d", 2304, 64, "L"), ("s", 2048, 128, "U")]
    refs = {}
    for x, n, nb, uplo in cases:
        a = orc.gen_hpd(n, nb, NP[x])
        ref = a.copy(order="F")
        assert orc.cholesky_local(uplo, ref, nb) == 0
        refs[(x, n, nb, uplo)] = (a, ref)
    nfail = 0
    for rep in range(reps):
        for flag in ("1", "0"):
            for (x, n, nb, uplo) in cases:
                dt = NP[x]
                a, ref = refs[(x, n, nb, uplo)]
                poison()
                L.dlaf_initialize(0, None, 0, None)
                pkg.check(L.dlafb_mpi_init(0, 1, None), "mpi_init")
                L.dlaf_create_grid_from_blacs_info(ctx, L.dlafb_mpi_comm_world(), 1, 1, b"R")
                os.environ["DLAFB_POTRF_REDISTRIBUTE"] = flag
                keep = (np.tril if uplo == "L" else np.triu)(np.ones((n, n))) > 0
                a_in = np.where(keep, a, dt(-9.9))
                ld = n + 5
                h = np.full((n, ld), dt(-7.7))
                h[:, :n] = a_in.T
                desca = pkg.DESC9(1, ctx, n, n, nb, nb, 0, 0, ld)
                info = ctypes.c_int(-1)
                getattr(L, f"dlaf_p{x}potrf")(uplo.encode(), n, h.ctypes.data, 1, 1, desca, ctypes.byref(info))
                got = np.ascontiguousarray(h[:, :n].T)
                err = np.abs(np.where(keep, got - ref, 0))
                err = np.where(np.isnan(err), np.inf, err)
                tol = 50 * n * float(np.finfo(dt).eps) * np.sqrt(2.0) * max(1.0, np.abs(ref[keep]).max())
                bad = err.max() > tol or info.value != 0
                line = f"rep={rep} flag={flag} {x} n={n} nb={nb} {uplo} info={info.value} maxerr={err.max():.3e} tol={tol:.1e} {'FAIL' if bad else 'ok'}"
                if bad:
                    nfail += 1
                    t = 256
                    nt = -(-n // t)
                    m = [[err[i * t:(i + 1) * t, j * t:(j + 1) * t].max() > tol for j in range(nt)] for i in range(nt)]
                    line += "\n" + "\n".join("".join("X" if v else "." for v in row) for row in m)
                    line += f"\n nan_count={int(np.isnan(got).sum())}"
                print(line, flush=True)
                L.dlaf_free_grid(ctx)
                L.dlaf_finalize()
    print(f"flaky_probe failures={nfail}")


if __name__ == "__main__":
    main()
