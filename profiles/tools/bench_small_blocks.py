"""dlaf_pdpotrf for a caller with ScaLAPACK-sized blocks (nb = 64, 128): plain path against DLAFB_POTRF_REDISTRIBUTE=1
(redistribute to 512 on the device - factorize - redistribute back), host pointers, one GPU; and uplo = U against L at
nb = 512.   python profiles/tools/bench_small_blocks.py [n]"""
import ctypes
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    pkg = bench.load_package()
    L = pkg.lib()
    torch.cuda.set_device(0)
    L.dlaf_initialize(0, None, 0, None)
    pkg.check(L.dlafb_mpi_init(0, 1, None), "mpi_init")
    ctx = 61
    L.dlaf_create_grid_from_blacs_info(ctx, L.dlafb_mpi_comm_world(), 1, 1, b"R")
    stream = torch.cuda.current_stream().cuda_stream
    dev = torch.empty(n * n, dtype=torch.float64, device="cuda")
    host = torch.empty(n * n, dtype=torch.float64).pin_memory()
    fl = bench.flops_cholesky(n)

    def timed(uplo, nb, flag, device_ptr=False):
        os.environ["DLAFB_POTRF_REDISTRIBUTE"] = flag
        desca = pkg.DESC9(1, ctx, n, n, nb, nb, 0, 0, n)
        ts = []
        info = ctypes.c_int(0)
        for i in range(3):
            pkg.check(L.dlafb_gen_hpd_d(stream, n, nb, n, n, 0, 1, 0, 1, dev.data_ptr(), n, 9 + i), "gen")
            if not device_ptr:
                host.copy_(dev)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            L.dlaf_pdpotrf(uplo, n, dev.data_ptr() if device_ptr else host.data_ptr(), 1, 1, desca, ctypes.byref(info))
            torch.cuda.synchronize()
            ts.append((time.perf_counter() - t0) * 1e3)
            assert info.value == 0
        ms = min(ts[1:])
        return {"uplo": uplo.decode(), "nb": nb, "redistribute": flag, "pointer": "device" if device_ptr else "host",
                "ms": round(ms, 2), "gflops": round(fl / ms / 1e6, 1)}
    for nb in (64, 128):
        for flag in ("0", "1"):
            print(json.dumps(timed(b"L", nb, flag)), flush=True)
            print(json.dumps(timed(b"L", nb, flag, device_ptr=True)), flush=True)
    for uplo in (b"L", b"U"):
        print(json.dumps(timed(uplo, 512, "0")), flush=True)
        print(json.dumps(timed(uplo, 512, "0", device_ptr=True)), flush=True)
    L.dlaf_free_grid(ctx)
    L.dlaf_finalize()


if __name__ == "__main__":
    main()
