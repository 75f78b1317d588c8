"""-m gpu: the C++ front end EXECUTED on the device (round 1 only compiled it).

(i)  tests/cxx/run_cholesky_api.cpp: Matrix<T, CPU> with the reference's analytical A -> MatrixMirror<T, GPU, CPU> ->
     dlaf::cholesky_factorization<Backend::GPU, Device::GPU, T> (local and CommunicatorGrid overloads, device matrices,
     the pre-0.2.0 spelling) -> compared with the analytical factor at 4(m+1)*err for the sizes of
     test/unit/factorization/test_cholesky.cpp:54-62, four types, both uplo.
(ii) dla-future_b200/bin/miniapp_cholesky (the reference's benchmark driver, miniapp/miniapp_cholesky.cpp:107-198) run
     with the reference's flags; its result line (miniapp_cholesky.cpp:166-172: "[i] <t>s <g>GFlop/s <type><uplo>
     (n, n) (nb, nb) (P, Q) <threads> <backend>") and residual line are parsed.
"""
import glob
import os
import re
import shutil
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _env():
    env = dict(os.environ)
    extra = [os.path.dirname(p) for base in sys.path
             for p in glob.glob(os.path.join(base, "nvidia", "nccl", "lib", "libnccl.so.2"))]
    env["LD_LIBRARY_PATH"] = ":".join(extra + ["/usr/local/cuda/lib64", env.get("LD_LIBRARY_PATH", "")])
    return env


def test_cxx_api_known_answers_on_the_device(pkg, tmp_path):
    cuda_inc = "/usr/local/cuda/include"
    if shutil.which("g++") is None or not os.path.exists(os.path.join(cuda_inc, "cuda_runtime.h")):
        pytest.skip("needs g++ and the CUDA headers")
    libdir = os.path.dirname(pkg.LIB_PATH)
    exe = str(tmp_path / "run_cholesky_api")
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "include", "dlafb_mpi_shim"), "-I" + cuda_inc,
           os.path.join(ROOT, "tests", "cxx", "run_cholesky_api.cpp"), "-o", exe, "-L" + libdir, "-ldlaf_b200",
           "-Wl,-rpath," + libdir, "-L/usr/local/cuda/lib64", "-lcudart", "-Wl,-rpath,/usr/local/cuda/lib64",
           "-Wl,--allow-shlib-undefined"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([exe], capture_output=True, text=True, env=_env(), timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    m = re.search(r"checks=(\d+) failures=(\d+)", r.stdout)
    assert m and int(m.group(1)) >= 4 * 2 * (14 + 6) and int(m.group(2)) == 0, r.stdout


LINE = re.compile(r"^\[(\d+)\] ([0-9.eE+-]+)s ([0-9.eE+-]+)GFlop/s ([SDCZ])([LU]) \((\d+), (\d+)\) \((\d+), (\d+)\) "
                  r"\((\d+), (\d+)\) (\d+) (\w+)$")


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("uplo", ["lower", "upper"])
def test_miniapp_cholesky_runs_and_checks(x, uplo):
    exe = os.path.join(ROOT, "dla-future_b200", "bin", "miniapp_cholesky")
    if not os.path.exists(exe):
        pytest.skip("miniapp not built (make miniapp)")
    cmd = [exe, "--matrix-size", "2048", "--block-size", "256", "--type", x, "--uplo", uplo, "--check-result", "last",
           "--nruns", "2", "--nwarmups", "1", "--local"]
    r = subprocess.run(cmd, capture_output=True, text=True, env=_env(), timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    lines = [LINE.match(l) for l in r.stdout.splitlines()]
    runs = [m for m in lines if m]
    assert [int(m.group(1)) for m in runs] == [0, 1], r.stdout
    for m in runs:
        assert float(m.group(2)) > 0 and float(m.group(3)) > 0
        assert m.group(4) == x.upper() and m.group(5) == uplo[0].upper()
        assert m.group(6) == m.group(7) == "2048" and m.group(8) == m.group(9) == "256"
        assert (m.group(10), m.group(11)) == ("1", "1") and m.group(13) == "GPU"
    res = [l for l in r.stdout.splitlines() if "Max Diff / Max A:" in l]
    assert len(res) == 1 and not res[0].startswith(("ERROR", "Warning")), r.stdout
    assert float(res[0].split(":")[1]) >= 0


def test_miniapp_cholesky_grid_overload_single_rank():
    """The CommunicatorGrid overload (1 x 1 grid) and the CSV line (miniapp_cholesky.cpp:173-190)."""
    exe = os.path.join(ROOT, "dla-future_b200", "bin", "miniapp_cholesky")
    if not os.path.exists(exe):
        pytest.skip("miniapp not built (make miniapp)")
    cmd = [exe, "--matrix-size", "1536", "--block-size", "512", "--type", "d", "--check-result", "all", "--nruns", "1", "--nwarmups", "0",
           "--csv"]
    r = subprocess.run(cmd, capture_output=True, text=True, env=_env(), timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    assert any(LINE.match(l) for l in r.stdout.splitlines()), r.stdout
    assert any(l.startswith("CSVData-2, run, 0, time, ") and "matrixsize, 1536, blocksize, 512" in l
               for l in r.stdout.splitlines()), r.stdout
    assert sum("Max Diff / Max A:" in l and not l.startswith(("ERROR", "Warning")) for l in r.stdout.splitlines()) == 1
