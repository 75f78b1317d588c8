"""The C++ API classes matrix::LayoutInfo and matrix::Distribution (include/dlaf/matrix/) against the reference's own
known-answer tables (fixtures tests/golden/layout_info.json, distribution_constructor.json; extracted from
test/unit/matrix/test_layout_info.cpp:23-34, :91-99 and test/unit/matrix/test_distribution.cpp:40-104 by
tests/golden/extract_reference_tables.py). tests/cxx/check_matrix_api.cpp is compiled with the host compiler, linked
with the product library (Distribution's maps are the engine's exported index functions) and fed the rows; host logic
only, no GPU."""
import json
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def test_layout_and_distribution_tables(pkg, tmp_path):
    cuda_inc = "/usr/local/cuda/include"
    if shutil.which("g++") is None or not os.path.exists(os.path.join(cuda_inc, "cuda_runtime.h")):
        pytest.skip("needs g++ and the CUDA headers")
    libdir = os.path.dirname(pkg.LIB_PATH)
    if not os.path.exists(pkg.LIB_PATH):
        pytest.skip("library not built")
    exe = str(tmp_path / "check_matrix_api")
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "include", "dlafb_mpi_shim"), "-I" + cuda_inc,
           os.path.join(ROOT, "tests", "cxx", "check_matrix_api.cpp"), "-o", exe, "-L" + libdir, "-ldlaf_b200",
           "-Wl,-rpath," + libdir, "-L/usr/local/cuda/lib64", "-Wl,-rpath,/usr/local/cuda/lib64",
           "-Wl,--allow-shlib-undefined"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    lines = []
    lay = json.load(open(os.path.join(GOLDEN, "layout_info.json")))
    lines += ["L " + " ".join(map(str, row)) for row in lay["rows"]]
    lines += ["C " + " ".join(map(str, row)) for row in lay["comparison_rows"]]
    dist = json.load(open(os.path.join(GOLDEN, "distribution_constructor.json")))
    used = 0
    for row in dist["rows"]:
        t = dict(zip(dist["keys"], row))
        if t["block_size"] != t["tile_size"]:
            continue   # more than one tile per block: outside the path (factorization/cholesky.h:43,71)
        if t["offset"][0] % t["block_size"][0] or t["offset"][1] % t["block_size"][1]:
            continue   # unaligned offset: not supported
        src = [(t["src_rank"][a] + t["offset"][a] // t["block_size"][a]) % t["grid_size"][a] for a in (0, 1)]
        lines.append("D " + " ".join(map(str, t["size"] + t["block_size"] + t["rank"] + t["grid_size"] + src +
                                         t["global_tiles"] + t["local_tiles"] + t["local_size"])))
        used += 1
    assert used >= 16
    env = dict(os.environ)
    # libdlaf_b200.so needs libnccl.so.2: the copy bundled with torch, else the system one
    import glob
    import sys
    extra = [os.path.dirname(p) for base in sys.path for p in glob.glob(os.path.join(base, "nvidia", "nccl", "lib", "libnccl.so.2"))]
    env["LD_LIBRARY_PATH"] = ":".join(extra + [env.get("LD_LIBRARY_PATH", "")])
    r = subprocess.run([exe], input="\n".join(lines) + "\n", capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "failures=0" in r.stdout and "checks=0" not in r.stdout, r.stdout
