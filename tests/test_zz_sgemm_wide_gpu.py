"""-m gpu: the wide float contraction kernel (csrc/gemm_simt_wide.cuh, DLAFB_SGEMM_WIDE=1) against the default 64 x 64
kernel on the device. The switch is read once per process, so both variants run in child processes
(tests/wide_sgemm_probe.py) on identical seeded inputs: contractions through dlafb_gemm_nt_s (RECT mode: clipping, lower
mask, odd leading dimensions) must agree bit for bit - both kernels form every element as the same fmaf chain over k -
and a float factorization (LIST mode with diagonal tiles and tiled panel operands) must agree with the oracle within the
usual tolerance. CPU execution of the same kernel body: tests/test_sgemm_wide_emu.py.
(The file name sorts last on purpose: written after the round's GPU budget was spent.)"""
import os
import subprocess
import sys

import numpy as np
import pytest

import dlaf_oracle as orc

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def probe(tmp_path, wide):
    out = str(tmp_path / f"wide{wide}.npz")
    env = dict(os.environ)
    env["DLAFB_SGEMM_WIDE"] = wide
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "wide_sgemm_probe.py"), out], cwd=ROOT, env=env,
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    return np.load(out)


def test_wide_kernel_matches_default_kernel(tmp_path):
    base, wide = probe(tmp_path, "0"), probe(tmp_path, "1")
    gemms = sorted(k for k in base.files if k.startswith("gemm") and not k.endswith("_pad_ok"))
    assert len(gemms) >= 7
    for k in gemms:
        assert bool(wide[k + "_pad_ok"]) and bool(base[k + "_pad_ok"]), k
        assert np.array_equal(wide[k], base[k]), (k, float(np.abs(wide[k] - base[k]).max()))
    assert int(base["potrf_info"]) == 0 and int(wide["potrf_info"]) == 0
    n, nb = 1536, 256
    ref = orc.gen_hpd(n, nb, np.float32).copy(order="F")
    assert orc.cholesky_local("L", ref, nb) == 0
    tol = 50 * n * float(np.finfo(np.float32).eps) * np.sqrt(2.0)
    low = np.tril(np.ones((n, n), bool))
    for name, res in (("default", base["potrf"]), ("wide", wide["potrf"])):
        assert np.abs(res[low] - ref[low]).max() <= tol * max(1.0, float(np.abs(ref[low]).max())), name
