"""The pending device tests (tests/test_zz_*.py) must at least be right about their own data preparation, argument order
and tolerances: tests/dryrun_zz_with_emu.py runs their test functions with the CPU emulation / numpy / the oracle standing
in for the device library. In a child process: the stand-in replaces the `gpu_util` module."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_pending_device_tests_pass_against_the_stand_ins():
    so = os.path.join(ROOT, "tests", "emu", "_build", "libsolver_emu.so")
    if not os.path.exists(so):
        subprocess.check_call(["make", "emu"], cwd=ROOT)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "dryrun_zz_with_emu.py")], cwd=ROOT,
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    assert "dry run ok" in r.stdout


def test_device_only_scripts_have_no_unbound_names():
    """tests/check_names.py: every name loaded in a script that only runs on a GPU box is bound somewhere in it."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "check_names.py")], cwd=ROOT, capture_output=True,
                       text=True, timeout=120)
    assert r.returncode == 0, r.stdout[-3000:]


import pytest  # noqa: E402


@pytest.mark.parametrize("script", ["dist_trsolve_check", "dist_g2s_check", "dist_redistribute_check", "dist_submatrix_check"])
def test_multi_gpu_scripts_run_over_gloo_with_the_emulation(script):
    """tests/dryrun_dist_scripts.py: the multi-GPU parity scripts as real two-process jobs on the CPU - gloo instead of
    NCCL, the product's per-rank templates executed by the emulation - so that their scatter / compare / bookkeeping logic
    is exercised before it costs a multi-GPU session."""
    import socket
    so = os.path.join(ROOT, "tests", "emu", "_build", "libsolver_emu.so")
    if not os.path.exists(so):
        subprocess.check_call(["make", "emu"], cwd=ROOT)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    env = dict(os.environ, DLAFB_CHECK_SMALL="1", DLAFB_DIST_QUICK="1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
                        "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "dryrun_dist_scripts.py"),
                        script], cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)
    tail = (r.stdout + r.stderr)[-3000:]
    assert r.returncode == 0, tail
    assert "total_failures=0" in r.stdout, tail
