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
