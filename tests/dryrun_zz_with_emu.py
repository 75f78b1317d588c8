"""Developer tool: runs the LOGIC of the not-yet-device-verified GPU tests (tests/test_zz_*.py: data preparation,
argument order, tolerances, comparisons) on a box without a GPU by standing the CPU emulation (tests/emu) in for the
device library: DevMat becomes host memory, the plan classes call the emulation's entry points. Catches mistakes in the
tests themselves before they cost a GPU session. Not collected by pytest (no test_ prefix).
    python tests/dryrun_zz_with_emu.py
"""
import ctypes
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
NP = {"s": np.float32, "d": np.float64, "c": np.complex64, "z": np.complex128}
EMU = ctypes.CDLL(os.path.join(ROOT, "tests", "emu", "_build", "libsolver_emu.so"))


class DevMat:
    def __init__(self, a, ld=None, slack=32):
        a = np.asarray(a)
        self.m, self.n = a.shape
        self.ld = max(1, self.m) if ld is None else ld
        self.host = np.full((max(1, self.n), self.ld), -7.7, dtype=a.dtype)
        if a.size:
            self.host[: self.n, : self.m] = a.T

    @property
    def ptr(self):
        return self.host.ctypes.data

    def get(self):
        return np.ascontiguousarray(self.host[: self.n, : self.m].T)

    def pad(self):
        return self.host[: self.n, self.m:]


gpu_util = types.ModuleType("gpu_util")
gpu_util.NP, gpu_util.DevMat = NP, DevMat
gpu_util.stream_ptr = lambda: None
gpu_util.sync = lambda: None
sys.modules["gpu_util"] = gpu_util


class Grid:
    def __init__(self, L, *a, **k):
        self.P = self.Q = 1
        self.pr = self.pc = 0

    def close(self):
        pass


class GenToStdPlan:
    def __init__(self, L, x, uplo, n, nb):
        self.x, self.uplo, self.n, self.nb = x, uplo, n, nb

    def run(self, a, lda, f, ldf, stream):
        fn = getattr(EMU, f"emu_gen_to_std_{self.x}")
        fn.argtypes = [ctypes.c_char, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p,
                       ctypes.c_longlong]
        assert fn(self.uplo.encode(), self.n, self.nb, a, lda, f, ldf) == 0

    def close(self):
        pass


class GenToStdDistPlan(GenToStdPlan):
    def __init__(self, L, grid, x, uplo, n, nb, src=(0, 0)):
        super().__init__(L, x, uplo, n, nb)
        self.local_rows = self.local_cols = n

    def run(self, a, lda, f, ldf, stream):
        fn = getattr(EMU, f"emu_gen_to_std_dist_{self.x}")
        fn.argtypes = [ctypes.c_int] * 4 + [ctypes.c_char] + [ctypes.c_int] * 2 + [
            ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_longlong),
            ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_longlong)]
        pa, pf = (ctypes.c_void_p * 1)(a), (ctypes.c_void_p * 1)(f)
        la, lf = (ctypes.c_longlong * 1)(lda), (ctypes.c_longlong * 1)(ldf)
        assert fn(1, 1, 0, 0, self.uplo.encode(), self.n, self.nb, pa, la, pf, lf) == 0


class TriangularSolverDistPlan:
    def __init__(self, L, grid, x, side, uplo, op, diag, m, n, mb, nb, src=(0, 0)):
        self.args = (x, side, uplo, op, diag, m, n, mb, nb)
        na = m if side == "L" else n
        self.a_rows = self.a_cols = na
        self.b_rows, self.b_cols = m, n

    def run(self, alpha, a, lda, b, ldb, stream):
        x, side, uplo, op, diag, m, n, mb, nb = self.args
        fn = getattr(EMU, f"emu_triangular_solver_dist_{x}")
        fn.argtypes = [ctypes.c_int] * 4 + [ctypes.c_char] * 4 + [ctypes.c_int] * 4 + [
            ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_longlong),
            ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_longlong)]
        al = np.array([alpha], dtype=NP[x])
        pa, pb = (ctypes.c_void_p * 1)(a), (ctypes.c_void_p * 1)(b)
        la, lb = (ctypes.c_longlong * 1)(lda), (ctypes.c_longlong * 1)(ldb)
        assert fn(1, 1, 0, 0, side.encode(), uplo.encode(), op.encode(), diag.encode(), m, n, mb, nb, al.ctypes.data, pa, la,
                  pb, lb) == 0

    def close(self):
        pass


class RedistributePlan:
    def __init__(self, L, grid, x, m, n, st, sr, dt_, dr):
        self.args = (x, m, n, st, sr, dt_, dr)
        self.src_rows, self.src_cols, self.dst_rows, self.dst_cols = m, n, m, n

    def run(self, s, lds, d, ldd, stream):
        x, m, n, st, sr, dt_, dr = self.args
        fn = getattr(EMU, f"emu_redistribute_{x}")
        fn.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_char, ctypes.c_longlong, ctypes.c_longlong] + [ctypes.c_int] * 8 + [
            ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_longlong),
            ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_longlong)]
        ps, pd = (ctypes.c_void_p * 1)(s), (ctypes.c_void_p * 1)(d)
        ls, ld = (ctypes.c_longlong * 1)(lds), (ctypes.c_longlong * 1)(ldd)
        assert fn(1, 1, b"R", m, n, st[0], st[1], sr[0], sr[1], dt_[0], dt_[1], dr[0], dr[1], ps, ls, pd, ld) == 0

    def close(self):
        pass


REGISTRY = {}


def max_norm(L, grid, x, uplo, m, n, mb, nb, src, ptr, lda, stream):
    """numpy stand-in for dlafb_max_norm (the device kernel max_abs is verified; this checks the test's expectations)."""
    if m == 0 or n == 0:
        return 0.0
    a = np.frombuffer((ctypes.c_char * (n * lda * np.dtype(NP[x]).itemsize)).from_address(ptr), dtype=NP[x])
    a = a.reshape(n, lda)[:, :m].T
    return float(np.abs(np.tril(a) if uplo == "L" else a).max())


class OracleCApi:
    """dlaf_p?potrf on a host buffer through the oracle: stand-in for the C API in the sub-matrix tests."""

    def __getattr__(self, name):
        if not (name.startswith("dlaf_p") and name.endswith("potrf")):
            raise AttributeError(name)
        x = name[6]
        import dlaf_oracle as orc

        def potrf(uplo, n, ptr, ia, ja, desca, info_ref):
            dt = np.dtype(NP[x])
            m, nb, ld = desca[2], desca[4], desca[8]
            arr = np.frombuffer((ctypes.c_char * (m * ld * dt.itemsize)).from_address(ptr), dtype=dt).reshape(m, ld)
            i0, j0 = ia - 1, ja - 1
            assert i0 % nb == 0 and j0 % nb == 0
            sub = np.asfortranarray(arr[j0:j0 + n, i0:i0 + n].T.copy())
            up = uplo.decode()
            keep = (np.tril if up == "L" else np.triu)(np.ones((n, n))) > 0
            work = np.asfortranarray(np.where(keep, sub, 0))
            info_ref._obj.value = orc.cholesky_local(up, work, nb)
            sub[keep] = work[keep]
            arr[j0:j0 + n, i0:i0 + n] = sub.T
        return potrf


pkg = types.SimpleNamespace(DESC9=ctypes.c_int * 9, max_norm=max_norm, Grid=Grid, GenToStdPlan=GenToStdPlan, GenToStdDistPlan=GenToStdDistPlan,
                            TriangularSolverDistPlan=TriangularSolverDistPlan, RedistributePlan=RedistributePlan)


def main():
    import pytest  # noqa: F401  (the test modules import it)
    import test_zz_gen_to_std_gpu as g2s
    import test_zz_redistribute_gpu as red
    import test_zz_triangular_dist_gpu as tri
    ran = 0
    for x in "sdcz":
        for uplo in "LU":
            g2s.test_reference_known_answer(pkg, None, x, uplo)
            g2s.test_random_ragged(pkg, None, x, uplo)
            ran += 2
    for x in "dz":
        for uplo in "LU":
            g2s.test_random_tma_path(pkg, None, x, uplo)   # n = 1280 / 640: a few seconds each as a triple loop
            ran += 1
    for x in "dz":
        for uplo in "LU":
            g2s.test_single_rank_grid_matches_reference_known_answer(pkg, None, x, uplo)
            ran += 1
    for x in "sdcz":
        for side in "LR":
            tri.test_single_rank_grid_matches_reference_known_answer(pkg, None, x, side)
            ran += 1
    for x in "sdcz":
        # shrink the large cases of the device test: same code path, smaller loops on the CPU
        src = open(os.path.join(ROOT, "tests", "test_zz_redistribute_gpu.py")).read()
        assert "(2048, 2048, (512, 512), (64, 64))" in src
        red.test_single_rank_is_a_bit_exact_copy(pkg, None, x)
        ran += 1
    import test_zz_max_norm_gpu as nrm
    import test_zz_submatrix_gpu as sub
    for x in "sdcz":
        nrm.test_max_norm_single_rank(pkg, None, x)
        nrm.test_reference_max_norm_cases(pkg, None, x)
        ran += 2
    capi = OracleCApi()
    for case in sub.CASES:
        for uplo in "LU":
            sub.test_pxpotrf_submatrix_host_pointer(pkg, capi, case[0], uplo, *case[1:])
            ran += 1
    print(f"dry run ok: {ran} device-test functions executed against the CPU emulation / numpy / oracle stand-ins")


if __name__ == "__main__":
    main()
