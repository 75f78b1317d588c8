"""ctypes binding of libdlaf_b200.so (the C ABI in include/dlafb_kernels.h, include/dlafb_engine.h,
include/dlafb_solver.h, include/dlafb_gen_to_std.h and include/dlaf_c/*.h). Used by tests/ and bench.py; the product itself is the shared library.

The directory name contains a hyphen, so import it with `load_package()` from tests/conftest.py /
bench.py (importlib), under the module name `dlaf_b200`.

There is no CPU fallback: if the library is missing, `lib()` raises.
"""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libdlaf_b200.so")

TYPES = "sdcz"
_LIB = None

c_i64 = ctypes.c_int64
c_vp = ctypes.c_void_p
c_int = ctypes.c_int
c_char = ctypes.c_char
c_dbl = ctypes.c_double
c_u64 = ctypes.c_uint64
c_sz = ctypes.c_size_t

TILE_API = {
    "dlafb_potrf_tile_{x}": (c_int, [c_vp, c_char, c_int, c_vp, c_i64, c_vp, c_int, c_vp, c_int]),
    "dlafb_trsm_panel_{x}": (c_int, [c_vp, c_int, c_vp, c_i64, c_vp, c_int, c_vp, c_i64, c_int]),
    "dlafb_trsm_panel_packed_{x}": (c_int, [c_vp, c_int, c_vp, c_i64, c_vp, c_int, c_vp, c_i64, c_int, c_vp, c_int]),
    "dlafb_gemm_nt_{x}": (c_int, [c_vp, c_int, c_int, c_int, c_dbl, c_vp, c_i64, c_vp, c_i64, c_dbl, c_vp, c_i64,
                                  c_int]),
    "dlafb_transpose_{x}": (c_int, [c_vp, c_int, c_int, c_vp, c_i64, c_vp, c_i64, c_int, c_char]),
    "dlafb_laset_{x}": (c_int, [c_vp, c_char, c_int, c_int, c_dbl, c_dbl, c_dbl, c_dbl, c_vp, c_i64]),
    "dlafb_lacpy_{x}": (c_int, [c_vp, c_char, c_int, c_int, c_vp, c_i64, c_vp, c_i64]),
    "dlafb_max_abs_{x}": (c_int, [c_vp, c_char, c_int, c_int, c_vp, c_i64, c_vp]),
    "dlafb_gen_hpd_{x}": (c_int, [c_vp, c_i64, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_vp, c_i64, c_u64]),
}

PLAIN_API = {
    "dlafb_inner_block": (c_int, [c_char, c_vp, c_i64, c_int]),
    "dlafb_potrf_work_elems": (c_sz, [c_char, c_int, c_int]),
    "dlafb_potrf_tile_noalloc_d": (c_int, [c_vp, c_char, c_int, c_vp, c_i64, c_vp, c_int]),
}


def declared_symbols():
    """Every symbol the public headers declare (checked by the CPU test-suite against the built library)."""
    from . import _c_api, _engine_api
    return ([k.format(x=x) for k in TILE_API for x in TYPES] + list(PLAIN_API) + _engine_api.symbols() +
            _c_api.symbols())


def _preload_nccl():
    """libdlaf_b200.so needs libnccl.so.2. When torch is installed its bundled NCCL must be the one in the process
    (libtorch_cuda.so needs symbols newer than the system copy has), so load that one first."""
    import glob
    import sys
    for base in sys.path:
        for cand in glob.glob(os.path.join(base, "nvidia", "nccl", "lib", "libnccl.so.2")):
            try:
                ctypes.CDLL(cand, mode=ctypes.RTLD_GLOBAL)
                return
            except OSError:
                pass


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} not built: run `make` (or __graft_entry__.build()); no CPU fallback")
        _preload_nccl()
        L = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        for k, (res, args) in TILE_API.items():
            for x in TYPES:
                f = getattr(L, k.format(x=x))
                f.restype, f.argtypes = res, args
        for k, (res, args) in PLAIN_API.items():
            f = getattr(L, k)
            f.restype, f.argtypes = res, args
        from . import _c_api, _engine_api
        _engine_api.bind(L)
        _c_api.bind(L)
        _LIB = L
    return _LIB


from ._c_api import DESC9, DLAFDescriptor  # noqa: E402,F401
from ._engine_api import (CholeskyPlan, GenToStdDistPlan, GenToStdPlan, Grid, RedistributePlan,  # noqa: E402,F401
                          StepInfo, TriangularSolverDistPlan, TriangularSolverPlan, max_norm, schedule_step)


def check(err, what=""):
    if err != 0:
        raise RuntimeError(f"dlafb call failed ({what}): error {err}")
