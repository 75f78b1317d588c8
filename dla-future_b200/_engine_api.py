"""ctypes signatures of include/dlafb_engine.h, include/dlafb_solver.h, include/dlafb_gen_to_std.h and include/dlafb_matrix.h plus thin Python
wrappers (Grid, CholeskyPlan, TriangularSolverPlan, GenToStdPlan) used by the tests and by bench.py. No arithmetic
happens here."""
import ctypes

c_i64 = ctypes.c_int64
c_vp = ctypes.c_void_p
c_int = ctypes.c_int
c_char = ctypes.c_char
P_int = ctypes.POINTER(ctypes.c_int)
P_i64 = ctypes.POINTER(ctypes.c_int64)
P_dbl = ctypes.POINTER(ctypes.c_double)


class StepInfo(ctypes.Structure):
    _fields_ = [("k", c_i64), ("own_diag", c_int), ("in_col", c_int), ("diag_root", c_int), ("row_root", c_int),
                ("lk_row", c_int), ("lk_col", c_int), ("il0", c_int), ("jl0", c_int), ("il_end", c_int),
                ("jl_end", c_int), ("n_colb", c_int), ("n_hp", c_int), ("n_np", c_int)]


ENGINE_API = {
    "dlafb_nccl_unique_id": (c_int, [c_vp]),
    "dlafb_grid_create": (c_int, [ctypes.POINTER(c_vp), c_vp, c_int, c_int, c_int, c_int, c_char]),
    "dlafb_grid_create_local": (c_int, [ctypes.POINTER(c_vp)]),
    "dlafb_grid_destroy": (c_int, [c_vp]),
    "dlafb_grid_info": (c_int, [c_vp, P_int, P_int, P_int, P_int, P_int, ctypes.c_char_p]),
    "dlafb_grid_barrier": (c_int, [c_vp]),
    "dlafb_grid_allreduce_max": (c_int, [c_vp, P_dbl]),
    "dlafb_dist_nr_tiles": (c_i64, [c_i64, c_i64]),
    "dlafb_dist_rank_global_tile": (c_int, [c_i64, c_int, c_int]),
    "dlafb_dist_local_tile_from_global": (c_i64, [c_i64, c_int]),
    "dlafb_dist_next_local_tile_from_global": (c_i64, [c_i64, c_int, c_int, c_int]),
    "dlafb_dist_global_tile_from_local": (c_i64, [c_i64, c_int, c_int, c_int]),
    "dlafb_dist_local_nr_tiles": (c_i64, [c_i64, c_i64, c_int, c_int, c_int]),
    "dlafb_dist_local_size": (c_i64, [c_i64, c_i64, c_int, c_int, c_int]),
    "dlafb_schedule_step": (c_int, [c_i64, c_i64, c_int, c_int, c_int, c_int, c_int, c_int, c_i64,
                                    ctypes.POINTER(StepInfo), P_int, P_int, P_int]),
    "dlafb_cholesky_plan_create": (c_int, [ctypes.POINTER(c_vp), c_vp, c_char, c_char, c_i64, c_i64, c_int, c_int]),
    "dlafb_cholesky_plan_run": (c_int, [c_vp, c_vp, c_i64, c_vp]),
    "dlafb_cholesky_plan_info": (c_int, [c_vp, P_int]),
    "dlafb_cholesky_plan_destroy": (c_int, [c_vp]),
    "dlafb_cholesky_plan_local_size": (c_int, [c_vp, P_i64, P_i64]),
    "dlafb_cholesky_plan_run_host": (c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, c_vp]),
    "dlafb_cholesky_plan_set_profiling": (c_int, [c_vp, c_int]),
    "dlafb_cholesky_plan_profile": (c_int, [c_vp, P_dbl, P_dbl, P_int]),
    "dlafb_cholesky_plan_timeline": (c_int, [c_vp, P_dbl, ctypes.c_longlong]),
    "dlafb_cholesky_plan_exchange_mode": (c_int, [c_vp]),
    "dlafb_cholesky_plan_residual": (c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, c_vp, P_dbl, P_dbl]),
    "dlafb_launch_count": (ctypes.c_uint64, []),
    "dlafb_capi_cached_plans": (c_int, []),
    "dlafb_cholesky_plan_acquire": (c_int, [ctypes.POINTER(c_vp), c_vp, c_char, c_char, c_i64, c_i64, c_int, c_int]),
    "dlafb_cholesky_plan_release": (c_int, [c_vp]),
    "dlafb_cholesky_plan_cache_drop": (c_int, [c_vp]),
    "dlafb_cholesky_plan_cache_size": (c_int, []),
    # include/dlafb_solver.h
    "dlafb_triangular_solver_plan_create": (c_int, [ctypes.POINTER(c_vp), c_char, c_char, c_char, c_char, c_char, c_i64,
                                                    c_i64, c_i64]),
    "dlafb_triangular_solver_plan_run": (c_int, [c_vp, c_vp, c_vp, c_i64, c_vp, c_i64, c_vp]),
    "dlafb_triangular_solver_plan_workspace_bytes": (ctypes.c_size_t, [c_vp]),
    "dlafb_triangular_solver_plan_destroy": (c_int, [c_vp]),
    "dlafb_triangular_solver_dist_plan_create": (c_int, [ctypes.POINTER(c_vp), c_vp, c_char, c_char, c_char, c_char,
                                                         c_char, c_i64, c_i64, c_i64, c_i64, c_int, c_int]),
    "dlafb_triangular_solver_dist_plan_local_sizes": (c_int, [c_vp, P_i64, P_i64, P_i64, P_i64]),
    "dlafb_triangular_solver_dist_plan_run": (c_int, [c_vp, c_vp, c_vp, c_i64, c_vp, c_i64, c_vp]),
    "dlafb_triangular_solver_dist_plan_workspace_bytes": (ctypes.c_size_t, [c_vp]),
    "dlafb_triangular_solver_dist_plan_destroy": (c_int, [c_vp]),
    # include/dlafb_gen_to_std.h
    "dlafb_gen_to_std_plan_create": (c_int, [ctypes.POINTER(c_vp), c_char, c_char, c_i64, c_i64]),
    "dlafb_gen_to_std_plan_run": (c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, c_vp]),
    "dlafb_gen_to_std_plan_workspace_bytes": (ctypes.c_size_t, [c_vp]),
    "dlafb_gen_to_std_plan_destroy": (c_int, [c_vp]),
    "dlafb_gen_to_std_dist_plan_create": (c_int, [ctypes.POINTER(c_vp), c_vp, c_char, c_char, c_i64, c_i64, c_int,
                                                  c_int]),
    "dlafb_gen_to_std_dist_plan_local_size": (c_int, [c_vp, P_i64, P_i64]),
    "dlafb_gen_to_std_dist_plan_run": (c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, c_vp]),
    "dlafb_gen_to_std_dist_plan_workspace_bytes": (ctypes.c_size_t, [c_vp]),
    "dlafb_gen_to_std_dist_plan_destroy": (c_int, [c_vp]),
    # include/dlafb_matrix.h
    "dlafb_redistribute_plan_create": (c_int, [ctypes.POINTER(c_vp), c_vp, c_char, c_i64, c_i64, c_i64, c_i64, c_int,
                                               c_int, c_i64, c_i64, c_int, c_int]),
    "dlafb_redistribute_plan_local_sizes": (c_int, [c_vp, P_i64, P_i64, P_i64, P_i64]),
    "dlafb_redistribute_plan_run": (c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, c_vp]),
    "dlafb_redistribute_plan_destroy": (c_int, [c_vp]),
    "dlafb_max_norm": (c_int, [c_vp, c_char, c_char, c_i64, c_i64, c_i64, c_i64, c_int, c_int, c_vp, c_i64, c_vp, P_dbl]),
}


def symbols():
    return list(ENGINE_API)


def bind(L):
    for k, (res, args) in ENGINE_API.items():
        f = getattr(L, k)
        f.restype, f.argtypes = res, args


def _ck(err, what):
    if err != 0:
        raise RuntimeError(f"{what} failed with error {err}")


def schedule_step(L, n, nb, P, Q, pr, pc, sr, sc, k):
    """Returns (StepInfo, colb[(jl, g, root, src_il)], hp[(row, col, diag)], np[...]) of step k."""
    info = StepInfo()
    _ck(L.dlafb_schedule_step(n, nb, P, Q, pr, pc, sr, sc, k, ctypes.byref(info), None, None, None), "schedule")
    colb = (c_int * (4 * max(1, info.n_colb)))()
    hp = (c_int * (3 * max(1, info.n_hp)))()
    np_ = (c_int * (3 * max(1, info.n_np)))()
    _ck(L.dlafb_schedule_step(n, nb, P, Q, pr, pc, sr, sc, k, ctypes.byref(info), colb, hp, np_), "schedule")
    cl = [tuple(colb[4 * i:4 * i + 4]) for i in range(info.n_colb)]
    hl = [tuple(hp[3 * i:3 * i + 3]) for i in range(info.n_hp)]
    nl = [tuple(np_[3 * i:3 * i + 3]) for i in range(info.n_np)]
    return info, cl, hl, nl


class Grid:
    """comm::CommunicatorGrid stand-in. `unique_id` (128 bytes) must be the same on every rank."""

    def __init__(self, L, P=1, Q=1, rank=0, size=1, unique_id=None, ordering="R"):
        self.L = L
        self.h = c_vp()
        if size == 1:
            _ck(L.dlafb_grid_create_local(ctypes.byref(self.h)), "grid_create_local")
        else:
            buf = ctypes.create_string_buffer(bytes(unique_id), 128)
            _ck(L.dlafb_grid_create(ctypes.byref(self.h), buf, rank, size, P, Q, ordering.encode()), "grid_create")
        p, q, r, c, w = c_int(), c_int(), c_int(), c_int(), c_int()
        o = ctypes.create_string_buffer(1)
        _ck(L.dlafb_grid_info(self.h, p, q, r, c, w, o), "grid_info")
        self.P, self.Q, self.pr, self.pc, self.rank = p.value, q.value, r.value, c.value, w.value

    def barrier(self):
        _ck(self.L.dlafb_grid_barrier(self.h), "barrier")

    def allreduce_max(self, v):
        d = ctypes.c_double(v)
        _ck(self.L.dlafb_grid_allreduce_max(self.h, ctypes.byref(d)), "allreduce_max")
        return d.value

    def close(self):
        if self.h:
            self.L.dlafb_grid_destroy(self.h)
            self.h = c_vp()


class CholeskyPlan:
    def __init__(self, L, grid, type_, uplo, n, nb, src=(0, 0)):
        self.L = L
        self.h = c_vp()
        _ck(L.dlafb_cholesky_plan_create(ctypes.byref(self.h), grid.h, type_.encode(), uplo.encode(), n, nb, src[0],
                                         src[1]), "plan_create")
        r, c = c_i64(), c_i64()
        _ck(L.dlafb_cholesky_plan_local_size(self.h, r, c), "local_size")
        self.local_rows, self.local_cols = r.value, c.value
        self.nt = (n + nb - 1) // nb if n > 0 else 0

    def run(self, ptr, lda, stream):
        _ck(self.L.dlafb_cholesky_plan_run(self.h, ptr, lda, stream), "plan_run")

    def info(self):
        v = c_int()
        _ck(self.L.dlafb_cholesky_plan_info(self.h, ctypes.byref(v)), "plan_info")
        return v.value

    def set_profiling(self, on=True):
        _ck(self.L.dlafb_cholesky_plan_set_profiling(self.h, int(on)), "set_profiling")

    def profile(self):
        ms, fl, n = ctypes.c_double(), ctypes.c_double(), c_int()
        _ck(self.L.dlafb_cholesky_plan_profile(self.h, ctypes.byref(ms), ctypes.byref(fl), ctypes.byref(n)), "profile")
        return ms.value, fl.value, n.value

    def residual(self, a0_ptr, lda, l_ptr, ldl, stream):
        """(max|A0|, max|A0 - L L^H|) over the grid; a0 is overwritten. Collective."""
        a, d = ctypes.c_double(), ctypes.c_double()
        _ck(self.L.dlafb_cholesky_plan_residual(self.h, a0_ptr, lda, l_ptr, ldl, stream, ctypes.byref(a),
                                                ctypes.byref(d)), "residual")
        return a.value, d.value

    def exchange_mode(self):
        return {0: "local", 1: "nccl", 2: "p2p"}.get(self.L.dlafb_cholesky_plan_exchange_mode(self.h), "?")

    def timeline(self):
        """Rows of 8 times (ms since the start of the last profiled run), one row per step; see dlafb_engine.h."""
        buf = (ctypes.c_double * (8 * max(self.nt, 1)))()
        _ck(self.L.dlafb_cholesky_plan_timeline(self.h, buf, len(buf)), "timeline")
        return [list(buf[8 * k:8 * k + 8]) for k in range(self.nt)]

    def close(self):
        if self.h:
            self.L.dlafb_cholesky_plan_destroy(self.h)
            self.h = c_vp()


_CTYPE = {"s": ctypes.c_float * 1, "d": ctypes.c_double * 1, "c": ctypes.c_float * 2, "z": ctypes.c_double * 2}


class TriangularSolverPlan:
    """dlaf::triangular_solver on one GPU (include/dlafb_solver.h): op(A) X = alpha B or X op(A) = alpha B."""

    def __init__(self, L, type_, side, uplo, op, diag, m, n, nb):
        self.L = L
        self.type = type_
        self.h = c_vp()
        _ck(L.dlafb_triangular_solver_plan_create(ctypes.byref(self.h), type_.encode(), side.encode(), uplo.encode(),
                                                  op.encode(), diag.encode(), m, n, nb), "triangular_solver_plan_create")

    def run(self, alpha, a_ptr, lda, b_ptr, ldb, stream):
        al = complex(alpha)
        buf = _CTYPE[self.type](al.real, al.imag) if self.type in "cz" else _CTYPE[self.type](al.real)
        _ck(self.L.dlafb_triangular_solver_plan_run(self.h, ctypes.addressof(buf), a_ptr, lda, b_ptr, ldb, stream),
            "triangular_solver_plan_run")

    def workspace_bytes(self):
        return self.L.dlafb_triangular_solver_plan_workspace_bytes(self.h)

    def close(self):
        if self.h:
            self.L.dlafb_triangular_solver_plan_destroy(self.h)
            self.h = c_vp()


class TriangularSolverDistPlan:
    """dlaf::triangular_solver(grid, ...) (include/dlafb_solver.h): collective over the grid, local blocks in and out."""

    def __init__(self, L, grid, type_, side, uplo, op, diag, m, n, mb, nb, src=(0, 0)):
        self.L = L
        self.type = type_
        self.h = c_vp()
        _ck(L.dlafb_triangular_solver_dist_plan_create(ctypes.byref(self.h), grid.h, type_.encode(), side.encode(),
                                                       uplo.encode(), op.encode(), diag.encode(), m, n, mb, nb, src[0],
                                                       src[1]), "triangular_solver_dist_plan_create")
        v = [c_i64() for _ in range(4)]
        _ck(L.dlafb_triangular_solver_dist_plan_local_sizes(self.h, *v), "triangular_solver_dist_plan_local_sizes")
        self.a_rows, self.a_cols, self.b_rows, self.b_cols = (x.value for x in v)

    def run(self, alpha, a_ptr, lda, b_ptr, ldb, stream):
        al = complex(alpha)
        buf = _CTYPE[self.type](al.real, al.imag) if self.type in "cz" else _CTYPE[self.type](al.real)
        _ck(self.L.dlafb_triangular_solver_dist_plan_run(self.h, ctypes.addressof(buf), a_ptr, lda, b_ptr, ldb, stream),
            "triangular_solver_dist_plan_run")

    def workspace_bytes(self):
        return self.L.dlafb_triangular_solver_dist_plan_workspace_bytes(self.h)

    def close(self):
        if self.h:
            self.L.dlafb_triangular_solver_dist_plan_destroy(self.h)
            self.h = c_vp()


class GenToStdPlan:
    """generalized_to_standard on one GPU (include/dlafb_gen_to_std.h): A <- inv(L) A inv(L)^H / inv(U)^H A inv(U)."""

    def __init__(self, L, type_, uplo, n, nb):
        self.L = L
        self.h = c_vp()
        _ck(L.dlafb_gen_to_std_plan_create(ctypes.byref(self.h), type_.encode(), uplo.encode(), n, nb),
            "gen_to_std_plan_create")

    def run(self, a_ptr, lda, f_ptr, ldf, stream):
        _ck(self.L.dlafb_gen_to_std_plan_run(self.h, a_ptr, lda, f_ptr, ldf, stream), "gen_to_std_plan_run")

    def workspace_bytes(self):
        return self.L.dlafb_gen_to_std_plan_workspace_bytes(self.h)

    def close(self):
        if self.h:
            self.L.dlafb_gen_to_std_plan_destroy(self.h)
            self.h = c_vp()


class GenToStdDistPlan:
    """generalized_to_standard(grid, ...) (include/dlafb_gen_to_std.h): collective over the grid, local blocks in and out."""

    def __init__(self, L, grid, type_, uplo, n, nb, src=(0, 0)):
        self.L = L
        self.h = c_vp()
        _ck(L.dlafb_gen_to_std_dist_plan_create(ctypes.byref(self.h), grid.h, type_.encode(), uplo.encode(), n, nb,
                                                src[0], src[1]), "gen_to_std_dist_plan_create")
        r, c = c_i64(), c_i64()
        _ck(L.dlafb_gen_to_std_dist_plan_local_size(self.h, r, c), "gen_to_std_dist_plan_local_size")
        self.local_rows, self.local_cols = r.value, c.value

    def run(self, a_ptr, lda, f_ptr, ldf, stream):
        _ck(self.L.dlafb_gen_to_std_dist_plan_run(self.h, a_ptr, lda, f_ptr, ldf, stream), "gen_to_std_dist_plan_run")

    def workspace_bytes(self):
        return self.L.dlafb_gen_to_std_dist_plan_workspace_bytes(self.h)

    def close(self):
        if self.h:
            self.L.dlafb_gen_to_std_dist_plan_destroy(self.h)
            self.h = c_vp()


class RedistributePlan:
    """matrix::copy(src, dst, grid) (include/dlafb_matrix.h): dst <- src between two block-cyclic layouts. Collective."""

    def __init__(self, L, grid, type_, m, n, src_tile, src_rank, dst_tile, dst_rank):
        self.L = L
        self.h = c_vp()
        _ck(L.dlafb_redistribute_plan_create(ctypes.byref(self.h), grid.h, type_.encode(), m, n, src_tile[0], src_tile[1],
                                             src_rank[0], src_rank[1], dst_tile[0], dst_tile[1], dst_rank[0],
                                             dst_rank[1]), "redistribute_plan_create")
        v = [c_i64() for _ in range(4)]
        _ck(L.dlafb_redistribute_plan_local_sizes(self.h, *v), "redistribute_plan_local_sizes")
        self.src_rows, self.src_cols, self.dst_rows, self.dst_cols = (x.value for x in v)

    def run(self, src_ptr, ld_src, dst_ptr, ld_dst, stream):
        _ck(self.L.dlafb_redistribute_plan_run(self.h, src_ptr, ld_src, dst_ptr, ld_dst, stream), "redistribute_plan_run")

    def close(self):
        if self.h:
            self.L.dlafb_redistribute_plan_destroy(self.h)
            self.h = c_vp()


def max_norm(L, grid, type_, uplo, m, n, mb, nb, src, a_ptr, lda, stream):
    """dlaf::auxiliary::max_norm over the grid (include/dlafb_matrix.h). Collective; every rank gets the value."""
    out = ctypes.c_double()
    _ck(L.dlafb_max_norm(grid.h, type_.encode(), uplo.encode(), m, n, mb, nb, src[0], src[1], a_ptr, lda, stream,
                         ctypes.byref(out)), "max_norm")
    return out.value
