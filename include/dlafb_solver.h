/* C ABI of the triangular solver, on one GPU and distributed over a grid of GPUs: op(A) X = alpha B (side 'L') or X op(A) = alpha B (side 'R'), with X
 * overwriting B. Row 2 of the "next" list of the hot-path scope: the immediate consumer of the Cholesky factor.
 *
 * Reference interface this stands in for (paths relative to the DLA-Future tree):
 *   dlaf::triangular_solver<Backend::GPU, Device::GPU, T>(side, uplo, op, diag, alpha, mat_a, mat_b)
 *       include/dlaf/solver/triangular.h:25-101 (local overload), algorithms include/dlaf/solver/triangular/impl.h:235-470
 *   explicit instantiations for float, double, complex<float>, complex<double>: include/dlaf/solver/triangular/api.h
 * (the reference has no C entry point for this operation; include/dlaf/solver/triangular.h of this repository is the
 * C++ front end over these functions).
 *
 * Return values as in dlafb_engine.h: 0 on success, CUDA errors as positive cudaError_t, argument errors as -1. */
#ifndef DLAFB_SOLVER_H
#define DLAFB_SOLVER_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dlafb_trsolve_plan* dlafb_trsolve_plan_t;

/* type 's','d','c','z'; side 'L'/'R'; uplo 'L'/'U'; op 'N'/'T'/'C'; diag 'N'/'U'. B is m x n; A is m x m (side 'L')
 * or n x n (side 'R') with square tiles of nb. Allocates the device workspace on the current device: the inverted
 * diagonal tiles of A, one panel of B, and for the op values that the contraction kernel cannot read in place
 * (side 'L' with 'T'/'C', side 'R' with 'N') a transposed copy of A. */
int dlafb_triangular_solver_plan_create(dlafb_trsolve_plan_t* plan, char type, char side, char uplo, char op, char diag,
                                        int64_t m, int64_t n, int64_t nb);
/* Solves in place. `alpha` points to one HOST element of the plan's type; `a` and `b` are device pointers,
 * column-major, lda >= order of A, ldb >= m. Only the uplo triangle of `a` is read (and not its diagonal when
 * diag = 'U'). Asynchronous: all work is enqueued on `stream` (cudaStream_t) in order. */
int dlafb_triangular_solver_plan_run(dlafb_trsolve_plan_t plan, const void* alpha, const void* a, int64_t lda, void* b,
                                     int64_t ldb, void* stream);
size_t dlafb_triangular_solver_plan_workspace_bytes(dlafb_trsolve_plan_t plan);
/* The caller synchronises with the last run first. */
int dlafb_triangular_solver_plan_destroy(dlafb_trsolve_plan_t plan);

/* ---- distributed over a P x Q grid (dlafb_engine.h: dlafb_grid_t), one process per GPU ----
 * Reference: dlaf::triangular_solver(grid, side, uplo, op, diag, alpha, mat_a, mat_b),
 * include/dlaf/solver/triangular.h:103-178, algorithms include/dlaf/solver/triangular/impl.h:475-1195.
 * A (square tiles of mb for side 'L', of nb for side 'R') and B (mb x nb tiles) are distributed 2D block-cyclically
 * over the grid with the same source rank (src_r, src_c). Collective over the grid: every rank creates the plan and
 * calls run with its local blocks (column-major device pointers; a rank whose local block is empty passes any
 * pointer). The grid must outlive the plan. */
typedef struct dlafb_trsolve_dist_plan* dlafb_trsolve_dist_plan_t;
#ifndef DLAFB_ENGINE_H
typedef struct dlafb_grid* dlafb_grid_t;
#endif

int dlafb_triangular_solver_dist_plan_create(dlafb_trsolve_dist_plan_t* plan, dlafb_grid_t grid, char type, char side,
                                             char uplo, char op, char diag, int64_t m, int64_t n, int64_t mb, int64_t nb,
                                             int src_r, int src_c);
/* Local block dimensions of this rank: of A and of B. */
int dlafb_triangular_solver_dist_plan_local_sizes(dlafb_trsolve_dist_plan_t plan, int64_t* a_rows, int64_t* a_cols,
                                                  int64_t* b_rows, int64_t* b_cols);
/* Solves in place. `alpha`: one HOST element of the plan's type. Asynchronous: local work and the NCCL broadcasts /
 * reductions along the grid's rows and columns are enqueued on `stream` in order. Collective. */
int dlafb_triangular_solver_dist_plan_run(dlafb_trsolve_dist_plan_t plan, const void* alpha, const void* a, int64_t lda,
                                          void* b, int64_t ldb, void* stream);
size_t dlafb_triangular_solver_dist_plan_workspace_bytes(dlafb_trsolve_dist_plan_t plan);
int dlafb_triangular_solver_dist_plan_destroy(dlafb_trsolve_dist_plan_t plan);

#ifdef __cplusplus
}
#endif
#endif
