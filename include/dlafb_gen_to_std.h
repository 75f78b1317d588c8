/* C ABI of the reduction of a Hermitian definite generalized eigenproblem to standard form on one GPU:
 *   uplo 'L':  A <- inv(L) A inv(L)^H        uplo 'U':  A <- inv(U)^H A inv(U)
 * with L / U the Cholesky factor of B (the output of dlaf_p?potrf / dlafb_cholesky_plan_run). Row 3 of the "next"
 * list of the hot-path scope: the one in-tree algorithm that consumes the Cholesky factor inside a larger one.
 *
 * Reference interface this stands in for (paths relative to the DLA-Future tree):
 *   dlaf::eigensolver::internal::generalized_to_standard<Backend::GPU, Device::GPU, T>(uplo, mat_a, mat_b)
 *       include/dlaf/eigensolver/gen_to_std.h:23-77 (local overload),
 *       algorithms include/dlaf/eigensolver/gen_to_std/impl.h:222-283 (call_L), :508-569 (call_U)
 *   explicit instantiations for float, double, complex<float>, complex<double>: include/dlaf/eigensolver/gen_to_std/api.h
 * (the reference reaches it from C only through dlaf_hermitian_generalized_eigensolver_*; there is no separate C
 * entry point. include/dlaf/eigensolver/gen_to_std.h of this repository is the C++ front end over these functions.)
 *
 * Return values as in dlafb_engine.h: 0 on success, CUDA errors as positive cudaError_t, argument errors as -1. */
#ifndef DLAFB_GEN_TO_STD_H
#define DLAFB_GEN_TO_STD_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dlafb_gen_to_std_plan* dlafb_gen_to_std_plan_t;

/* type 's','d','c','z'; uplo 'L'/'U'; A and the factor are n x n with square tiles of nb. Allocates the device
 * workspace on the current device: the inverted diagonal tiles of the factor, two nb x nb tiles, two n x 2nb panels
 * and, for uplo 'U', lower-triangular images of A and of the factor (2 n^2 elements). */
int dlafb_gen_to_std_plan_create(dlafb_gen_to_std_plan_t* plan, char type, char uplo, int64_t n, int64_t nb);
/* In place on `a`. `a` and `factor` are device pointers, column-major, lda, ldf >= n. Only the uplo triangles are
 * read, only the uplo triangle of `a` is written, `factor` is never written. Asynchronous: all work is enqueued on
 * `stream` (cudaStream_t) in order. */
int dlafb_gen_to_std_plan_run(dlafb_gen_to_std_plan_t plan, void* a, int64_t lda, const void* factor, int64_t ldf,
                              void* stream);
size_t dlafb_gen_to_std_plan_workspace_bytes(dlafb_gen_to_std_plan_t plan);
/* The caller synchronises with the last run first. */
int dlafb_gen_to_std_plan_destroy(dlafb_gen_to_std_plan_t plan);

/* ---- distributed over a P x Q grid (dlafb_engine.h: dlafb_grid_t), one process per GPU ----
 * Reference: generalized_to_standard(grid, uplo, mat_a, mat_b), include/dlaf/eigensolver/gen_to_std.h:79-140,
 * algorithms include/dlaf/eigensolver/gen_to_std/impl.h:286-505 (call_L), :572-771 (call_U).
 * A and the factor are distributed 2D block-cyclically with nb x nb tiles and source rank (src_r, src_c). Collective:
 * every rank creates the plan and calls run with its local blocks. The grid must outlive the plan. */
typedef struct dlafb_gen_to_std_dist_plan* dlafb_gen_to_std_dist_plan_t;
#ifndef DLAFB_ENGINE_H
typedef struct dlafb_grid* dlafb_grid_t;
#endif

int dlafb_gen_to_std_dist_plan_create(dlafb_gen_to_std_dist_plan_t* plan, dlafb_grid_t grid, char type, char uplo,
                                      int64_t n, int64_t nb, int src_r, int src_c);
/* Local block dimensions of this rank (the same for A and for the factor). */
int dlafb_gen_to_std_dist_plan_local_size(dlafb_gen_to_std_dist_plan_t plan, int64_t* rows, int64_t* cols);
/* In place on the local block `a`; `factor` is never written. Asynchronous: local work and the NCCL broadcasts along
 * the grid's rows and columns are enqueued on `stream` in order. Collective. */
int dlafb_gen_to_std_dist_plan_run(dlafb_gen_to_std_dist_plan_t plan, void* a, int64_t lda, const void* factor,
                                   int64_t ldf, void* stream);
size_t dlafb_gen_to_std_dist_plan_workspace_bytes(dlafb_gen_to_std_dist_plan_t plan);
int dlafb_gen_to_std_dist_plan_destroy(dlafb_gen_to_std_dist_plan_t plan);

#ifdef __cplusplus
}
#endif
#endif
