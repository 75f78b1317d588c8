/* C ABI of the matrix redistribution: dst <- src between two 2D block-cyclic layouts of the same global matrix on the
 * same process grid (another block size and / or another source rank).
 *
 * Reference interface this stands in for (paths relative to the DLA-Future tree):
 *   dlaf::matrix::copy(Matrix<const T, Source>& src, Matrix<T, Destination>& dst, comm::CommunicatorGrid& grid)
 *       include/dlaf/matrix/copy.h:118-231 ("Copy of a matrix performing data reshuffling", CHANGELOG 0.2.0 #933)
 * with the reference's preconditions (:124-131): same size, same grid, and in each dimension the smaller block size
 * divides the larger one. include/dlaf/matrix/copy.h of this repository is the C++ front end over these functions.
 *
 * Return values as in dlafb_engine.h: 0 on success, CUDA errors as positive cudaError_t, NCCL errors as 1000 +
 * ncclResult_t, argument errors as -1. */
#ifndef DLAFB_MATRIX_H
#define DLAFB_MATRIX_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dlafb_redistribute_plan* dlafb_redistribute_plan_t;
#ifndef DLAFB_ENGINE_H
typedef struct dlafb_grid* dlafb_grid_t;
#endif

/* type 's','d','c','z'; the m x n matrix has (mb_src x nb_src) tiles with first tile on rank (src_r_src, src_c_src) in
 * `src` and (mb_dst x nb_dst) tiles from (src_r_dst, src_c_dst) in `dst`. Host index algebra only; the device lists and
 * exchange buffers are created by the first run. The grid must outlive the plan. */
int dlafb_redistribute_plan_create(dlafb_redistribute_plan_t* plan, dlafb_grid_t grid, char type, int64_t m, int64_t n,
                                   int64_t mb_src, int64_t nb_src, int src_r_src, int src_c_src, int64_t mb_dst,
                                   int64_t nb_dst, int src_r_dst, int src_c_dst);
/* Local block dimensions of this rank in the two layouts. */
int dlafb_redistribute_plan_local_sizes(dlafb_redistribute_plan_t plan, int64_t* src_rows, int64_t* src_cols,
                                        int64_t* dst_rows, int64_t* dst_cols);
/* dst <- src (local blocks, column-major device pointers). Collective over the grid. Asynchronous on `stream`: one
 * gather launch, one grouped NCCL send / receive per pair of ranks, one scatter launch. The first call with a new pair
 * of leading dimensions builds and uploads the rectangle lists (synchronises the stream once). */
int dlafb_redistribute_plan_run(dlafb_redistribute_plan_t plan, const void* src, int64_t ld_src, void* dst,
                                int64_t ld_dst, void* stream);
/* The caller synchronises with the last run first. */
int dlafb_redistribute_plan_destroy(dlafb_redistribute_plan_t plan);

/* Max norm of a distributed matrix, max |a_ij| over the general matrix (uplo 'G', any m x n with mb x nb tiles) or over
 * the lower triangle of a square matrix with square tiles (uplo 'L': valid for symmetric / Hermitian / lower-triangular
 * data). Reference: dlaf::auxiliary::max_norm(grid, rank, uplo, A), include/dlaf/auxiliary/norm.h:35-58 - implemented
 * there for the CPU backend only (auxiliary/norm/mc.h:35-149); 0 for an empty matrix (:41-43). Collective over the grid;
 * synchronises; every rank gets the result (the reference delivers it to one rank). NaN entries are reported as +inf. */
int dlafb_max_norm(dlafb_grid_t grid, char type, char uplo, int64_t m, int64_t n, int64_t mb, int64_t nb, int src_r,
                   int src_c, const void* a, int64_t lda, void* stream, double* result);

#ifdef __cplusplus
}
#endif
#endif
