/* C ABI of the host-side Cholesky engine: process grid (NCCL over NVLink), 2D block-cyclic index algebra,
 * per-step schedule and the look-ahead scheduler over CUDA streams. Host orchestration is C++; this header is
 * what the C API (include/dlaf_c), the C++ API (include/dlaf) and the Python test/bench harness bind.
 *
 * Reference interfaces these stand in for (paths relative to the DLA-Future tree):
 *   dlafb_grid_*            comm::CommunicatorGrid   include/dlaf/communication/communicator_grid.h:37-153,
 *                                                    src/communication/communicator_grid.cpp:19-68
 *   dlafb_dist_*            matrix::Distribution     include/dlaf/matrix/util_distribution.h:85-175
 *   dlafb_schedule_*        Cholesky<..>::call_L     include/dlaf/factorization/cholesky/impl.h:191-313 and
 *                           comm::broadcast(panel)   include/dlaf/communication/broadcast_panel.h:123-213
 *   dlafb_cholesky_plan_*   dlaf::cholesky_factorization<Backend::GPU>  include/dlaf/factorization/cholesky.h:40,:68
 *
 * Return values: 0 on success; CUDA errors as positive cudaError_t; NCCL errors as 1000 + ncclResult_t;
 * argument errors as -1. */
#ifndef DLAFB_ENGINE_H
#define DLAFB_ENGINE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dlafb_grid* dlafb_grid_t;
typedef struct dlafb_chol_plan* dlafb_chol_plan_t;

/* ---- process grid ---- */
/* 128-byte NCCL unique id, created on one rank and handed to every rank by the launcher (torch.distributed,
 * a file, MPI ...). */
int dlafb_nccl_unique_id(void* out128);
/* Collective over all world_size = P*Q ranks. ordering 'R' (row-major) or 'C' (column-major) maps the linear rank
 * to grid coordinates as comm::CommunicatorGrid does. Uses the current CUDA device. */
int dlafb_grid_create(dlafb_grid_t* grid, const void* unique_id128, int world_rank, int world_size, int P, int Q,
                      char ordering);
/* 1 x 1 grid without NCCL. */
int dlafb_grid_create_local(dlafb_grid_t* grid);
int dlafb_grid_destroy(dlafb_grid_t grid);
int dlafb_grid_info(dlafb_grid_t grid, int* P, int* Q, int* pr, int* pc, int* world_rank, char* ordering);
/* Grid-wide barrier (device all-reduce on the world communicator + stream synchronisation). */
int dlafb_grid_barrier(dlafb_grid_t grid);
/* max over all ranks of a host double (used for max-over-ranks timing and the distributed max norm). */
int dlafb_grid_allreduce_max(dlafb_grid_t grid, double* value);

/* ---- 2D block-cyclic index algebra along one axis (n elements, block nb, P ranks, source rank src) ---- */
int64_t dlafb_dist_nr_tiles(int64_t n, int64_t nb);
int dlafb_dist_rank_global_tile(int64_t g, int P, int src);
int64_t dlafb_dist_local_tile_from_global(int64_t g, int P);
int64_t dlafb_dist_next_local_tile_from_global(int64_t g, int P, int rank, int src);
int64_t dlafb_dist_global_tile_from_local(int64_t l, int P, int rank, int src);
int64_t dlafb_dist_local_nr_tiles(int64_t n, int64_t nb, int P, int rank, int src);
int64_t dlafb_dist_local_size(int64_t n, int64_t nb, int P, int rank, int src);

/* ---- per-step schedule of the Lower factorization as seen from rank (pr, pc): host logic only, no GPU ---- */
typedef struct {
  int64_t k;
  int own_diag, in_col;
  int diag_root, row_root;
  int lk_row, lk_col;
  int il0, jl0, il_end, jl_end;
  int n_colb, n_hp, n_np;
} dlafb_step_info;
/* colb: n_colb x {jl, g, root, src_il}; hp / np: n x {row, col, diag}. Pass NULL arrays to query the counts. */
int dlafb_schedule_step(int64_t n, int64_t nb, int P, int Q, int pr, int pc, int src_r, int src_c, int64_t k,
                        dlafb_step_info* info, int* colb, int* hp, int* np);

/* ---- the factorization ---- */
/* type 's','d','c','z'; uplo 'L' or 'U'; (src_r, src_c) = rank owning the first tile. Collective over the grid. */
int dlafb_cholesky_plan_create(dlafb_chol_plan_t* plan, dlafb_grid_t grid, char type, char uplo, int64_t n,
                               int64_t nb, int src_r, int src_c);
/* Factorizes the local block `a` (device pointer, column-major, leading dimension lda >= local rows) in place.
 * Asynchronous: work starts after everything already enqueued on `stream` (cudaStream_t) and `stream` waits for
 * the factorization before running anything enqueued later. Collective over the grid. */
int dlafb_cholesky_plan_run(dlafb_chol_plan_t plan, void* a, int64_t lda, void* stream);
/* Host-mirrored run (what the reference's C API does with matrix::MatrixMirror, src/c_api/factorization/cholesky.h:48-55):
 * copies the referenced triangle of the host block `host` (ld `ldh`) into the device mirror `dev` (ld `ldd`, same
 * local shape, allocated by the caller), factorizes it there and copies the factor back. Only the referenced triangle
 * travels; for uplo = L every column panel leaves for the host as soon as it is final, overlapping the remaining
 * steps. Asynchronous on `stream`; `host` should be pinned for the copies to overlap. */
int dlafb_cholesky_plan_run_host(dlafb_chol_plan_t plan, void* host, int64_t ldh, void* dev, int64_t ldd, void* stream);
/* Synchronises with the last run and returns its info: 0, or the 1-based global column of the first
 * non-positive pivot (ScaLAPACK p?potrf convention: the same value on every rank, so on a grid this call is
 * collective - one small all-reduce). */
int dlafb_cholesky_plan_info(dlafb_chol_plan_t plan, int* info);
int dlafb_cholesky_plan_destroy(dlafb_chol_plan_t plan);
/* Cached plans (what the C++ front end uses: creating a plan is a collective with allocations and peer mappings,
 * destroying one synchronises the device). `acquire` returns the plan the library keeps for (grid, type, uplo, n,
 * nb, source rank), creating it on first use; `release` says the caller is done with it (after the run has been
 * waited for). Idle plans beyond a small limit are destroyed least recently used first; `cache_drop(grid)` destroys
 * the plans of one grid (dlafb_grid_destroy calls it), `cache_drop(NULL)` all of them. Collective like create. */
int dlafb_cholesky_plan_acquire(dlafb_chol_plan_t* plan, dlafb_grid_t grid, char type, char uplo, int64_t n, int64_t nb,
                                int src_r, int src_c);
int dlafb_cholesky_plan_release(dlafb_chol_plan_t plan);
int dlafb_cholesky_plan_cache_drop(dlafb_grid_t grid);
int dlafb_cholesky_plan_cache_size(void);
/* Local block dimensions of this rank for the plan's matrix. */
int dlafb_cholesky_plan_local_size(dlafb_chol_plan_t plan, int64_t* rows, int64_t* cols);

/* Optional timing of the bulk trailing-update launches (CUDA events on the stream they run on). After a run,
 * dlafb_cholesky_plan_profile returns their summed duration, their algorithmic flops and their number. */
int dlafb_cholesky_plan_set_profiling(dlafb_chol_plan_t plan, int on);
int dlafb_cholesky_plan_profile(dlafb_chol_plan_t plan, double* update_ms, double* update_flops, int* launches);
/* Timeline of the last profiled run: 8 times per step (ms since the run started on the caller's stream): step start,
 * diagonal tile factored, panel solved and packed, row broadcast done, column broadcasts done, look-ahead column
 * updated (all on the high-priority stream), bulk update begin / end. -1 where not recorded. out holds >= 8*nt. */
int dlafb_cholesky_plan_timeline(dlafb_chol_plan_t plan, double* out, long long cap);

/* Residual check of a factorization on the device, distributed like the matrix (replaces the reference miniapp's
 * CPU checker, miniapp/miniapp_cholesky.cpp:243-442: max|A - L L^H| / max|A| <= n*eps). `a0` = this rank's block
 * of the ORIGINAL matrix (overwritten with the residual), `l` = the factor. Collective over the grid; uplo = L plans
 * only (returns 2 otherwise). Returns the two maxima over the whole grid, taken over the referenced triangle.
 * Synchronises. */
int dlafb_cholesky_plan_residual(dlafb_chol_plan_t plan, void* a0, int64_t lda, const void* l, int64_t ldl, void* stream,
                                 double* max_abs_a, double* max_abs_diff);

/* How the panels travel between ranks: 0 = single rank, 1 = ncclBroadcast on row / column communicators,
 * 2 = copies into the peers' memory over NVLink (CUDA IPC, copy engines) with stream-ordered step counters; chosen at
 * plan creation (2 when every rank of the grid can map its row and column peers; DLAFB_P2P=0 forces 1). */
int dlafb_cholesky_plan_exchange_mode(dlafb_chol_plan_t plan);

/* Factorization plans the reference-facing C API (include/dlaf_c) currently keeps alive: at most DLAFB_PLAN_CACHE
 * (default 4), least recently used first out. */
int dlafb_capi_cached_plans(void);

/* Kernels launched by this library so far in this process. */
uint64_t dlafb_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif
