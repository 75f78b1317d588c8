/* TEST INFRASTRUCTURE: the C-ABI boundary must be usable from plain C. Includes every public C header and takes the
 * address of one function of each (syntax check only, never run). */
#include <dlaf_c/desc.h>
#include <dlaf_c/factorization/cholesky.h>
#include <dlaf_c/grid.h>
#include <dlaf_c/init.h>
#include <dlaf_c/utils.h>
#include <dlafb_engine.h>
#include <dlafb_gen_to_std.h>
#include <dlafb_kernels.h>
#include <dlafb_matrix.h>
#include <dlafb_solver.h>

typedef void (*any_fn)(void);
any_fn c_api_symbols[] = {
    (any_fn)dlaf_initialize,
    (any_fn)dlaf_finalize,
    (any_fn)dlaf_create_grid,
    (any_fn)dlaf_free_grid,
    (any_fn)make_dlaf_descriptor,
    (any_fn)dlaf_cholesky_factorization_d,
    (any_fn)dlaf_pzpotrf,
    (any_fn)dlafb_grid_create,
    (any_fn)dlafb_cholesky_plan_run,
    (any_fn)dlafb_potrf_tile_d,
    (any_fn)dlafb_gemm_nt_z,
    (any_fn)dlafb_triangular_solver_plan_run,
    (any_fn)dlafb_triangular_solver_dist_plan_run,
    (any_fn)dlafb_gen_to_std_plan_run,
    (any_fn)dlafb_gen_to_std_dist_plan_run,
    (any_fn)dlafb_redistribute_plan_run,
    (any_fn)dlafb_max_norm,
};

int use_descriptor(void) {
  struct DLAF_descriptor d = {8, 8, 4, 4, 0, 0, 0, 0, 8};
  int desca[9] = {1, 0, 8, 8, 4, 4, 0, 0, 8};
  struct DLAF_descriptor e = make_dlaf_descriptor(8, 8, 5, 5, desca) /* 1-based ia, ja: sub-matrix at tile (1, 1) */;
  return d.m + e.i;
}

/* The call sequence of a C / ScaLAPACK-style caller, compiled and linked, never run here. */
int factorize(int rank, int size, const void* nccl_id128, int nprow, int npcol, int n, int nb, double* a_local, int lld) {
  int info = -1;
  int desca[9];
  int ctx;
  if (dlafb_mpi_init(rank, size, nccl_id128) != 0)
    return -1;
  dlaf_initialize(0, 0, 0, 0);
  ctx = dlaf_create_grid(MPI_COMM_WORLD, nprow, npcol, 'R');
  {
    struct DLAF_descriptor d = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    d.m = d.n = n;
    d.mb = d.nb = nb;
    d.ld = lld;
    info = dlaf_cholesky_factorization_d(ctx, 'L', a_local, d);
  }
  /* ScaLAPACK style: the context number doubles as the BLACS context; sub-matrix A(nb+1:n, nb+1:n) */
  dlaf_create_grid_from_blacs_info(7, MPI_COMM_WORLD, nprow, npcol, grid_ordering(MPI_COMM_WORLD, nprow, npcol, 0, 0));
  desca[0] = 1, desca[1] = 7, desca[2] = n, desca[3] = n, desca[4] = nb, desca[5] = nb, desca[6] = 0, desca[7] = 0, desca[8] = lld;
  dlaf_pdpotrf('L', n - nb, a_local, nb + 1, nb + 1, desca, &info);
  dlaf_free_grid(7);
  dlaf_free_grid(ctx);
  dlaf_finalize();
  dlafb_mpi_finalize();
  return info;
}

int main(int argc, char** argv) {
  (void)argv;
  if (argc > 1000)  /* never */
    return factorize(0, 1, 0, 1, 1, 8, 4, 0, 8);
  return use_descriptor() == 12 && c_api_symbols[0] ? 0 : 1;
}
