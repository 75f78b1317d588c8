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
  struct DLAF_descriptor e = make_dlaf_descriptor(8, 8, 4, 4, desca);
  return d.m + e.i;
}

int main(void) { return use_descriptor() == 12 && c_api_symbols[0] ? 0 : 1; }
