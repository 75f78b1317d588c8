/* Process-grid registry of the C API (reference: include/dlaf_c/grid.h:31-74, src/c_api/grid.cpp:26-92).
 * MPI_Comm comes from <mpi.h> = the single-node shim in include/dlafb_mpi_shim/mpi.h (ranks = processes, one per GPU,
 * bootstrapped with an NCCL unique id; see that header for applications that use a real MPI). */
#pragma once
#include <mpi.h>

#include "dlaf_c/utils.h"

/* Creates a nprow x npcol grid over comm ('R' row-major / 'C' column-major rank order); returns the context,
 * counting down from INT_MAX as the reference does. Collective. */
DLAF_EXTERN_C int dlaf_create_grid(MPI_Comm comm, int nprow, int npcol, char order) DLAF_NOEXCEPT;
DLAF_EXTERN_C void dlaf_free_grid(int context) DLAF_NOEXCEPT;
/* 'R' or 'C' according to how (myprow, mypcol) relates to the rank in comm. */
DLAF_EXTERN_C char grid_ordering(MPI_Comm comm, int nprow, int npcol, int myprow, int mypcol) DLAF_NOEXCEPT;
/* Stand-in for dlaf_create_grid_from_blacs(blacs_ctxt) when no BLACS is linked: registers a grid under the
 * caller's BLACS context number so that dlaf_p?potrf (which reads the context from desca[1]) finds it. */
DLAF_EXTERN_C void dlaf_create_grid_from_blacs_info(int blacs_ctxt, MPI_Comm comm, int nprow, int npcol,
                                                    char order) DLAF_NOEXCEPT;
