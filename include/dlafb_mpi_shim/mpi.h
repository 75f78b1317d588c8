/* Minimal single-node stand-in for <mpi.h>, used when no MPI is installed (this image has none).
 * It keeps the MPI_Comm-typed signatures of the reference's C and C++ APIs compiling and gives them meaning for
 * the one-process-per-GPU launch this build targets: a communicator is {rank, size, NCCL unique id}; the
 * launcher (torchrun + torch.distributed, or any other) distributes the unique id and calls dlafb_mpi_init.
 * An application that uses a real MPI builds the library with -DDLAFB_WITH_MPI and its own <mpi.h> on the include
 * path instead of this directory: dlaf_create_grid / grid_ordering / dlaf_create_grid_from_blacs_info then take the real
 * communicator (rank and size from MPI, the NCCL id over MPI_Bcast, src/c_api.cu). That build is compile-checked
 * against a stub <mpi.h> only (tests/test_cxx_headers.py): this image has no MPI to run it with. */
#ifndef DLAFB_MPI_SHIM_H
#define DLAFB_MPI_SHIM_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dlafb_mpi_comm* MPI_Comm;
#define MPI_SUCCESS 0
#define MPI_COMM_NULL ((MPI_Comm)0)
/* The world communicator set up by dlafb_mpi_init. */
MPI_Comm dlafb_mpi_comm_world(void);
#define MPI_COMM_WORLD (dlafb_mpi_comm_world())

/* rank / size of this process, nccl_unique_id128: same 128 bytes on every rank (from dlafb_nccl_unique_id),
 * may be NULL when size == 1. */
int dlafb_mpi_init(int rank, int size, const void* nccl_unique_id128);
int dlafb_mpi_finalize(void);

int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Comm_size(MPI_Comm comm, int* size);

#ifdef __cplusplus
}
#endif
#endif
