/* Minimal single-node stand-in for <mpi.h>, used when no MPI is installed (this image has none).
 * It keeps the MPI_Comm-typed signatures of the reference's C and C++ APIs compiling and gives them meaning for
 * the one-process-per-GPU launch this build targets: a communicator is {rank, size, NCCL unique id}; the
 * launcher (torchrun + torch.distributed, or any other) distributes the unique id and calls dlafb_mpi_init.
 * An application that uses a real MPI bootstraps this handle with its own rank / size and an id it broadcasts with
 * MPI_Bcast, from a translation unit that includes this header and not the real <mpi.h> (both define MPI_Comm);
 * passing real MPI communicators into the library is not supported in this build. */
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
