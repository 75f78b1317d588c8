/* TEST INFRASTRUCTURE: declarations of the few MPI functions the -DDLAFB_WITH_MPI build of src/c_api.cu uses, with
 * MPICH-style handle types (MPI_Comm is an int), so that the real-MPI bootstrap can be COMPILE-checked on an image
 * without MPI (tests/test_cxx_headers.py). Nothing is ever linked or run against this file. */
#ifndef DLAFB_TEST_MPI_STUB_H
#define DLAFB_TEST_MPI_STUB_H
#ifdef __cplusplus
extern "C" {
#endif
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
#define MPI_SUCCESS 0
#define MPI_COMM_WORLD ((MPI_Comm)0x44000000)
#define MPI_COMM_NULL ((MPI_Comm)0x04000000)
#define MPI_BYTE ((MPI_Datatype)0x4c00010d)
#define MPI_INT ((MPI_Datatype)0x4c000405)
#define MPI_MIN ((MPI_Op)0x58000002)
int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Comm_size(MPI_Comm comm, int* size);
int MPI_Barrier(MPI_Comm comm);
int MPI_Bcast(void* buffer, int count, MPI_Datatype datatype, int root, MPI_Comm comm);
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype datatype, MPI_Op op, MPI_Comm comm);
#ifdef __cplusplus
}
#endif
#endif
