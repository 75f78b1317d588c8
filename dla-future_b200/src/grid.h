// Process grid: P x Q ranks, one GPU per rank, row / column / world NCCL communicators over NVLink.
// Stands in for comm::CommunicatorGrid (include/dlaf/communication/communicator_grid.h:37-153): the reference
// splits an MPI communicator three ways (src/communication/communicator_grid.cpp:36-38); here the split is
// ncclCommSplit and collectives are enqueued on CUDA streams in program order, which replaces the reference's
// communicator pipelines (include/dlaf/communication/communicator_pipeline.h:41-148).
#pragma once
#include <cuda_runtime.h>
#include <nccl.h>

struct dlafb_grid {
  int P = 1, Q = 1, pr = 0, pc = 0;
  int world_rank = 0, world_size = 1;
  char ordering = 'R';
  ncclComm_t world = nullptr, row = nullptr, col = nullptr;
  bool owns_world = true;
  cudaStream_t stream = nullptr;  // for barriers / small reductions
  double* scratch = nullptr;
};

extern "C" int dlafb_grid_create_from_nccl(dlafb_grid** out, ncclComm_t world, int owns_world, int world_rank,
                                           int world_size, int P, int Q, char ordering);
