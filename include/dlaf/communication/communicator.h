// comm::Communicator: thin value wrapper of an MPI_Comm (reference: include/dlaf/communication/communicator.h,
// src/communication/communicator_impl.cpp). <mpi.h> is the single-node shim include/dlafb_mpi_shim/mpi.h.
#pragma once
#include <mpi.h>

#include "dlaf/common/index2d.h"

namespace dlaf::comm {

class Communicator {
public:
  Communicator() : comm_(MPI_COMM_NULL) {}
  Communicator(MPI_Comm mpi_communicator) : comm_(mpi_communicator) {
    if (comm_ != MPI_COMM_NULL) {
      MPI_Comm_rank(comm_, &rank_);
      MPI_Comm_size(comm_, &size_);
    }
  }
  operator MPI_Comm() const noexcept { return comm_; }
  IndexT_MPI rank() const noexcept { return rank_; }
  IndexT_MPI size() const noexcept { return size_; }

private:
  MPI_Comm comm_;
  IndexT_MPI rank_ = 0, size_ = 1;
};

}  // namespace dlaf::comm
