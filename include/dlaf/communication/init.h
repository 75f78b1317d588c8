// comm::mpi_init: RAII start-up of the communication layer, as the reference's miniapps and tests do first in main()
// (`dlaf::comm::mpi_init mpi_initter(argc, argv);`, include/dlaf/communication/init.h). There it calls MPI_Init_thread
// with MPI_THREAD_MULTIPLE; here there is no MPI: one process per GPU, the launcher provides RANK / WORLD_SIZE /
// LOCAL_RANK (torchrun style), and the processes share the 128-byte NCCL id through the file named by DLAFB_ID_FILE
// (default /tmp/dlafb_nccl_id; written by rank 0, removed once every rank has joined). Selects the CUDA device
// LOCAL_RANK. After construction MPI_COMM_WORLD (the shim's world handle) is valid.
#pragma once
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <string>

#include <cuda_runtime.h>
#include <mpi.h>

#include "dlafb_engine.h"

namespace dlaf::comm {

class mpi_init {
public:
  mpi_init(int /*argc*/, char** /*argv*/) {
    rank_ = std::getenv("RANK") ? std::atoi(std::getenv("RANK")) : 0;
    size_ = std::getenv("WORLD_SIZE") ? std::atoi(std::getenv("WORLD_SIZE")) : 1;
    const int local = std::getenv("LOCAL_RANK") ? std::atoi(std::getenv("LOCAL_RANK")) : rank_;
    cudaSetDevice(local);
    unsigned char id[128] = {0};
    const char* path = std::getenv("DLAFB_ID_FILE") ? std::getenv("DLAFB_ID_FILE") : "/tmp/dlafb_nccl_id";
    if (size_ > 1) {
      if (rank_ == 0) {
        const std::string tmp = std::string(path) + ".tmp";
        dlafb_nccl_unique_id(id);
        FILE* f = std::fopen(tmp.c_str(), "wb");
        if (f) {
          std::fwrite(id, 1, 128, f);
          std::fclose(f);
          std::rename(tmp.c_str(), path);
        }
      }
      else {
        bool ok = false;
        for (int tries = 0; tries < 600 && !ok; ++tries) {
          FILE* f = std::fopen(path, "rb");
          ok = f && std::fread(id, 1, 128, f) == 128;
          if (f)
            std::fclose(f);
          if (!ok)
            usleep(100000);
        }
      }
    }
    if (dlafb_mpi_init(rank_, size_, size_ > 1 ? id : nullptr) != 0) {
      std::fprintf(stderr, "[dlaf_b200] fatal: communicator bootstrap failed (rank %d of %d)\n", rank_, size_);
      std::abort();
    }
    if (size_ > 1 && rank_ == 0)
      std::remove(path);  // every rank has joined the communicator once dlafb_mpi_init returns
  }
  ~mpi_init() { dlafb_mpi_finalize(); }
  mpi_init(const mpi_init&) = delete;
  mpi_init& operator=(const mpi_init&) = delete;
  int rank() const noexcept { return rank_; }
  int size() const noexcept { return size_; }

private:
  int rank_ = 0, size_ = 1;
};

}  // namespace dlaf::comm
