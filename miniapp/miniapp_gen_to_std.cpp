// miniapp_gen_to_std: benchmark driver of the generalized-to-standard reduction with the reference's flags and output
// lines (miniapp/miniapp_gen_to_std.cpp:79-186): A Hermitian, B = a triangular factor with a strictly positive diagonal
// (the reference fills it with set_random_hermitian_positive_definite and reads one triangle, :99-109), timed region =
// the reduction only, device-resident operands, barriers on both sides, GFlop/s = total_ops<T>(n^3/2, n^3/2) / t.
//
//   miniapp_gen_to_std --matrix-size 32768 --block-size 512 --uplo L --type d --nruns 3
//
// One process per GPU for --grid-rows x --grid-cols > 1. --check-result is not implemented, as in the reference (:63-67).
#include <chrono>
#include <iostream>

#include <dlaf/eigensolver/gen_to_std.h>

#include <dlaf/communication/init.h>

#include "miniapp_common.h"

using namespace dlaf;

struct Options : miniapp::CommonOptions {
  SizeType m = 4096, mb = 256;
  blas::Uplo uplo = blas::Uplo::Lower;
};

static Options parse(int argc, char** argv) {
  Options o;
  for (int i = 1; i < argc; ++i) {
    const char* v;
    if ((v = miniapp::option_value(argc, argv, i, "matrix-size"))) o.m = std::atoll(v);
    else if ((v = miniapp::option_value(argc, argv, i, "block-size"))) o.mb = std::atoll(v);
    else if ((v = miniapp::option_value(argc, argv, i, "uplo"))) o.uplo = (v[0] == 'U' || v[0] == 'u') ? blas::Uplo::Upper : blas::Uplo::Lower;
    else if (std::string(argv[i]) == "--help") {
      std::puts("miniapp_gen_to_std --matrix-size N --block-size NB --uplo L|U --grid-rows P --grid-cols Q --nruns R "
                "--nwarmups W --type s|d|c|z [--csv] [--local]");
      std::exit(0);
    }
    else
      miniapp::parse_common(argc, argv, i, o, "miniapp_gen_to_std");
  }
  if (o.check != "none") {
    std::cerr << "Warning! At the moment result checking it is not implemented." << std::endl;
    o.check = "none";
  }
  return o;
}

template <class T>
static void run(const Options& opts, int rank) {
  comm::Communicator world(MPI_COMM_WORLD);
  comm::CommunicatorGrid comm_grid(world, opts.grid_rows, opts.grid_cols, common::Ordering::ColumnMajor);
  const GlobalElementSize size(opts.m, opts.m);
  const TileElementSize block(opts.mb, opts.mb);
  Matrix<T, Device::GPU> a_ref(size, block, comm_grid), factor(size, block, comm_grid), a(size, block, comm_grid);
  miniapp::generate(a_ref, 13);
  miniapp::generate(factor, 17);
  const char ul = opts.uplo == blas::Uplo::Lower ? 'L' : 'U';
  for (int64_t run_index = -opts.nwarmups; run_index < opts.nruns; ++run_index) {
    if (rank == 0 && run_index >= 0)
      std::cout << "[" << run_index << "]" << std::endl;
    const auto ls = a.distribution().local_size();
    if (ls.rows() > 0 && ls.cols() > 0)
      cudaMemcpy2DAsync(a.ptr(), a.ld() * sizeof(T), a_ref.ptr(), a_ref.ld() * sizeof(T), ls.rows() * sizeof(T),
                        ls.cols(), cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(a.stream()));
    a_ref.waitLocalTiles();
    factor.waitLocalTiles();
    a.waitLocalTiles();
    comm_grid.barrier();
    const auto t0 = std::chrono::steady_clock::now();
    if (opts.local)
      eigensolver::internal::generalized_to_standard<Backend::GPU, Device::GPU, T>(opts.uplo, a, factor);
    else
      eigensolver::internal::generalized_to_standard<Backend::GPU, Device::GPU, T>(comm_grid, opts.uplo, a, factor);
    a.waitLocalTiles();
    factor.waitLocalTiles();
    comm_grid.barrier();
    const double elapsed_time = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (rank == 0 && run_index >= 0) {
      const double n = (double)opts.m;
      const double add_mul = n * n * n / 2;
      const double gigaflops = dlaf::total_ops<T>(add_mul, add_mul) / elapsed_time / 1e9;
      std::cout << "[" << run_index << "]"
                << " " << elapsed_time << "s"
                << " " << gigaflops << "GFlop/s"
                << " " << miniapp::type_name(opts.type) << ul << " " << a.size() << " " << a.block_size() << " "
                << comm_grid.size() << " " << 1 << " GPU" << std::endl;
      if (opts.csv)
        std::cout << "CSVData-2, run, " << run_index << ", time, " << elapsed_time << ", GFlops, " << gigaflops
                  << ", type, " << miniapp::type_name(opts.type) << ", uplo, " << ul << ", matrixsize, " << opts.m
                  << ", blocksize, " << opts.mb << ", comm_rows, " << comm_grid.size().rows() << ", comm_cols, "
                  << comm_grid.size().cols() << ", threads, 1, backend, GPU, " << std::endl;
    }
  }
}

int main(int argc, char** argv) {
  const Options opts = parse(argc, argv);
  dlaf::comm::mpi_init mpi_initter(argc, argv);
  const int rank = mpi_initter.rank(), size = mpi_initter.size();
  if (opts.grid_rows * opts.grid_cols != size) {
    std::fprintf(stderr, "miniapp_gen_to_std: grid %dx%d needs %d ranks, WORLD_SIZE is %d\n", opts.grid_rows,
                 opts.grid_cols, opts.grid_rows * opts.grid_cols, size);
    return 1;
  }
  {
    dlaf::ScopedInitializer init(argc, argv);
    switch (opts.type) {
      case 's': run<float>(opts, rank); break;
      case 'd': run<double>(opts, rank); break;
      case 'c': run<std::complex<float>>(opts, rank); break;
      case 'z': run<std::complex<double>>(opts, rank); break;
      default: std::fprintf(stderr, "unknown --type %c\n", opts.type); return 1;
    }
  }
  return 0;
}
