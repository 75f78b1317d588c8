// miniapp_triangular_solver: benchmark driver of the triangular solver with the reference's flags and output lines
// (miniapp/miniapp_triangular_solver.cpp:88-206, :218-246): op(A) X = 2 B / X op(A) = 2 B, timed region = the solve
// only, device-resident operands, barriers on both sides, GFlop/s = total_ops<T>(m n k / 2, m n k / 2) / t with k the
// order of A.
//
//   miniapp_triangular_solver --m 32768 --n 32768 --mb 512 --nb 512 --side L --uplo L --op N --diag N --type d --nruns 3
//
// One process per GPU for --grid-rows x --grid-cols > 1 (RANK / WORLD_SIZE / LOCAL_RANK, see miniapp_cholesky.cpp).
// As in the reference, --check-result is not implemented for this miniapp (:74-78).
#include <chrono>
#include <iostream>

#include <dlaf/solver/triangular.h>

#include <dlaf/communication/init.h>

#include "miniapp_common.h"

using namespace dlaf;

struct Options : miniapp::CommonOptions {
  SizeType m = 4096, n = 512, mb = 256, nb = 512;  // miniapp_triangular_solver.cpp:231-234
  blas::Side side = blas::Side::Left;
  blas::Uplo uplo = blas::Uplo::Lower;
  blas::Op op = blas::Op::NoTrans;
  blas::Diag diag = blas::Diag::NonUnit;
};

static Options parse(int argc, char** argv) {
  Options o;
  for (int i = 1; i < argc; ++i) {
    const char* v;
    if ((v = miniapp::option_value(argc, argv, i, "m"))) o.m = std::atoll(v);
    else if ((v = miniapp::option_value(argc, argv, i, "n"))) o.n = std::atoll(v);
    else if ((v = miniapp::option_value(argc, argv, i, "mb"))) o.mb = std::atoll(v);
    else if ((v = miniapp::option_value(argc, argv, i, "nb"))) o.nb = std::atoll(v);
    else if ((v = miniapp::option_value(argc, argv, i, "side"))) o.side = (v[0] == 'R' || v[0] == 'r') ? blas::Side::Right : blas::Side::Left;
    else if ((v = miniapp::option_value(argc, argv, i, "uplo"))) o.uplo = (v[0] == 'U' || v[0] == 'u') ? blas::Uplo::Upper : blas::Uplo::Lower;
    else if ((v = miniapp::option_value(argc, argv, i, "op")))
      o.op = (v[0] == 'T' || v[0] == 't') ? blas::Op::Trans : (v[0] == 'C' || v[0] == 'c') ? blas::Op::ConjTrans : blas::Op::NoTrans;
    else if ((v = miniapp::option_value(argc, argv, i, "diag"))) o.diag = (v[0] == 'U' || v[0] == 'u') ? blas::Diag::Unit : blas::Diag::NonUnit;
    else if (std::string(argv[i]) == "--help") {
      std::puts("miniapp_triangular_solver --m M --n N --mb MB --nb NB --side L|R --uplo L|U --op N|T|C --diag N|U "
                "--grid-rows P --grid-cols Q --nruns R --nwarmups W --type s|d|c|z [--csv] [--local]");
      std::exit(0);
    }
    else
      miniapp::parse_common(argc, argv, i, o, "miniapp_triangular_solver");
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
  const bool left = opts.side == blas::Side::Left;
  const SizeType k = left ? opts.m : opts.n, kb = left ? opts.mb : opts.nb;
  Matrix<T, Device::GPU> a(GlobalElementSize(k, k), TileElementSize(kb, kb), comm_grid);
  miniapp::generate(a, 7);
  Matrix<T, Device::GPU> b_ref(GlobalElementSize(opts.m, opts.n), TileElementSize(opts.mb, opts.nb), comm_grid);
  miniapp::generate(b_ref, 11);
  Matrix<T, Device::GPU> b(GlobalElementSize(opts.m, opts.n), TileElementSize(opts.mb, opts.nb), comm_grid);
  const T alpha = 2.0;
  const double m = (double)opts.m, n = (double)opts.n;
  const double add_mul = n * m * (left ? m : n) / 2;
  const double total = dlaf::total_ops<T>(add_mul, add_mul);
  const char sd = left ? 'L' : 'R', ul = opts.uplo == blas::Uplo::Lower ? 'L' : 'U';
  const char op = opts.op == blas::Op::NoTrans ? 'N' : opts.op == blas::Op::Trans ? 'T' : 'C';
  const char dg = opts.diag == blas::Diag::Unit ? 'U' : 'N';

  for (int64_t run_index = -opts.nwarmups; run_index < opts.nruns; ++run_index) {
    if (rank == 0 && run_index >= 0)
      std::cout << "[" << run_index << "]" << std::endl;
    const auto ls = b.distribution().local_size();
    if (ls.rows() > 0 && ls.cols() > 0)
      cudaMemcpy2DAsync(b.ptr(), b.ld() * sizeof(T), b_ref.ptr(), b_ref.ld() * sizeof(T), ls.rows() * sizeof(T),
                        ls.cols(), cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(b.stream()));
    a.waitLocalTiles();
    b_ref.waitLocalTiles();
    b.waitLocalTiles();
    comm_grid.barrier();
    const auto t0 = std::chrono::steady_clock::now();
    if (opts.local)
      dlaf::triangular_solver<Backend::GPU, Device::GPU, T>(opts.side, opts.uplo, opts.op, opts.diag, alpha, a, b);
    else
      dlaf::triangular_solver<Backend::GPU, Device::GPU, T>(comm_grid, opts.side, opts.uplo, opts.op, opts.diag, alpha,
                                                            a, b);
    a.waitLocalTiles();
    b.waitLocalTiles();
    comm_grid.barrier();
    const double elapsed_time = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (rank == 0 && run_index >= 0) {
      const double gigaflops = total / elapsed_time / 1e9;
      std::cout << "[" << run_index << "]"
                << " " << elapsed_time << "s"
                << " " << gigaflops << "GFlop/s"
                << " " << miniapp::type_name(opts.type) << sd << ul << op << dg << " " << b.size() << " "
                << b.block_size() << " " << comm_grid.size() << " " << 1 << " GPU" << std::endl;
      if (opts.csv)
        std::cout << "CSVData-2, run, " << run_index << ", time, " << elapsed_time << ", GFlops, " << gigaflops
                  << ", type, " << miniapp::type_name(opts.type) << ", size, " << sd << ", uplo, " << ul << ", op, " << op
                  << ", diag, " << dg << ", matrixsize, " << opts.m << ", blocksize, " << opts.mb << ", comm_rows, "
                  << comm_grid.size().rows() << ", comm_cols, " << comm_grid.size().cols() << ", threads, 1"
                  << ", backend, GPU, " << std::endl;
    }
  }
}

int main(int argc, char** argv) {
  const Options opts = parse(argc, argv);
  dlaf::comm::mpi_init mpi_initter(argc, argv);
  const int rank = mpi_initter.rank(), size = mpi_initter.size();
  if (opts.grid_rows * opts.grid_cols != size) {
    std::fprintf(stderr, "miniapp_triangular_solver: grid %dx%d needs %d ranks, WORLD_SIZE is %d\n", opts.grid_rows,
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
