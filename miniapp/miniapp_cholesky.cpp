// miniapp_cholesky: benchmark driver of the Cholesky path with the reference's flags and output lines
// (miniapp/miniapp_cholesky.cpp:107-235, miniapp/include/dlaf/miniapp/options.h): timed region = factorization only,
// device-resident input, barriers on both sides, GFlop/s = total_ops<T>(n^3/6, n^3/6) / t.
//
//   miniapp_cholesky --matrix-size 32768 --block-size 512 --type d --nruns 5 --nwarmups 1 --check-result last
//
// Multi-rank runs: one process per GPU. Without MPI the launcher provides RANK / WORLD_SIZE / LOCAL_RANK (torchrun
// style) and the ranks share the NCCL id through the file named by DLAFB_ID_FILE (default /tmp/dlafb_nccl_id).
// Differences to the reference, all outside the timed region: the input is generated on the device by the
// counter-based generator (same distribution of values as set_random_hermitian_positive_definite), and the
// residual check runs on the GPU; multi-rank runs use the engine's distributed checker (uplo = lower).
#include <unistd.h>

#include <chrono>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <limits>
#include <string>

#include <dlaf/factorization/cholesky.h>
#include <dlaf/init.h>
#include <dlaf/matrix/matrix.h>
#include <dlafb_kernels.h>

using namespace dlaf;

struct Options {
  SizeType m = 4096, mb = 256;  // miniapp_cholesky.cpp:223-224
  int grid_rows = 1, grid_cols = 1;
  int64_t nruns = 1, nwarmups = 1;  // options.h:252-255
  char type = 'd';
  blas::Uplo uplo = blas::Uplo::Lower;
  std::string check = "none";
  bool csv = false, local = false;
};

static Options parse(int argc, char** argv) {
  Options o;
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    auto val = [&](const char* name) -> const char* {
      const std::string n = std::string("--") + name;
      if (a == n && i + 1 < argc)
        return argv[++i];
      if (a.rfind(n + "=", 0) == 0)
        return argv[i] + n.size() + 1;
      return nullptr;
    };
    const char* v;
    if ((v = val("matrix-size"))) o.m = std::atoll(v);
    else if ((v = val("block-size"))) o.mb = std::atoll(v);
    else if ((v = val("grid-rows"))) o.grid_rows = std::atoi(v);
    else if ((v = val("grid-cols"))) o.grid_cols = std::atoi(v);
    else if ((v = val("nruns"))) o.nruns = std::atoll(v);
    else if ((v = val("nwarmups"))) o.nwarmups = std::atoll(v);
    else if ((v = val("type"))) o.type = v[0];
    else if ((v = val("uplo"))) o.uplo = (v[0] == 'u' || v[0] == 'U') ? blas::Uplo::Upper : blas::Uplo::Lower;
    else if ((v = val("check-result"))) o.check = v;
    else if ((v = val("backend"))) {
      if (std::string(v) == "mc") {
        std::fprintf(stderr, "miniapp_cholesky: this build has no MC (CPU) backend\n");
        std::exit(2);
      }
    }
    else if (a == "--csv") o.csv = true;
    else if (a == "--local") o.local = true;
    else if (a == "--help") {
      std::puts("miniapp_cholesky --matrix-size N --block-size NB --grid-rows P --grid-cols Q --nruns R --nwarmups W "
                "--type s|d|c|z --uplo lower|upper --check-result none|last|all [--csv] [--local]");
      std::exit(0);
    }
  }
  return o;
}

static void bootstrap(int& rank, int& size) {
  rank = std::getenv("RANK") ? std::atoi(std::getenv("RANK")) : 0;
  size = std::getenv("WORLD_SIZE") ? std::atoi(std::getenv("WORLD_SIZE")) : 1;
  const int local = std::getenv("LOCAL_RANK") ? std::atoi(std::getenv("LOCAL_RANK")) : rank;
  cudaSetDevice(local);
  unsigned char id[128] = {0};
  const char* path = std::getenv("DLAFB_ID_FILE") ? std::getenv("DLAFB_ID_FILE") : "/tmp/dlafb_nccl_id";
  if (size > 1) {
    if (rank == 0) {
      const std::string tmp = std::string(path) + ".tmp";
      dlafb_nccl_unique_id(id);
      FILE* f = std::fopen(tmp.c_str(), "wb");
      std::fwrite(id, 1, 128, f);
      std::fclose(f);
      std::rename(tmp.c_str(), path);
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
  if (dlafb_mpi_init(rank, size, size > 1 ? id : nullptr) != 0) {
    std::fprintf(stderr, "miniapp_cholesky: communicator bootstrap failed\n");
    std::exit(1);
  }
  if (size > 1 && rank == 0)
    std::remove(path);  // every rank has joined the communicator once dlafb_mpi_init returns
}

// element-type overloads of the C ABI helpers used by the generator and the checker
#define MINIAPP_OVERLOADS(T, CT, X)                                                                            \
  static int k_gen(void* s, int64_t n, int nb, int r, int c, int rf, int rs, int cf, int cs, T* a, int64_t ld, \
                   uint64_t seed) {                                                                            \
    return dlafb_gen_hpd_##X(s, n, nb, r, c, rf, rs, cf, cs, (CT*)a, ld, seed);                                \
  }                                                                                                            \
  static int k_gemm(void* s, int m, int n, int k, double al, const T* a, int64_t lda, const T* b, int64_t ldb, \
                    double be, T* c, int64_t ldc, int lower) {                                                 \
    return dlafb_gemm_nt_##X(s, m, n, k, al, (const CT*)a, lda, (const CT*)b, ldb, be, (CT*)c, ldc, lower);    \
  }                                                                                                            \
  static int k_laset0(void* s, char u, int m, int n, T* a, int64_t ld) {                                       \
    return dlafb_laset_##X(s, u, m, n, 0, 0, 0, 0, (CT*)a, ld);                                                \
  }                                                                                                            \
  static int k_maxabs(void* s, char u, int m, int n, const T* a, int64_t ld, double* out) {                    \
    return dlafb_max_abs_##X(s, u, m, n, (const CT*)a, ld, out);                                               \
  }                                                                                                            \
  static int k_transpose(void* s, int m, int n, const T* in, int64_t ldi, T* out, int64_t ldo) {               \
    return dlafb_transpose_##X(s, m, n, (const CT*)in, ldi, (CT*)out, ldo, 1, 'G');                            \
  }
MINIAPP_OVERLOADS(float, float, s)
MINIAPP_OVERLOADS(double, double, d)
MINIAPP_OVERLOADS(std::complex<float>, void, c)
MINIAPP_OVERLOADS(std::complex<double>, void, z)

template <class T>
static void generate(Matrix<T, Device::GPU>& m, uint64_t seed) {
  const auto& d = m.distribution();
  const auto ls = d.local_size();
  k_gen(m.stream(), d.size().rows(), (int)d.block_size().rows(), (int)ls.rows(), (int)ls.cols(),
        (int)d.template global_tile_from_local_tile<Coord::Row>(0), d.grid_size().rows(),
        (int)d.template global_tile_from_local_tile<Coord::Col>(0), d.grid_size().cols(), m.ptr(), m.ld(), seed);
}

// max|A - L L^H| / max|A| over the referenced triangle, thresholds eps*n (warning) and 100*eps*n (error):
// miniapp_cholesky.cpp:408-442. GPU version for single-rank runs; `a` is consumed.
template <class T>
static void check_cholesky(Matrix<T, Device::GPU>& a, Matrix<T, Device::GPU>& fac, blas::Uplo uplo) {
  const int n = (int)a.size().rows();
  if (n == 0)
    return;
  fac.waitLocalTiles();
  a.waitLocalTiles();
  void* s = a.stream();
  T *pa = a.ptr(), *pl = fac.ptr();
  int64_t lda = a.ld(), ldl = fac.ld();
  Matrix<T, Device::GPU> ta(a.distribution()), tl(a.distribution());
  if (uplo == blas::Uplo::Upper) {  // U^H U on the upper triangle == L L^H on the lower triangle of the transposes
    k_transpose(s, n, n, pa, lda, ta.ptr(), ta.ld());
    k_transpose(s, n, n, pl, ldl, tl.ptr(), tl.ld());
    pa = ta.ptr(), pl = tl.ptr(), lda = ta.ld(), ldl = tl.ld();
  }
  double* dv = nullptr;
  cudaMalloc(&dv, 2 * sizeof(double));
  cudaMemsetAsync(dv, 0, 2 * sizeof(double), a.stream());
  k_maxabs(s, 'L', n, n, pa, lda, dv);
  if (n > 1)
    k_laset0(s, 'U', n - 1, n - 1, pl + ldl, ldl);  // strict upper triangle of the factor <- 0
  k_gemm(s, n, n, n, -1.0, pl, ldl, pl, ldl, 1.0, pa, lda, 1);
  k_maxabs(s, 'L', n, n, pa, lda, dv + 1);
  double hv[2];
  cudaMemcpyAsync(hv, dv, sizeof(hv), cudaMemcpyDeviceToHost, a.stream());
  cudaStreamSynchronize(a.stream());
  cudaFree(dv);
  const double diff = hv[1] / hv[0];
  const double eps = std::numeric_limits<BaseType<T>>::epsilon();
  if (diff > 100 * eps * n)
    std::cout << "ERROR: ";
  else if (diff > eps * n)
    std::cout << "Warning: ";
  std::cout << "Max Diff / Max A: " << diff << std::endl;
}

template <class T>
static void run(const Options& opts, int rank, int size) {
  comm::Communicator world(MPI_COMM_WORLD);
  comm::CommunicatorGrid comm_grid(world, opts.grid_rows, opts.grid_cols, common::Ordering::ColumnMajor);
  const GlobalElementSize matrix_size(opts.m, opts.m);
  const TileElementSize block_size(opts.mb, opts.mb);
  const char* type_name = opts.type == 's' ? "S" : opts.type == 'd' ? "D" : opts.type == 'c' ? "C" : "Z";

  for (int64_t run_index = -opts.nwarmups; run_index < opts.nruns; ++run_index) {
    if (rank == 0 && run_index >= 0)
      std::cout << "[" << run_index << "]" << std::endl;
    Matrix<T, Device::GPU> matrix(matrix_size, block_size, comm_grid);
    generate(matrix, 42);
    matrix.waitLocalTiles();
    comm_grid.barrier();

    const auto t0 = std::chrono::steady_clock::now();
    if (opts.local)
      dlaf::cholesky_factorization<Backend::GPU, Device::GPU, T>(opts.uplo, matrix);
    else
      dlaf::cholesky_factorization<Backend::GPU, Device::GPU, T>(comm_grid, opts.uplo, matrix);
    matrix.waitLocalTiles();
    comm_grid.barrier();
    const double elapsed_time = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    const double n = (double)opts.m;
    const double add_mul = n * n * n / 6;
    const double gigaflops = dlaf::total_ops<T>(add_mul, add_mul) / elapsed_time / 1e9;
    if (rank == 0 && run_index >= 0) {
      std::cout << "[" << run_index << "]"
                << " " << elapsed_time << "s"
                << " " << gigaflops << "GFlop/s"
                << " " << type_name << (opts.uplo == blas::Uplo::Lower ? "L" : "U") << " " << matrix.size() << " "
                << matrix.block_size() << " " << comm_grid.size() << " " << 1 << " GPU" << std::endl;
      if (opts.csv)
        std::cout << "CSVData-2, run, " << run_index << ", time, " << elapsed_time << ", GFlops, " << gigaflops
                  << ", type, " << type_name << ", UpLo, " << (opts.uplo == blas::Uplo::Lower ? "L" : "U")
                  << ", matrixsize, " << opts.m << ", blocksize, " << opts.mb << ", comm_rows, "
                  << comm_grid.size().rows() << ", comm_cols, " << comm_grid.size().cols() << ", threads, 1"
                  << ", backend, GPU, " << std::endl;
    }
    const bool do_check = (opts.check == "all") || (opts.check == "last" && run_index == opts.nruns - 1);
    if (do_check) {
      if (opts.m % opts.mb != 0) {
        if (rank == 0)
          std::cout << "Warning! At the moment result checking works just with block sizes dividing matrix size."
                    << std::endl;  // miniapp_cholesky.cpp:90-95
      }
      else if (size > 1 && opts.uplo == blas::Uplo::Upper) {
        if (rank == 0)
          std::cout << "check-result: the distributed residual check covers uplo = lower only" << std::endl;
      }
      else if (size > 1) {
        // distributed, on the device: A - L L^H through the engine's own exchange and update kernels
        Matrix<T, Device::GPU> original(matrix_size, block_size, comm_grid);
        generate(original, 42);
        original.waitLocalTiles();
        dlafb_chol_plan_t plan = nullptr;
        double amax = 0, dmax = 0;
        int e = dlafb_cholesky_plan_create(&plan, comm_grid.handle(), dlaf::internal::TypeCode<T>::value, 'L', opts.m,
                                           opts.mb, 0, 0);
        if (e == 0)
          e = dlafb_cholesky_plan_residual(plan, original.ptr(), original.ld(), matrix.ptr(), matrix.ld(),
                                           original.stream(), &amax, &dmax);
        if (plan)
          dlafb_cholesky_plan_destroy(plan);
        if (rank == 0) {
          const double diff = amax > 0 ? dmax / amax : 0.0;
          const double eps = std::numeric_limits<BaseType<T>>::epsilon();
          if (e != 0)
            std::cout << "ERROR: residual check failed to run (" << e << ")" << std::endl;
          else {
            if (diff > 100 * eps * opts.m)
              std::cout << "ERROR: ";
            else if (diff > eps * opts.m)
              std::cout << "Warning: ";
            std::cout << "Max Diff / Max A: " << diff << std::endl;
          }
        }
      }
      else {
        Matrix<T, Device::GPU> original(matrix_size, block_size, comm_grid);
        generate(original, 42);
        check_cholesky(original, matrix, opts.uplo);
      }
    }
  }
}

int main(int argc, char** argv) {
  const Options opts = parse(argc, argv);
  int rank = 0, size = 1;
  bootstrap(rank, size);
  if (opts.grid_rows * opts.grid_cols != size) {
    std::fprintf(stderr, "miniapp_cholesky: grid %dx%d needs %d ranks, WORLD_SIZE is %d\n", opts.grid_rows,
                 opts.grid_cols, opts.grid_rows * opts.grid_cols, size);
    return 1;
  }
  {
    dlaf::ScopedInitializer init(argc, argv);
    switch (opts.type) {
      case 's': run<float>(opts, rank, size); break;
      case 'd': run<double>(opts, rank, size); break;
      case 'c': run<std::complex<float>>(opts, rank, size); break;
      case 'z': run<std::complex<double>>(opts, rank, size); break;
      default: std::fprintf(stderr, "unknown --type %c\n", opts.type); return 1;
    }
  }
  dlafb_mpi_finalize();
  return 0;
}
