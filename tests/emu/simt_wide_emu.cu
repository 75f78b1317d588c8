// TEST INFRASTRUCTURE ONLY. CPU execution of the wide float contraction kernel (dla-future_b200/csrc/
// gemm_simt_wide.cuh) for boxes without a GPU: the product's own kernel body, one host thread per CUDA thread of a
// block, a real barrier for __syncthreads and a real shared-memory slice per block - so the loader mapping, the
// register prefetch / barrier protocol (a missing barrier is a data race here as on the device; run under the barrier
// it either deadlocks or corrupts the result), the 8 x 8 ownership map, clipping, the lower-triangle mask, LIST mode
// with tiled operands and the block-count arithmetic of the launcher are all exercised. Built host-only by nvcc.
#include <condition_variable>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

#include "gemm_simt_wide.cuh"

namespace {
using namespace dlafb;

struct Barrier {
  std::mutex mu;
  std::condition_variable cv;
  int size, arrived = 0;
  unsigned long long gen = 0;
  explicit Barrier(int n) : size(n) {}
  void wait() {
    std::unique_lock<std::mutex> lk(mu);
    const unsigned long long g = gen;
    if (++arrived == size) {
      arrived = 0;
      ++gen;
      cv.notify_all();
    }
    else
      cv.wait(lk, [&] { return gen != g; });
  }
};

struct HostCtx {
  int tid_, bx_, by_;
  float* sm;
  Barrier* bar;
  int tid() const { return tid_; }
  int bx() const { return bx_; }
  int by() const { return by_; }
  float* smem() const { return sm; }
  void sync() const { bar->wait(); }
};

// Runs the grid block after block; every block with kWideThreads host threads.
void run_grid(const SimtParams<float>& p, unsigned nx, unsigned ny) {
  for (unsigned by = 0; by < ny; ++by)
    for (unsigned bx = 0; bx < nx; ++bx) {
      alignas(16) static float sm[kWideSmemFloats];
      for (int i = 0; i < kWideSmemFloats; ++i)
        sm[i] = std::numeric_limits<float>::quiet_NaN();
      Barrier bar(kWideThreads);
      std::vector<std::thread> th;
      th.reserve(kWideThreads);
      for (int t = 0; t < kWideThreads; ++t)
        th.emplace_back([&, t] {
          HostCtx c{t, (int)bx, (int)by, sm, &bar};
          gemm_simt_wide_body(p, c);
        });
      for (auto& x : th)
        x.join();
    }
}

}  // namespace

extern "C" {

// RECT mode as dlafb::gemm_nt<float> issues it (tile_ops.cu): blocks counted in 64 x 64 units by the caller, converted
// with the launcher's rule (gemm_simt.cu: launch_wide).
int emu_sgemm_wide_rect(int M, int N, int K, float alpha, const float* A, long long lda, const float* B, long long ldb,
                        float beta, float* C, long long ldc, int lower) {
  SimtParams<float> p{};
  p.C = C;
  p.ldc = ldc;
  p.A = A;
  p.lda = lda;
  p.B = B;
  p.ldb = ldb;
  p.m_total = M;
  p.n_total = N;
  p.K = K;
  p.alpha = alpha;
  p.beta = beta;
  p.lower = lower;
  const unsigned nbx = (unsigned)ceil_div(M, kSimtTile), nby = (unsigned)ceil_div(N, kSimtTile);
  run_grid(p, (nbx + 1) / 2, (nby + 1) / 2);
  return 0;
}

// LIST mode as the factorization engine issues it for s (cholesky_engine.cu: update with tile lists): C -= A B^T on
// the listed nb x nb tiles, entry = (first row, first col | diag << 31); operands either the local block itself
// (lda = ldb = ld, not tiled) or panels stored as contiguous nb x nb tiles.
int emu_sgemm_wide_list(int m_total, int n_total, int nb, float alpha, float beta, const float* A, long long lda,
                        int a_tiled, const float* B, long long ldb, int b_tiled, float* C, long long ldc,
                        const int* list_xy, int count) {
  std::vector<int2> list(count);
  for (int i = 0; i < count; ++i) {
    list[i].x = list_xy[2 * i];
    list[i].y = list_xy[2 * i + 1];
  }
  SimtParams<float> p{};
  p.C = C;
  p.ldc = ldc;
  p.A = A;
  p.lda = lda;
  p.B = B;
  p.ldb = ldb;
  p.m_total = m_total;
  p.n_total = n_total;
  p.K = nb;
  p.alpha = alpha;
  p.beta = beta;
  p.list = list.data();
  p.sub = (nb + kSimtTile - 1) / kSimtTile;
  p.tile = nb;
  p.a_tiled = a_tiled;
  p.b_tiled = b_tiled;
  const unsigned nblocks64 = (unsigned)(count * p.sub * p.sub);
  const unsigned items = nblocks64 / (unsigned)(p.sub * p.sub);
  const unsigned sub = (unsigned)((nb + kWideTile - 1) / kWideTile);
  run_grid(p, items * sub * sub, 1);
  return 0;
}

// The arithmetic every element must reproduce bit for bit: acc = fmaf(a_ik, b_jk, acc) over k = 0, 1, ...; then
// alpha * acc (+ beta * c). Compiled in this translation unit with the same floating-point contraction setting.
void emu_sgemm_scalar(int M, int N, int K, float alpha, const float* A, long long lda, const float* B, long long ldb,
                      float beta, float* C, long long ldc, int lower) {
  for (int j = 0; j < N; ++j)
    for (int i = lower ? j : 0; i < M; ++i) {
      float acc = 0.f;
      for (int k = 0; k < K; ++k)
        acc = fmaf(A[i + k * lda], B[j + k * ldb], acc);
      float v = acc * alpha;
      if (beta != 0.f)
        v = v + C[i + j * ldc] * beta;
      C[i + j * ldc] = v;
    }
}

}  // extern "C"
