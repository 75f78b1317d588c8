// Generic register-tiled contraction kernel (see gemm_simt.cuh).
#include "gemm_simt.cuh"

#include <cstdlib>
#include <type_traits>

#include "gemm_simt_wide.cuh"

namespace dlafb {
namespace {

// ---- wide float variant (gemm_simt_wide.cuh) ----
struct WideDeviceCtx {
  float* sm;
  __device__ int tid() const { return threadIdx.x; }
  __device__ int bx() const { return blockIdx.x; }
  __device__ int by() const { return blockIdx.y; }
  __device__ float* smem() const { return sm; }
  __device__ void sync() const { __syncthreads(); }
};

__global__ void __launch_bounds__(kWideThreads, 2) gemm_simt_wide_kernel(const SimtParams<float> p) {
  __shared__ __align__(16) float sm[kWideSmemFloats];
  WideDeviceCtx c{sm};
  gemm_simt_wide_body(p, c);
}

// float contractions run on the 128 x 128 kernel; DLAFB_SGEMM_WIDE=0 keeps the 64 x 64 one (read once).
bool wide_enabled() {
  static const bool on = [] {
    const char* e = getenv("DLAFB_SGEMM_WIDE");
    return !(e && e[0] == '0');
  }();
  return on;
}

template <class T>
bool launch_wide(cudaStream_t, const SimtParams<T>&, unsigned, unsigned, cudaError_t*) {
  return false;
}
template <>
bool launch_wide<float>(cudaStream_t stream, const SimtParams<float>& p, unsigned nblocks_x, unsigned nblocks_y,
                        cudaError_t* err) {
  if (!wide_enabled())
    return false;
  // the callers count 64 x 64 blocks: RECT mode ceil(M / 64) x ceil(N / 64); LIST mode items * sub^2, sub = ceil(tile / 64)
  unsigned nx, ny;
  if (p.list != nullptr) {
    if (p.sub <= 0)
      return false;
    const unsigned items = nblocks_x / (unsigned)(p.sub * p.sub);
    const unsigned sub = (unsigned)((p.tile + kWideTile - 1) / kWideTile);
    nx = items * sub * sub;
    ny = 1;
  }
  else {
    nx = (nblocks_x + 1) / 2;
    ny = (nblocks_y + 1) / 2;
  }
  gemm_simt_wide_kernel<<<dim3(nx, ny), kWideThreads, 0, stream>>>(p);
  ++g_launches;
  *err = cudaGetLastError();
  return true;
}

constexpr int kSK = 8;  // K-slice

template <class T>
__global__ void __launch_bounds__(256) gemm_simt_kernel(const SimtParams<T> p) {
  using R = typename Tr<T>::Real;
  int trow, tcol, row0, col0, lower, m_lim = p.m_total, n_lim = p.n_total;
  bool b_from_a = false;
  int b_row_a = 0;
  if (p.list != nullptr) {
    const int nsub = p.sub * p.sub;
    const int item = blockIdx.x / nsub;
    const int s = blockIdx.x - item * nsub;
    const int2 it = p.list[item];
    lower = it.y < 0;
    const int c0 = it.y & 0x7fffffff;
    trow = (s % p.sub) * kSimtTile;
    tcol = (s / p.sub) * kSimtTile;
    row0 = it.x + trow;
    col0 = c0 + tcol;
    b_from_a = lower;  // diagonal tile of the list: HERK, both operands from the row panel
    b_row_a = it.x + tcol + p.a_row_off;
    m_lim = min(m_lim, it.x + p.tile);
    n_lim = min(n_lim, c0 + p.tile);
  }
  else {
    lower = p.lower;
    trow = blockIdx.x * kSimtTile;
    tcol = blockIdx.y * kSimtTile;
    row0 = p.rect_row0 + trow;
    col0 = p.rect_col0 + tcol;
  }
  if (row0 >= m_lim || col0 >= n_lim)
    return;
  if (lower && trow + kSimtTile - 1 < tcol)
    return;

  __shared__ T As[kSK][kSimtTile + 1];
  __shared__ T Bs[kSK][kSimtTile + 1];

  const int tx = threadIdx.x & 15;
  const int ty = threadIdx.x >> 4;
  const int lr = threadIdx.x & 63;  // loader: row inside the block
  const int lk = threadIdx.x >> 6;  // loader: k inside the slice (0..3, +4)

  T acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
      acc[i][j] = Tr<T>::zero();

  const bool a_ok = row0 + lr < m_lim;
  const bool b_ok = col0 + lr < n_lim;
  auto operand = [&](const T* base, int row, int tiled) {
    return tiled ? base + (long long)(row / p.tile) * p.tile * p.tile + row % p.tile : base + row;
  };
  const T* ap = operand(p.A, row0 + lr + p.a_row_off, p.a_tiled);
  const T* bp = b_from_a ? operand(p.A, b_row_a + lr, p.a_tiled) : operand(p.B, col0 + lr + p.b_row_off, p.b_tiled);
  const long long ldb = b_from_a ? p.lda : p.ldb;

  for (int k0 = 0; k0 < p.K; k0 += kSK) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int k = k0 + lk + 4 * h;
      As[lk + 4 * h][lr] = (a_ok && k < p.K) ? ap[(long long)k * p.lda] : Tr<T>::zero();
      Bs[lk + 4 * h][lr] = (b_ok && k < p.K) ? bp[(long long)k * ldb] : Tr<T>::zero();
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kSK; ++k) {
      T a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i)
        a[i] = As[k][tx + 16 * i];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        b[j] = Tr<T>::conj(Bs[k][ty + 16 * j]);
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
          acc[i][j] = Tr<T>::fma(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

  const R alpha = p.alpha, beta = p.beta;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int cl = ty + 16 * j;
    const int c = col0 + cl;
    if (c >= n_lim)
      continue;
    T* colp = p.C + (long long)c * p.ldc;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int rl = tx + 16 * i;
      const int rr = row0 + rl;
      if (rr >= m_lim)
        continue;
      if (lower && (trow + rl < tcol + cl))
        continue;
      T v = Tr<T>::scale(acc[i][j], alpha);
      if (beta != R(0))
        v = Tr<T>::add(v, Tr<T>::scale(colp[rr], beta));
      if (Tr<T>::is_complex && lower && (trow + rl == tcol + cl))
        v = Tr<T>::from_real(Tr<T>::real(v));  // HERK keeps the diagonal real (blas herk semantics)
      colp[rr] = v;
    }
  }
}

}  // namespace

template <class T>
cudaError_t launch_gemm_simt(cudaStream_t stream, const SimtParams<T>& p, unsigned nblocks_x, unsigned nblocks_y) {
  if (nblocks_x == 0 || nblocks_y == 0)
    return cudaSuccess;
  cudaError_t werr = cudaSuccess;
  if (launch_wide<T>(stream, p, nblocks_x, nblocks_y, &werr))
    return werr;
  gemm_simt_kernel<T><<<dim3(nblocks_x, nblocks_y), 256, 0, stream>>>(p);
  ++g_launches;
  return cudaGetLastError();
}

template cudaError_t launch_gemm_simt<float>(cudaStream_t, const SimtParams<float>&, unsigned, unsigned);
template cudaError_t launch_gemm_simt<double>(cudaStream_t, const SimtParams<double>&, unsigned, unsigned);
template cudaError_t launch_gemm_simt<float2>(cudaStream_t, const SimtParams<float2>&, unsigned, unsigned);
template cudaError_t launch_gemm_simt<double2>(cudaStream_t, const SimtParams<double2>&, unsigned, unsigned);

}  // namespace dlafb
