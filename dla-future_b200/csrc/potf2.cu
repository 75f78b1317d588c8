// Diagonal-block Cholesky: a latency-bound single-CTA kernel that keeps the whole block in shared memory.
// Replaces cusolverDn{S,D,C,Z}potrf on the diagonal tile (reference: include/dlaf/lapack/tile.h:578-606,
// called from include/dlaf/factorization/cholesky/impl.h:46-53 and :98-105).
//
// One CTA of 512 threads, block order n <= Potf2Limits<T>::nmax, inner step BS (32, or 16 for complex<double>):
//   (a) BS x BS diagonal sub-block: ONE warp, one matrix row per lane, the row lives in registers and the
//       column scale / rank-1 updates travel by warp shuffle (no shared-memory traffic, no block barrier);
//   (b) the rows below: one row per thread, right-looking substitution against the sub-block;
//   (c) trailing update of the remaining lower triangle with register-tiled accumulators.
// The same CTA then forms W = inv(L) (needed by the panel solve, which multiplies by W^H with the DMMA
// contraction kernel instead of substituting) with the strict upper triangle of the shared buffer as scratch.
// `info` follows potrf: index (1-based, offset by info_base) of the first non-positive pivot, first failure wins.
#include "tile_ops.cuh"

namespace dlafb {
namespace {

constexpr int kPotf2Threads = 512;

template <class T, int BS, int NMAX>
__global__ void __launch_bounds__(kPotf2Threads, 1)
    potf2_kernel(int upper, int n, T* __restrict__ a, long long lda, T* __restrict__ winv, long long ldw,
                 int* info, int info_base, int mode) {
  using R = typename Tr<T>::Real;
  extern __shared__ __align__(16) unsigned char smraw[];
  const int lds = ((n + 15) & ~15) + 8;  // == 8 (mod 16): spreads the 4 row-parts of the inversion over the banks
  T* S = reinterpret_cast<T*>(smraw);
  R* dinv = reinterpret_cast<R*>(S + (size_t)lds * n);
  __shared__ int fail;

  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int warp = tid >> 5;
  if (tid == 0)
    fail = 0;

  // ---- load the referenced triangle as a lower-triangular block (upper: conjugate transpose) ----
  for (int idx = tid; idx < n * n; idx += kPotf2Threads) {
    if (!upper) {
      const int i = idx % n, j = idx / n;
      if (i >= j)
        S[i + j * lds] = a[i + j * lda];
    }
    else {
      const int j = idx % n, i = idx / n;
      if (i >= j)
        S[i + j * lds] = Tr<T>::conj(a[j + i * lda]);
    }
  }
  __syncthreads();

  if (mode & 1) {
    for (int jb = 0; jb < n; jb += BS) {
      const int bs = min(BS, n - jb);
      // (a) diagonal sub-block in one warp
      if (warp == 0) {
        T v[BS];
#pragma unroll
        for (int c = 0; c < BS; ++c)
          v[c] = (lane < bs && c <= lane) ? S[(jb + lane) + (jb + c) * lds] : Tr<T>::zero();
        int bad = 0;
#pragma unroll
        for (int j = 0; j < BS; ++j) {
          if (j < bs && !bad) {
            const R d = Tr<T>::real(shfl(v[j], j));
            if (!(d > R(0))) {
              bad = j + 1;
            }
            else {
              const R s = sqrt(d);
              const R is = R(1) / s;
              if (lane == j)
                v[j] = Tr<T>::from_real(s);
              else if (lane > j)
                v[j] = Tr<T>::scale(v[j], is);
              if (lane == 0)
                dinv[jb + j] = is;
#pragma unroll
              for (int c = j + 1; c < BS; ++c) {
                const T lc = shfl(v[j], c);
                if (lane >= c)
                  v[c] = Tr<T>::fnma_conj(v[j], lc, v[c]);
              }
            }
          }
        }
#pragma unroll
        for (int c = 0; c < BS; ++c)
          if (lane < bs && c <= lane)
            S[(jb + lane) + (jb + c) * lds] = v[c];
        if (bad && lane == 0)
          fail = jb + bad;
      }
      __syncthreads();
      if (fail)
        break;
      const int nn = n - jb - bs;  // rows below the sub-block
      if (nn == 0)
        break;
      // (b) rows below: X * L^H = B, one row per thread
      if (tid < nn) {
        const int i = jb + bs + tid;
        T x[BS];
#pragma unroll
        for (int c = 0; c < BS; ++c)
          x[c] = c < bs ? S[i + (jb + c) * lds] : Tr<T>::zero();
#pragma unroll
        for (int c = 0; c < BS; ++c) {
          if (c < bs) {
            x[c] = Tr<T>::scale(x[c], dinv[jb + c]);
#pragma unroll
            for (int c2 = c + 1; c2 < BS; ++c2)
              if (c2 < bs)
                x[c2] = Tr<T>::fnma_conj(x[c], S[(jb + c2) + (jb + c) * lds], x[c2]);
          }
        }
#pragma unroll
        for (int c = 0; c < BS; ++c)
          if (c < bs)
            S[i + (jb + c) * lds] = x[c];
      }
      __syncthreads();
      // (c) trailing lower triangle: S[i][j] -= sum_l X[i][l] conj(X[j][l]); thread (tx, ty) owns
      //     rows tx + 32*a and columns ty + 16*b
      {
        constexpr int RA = (NMAX + 31) / 32;
        constexpr int CB = (NMAX + 15) / 16;
        const int tx = tid & 31, ty = tid >> 5;
        const int r0 = jb + bs;
        T acc[RA][CB];
#pragma unroll
        for (int ia = 0; ia < RA; ++ia)
#pragma unroll
          for (int ib = 0; ib < CB; ++ib)
            acc[ia][ib] = Tr<T>::zero();
        for (int l = 0; l < bs; ++l) {
          const T* col = S + (jb + l) * lds + r0;
          T xa[RA], xb[CB];
#pragma unroll
          for (int ia = 0; ia < RA; ++ia)
            xa[ia] = (tx + 32 * ia < nn) ? col[tx + 32 * ia] : Tr<T>::zero();
#pragma unroll
          for (int ib = 0; ib < CB; ++ib)
            xb[ib] = (ty + 16 * ib < nn) ? col[ty + 16 * ib] : Tr<T>::zero();
#pragma unroll
          for (int ia = 0; ia < RA; ++ia)
#pragma unroll
            for (int ib = 0; ib < CB; ++ib)
              acc[ia][ib] = Tr<T>::fnma_conj(xa[ia], xb[ib], acc[ia][ib]);
        }
#pragma unroll
        for (int ia = 0; ia < RA; ++ia)
#pragma unroll
          for (int ib = 0; ib < CB; ++ib) {
            const int i = tx + 32 * ia, j = ty + 16 * ib;
            if (i < nn && j < nn && i >= j) {
              T* dst = S + (r0 + i) + (r0 + j) * lds;
              *dst = Tr<T>::add(*dst, acc[ia][ib]);
            }
          }
      }
      __syncthreads();
    }
    if (fail) {
      if (tid == 0 && info != nullptr)
        atomicCAS(info, 0, info_base + fail);
      // leave the tile as potrf would on failure: nothing further is defined; keep the input untouched
      return;
    }
    // ---- store L (or U = L^H) ----
    for (int idx = tid; idx < n * n; idx += kPotf2Threads) {
      if (!upper) {
        const int i = idx % n, j = idx / n;
        if (i >= j)
          a[i + j * lda] = S[i + j * lds];
      }
      else {
        const int j = idx % n, i = idx / n;
        if (i >= j)
          a[j + i * lda] = Tr<T>::conj(S[i + j * lds]);
      }
    }
  }
  else {
    for (int j = tid; j < n; j += kPotf2Threads)
      dinv[j] = R(1) / Tr<T>::real(S[j + j * lds]);
    __syncthreads();
  }

  if (!(mode & 2) || winv == nullptr)
    return;

  // ---- W = inv(L). Column t of W is built by 4 cooperating lanes (row-parts p = r mod 4); the running sums
  //      acc(r,t) = sum_{l<r} L[r][l] W[l][t] sit in the strict upper triangle at S[t + r*lds]. ----
  {
    const int t = tid >> 2;
    const int p = tid & 3;
    if (t < n)
      for (int r = t + 1 + p; r < n; r += 4)
        S[t + r * lds] = Tr<T>::zero();
    __syncthreads();
    const int t0 = (warp * 32) >> 2;  // smallest column handled by this warp
    for (int i = t0; i < n; ++i) {
      T w = Tr<T>::zero();
      const bool act = (t < n) && (i >= t);
      if (act && (i & 3) == p) {
        if (i == t)
          w = Tr<T>::from_real(dinv[t]);
        else {
          w = Tr<T>::scale(S[t + i * lds], -dinv[i]);
          S[t + i * lds] = w;
        }
      }
      w = shfl(w, (lane & ~3) | (i & 3));
      if (act) {
        for (int r = i + 1 + ((p - (i + 1)) & 3); r < n; r += 4)
          S[t + r * lds] = Tr<T>::fma(S[r + i * lds], w, S[t + r * lds]);
      }
    }
    __syncthreads();
    // write W (lower, with explicit zeros above the diagonal: the panel solve multiplies by the full block)
    for (int idx = tid; idx < n * n; idx += kPotf2Threads) {
      const int i = idx % n, j = idx / n;
      T v;
      if (i > j)
        v = S[j + i * lds];
      else if (i == j)
        v = Tr<T>::from_real(dinv[i]);
      else
        v = Tr<T>::zero();
      winv[i + j * ldw] = v;
    }
  }
}

template <class T>
size_t potf2_smem_bytes(int n) {
  const int lds = ((n + 15) & ~15) + 8;
  return (size_t)lds * n * sizeof(T) + (size_t)n * sizeof(typename Tr<T>::Real) + 16;
}

}  // namespace

template <class T>
cudaError_t launch_potf2(cudaStream_t stream, char uplo, int n, T* a, long long lda, T* winv, long long ldw,
                         int* info, int info_base, int mode) {
  constexpr int BS = Potf2Limits<T>::bs;
  constexpr int NMAX = Potf2Limits<T>::nmax;
  if (n <= 0)
    return cudaSuccess;
  if (n > NMAX)
    return cudaErrorInvalidValue;
  auto kern = potf2_kernel<T, BS, NMAX>;
  static cudaError_t attr =
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)potf2_smem_bytes<T>(NMAX));
  if (attr != cudaSuccess)
    return attr;
  kern<<<1, kPotf2Threads, potf2_smem_bytes<T>(n), stream>>>(uplo == 'U' || uplo == 'u', n, a, lda, winv, ldw, info,
                                                             info_base, mode);
  ++g_launches;
  return cudaGetLastError();
}

template cudaError_t launch_potf2<float>(cudaStream_t, char, int, float*, long long, float*, long long, int*, int, int);
template cudaError_t launch_potf2<double>(cudaStream_t, char, int, double*, long long, double*, long long, int*, int,
                                          int);
template cudaError_t launch_potf2<float2>(cudaStream_t, char, int, float2*, long long, float2*, long long, int*, int,
                                          int);
template cudaError_t launch_potf2<double2>(cudaStream_t, char, int, double2*, long long, double2*, long long, int*,
                                           int, int);

}  // namespace dlafb
