// Diagonal-block Cholesky: a latency-bound single-CTA kernel that keeps the whole block in shared memory.
// Replaces cusolverDn{S,D,C,Z}potrf on the diagonal tile (reference: include/dlaf/lapack/tile.h:578-606,
// called from include/dlaf/factorization/cholesky/impl.h:46-53 and :98-105) and also forms W = inv(L): the panel
// solve multiplies by W^H with the DMMA contraction kernel instead of substituting.
//
// One CTA of 512 threads, block order n <= Potf2Limits<T>::nmax (128; 64 for complex<double>), panels of 8 columns:
//   panel:  one thread per row keeps its row of the panel in registers; threads [NMAX, 2*NMAX) own the rows of an
//           identity appended below A, which the same column operations turn into inv(L)^H, so the inverse adds no
//           dependent chain of its own. Every participating warp factors the 8 x 8 diagonal sub-block itself
//           (lane l keeps row l % 8, columns travel by warp shuffle, the diagonal is tracked by every lane), so the
//           column-to-column chain is rsqrt -> scale -> FMA: no block barrier, no shared-memory round trip;
//   update: trailing columns (A rows below and E rows above) with 4 x CBN register tiles, all 512 threads,
//           instantiated per column count; loads and stores are select/predicate only.
// `info` follows potrf: index (1-based, offset by info_base) of the first non-positive pivot, first failure wins.
//
// History (round 1, profiles/r01/critical_path_v*.json): v1 factored the sub-block in ONE warp and solved the rows
// below with one thread per row, then inverted separately: 252 us for n = 128. A 32-column, fully unrolled register
// variant: 131 us (FP64 dependent latency is 23 cycles on B200, issue 2.2 cycles per warp instruction:
// profiles/r01/latency_b200.txt). This version: 75 us; the 512 x 512 tile built from it: 1170 -> 362 us.
#include "tile_ops.cuh"

#include <cstdlib>

namespace dlafb {
namespace {

constexpr int kPotf2Threads = 512;

#ifdef DLAFB_TIMING
__device__ long long g_potf2_clk[64];
#define STAMP(i)          \
  if (threadIdx.x == 0)   \
  g_potf2_clk[i] = clock64()
#define STAMPT(i, t)       \
  if (threadIdx.x == (t))  \
  g_potf2_clk[i] = clock64()
#else
#define STAMP(i)
#define STAMPT(i, t)
#endif

template <class T, int NMAX>
struct Potf2Smem {
  using R = typename Tr<T>::Real;
  static constexpr int lds = NMAX + 9;    // odd: walking along a row or along a column is bank-conflict free
  static constexpr int ncol = NMAX + 1;   // one dummy column absorbs the out-of-range lanes of the trailing update
  static constexpr size_t bytes(int bs) {
    return ((size_t)lds * ncol + (size_t)bs * NMAX) * sizeof(T) + (size_t)NMAX * sizeof(R) + 16;
  }
};

// Trailing update of one warp's share: rows tx + 32*ia (all RA strips), columns col0 + 16*ib for ib < CBN, over the
// panel X = Xp (rows < r0 hold E, rows >= r0 hold L):  S[rw][col] -= sum_l X[rw][l] conj(X[col][l]), kept only for
// A rows rw >= col and, when inverting, E rows rw < r0 (same formula, same address). Straight-line code: no branch
// in the l loop, and the read-modify-write adds a selected zero instead of branching per element.
template <class T, int BS, int NMAX, int CBN>
__device__ __forceinline__ void potf2_trailing(T* __restrict__ S, const T* __restrict__ Xp, int n, int r0, int bs,
                                               bool factor, bool inv, int tx, int col0, int cend) {
  constexpr int RA = NMAX / 32;
  constexpr int LDS = Potf2Smem<T, NMAX>::lds;
  T acc[RA][CBN];
  int cidx[CBN];
#pragma unroll
  for (int ib = 0; ib < CBN; ++ib) {
    cidx[ib] = min(col0 + 16 * ib, NMAX - 1);
#pragma unroll
    for (int ia = 0; ia < RA; ++ia)
      acc[ia][ib] = Tr<T>::zero();
  }
#pragma unroll 2
  for (int l = 0; l < bs; ++l) {
    const T* xp = Xp + l * NMAX;
    T xa[RA], xb[CBN];
#pragma unroll
    for (int ia = 0; ia < RA; ++ia)
      xa[ia] = xp[tx + 32 * ia];
#pragma unroll
    for (int ib = 0; ib < CBN; ++ib)
      xb[ib] = xp[cidx[ib]];
#pragma unroll
    for (int ia = 0; ia < RA; ++ia)
#pragma unroll
      for (int ib = 0; ib < CBN; ++ib)
        acc[ia][ib] = Tr<T>::fnma_conj(xa[ia], xb[ib], acc[ia][ib]);
  }
#pragma unroll
  for (int ib = 0; ib < CBN; ++ib) {
    const int col = col0 + 16 * ib;
    const bool colok = col < cend;  // cend <= n: the columns this update is limited to
    T* dcol = S + (colok ? col : NMAX) * LDS;
#pragma unroll
    for (int ia = 0; ia < RA; ++ia) {
      const int rw = tx + 32 * ia;
      const bool keep = colok && rw < n && ((factor && rw >= col) || (inv && rw < r0));
      dcol[rw] = Tr<T>::add(dcol[rw], keep ? acc[ia][ib] : Tr<T>::zero());
    }
  }
}

template <class T, int BS, int NMAX, bool FACTOR>
__global__ void __launch_bounds__(kPotf2Threads, 1)
    potf2_kernel(int upper, int n, T* __restrict__ a, long long lda, T* __restrict__ winv, long long ldw,
                 int* info, int info_base) {
  using R = typename Tr<T>::Real;
  using Sm = Potf2Smem<T, NMAX>;
  static_assert((NMAX & (NMAX - 1)) == 0 && kPotf2Threads % NMAX == 0 && 2 * NMAX <= kPotf2Threads, "bad NMAX");
  static_assert(BS == 8, "lane l keeps row l % 8 of the sub-block");
  extern __shared__ __align__(16) unsigned char smraw[];
  constexpr int LDS = Sm::lds;
  T* S = reinterpret_cast<T*>(smraw);             // lower: A -> L; strict upper: E = inv(L)^H (when inverting)
  T* Xp = S + (size_t)LDS * Sm::ncol;             // current panel, BS x NMAX: rows < r0 from E, rows >= r0 from A
  R* dinv = reinterpret_cast<R*>(Xp + BS * NMAX);  // 1 / L_jj
  int* sfail = reinterpret_cast<int*>(dinv + NMAX);

  const int tid = threadIdx.x;
  const bool inv = winv != nullptr;
  const int li = tid & (NMAX - 1);  // element loops without integer division: row li, columns lj0 + k*lstep
  const int lj0 = tid / NMAX;
  constexpr int lstep = kPotf2Threads / NMAX;
  const T one = Tr<T>::from_real(R(1));
  STAMP(0);

  // ---- load the referenced triangle as a lower-triangular block (upper: conjugate transpose); zero elsewhere ----
  for (int j = lj0; j < n; j += lstep) {
    T v = Tr<T>::zero();
    if (li < n && li >= j)
      v = upper ? Tr<T>::conj(a[j + li * lda]) : a[li + j * lda];
    S[li + j * LDS] = v;
  }
  __syncthreads();
  STAMP(1);

  // Column panels of width BS, one thread per row, the row of the panel in registers:
  //   threads [0, NMAX)        row = tid of A (rows of the BS x BS diagonal sub-block and the rows below it)
  //   threads [NMAX, 2*NMAX)   row e = tid - NMAX of E: the identity appended below A, which the same column
  //                            operations turn into inv(L)^H (X * L^H = I solved panel by panel)
  // Every one of these warps ALSO factors the sub-block itself (lane l keeps row l % BS of it, columns travel by
  // warp shuffle), so the column-to-column dependent chain has no block barrier and no shared-memory round trip:
  //   rsqrt(pivot) -> scale the (already shuffled, unscaled) column -> update the locally tracked diagonal -> next
  // (23-cycle FP64 latency, 55-80 for rsqrt: profiles/r01/latency_b200.txt). All copies run the same operations in
  // the same order and agree bit for bit. Loads and stores are select/predicate only: divergent branches cost more
  // than the arithmetic here.
  const bool roleA = tid < NMAX;
  const int row = roleA ? tid : tid - NMAX;
  int fail = 0;
  for (int jb = 0; jb < n; jb += BS) {
    const int bs = min(BS, n - jb);
    const int r0 = jb + bs;
    if (tid < 2 * NMAX) {
      T* Sp = S + jb * LDS;    // the panel's columns
      const int t = row - jb;  // index inside the diagonal sub-block (when 0 <= t < bs)
      const int r = tid & (BS - 1);
      const bool act = roleA ? (row >= jb && row < n) : (inv && row < r0);
      T Dr[BS];  // row r of the sub-block (identity-padded beyond bs)
      T dg[BS];  // its diagonal, tracked by every lane
      T x[BS];   // this thread's row of the panel
#pragma unroll
      for (int c = 0; c < BS; ++c) {
        const T dv = Sp[(jb + r) + c * LDS];
        const T gv = Sp[(jb + c) + c * LDS];
        const T xv = Sp[row + c * LDS];
        Dr[c] = (r >= c && r < bs) ? dv : (r == c ? one : Tr<T>::zero());
        dg[c] = c < bs ? gv : one;
        const bool in = act && c < bs;
        const bool keep = roleA ? (in && (t >= bs || c <= t)) : (in && t < 0);
        x[c] = keep ? xv : ((!roleA && in && t == c) ? one : Tr<T>::zero());
      }
#pragma unroll
      for (int j = 0; j < BS; ++j) {
        const R d = Tr<T>::real(dg[j]);
        fail = (!(d > R(0)) && fail == 0) ? jb + j + 1 : fail;
        T lraw[BS];
#pragma unroll
        for (int c = j + 1; c < BS; ++c)
          lraw[c] = shfl(Dr[j], (tid & 24) + c);  // element (c, j) from the lane that keeps row c
        if (FACTOR) {
          const R is = Tr<T>::rsqrt_(d);
          if (tid == j && j < bs)
            dinv[jb + j] = is;
          T xs = Tr<T>::scale(x[j], is);
          if (roleA && t == j)
            xs = Tr<T>::from_real(Tr<T>::real(xs));  // the diagonal is real
          x[j] = xs;
          const T drj = Tr<T>::scale(Dr[j], is);
#pragma unroll
          for (int c = j + 1; c < BS; ++c) {
            const T lc = Tr<T>::scale(lraw[c], is);
            dg[c] = Tr<T>::fnma_conj(lc, lc, dg[c]);
            Dr[c] = Tr<T>::fnma_conj(drj, lc, Dr[c]);
            x[c] = Tr<T>::fnma_conj(xs, lc, x[c]);
          }
        }
        else {  // L is given: only the E rows are solved (A rows pass through unchanged)
          const R is = R(1) / d;
          if (tid == j && j < bs)
            dinv[jb + j] = is;
          const T xs = roleA ? x[j] : Tr<T>::scale(x[j], is);
          x[j] = xs;
#pragma unroll
          for (int c = j + 1; c < BS; ++c)
            x[c] = roleA ? x[c] : Tr<T>::fnma_conj(xs, lraw[c], x[c]);
        }
      }
      if (tid == 0)
        *sfail = fail;
      STAMP(4 + 3 * (jb / BS));
      // write the panel back: into S (final values) and into Xp (operand of the trailing update)
#pragma unroll
      for (int c = 0; c < BS; ++c) {
        const bool in = c < bs;
        const bool ps = in && act && (roleA ? (FACTOR && (t >= bs || c <= t)) : (row < jb + c));
        const bool px = in && (roleA ? row >= r0 : row < r0);
        if (ps)
          Sp[row + c * LDS] = x[c];
        if (px)
          Xp[c * NMAX + row] = x[c];
      }
    }
    __syncthreads();
    fail = *sfail;
    if (fail)
      break;  // uniform
    STAMP(5 + 3 * (jb / BS));
    if (r0 >= n)
      break;
    {  // warp w owns the columns r0 + w + 16*ib; instantiated per column count so that no lane tests inside
      const int tx = tid & 31, col0 = r0 + (tid >> 5);
      const int nib = col0 < n ? (n - col0 + 15) >> 4 : 0;
      constexpr int CB = (NMAX - BS + 15) / 16;
      if (CB > 6 && nib > 6)
        potf2_trailing<T, BS, NMAX, (CB > 6 ? 8 : CB)>(S, Xp, n, r0, bs, FACTOR, inv, tx, col0, n);
      else if (CB > 4 && nib > 4)
        potf2_trailing<T, BS, NMAX, (CB > 4 ? 6 : CB)>(S, Xp, n, r0, bs, FACTOR, inv, tx, col0, n);
      else if (nib > 2)
        potf2_trailing<T, BS, NMAX, 4>(S, Xp, n, r0, bs, FACTOR, inv, tx, col0, n);
      else if (nib > 0)
        potf2_trailing<T, BS, NMAX, 2>(S, Xp, n, r0, bs, FACTOR, inv, tx, col0, n);
    }
    __syncthreads();
    STAMP(6 + 3 * (jb / BS));
  }
  STAMP(60);
  if (fail) {
    if (tid == 0 && info != nullptr)
      atomicCAS(info, 0, info_base + fail);
    return;  // like potrf on failure nothing further is defined: the tile is left as it was
  }
  // ---- store L (or U = L^H) ----
  if (FACTOR && li < n)
    for (int j = lj0; j <= li; j += lstep) {
      if (upper)
        a[j + li * lda] = Tr<T>::conj(S[li + j * LDS]);
      else
        a[li + j * lda] = S[li + j * LDS];
    }
  STAMP(61);
  // ---- W = inv(L) = E^H, with explicit zeros above the diagonal (the panel solve multiplies by the full block).
  //      Lanes run along i: coalesced global writes, shared reads with the odd stride LDS. ----
  if (inv && li < n)
    for (int c = lj0; c < n; c += lstep) {
      T v = Tr<T>::zero();  // W(li, c)
      if (c < li)
        v = Tr<T>::conj(S[c + li * LDS]);
      else if (c == li)
        v = Tr<T>::from_real(dinv[li]);
      winv[li + c * ldw] = v;
    }
  STAMP(62);
}

#include "potf2_d128.cuh"

}  // namespace

// factor + inverse of a full 128 x 128 lower double block: the delayed-update kernel (DLAFB_POTF2_FAST=0: the generic one)
template <class T>
bool launch_potf2_fast(cudaStream_t, int, int, T*, long long, T*, long long, int*, int, int) {
  return false;
}
template <>
bool launch_potf2_fast<double>(cudaStream_t stream, int upper, int n, double* a, long long lda, double* w,
                               long long ldw, int* info, int info_base, int mode) {
  static const bool on = [] {
    const char* e = getenv("DLAFB_POTF2_FAST");
    return !(e && e[0] == '0');
  }();
  if (!on || upper || n != kP128 || mode != 3 || w == nullptr)
    return false;
  static cudaError_t attr =
      cudaFuncSetAttribute(potf2_d128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kP128Smem);
  if (attr != cudaSuccess)
    return false;
  potf2_d128_kernel<<<1, kPotf2Threads, kP128Smem, stream>>>(a, lda, w, ldw, info, info_base);
  ++g_launches;
  return true;
}

template <class T>
cudaError_t launch_potf2(cudaStream_t stream, char uplo, int n, T* a, long long lda, T* winv, long long ldw,
                         int* info, int info_base, int mode) {
  constexpr int BS = Potf2Limits<T>::bs;
  constexpr int NMAX = Potf2Limits<T>::nmax;
  constexpr size_t smem = Potf2Smem<T, NMAX>::bytes(BS);
  if (n <= 0)
    return cudaSuccess;
  if (n > NMAX)
    return cudaErrorInvalidValue;
  auto kern_f = potf2_kernel<T, BS, NMAX, true>;
  auto kern_i = potf2_kernel<T, BS, NMAX, false>;
  static cudaError_t attr_f = cudaFuncSetAttribute(kern_f, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  static cudaError_t attr_i = cudaFuncSetAttribute(kern_i, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (attr_f != cudaSuccess || attr_i != cudaSuccess)
    return attr_f != cudaSuccess ? attr_f : attr_i;
  T* w = (mode & 2) ? winv : nullptr;
  const int up = uplo == 'U' || uplo == 'u';
  if (launch_potf2_fast<T>(stream, up, n, a, lda, w, ldw, info, info_base, mode))
    return cudaGetLastError();
  if (mode & 1)
    kern_f<<<1, kPotf2Threads, smem, stream>>>(up, n, a, lda, w, ldw, info, info_base);
  else if (w != nullptr)
    kern_i<<<1, kPotf2Threads, smem, stream>>>(up, n, a, lda, w, ldw, info, info_base);
  else
    return cudaSuccess;
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
