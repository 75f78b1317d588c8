// Tile-level operations of the Cholesky path, built from the three kernels (potf2, DMMA/SIMT contraction) and a
// few element-wise helpers. Counterparts in the reference: tile::potrf (include/dlaf/lapack/tile.h:602-606),
// tile::trsm / herk / gemm (include/dlaf/blas/tile.h:370-470), tile::laset (lapack/tile.h:535-541).
#pragma once
#include <cuda_runtime.h>

#include "common.cuh"

namespace dlafb {

template <class T>
struct Potf2Limits {
  static constexpr int bs = 8;
  static constexpr int nmax = 128;
};
template <>
struct Potf2Limits<double2> {
  static constexpr int bs = 8;
  static constexpr int nmax = 64;
};

// mode bit 0: factorize in place; bit 1: write W = inv(L) (n x n, zeros above the diagonal) to winv.
template <class T>
cudaError_t launch_potf2(cudaStream_t stream, char uplo, int n, T* a, long long lda, T* winv, long long ldw,
                         int* info, int info_base, int mode);

// C (M x N) <- beta*C + alpha * A (M x K) * B (N x K)^H; lower != 0 touches only i >= j.
// Picks the TMA+DMMA kernel for aligned double operands, the generic kernel otherwise.
// C may alias A when N <= in-place limit of the chosen kernel (128 DMMA, 64 generic); checked.
template <class T>
cudaError_t gemm_nt(cudaStream_t stream, int M, int N, int K, typename Tr<T>::Real alpha, const T* A, long long lda,
                    const T* B, long long ldb, typename Tr<T>::Real beta, T* C, long long ldc, int lower);

// Inner block size used by potrf_tile / trsm_panel for this operand (128 on the DMMA path, 64 otherwise).
template <class T>
int inner_block(const T* a, long long lda, int nb);

// Elements of workspace potrf_tile needs for an n x n tile (inverted diagonal blocks, + n*n for uplo = U).
template <class T>
size_t potrf_work_elems(int n, int ib);

// Blocked Cholesky of one n x n tile. On return work[j*ib*ib ...] (ld = ib) holds inv(L_jj) of the j-th ib block.
template <class T>
cudaError_t potrf_tile(cudaStream_t stream, char uplo, int n, T* a, long long lda, T* work, int ib, int* info,
                       int info_base);

// Panel solve, Lower form: B (m x nk) <- B * L^-H (Right, Lower, ConjTrans, NonUnit; impl.h:55-67) given the
// inverted diagonal blocks W of L (as left by potrf_tile).
template <class T>
cudaError_t trsm_panel(cudaStream_t stream, int nk, const T* L, long long ldl, const T* W, int ib, T* B,
                       long long ldb, int m);

// The same solve followed by the copy of the solved panel into tile layout (pack_tiles below: row i of B -> slot
// i / tile of `packed`), which is how the distributed factorization hands the panel to the exchange. One launch on the
// fused path.
template <class T>
cudaError_t trsm_panel_packed(cudaStream_t stream, int nk, const T* L, long long ldl, const T* W, int ib, T* B,
                              long long ldb, int m, T* packed, int tile);

// One-kernel panel solve for double (csrc/trsm_fused.cu). Returns false when the case is not the fused kernel's; the
// caller then takes the multi-launch path. `packed` may be null.
bool trsm_panel_fused_d(cudaStream_t stream, int nk, const double* L, long long ldl, const double* W, int ib, double* B,
                        long long ldb, int m, double* packed, int tile, cudaError_t* err);

// out (n x m) <- op(in (m x n)), op = transpose (+ conjugate when conj != 0); tri: 'G' all, 'L' / 'U' only that
// triangle of OUT is written.
template <class T>
cudaError_t transpose(cudaStream_t stream, int m, int n, const T* in, long long ldin, T* out, long long ldout, int conj,
                      char tri);

// laset: off-diagonal <- alpha, diagonal <- beta on the uplo part ('G', 'L', 'U') of an m x n tile
// (reference kernel: src/lapack/gpu/laset.cu:66-120).
template <class T>
cudaError_t laset(cudaStream_t stream, char uplo, int m, int n, T alpha, T beta, T* a, long long lda);

// lacpy with uplo ('G', 'L', 'U').
template <class T>
cudaError_t lacpy(cudaStream_t stream, char uplo, int m, int n, const T* a, long long lda, T* b, long long ldb);

// Synthetic Hermitian positive definite matrix for timing runs (counter-based; NOT the reference's mt19937_64
// stream, see include/dlaf/util_matrix.h:499-501 for the property it mimics: |off-diag| <= 1, diag = u + 2n).
// Fills the rows x cols local block whose element (i, j) is global (grow[i], gcol[j]) through the index maps
// row_tile/col_tile: local tile l <-> global tile first + l*stride.
template <class T>
cudaError_t gen_hpd(cudaStream_t stream, long long n, int nb, int rows, int cols, int row_first, int row_stride,
                    int col_first, int col_stride, T* a, long long lda, unsigned long long seed);

// Copies the m x n column-major block src into tile layout: row i goes to slot i / tile (contiguous tile x tile
// column-major tiles) of dst; dst points at the slot of row 0.
template <class T>
cudaError_t pack_tiles(cudaStream_t stream, int m, int n, const T* src, long long ld, T* dst, int tile);

// max |a_ij| over the uplo part, atomically combined into *out (double, must be zeroed by the caller).
template <class T>
cudaError_t max_abs(cudaStream_t stream, char uplo, int m, int n, const T* a, long long lda, double* out);

}  // namespace dlafb
