// FP64 contraction kernel of the Cholesky path: C <- beta*C + alpha * A * B^T on 128x128 blocks of C.
// This is the kernel that replaces the reference's per-tile cublasDgemm / cublasDsyrk / cublasDtrsm calls
// (include/dlaf/blas/tile.h:370-382, :414-425, :458-470) for the trailing update, the panel solve
// (multiplication by inverted diagonal blocks) and the blocked diagonal-tile factorization.
//
// Data path: TMA (cp.async.bulk.tensor) -> 128B-swizzled shared memory -> DMMA m8n8k4 (mma.sync f64).
// A and B are column-major panels (rows x K); one pipeline stage holds K-slices of 16 columns of a
// 128-row block of each, laid out as 8 row-groups of [16 k-columns][16 rows] (128 B per k-column).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace dlafb {

constexpr int kBM = 128;      // rows of C per CTA
constexpr int kBN = 128;      // cols of C per CTA
constexpr int kBK = 16;       // K-slice per pipeline stage
constexpr int kStages = 4;    // TMA pipeline depth
constexpr int kGroupBytes = kBK * 128;             // one 16-row group of one K-slice
constexpr int kOperandStageBytes = kBM * kBK * 8;  // 16 KiB
constexpr int kStageBytes = 2 * kOperandStageBytes;
constexpr int kGemmSmemBytes = kStages * kStageBytes + 2 * kStages * 8 + 1024;  // + barriers + align slack
constexpr int kGemmThreads = 256;  // 8 DMMA warps; lane 0 of warp 0 also issues the TMA loads
constexpr int kPrefetch = kStages - 2;  // slices in flight ahead of the math (refills the stage freed two slices ago)

// Block of C per CTA. 128 x 128 is the throughput shape (88 % of the DMMA peak); one CTA needs 2.1 us per K-slice of
// 16 whatever the shape's area ratio, so the launches on the critical path (few blocks, K = 128..512) use the small
// shapes: 4x (64 x 64) or 4x (32 x 128, the in-place panel solve needs all N columns in one CTA) more CTAs.
enum class GemmShape : int { k128x128 = 0, k64x64 = 1, k32x128 = 2, k64x128 = 3 };
constexpr GemmShape kDefaultBulkShape = GemmShape::k64x128;  // two co-resident CTAs hide each other's prologue and epilogue
constexpr int shape_bm(GemmShape s) {
  return s == GemmShape::k128x128 ? 128 : (s == GemmShape::k64x64 || s == GemmShape::k64x128) ? 64 : 32;
}
constexpr int shape_bn(GemmShape s) {
  return s == GemmShape::k64x64 ? 64 : 128;
}
constexpr int gemm_smem_bytes(int bm, int bn, int stages = kStages) {
  // operand ring + half a staged block of C for the reduce epilogue (column stride bm*8 + 16) + barriers + align slack
  return stages * (bm + bn) * kBK * 8 + (bn / 2) * (bm * 8 + 16) + 2 * stages * 8 + 1024;
}

struct GemmParams {
  double* C;
  long long ldc;
  int m_total, n_total;  // C is clipped to [0, m_total) x [0, n_total)
  int K;                 // contraction length, multiple of kBK
  int a_col0, b_col0;    // first K column inside the A / B tensor maps
  int a_row_off, b_row_off;  // A row = C row + a_row_off, B row = C col + b_row_off
  double alpha, beta;
  // LIST mode (list != nullptr): one entry per nb x nb tile, x = first row, y = first col | diag<<31;
  // each tile is covered by (tile/BM)*(tile/BN) CTAs (blockIdx.x = item * that + block inside the tile).
  const int2* list;
  int tile;  // nb: list tiles are clipped to tile x tile; a multiple of BM and BN
  // RECT mode: blocks (blockIdx.x, blockIdx.y) of the rectangle at (rect_row0, rect_col0);
  // lower != 0 keeps only elements on or below the rectangle's own diagonal.
  int rect_row0, rect_col0;
  int lower;
  int ntiles, rect_nx;  // filled by the launcher: number of blocks, blocks per rectangle row
  int list_scale;       // 2 for the real view of complex<double> (list entries count complex elements), else 1
  // 1: rank-3 map (one TMA per operand slice), 8: rank-2 map (one TMA per 16-row group),
  // 4: rank-4 map of a panel stored as contiguous tile x tile tiles (one TMA per slice)
  int nbox_a, nbox_b;
};

// Builds the tensor map of a column-major FP64 matrix for the layout above. kind 3: rank-3 view
// (16 rows, cols, row-groups) so that a whole 128 x 16 slice arrives with ONE bulk copy; kind 2: plain
// rank-2 view with 16 x 16 boxes; kind 4: panel in tile layout (rows = ntiles*ld, ld = tile size, cols = ld).
// Returns 0 on success.
// box_rows = rows of the operand one CTA loads per slice (BM for the A map, BN for the B map of the shape used).
// kind 4: tiles of ld rows x tile_cols columns (tile_cols = 0: square). For complex<double> pass the real view:
// rows and ld doubled.
int make_operand_map(CUtensorMap* map, const double* base, long long rows, long long cols, long long ld,
                     int kind, int box_rows = 128, long long tile_cols = 0);

// Launches the kernel for `nblocks_x` x `nblocks_y` CTAs on `stream`. The maps' boxes must match `shape`.
cudaError_t launch_gemm_nt_dmma(cudaStream_t stream, const CUtensorMap& map_a, const CUtensorMap& map_b,
                                const GemmParams& p, unsigned nblocks_x, unsigned nblocks_y,
                                GemmShape shape = GemmShape::k128x128, bool cplx = false, unsigned max_ctas = 0);
// max_ctas > 0: at most that many CTAs, each looping over blocks b, b + grid, ... (persistent). The engine uses it
// with one CTA per SM minus one SM when the dependent chain of the factorization is the bottleneck: the
// high-priority kernels then always find a free half SM (and the single-CTA diagonal kernel a whole one).
int gemm_sm_count();

}  // namespace dlafb
