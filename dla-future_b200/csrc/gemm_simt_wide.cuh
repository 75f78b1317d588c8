// Wide variant of the generic contraction kernel for float: C <- beta*C + alpha * A * B^T on 128 x 128 blocks of C,
// same parameters, modes (RECT / LIST, lower, tiled operands, clipping) and results as gemm_simt.cu - bit for bit:
// every element is the same chain acc = fma(a_ik, b_jk, acc) over k = 0, 1, ... - but with four times the arithmetic
// per byte moved on chip: an 8 x 8 accumulator block per thread fed by two 16-byte shared-memory loads per operand and
// k (64 FMA per 4 LDS.128; the 64 x 64 kernel does 16 FMA per 8 scalar loads and is bound by them: 27 of 74 TFLOP/s),
// and the next K-slice travels from global memory into registers while the current one is multiplied.
//
// The body is a __host__ __device__ function of a context object (thread / block indices, the two shared-memory
// slices, the barrier), so that tests/emu/ runs exactly this code on the CPU with one host thread per CUDA thread of a
// block. The default for float since round 2 (38.2 vs 27.5 TFLOP/s for the n = 32768 factorization with the single-buffered
// first version, profiles/r02/session1_summary.txt); DLAFB_SGEMM_WIDE=0 selects the 64 x 64 kernel.
#pragma once
#include "gemm_simt.cuh"

// the host pass of this __host__ __device__ body does not know the pragma
#ifdef __CUDA_ARCH__
#define DLAFB_WIDE_UNROLL _Pragma("unroll")
#else
#define DLAFB_WIDE_UNROLL
#endif

namespace dlafb {

constexpr int kWideTile = 128;           // C block per CTA
constexpr int kWideSK = 8;               // K-slice
constexpr int kWideLd = kWideTile + 4;   // row stride of a slice in shared memory (floats); keeps 16-byte alignment
constexpr int kWideThreads = 256;
constexpr int kWideSmemFloats = 2 * 2 * kWideSK * kWideLd;  // two buffers of (A slice + B slice)

// Context on the device: see gemm_simt.cu. Context on the host: tests/emu/simt_wide_emu.cu.
//   int tid(), bx(), by();  float* smem();  void sync();

template <class Ctx>
__host__ __device__ inline void gemm_simt_wide_body(const SimtParams<float>& p, Ctx& c) {
  constexpr int TILE = kWideTile, SK = kWideSK, LD = kWideLd;
  int trow, tcol, row0, col0, lower, m_lim = p.m_total, n_lim = p.n_total;
  bool b_from_a = false;
  int b_row_a = 0;
  if (p.list != nullptr) {
    const int sub = (p.tile + TILE - 1) / TILE;
    const int nsub = sub * sub;
    const int item = c.bx() / nsub;
    const int s = c.bx() - item * nsub;
    const int2 it = p.list[item];
    lower = it.y < 0;
    const int c0 = it.y & 0x7fffffff;
    trow = (s % sub) * TILE;
    tcol = (s / sub) * TILE;
    row0 = it.x + trow;
    col0 = c0 + tcol;
    b_from_a = lower;  // diagonal tile of the list: both operands from the row panel
    b_row_a = it.x + tcol + p.a_row_off;
    m_lim = m_lim < it.x + p.tile ? m_lim : it.x + p.tile;
    n_lim = n_lim < c0 + p.tile ? n_lim : c0 + p.tile;
  }
  else {
    lower = p.lower;
    trow = c.bx() * TILE;
    tcol = c.by() * TILE;
    row0 = p.rect_row0 + trow;
    col0 = p.rect_col0 + tcol;
  }
  // whole-block exits: taken by every thread of the block alike, so no thread is left waiting at a barrier
  if (row0 >= m_lim || col0 >= n_lim)
    return;
  if (lower && trow + TILE - 1 < tcol)
    return;

  // two buffers of [A slice | B slice], each slice [SK][LD]: the next slice is written into the buffer that was read in
  // the previous iteration, so one barrier per slice is enough
  float* As = c.smem();
  float* Bs = As + SK * LD;
  constexpr int BUF = 2 * SK * LD;
  const int tid = c.tid();
  const int tx = tid & 15, ty = tid >> 4;
  const int lr = tid & (TILE - 1);  // loader: row inside the block
  const int lk = tid >> 7;          // loader: k inside the slice (0 or 1; + 2, 4, 6)

  float acc[8][8];
DLAFB_WIDE_UNROLL
  for (int i = 0; i < 8; ++i)
DLAFB_WIDE_UNROLL
    for (int j = 0; j < 8; ++j)
      acc[i][j] = 0.f;

  const bool a_ok = row0 + lr < m_lim;
  const bool b_ok = col0 + lr < n_lim;
  // first element of the operand row: a plain column-major panel, or a panel stored as contiguous tile x tile tiles
  const int arow = row0 + lr + p.a_row_off;
  const float* ap = p.a_tiled ? p.A + (long long)(arow / p.tile) * p.tile * p.tile + arow % p.tile : p.A + arow;
  const int brow = b_from_a ? b_row_a + lr : col0 + lr + p.b_row_off;
  const float* bbase = b_from_a ? p.A : p.B;
  const int btiled = b_from_a ? p.a_tiled : p.b_tiled;
  const float* bp = btiled ? bbase + (long long)(brow / p.tile) * p.tile * p.tile + brow % p.tile : bbase + brow;
  const long long lda = p.lda, ldb = b_from_a ? p.lda : p.ldb;

  float pa[4], pb[4];
  // slice k0 -> registers
  auto fetch = [&](int k0) {
DLAFB_WIDE_UNROLL
    for (int h = 0; h < 4; ++h) {
      const int k = k0 + lk + 2 * h;
      pa[h] = (a_ok && k < p.K) ? ap[(long long)k * lda] : 0.f;
      pb[h] = (b_ok && k < p.K) ? bp[(long long)k * ldb] : 0.f;
    }
  };
  auto stash = [&](int buf) {
DLAFB_WIDE_UNROLL
    for (int h = 0; h < 4; ++h) {
      As[buf * BUF + (lk + 2 * h) * LD + lr] = pa[h];
      Bs[buf * BUF + (lk + 2 * h) * LD + lr] = pb[h];
    }
  };

  fetch(0);
  stash(0);
  c.sync();
  int cur = 0;
  for (int k0 = 0; k0 < p.K; k0 += SK) {
    const bool more = k0 + SK < p.K;
    if (more)
      fetch(k0 + SK);  // global loads in flight during the multiplication below
    const float* Ac = As + cur * BUF;
    const float* Bc = Bs + cur * BUF;
DLAFB_WIDE_UNROLL
    for (int k = 0; k < SK; ++k) {
      // rows 4 tx + {0..3} and 64 + 4 tx + {0..3}; columns 4 ty + {0..3} and 64 + 4 ty + {0..3}
      const float4 a0 = *reinterpret_cast<const float4*>(Ac + k * LD + 4 * tx);
      const float4 a1 = *reinterpret_cast<const float4*>(Ac + k * LD + 64 + 4 * tx);
      const float4 b0 = *reinterpret_cast<const float4*>(Bc + k * LD + 4 * ty);
      const float4 b1 = *reinterpret_cast<const float4*>(Bc + k * LD + 64 + 4 * ty);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
DLAFB_WIDE_UNROLL
      for (int i = 0; i < 8; ++i)
DLAFB_WIDE_UNROLL
        for (int j = 0; j < 8; ++j)
          acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (more)
      stash(cur ^ 1);  // the buffer everyone finished reading before the previous barrier
    c.sync();
    cur ^= 1;
  }

  const float alpha = p.alpha, beta = p.beta;
DLAFB_WIDE_UNROLL
  for (int j = 0; j < 8; ++j) {
    const int cl = 4 * ty + (j & 3) + 64 * (j >> 2);
    const int cc = col0 + cl;
    if (cc >= n_lim)
      continue;
    float* colp = p.C + (long long)cc * p.ldc;
DLAFB_WIDE_UNROLL
    for (int i = 0; i < 8; ++i) {
      const int rl = 4 * tx + (i & 3) + 64 * (i >> 2);
      const int rr = row0 + rl;
      if (rr >= m_lim)
        continue;
      if (lower && (trow + rl < tcol + cl))
        continue;
      float v = acc[i][j] * alpha;
      if (beta != 0.f)
        v = v + colp[rr] * beta;
      colp[rr] = v;
    }
  }
}

}  // namespace dlafb
