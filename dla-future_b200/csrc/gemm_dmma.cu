// TMA + DMMA FP64 contraction kernel (see gemm_dmma.cuh). sm_100a only.
#include "gemm_dmma.cuh"

#include "common.cuh"

#include <cstdio>

namespace dlafb {
namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
  } while (!ok);
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2,
                                            uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, int c3,
                                            uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(bar)
      : "memory");
}
__device__ __forceinline__ double lds_f64(uint32_t addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_f64(uint32_t addr, double v) {
  asm volatile("st.shared.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory");
}
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

// BM x BN block of C per CTA; 8 warps as 2 (rows) x 4 (cols), each owning (BM/2) x (BN/4) of it as F x G DMMA
// fragments. 128 x 128 is the throughput shape; the smaller ones exist for the launches on the critical path of the
// factorization (few CTAs, each a serial chain of K/16 slices at 128 flop/clk/SM).
template <int BM, int BN>
__global__ void __launch_bounds__(kGemmThreads, 1)
    gemm_nt_dmma_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                        const GemmParams p) {
  // ---- which block of C ----
  constexpr int F = BM / 16, G = BN / 32;  // 8-row / 8-column fragments per warp
  constexpr int kABytes = BM * kBK * 8, kBBytes = BN * kBK * 8, kStageB = kABytes + kBBytes;
  static_assert(F >= 2 && G >= 2 && F % 2 == 0 && G % 2 == 0, "a warp owns whole 16-row groups of both operands");
  int trow, tcol, row0, col0, lower, m_lim = p.m_total, n_lim = p.n_total;
  int b_from_a = 0, b_row_a = 0;  // LIST mode, diagonal tile: both operands come from the row panel (HERK)
  if (p.list != nullptr) {
    const int sub_m = p.tile / BM, nsub = sub_m * (p.tile / BN);
    const int item = blockIdx.x / nsub;
    const int s = blockIdx.x - item * nsub;
    const int2 it = p.list[item];
    lower = it.y < 0;
    trow = (s % sub_m) * BM;
    tcol = (s / sub_m) * BN;
    row0 = it.x + trow;
    col0 = (it.y & 0x7fffffff) + tcol;
    b_from_a = lower;
    b_row_a = it.x + tcol + p.a_row_off;
    m_lim = min(m_lim, it.x + p.tile);
    n_lim = min(n_lim, (it.y & 0x7fffffff) + p.tile);
  }
  else {
    lower = p.lower;
    trow = blockIdx.x * BM;
    tcol = blockIdx.y * BN;
    row0 = p.rect_row0 + trow;
    col0 = p.rect_col0 + tcol;
  }
  if (row0 >= m_lim || col0 >= n_lim)
    return;
  if (lower && trow + BM - 1 < tcol)
    return;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_full = base + kStages * kStageB;
  const uint32_t bar_empty = bar_full + 8 * kStages;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 8);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();

  const int nk = p.K / kBK;

  // ---- pull this block of C into L2 while the contraction runs (the epilogue reads it) ----
  if (p.beta != 0.0 && p.beta != 1.0) {
    const int rows = min(BM, m_lim - row0);
    const int cols = min(BN, n_lim - col0);
    const int segs = (rows + 15) >> 4;  // 128-byte lines per column
    for (int i = threadIdx.x; i < cols * segs; i += kGemmThreads) {
      const int c = i / segs, sgm = i - c * segs;
      const double* ptr = p.C + (long long)(col0 + c) * p.ldc + row0 + sgm * 16;
      asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr));
    }
  }

  // ---- TMA producer role: lane 0 of warp 0 feeds the ring of stages, kPrefetch slices ahead of the math ----
  const bool producer = (threadIdx.x == 0);
  const int a_row = row0 + p.a_row_off;
  const int b_row = b_from_a ? b_row_a : col0 + p.b_row_off;
  const CUtensorMap* mb = b_from_a ? &map_a : &map_b;
  const int nbox_b = b_from_a ? p.nbox_a : p.nbox_b;
  const int b_col0 = b_from_a ? p.a_col0 : p.b_col0;
  auto issue = [&](int kt) {
    const int s = kt % kStages;
    const uint32_t ph = (kt / kStages) & 1;
    mbar_wait(bar_empty + 8 * s, ph ^ 1);
    const uint32_t full = bar_full + 8 * s;
    mbar_expect_tx(full, kStageB);
    const uint32_t sa = base + s * kStageB;
    const uint32_t sb = sa + kABytes;
    const int ka = p.a_col0 + kt * kBK;
    const int kb = b_col0 + kt * kBK;
    if (p.nbox_a == 1)
      tma_load_3d(sa, &map_a, 0, ka, a_row >> 4, full);
    else if (p.nbox_a == 4)
      tma_load_4d(sa, &map_a, 0, ka, (a_row % p.tile) >> 4, a_row / p.tile, full);
    else
      for (int g = 0; g < BM / 16; ++g)
        tma_load_2d(sa + g * kGroupBytes, &map_a, a_row + 16 * g, ka, full);
    if (nbox_b == 1)
      tma_load_3d(sb, mb, 0, kb, b_row >> 4, full);
    else if (nbox_b == 4)
      tma_load_4d(sb, mb, 0, kb, (b_row % p.tile) >> 4, b_row / p.tile, full);
    else
      for (int g = 0; g < BN / 16; ++g)
        tma_load_2d(sb + g * kGroupBytes, mb, b_row + 16 * g, kb, full);
  };
  if (producer)
    for (int kt = 0; kt < kPrefetch && kt < nk; ++kt)
      issue(kt);

  // ---- DMMA: 8 warps as 2 (rows) x 4 (cols); each owns a (BM/2) x (BN/4) block of C ----
  const int wm = warp & 1;
  const int wn = warp >> 1;
  const int r = lane >> 2;   // fragment row (A) / column (B)
  const int kk = lane & 3;   // fragment k index; MMA step s4 uses k-columns 8*(s4/2) + 2*kk + (s4&1)

  double acc[F][G][2];
#pragma unroll
  for (int f = 0; f < F; ++f)
#pragma unroll
    for (int g = 0; g < G; ++g)
      acc[f][g][0] = acc[f][g][1] = 0.0;

  // Swizzle<3,4,3> of the 128B rows: 16-byte chunk index ^= (k-column & 7). With the k permutation above
  // the 16 lanes of a half-warp hit 16 distinct 8-byte bank pairs.
  uint32_t xo[2][2];
#pragma unroll
  for (int fb = 0; fb < 2; ++fb)
#pragma unroll
    for (int sb = 0; sb < 2; ++sb)
      xo[fb][sb] = (((uint32_t)((fb << 2) | (r >> 1)) ^ (uint32_t)((kk << 1) | sb)) << 4) + ((r & 1) << 3);
  const uint32_t a_thr = (wm * (F / 2)) * kGroupBytes + (2 * kk) * 128;
  const uint32_t b_thr = kABytes + (wn * (G / 2)) * kGroupBytes + (2 * kk) * 128;

  for (int kt = 0; kt < nk; ++kt) {
    // refill the stage freed two slices ago: its empty barrier is (almost always) already complete
    if (producer && kt + kPrefetch < nk)
      issue(kt + kPrefetch);
    const int s = kt % kStages;
    const uint32_t ph = (kt / kStages) & 1;
    mbar_wait(bar_full + 8 * s, ph);
    const uint32_t st = base + s * kStageB;
#pragma unroll
    for (int s4 = 0; s4 < 4; ++s4) {
      const uint32_t koff = (8 * (s4 >> 1) + (s4 & 1)) * 128;
      double a[F], b[G];
#pragma unroll
      for (int f = 0; f < F; ++f)
        a[f] = lds_f64(st + a_thr + (f >> 1) * kGroupBytes + koff + xo[f & 1][s4 & 1]);
#pragma unroll
      for (int g = 0; g < G; ++g)
        b[g] = lds_f64(st + b_thr + (g >> 1) * kGroupBytes + koff + xo[g & 1][s4 & 1]);
#pragma unroll
      for (int f = 0; f < F; ++f)
#pragma unroll
        for (int g = 0; g < G; ++g)
          dmma884(acc[f][g][0], acc[f][g][1], a[f], b[g]);
    }
    __syncwarp();
    if (lane == 0)
      mbar_arrive(bar_empty + 8 * s);
  }

  // ---- epilogue: C <- beta*C + alpha*acc, clipped and (optionally) restricted to the lower triangle ----
  const double alpha = p.alpha, beta = p.beta;
  const bool full_block = !lower || (trow >= tcol + BN - 1);
  if (full_block && row0 + BM <= m_lim && col0 + BN <= n_lim) {
    // interior block: no predicates; the 16 loads of column group g+1 are in flight while group g is stored
    double* cb = p.C + (long long)(col0 + wn * (BN / 4) + 2 * kk) * p.ldc + (row0 + wm * (BM / 2) + r);
    const long long ldc = p.ldc;
    double* c00 = p.C + (long long)col0 * ldc + row0;
    if (beta == 1.0 && (reinterpret_cast<uintptr_t>(c00) & 15) == 0 && (ldc & 1) == 0) {
      // C += alpha*acc without reading C into the SM: the block is staged in shared memory (the operand ring is
      // free now) and added to global memory by bulk reduce copies, one per column; the adds execute in L2 and the
      // CTA only waits until its shared memory has been read. One rounding per element, exactly like fma(1, c, v).
      // Column stride BM*8 + 16 bytes: the 16 lanes of a half-warp (4 columns x 4 row pairs) hit 32 distinct banks.
      constexpr uint32_t kColStride = BM * 8 + 16;
      __syncthreads();  // every warp is done with the operand ring
      const uint32_t mine = base + (wn * (BN / 4) + 2 * kk) * kColStride + (wm * (BM / 2) + r) * 8;
#pragma unroll
      for (int g = 0; g < G; ++g)
#pragma unroll
        for (int e = 0; e < 2; ++e)
#pragma unroll
          for (int f = 0; f < F; ++f)
            sts_f64(mine + (g * 8 + e) * kColStride + f * 64, alpha * acc[f][g][e]);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncthreads();
      if (threadIdx.x < BN) {
        double* gp = c00 + (long long)threadIdx.x * ldc;
        asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f64 [%0], [%1], %2;" ::"l"(gp),
                     "r"(base + threadIdx.x * kColStride), "r"(BM * 8)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      }
      return;
    }
    if (beta != 0.0) {
      double cv[2][2][F];
#pragma unroll
      for (int e = 0; e < 2; ++e)
#pragma unroll
        for (int f = 0; f < F; ++f)
          cv[0][e][f] = __ldcg(cb + e * ldc + f * 8);
#pragma unroll
      for (int g = 0; g < G; ++g) {
        if (g + 1 < G) {
#pragma unroll
          for (int e = 0; e < 2; ++e)
#pragma unroll
            for (int f = 0; f < F; ++f)
              cv[(g + 1) & 1][e][f] = __ldcg(cb + ((g + 1) * 8 + e) * ldc + f * 8);
        }
#pragma unroll
        for (int e = 0; e < 2; ++e)
#pragma unroll
          for (int f = 0; f < F; ++f)
            __stcg(cb + (g * 8 + e) * ldc + f * 8, fma(beta, cv[g & 1][e][f], alpha * acc[f][g][e]));
      }
    }
    else {
#pragma unroll
      for (int g = 0; g < G; ++g)
#pragma unroll
        for (int e = 0; e < 2; ++e)
#pragma unroll
          for (int f = 0; f < F; ++f)
            __stcg(cb + (g * 8 + e) * ldc + f * 8, alpha * acc[f][g][e]);
    }
    return;
  }
#pragma unroll
  for (int g = 0; g < G; ++g) {
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int cl = wn * (BN / 4) + g * 8 + 2 * kk + e;  // column inside the CTA block
      const int c = col0 + cl;
      if (c >= n_lim)
        continue;
      double* colp = p.C + (long long)c * p.ldc;
#pragma unroll
      for (int f = 0; f < F; ++f) {
        const int rl = wm * (BM / 2) + f * 8 + r;
        const int rr = row0 + rl;
        if (rr >= m_lim)
          continue;
        if (!full_block && (trow + rl < tcol + cl))
          continue;
        double v = alpha * acc[f][g][e];
        if (beta != 0.0)
          v = fma(beta, colp[rr], v);
        colp[rr] = v;
      }
    }
  }
}

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      f = nullptr;
    return reinterpret_cast<EncodeTiledFn>(f);
  }();
  return fn;
}

}  // namespace

int make_operand_map(CUtensorMap* map, const double* base, long long rows, long long cols, long long ld,
                     int kind, int box_rows) {
  const cuuint32_t groups = (cuuint32_t)(box_rows / 16);
  EncodeTiledFn enc = encode_fn();
  if (!enc)
    return -1;
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld & 1) || rows <= 0 || cols <= 0)
    return -2;
  CUresult r;
  if (kind == 3) {
    cuuint64_t dims[3] = {16, (cuuint64_t)cols, (cuuint64_t)((rows + 15) / 16)};
    cuuint64_t strides[2] = {(cuuint64_t)ld * 8, 128};
    cuuint32_t box[3] = {16, (cuuint32_t)kBK, groups};
    cuuint32_t es[3] = {1, 1, 1};
    r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<double*>(base), dims, strides, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  else if (kind == 4) {
    // tile layout (contiguous ld x ld tiles, ld = tile size): rows = number of tiles * ld, cols = ld
    const cuuint64_t nb = (cuuint64_t)ld;
    cuuint64_t dims[4] = {16, nb, nb / 16, (cuuint64_t)(rows / ld)};
    cuuint64_t strides[3] = {nb * 8, 128, nb * nb * 8};
    cuuint32_t box[4] = {16, (cuuint32_t)kBK, groups, 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 4, const_cast<double*>(base), dims, strides, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  else {
    cuuint64_t dims[2] = {(cuuint64_t)rows, (cuuint64_t)cols};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 8};
    cuuint32_t box[2] = {16, (cuuint32_t)kBK};
    cuuint32_t es[2] = {1, 1};
    r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double*>(base), dims, strides, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  return r == CUDA_SUCCESS ? 0 : (int)r;
}

namespace {
template <int BM, int BN>
cudaError_t launch_shape(cudaStream_t stream, const CUtensorMap& map_a, const CUtensorMap& map_b, const GemmParams& p,
                         unsigned nx, unsigned ny) {
  constexpr int smem = gemm_smem_bytes(BM, BN);
  static cudaError_t attr =
      cudaFuncSetAttribute(gemm_nt_dmma_kernel<BM, BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (attr != cudaSuccess)
    return attr;
  gemm_nt_dmma_kernel<BM, BN><<<dim3(nx, ny), kGemmThreads, smem, stream>>>(map_a, map_b, p);
  ++g_launches;
  return cudaGetLastError();
}
}  // namespace

cudaError_t launch_gemm_nt_dmma(cudaStream_t stream, const CUtensorMap& map_a, const CUtensorMap& map_b,
                                const GemmParams& p, unsigned nblocks_x, unsigned nblocks_y, GemmShape shape) {
  if (nblocks_x == 0 || nblocks_y == 0)
    return cudaSuccess;
  switch (shape) {
    case GemmShape::k128x128:
      return launch_shape<128, 128>(stream, map_a, map_b, p, nblocks_x, nblocks_y);
    case GemmShape::k64x64:
      return launch_shape<64, 64>(stream, map_a, map_b, p, nblocks_x, nblocks_y);
    case GemmShape::k32x128:
      return launch_shape<32, 128>(stream, map_a, map_b, p, nblocks_x, nblocks_y);
  }
  return cudaErrorInvalidValue;
}

}  // namespace dlafb
