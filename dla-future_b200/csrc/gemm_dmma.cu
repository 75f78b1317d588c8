// TMA + DMMA FP64 contraction kernel (see gemm_dmma.cuh). sm_100a only.
#include "gemm_dmma.cuh"

#include "common.cuh"

#include <algorithm>
#include <cstdio>
#include <cstdlib>

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
//
// Persistent: CTA b works on blocks b, b + gridDim.x, ... The TMA producer (lane 0 of warp 0) keeps kPrefetch slices
// in flight ACROSS block boundaries, so the first slices of the next block land while this block's epilogue runs;
// measured fixed cost per block before: 5.2 us of 73.8 us at K = 512 (profiles/r01/update_kernel_alone.json).
struct BlockCoord {
  int trow, tcol;      // inside the tile (LIST) / rectangle (RECT)
  int row0, col0;      // in C
  int lower;           // keep only elements on or below the diagonal of the tile / rectangle
  int m_lim, n_lim;    // clip
  int b_from_a;        // LIST mode, diagonal tile: both operands come from the row panel (HERK)
  int b_row_a;
};

template <int BM, int BN>
__device__ __forceinline__ bool decode_block(const GemmParams& p, int id, BlockCoord& c) {
  c.m_lim = p.m_total;
  c.n_lim = p.n_total;
  c.b_from_a = 0;
  c.b_row_a = 0;
  if (p.list != nullptr) {
    const int sub_m = p.tile / BM, nsub = sub_m * (p.tile / BN);
    const int item = id / nsub;
    const int s = id - item * nsub;
    const int2 it = p.list[item];
    const int ir = it.x * p.list_scale, ic = (it.y & 0x7fffffff) * p.list_scale;  // list entries are in elements of T
    c.lower = it.y < 0;
    c.trow = (s % sub_m) * BM;
    c.tcol = (s / sub_m) * BN;
    c.row0 = ir + c.trow;
    c.col0 = ic + c.tcol;
    c.b_from_a = c.lower;
    c.b_row_a = ir + c.tcol + p.a_row_off;
    c.m_lim = min(c.m_lim, ir + p.tile);
    c.n_lim = min(c.n_lim, ic + p.tile);
  }
  else {
    const int by = id / p.rect_nx;
    c.lower = p.lower;
    c.trow = (id - by * p.rect_nx) * BM;
    c.tcol = by * BN;
    c.row0 = p.rect_row0 + c.trow;
    c.col0 = p.rect_col0 + c.tcol;
  }
  if (c.row0 >= c.m_lim || c.col0 >= c.n_lim)
    return false;
  if (c.lower && c.trow + BM - 1 < c.tcol)
    return false;
  return true;
}

// CPLX: the operands are the REAL VIEWS of complex<double> matrices (re/im interleaved along the rows: 2m x K doubles,
// leading dimension 2*ld), so the same main loop forms the four real products of C = A * B^H at once:
//   P[2i][2j] = Ar.Br   P[2i][2j+1] = Ar.Bi   P[2i+1][2j] = Ai.Br   P[2i+1][2j+1] = Ai.Bi
// and the epilogue folds them: Cr = P[2i][2j] + P[2i+1][2j+1], Ci = P[2i+1][2j] - P[2i][2j+1]. In the DMMA fragment
// a lane holds P[r][2kk], P[r][2kk+1]; its partner row r^1 is lane^4: one shuffle per fragment. All coordinates in
// GemmParams are in real-view units (rows and product columns doubled); C is addressed as its real view
// (2m rows, n columns, leading dimension p.ldc doubles).
template <int BM, int BN, int STAGES, int MINB, bool CPLX>
__global__ void __launch_bounds__(kGemmThreads, MINB)
    gemm_nt_dmma_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                        const GemmParams p) {
  constexpr int F = BM / 16, G = BN / 32;  // 8-row / 8-column fragments per warp
  constexpr int kABytes = BM * kBK * 8, kBBytes = BN * kBK * 8, kStageB = kABytes + kBBytes;
  constexpr uint32_t kColStride = BM * 8 + 16;  // staged C: the 16 lanes of a half-warp hit 32 distinct banks
  static_assert(F >= 2 && G >= 2 && F % 2 == 0 && G % 2 == 0, "a warp owns whole 16-row groups of both operands");

  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t stg = base + STAGES * kStageB;  // half a block of C (BN/2 columns) for the reduce epilogue
  const uint32_t bar_full = stg + (BN / 2) * kColStride;
  const uint32_t bar_empty = bar_full + 8 * STAGES;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 8);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();

  const int nk = p.K / kBK;

  // ---- TMA producer role: lane 0 of warp 0 feeds the ring of stages, (STAGES - 2) slices ahead of the math ----
  const bool producer = (threadIdx.x == 0);
  int pr_id = blockIdx.x, pr_kt = 0;
  uint32_t pr_g = 0;  // slices issued so far (stage = pr_g % STAGES)
  BlockCoord pc;
  bool pr_valid = false;
  auto pr_seek = [&]() {
    while (pr_id < p.ntiles && !decode_block<BM, BN>(p, pr_id, pc))
      pr_id += gridDim.x;
    pr_valid = pr_id < p.ntiles;
    pr_kt = 0;
  };
  auto issue = [&]() {
    if (!pr_valid)
      return;
    const int s = pr_g % STAGES;
    const uint32_t ph = (pr_g / STAGES) & 1;
    mbar_wait(bar_empty + 8 * s, ph ^ 1);
    const uint32_t full = bar_full + 8 * s;
    mbar_expect_tx(full, kStageB);
    const uint32_t sa = base + s * kStageB;
    const uint32_t sb = sa + kABytes;
    const int a_row = pc.row0 + p.a_row_off;
    const int b_row = pc.b_from_a ? pc.b_row_a : pc.col0 + p.b_row_off;
    const CUtensorMap* mb = pc.b_from_a ? &map_a : &map_b;
    const int nbox_b = pc.b_from_a ? p.nbox_a : p.nbox_b;
    const int ka = p.a_col0 + pr_kt * kBK;
    const int kb = (pc.b_from_a ? p.a_col0 : p.b_col0) + pr_kt * kBK;
    if (p.nbox_a == 1)
      tma_load_3d(sa, &map_a, 0, ka, a_row >> 4, full);
    else if (p.nbox_a == 4)
      tma_load_4d(sa, &map_a, 0, ka, (a_row % p.tile) >> 4, a_row / p.tile, full);
    else
      for (int g = 0; g < BM / 16; ++g)
        tma_load_2d(sa + g * kGroupBytes, &map_a, a_row + 16 * g, ka, full);
    // a diagonal tile takes B through the A map, whose box has BM rows: BN / BM copies
    const int bbox = pc.b_from_a ? BM : BN;
    if (nbox_b == 1)
      for (int i = 0; i < BN; i += bbox)
        tma_load_3d(sb + i * 128, mb, 0, kb, (b_row + i) >> 4, full);
    else if (nbox_b == 4)
      for (int i = 0; i < BN; i += bbox)
        tma_load_4d(sb + i * 128, mb, 0, kb, ((b_row + i) % p.tile) >> 4, (b_row + i) / p.tile, full);
    else
      for (int g = 0; g < BN / 16; ++g)
        tma_load_2d(sb + g * kGroupBytes, mb, b_row + 16 * g, kb, full);
    ++pr_g;
    if (++pr_kt == nk) {
      pr_id += gridDim.x;
      pr_seek();
    }
  };
  if (producer) {
    pr_seek();
    for (int i = 0; i < (STAGES - 2); ++i)
      issue();
  }

  // ---- DMMA: 8 warps as 2 (rows) x 4 (cols); each owns a (BM/2) x (BN/4) block of C ----
  const int wm = warp & 1;
  const int wn = warp >> 1;
  const int r = lane >> 2;   // fragment row (A) / column (B)
  const int kk = lane & 3;   // fragment k index; MMA step s4 uses k-columns 8*(s4/2) + 2*kk + (s4&1)

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

  uint32_t cg = 0;          // slices consumed so far
  bool stg_pending = false;  // bulk copies of the previous block may still be reading the staging buffer

  for (int id = blockIdx.x; id < p.ntiles; id += gridDim.x) {
    BlockCoord c;
    if (!decode_block<BM, BN>(p, id, c))
      continue;
    const int row0 = c.row0, col0 = c.col0, m_lim = c.m_lim, n_lim = c.n_lim;

    // pull this block of C into L2 while the contraction runs (only when the epilogue reads it)
    if (p.beta != 0.0 && p.beta != 1.0) {
      const int rows = min(BM, m_lim - row0);
      const int cols = min(BN, n_lim - col0);
      const int segs = (rows + 15) >> 4;  // 128-byte lines per column
      for (int i = threadIdx.x; i < cols * segs; i += kGemmThreads) {
        const int cc = i / segs, sgm = i - cc * segs;
        const double* ptr = p.C + (long long)(col0 + cc) * p.ldc + row0 + sgm * 16;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr));
      }
    }

    double acc[F][G][2];
#pragma unroll
    for (int f = 0; f < F; ++f)
#pragma unroll
      for (int g = 0; g < G; ++g)
        acc[f][g][0] = acc[f][g][1] = 0.0;

    for (int kt = 0; kt < nk; ++kt, ++cg) {
      // refill the stage freed two slices ago: its empty barrier is (almost always) already complete
      if (producer)
        issue();
      const int s = cg % STAGES;
      const uint32_t ph = (cg / STAGES) & 1;
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
    const bool full_block = !c.lower || (c.trow >= c.tcol + BN - 1);
    const long long ldc = p.ldc;
    if constexpr (CPLX) {
      // fold the four real products; this lane then owns ONE double per fragment: real-view row rr of C (even: real
      // part, odd: imaginary part of complex row rr/2), complex column (col0 + cl)/2
      const double sgn = (r & 1) ? -1.0 : 1.0;
      double* c00 = p.C + (long long)(col0 >> 1) * ldc + row0;
      const bool interior = full_block && row0 + BM <= m_lim && col0 + BN <= n_lim;
      if (interior && beta == 1.0 && (reinterpret_cast<uintptr_t>(c00) & 15) == 0 && (ldc & 1) == 0) {
        if (stg_pending && threadIdx.x < BN / 2)
          asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncthreads();  // staging buffer free
        const uint32_t mine = stg + (wn * (BN / 8) + kk) * kColStride + (wm * (BM / 2) + r) * 8;
#pragma unroll
        for (int g = 0; g < G; ++g)
#pragma unroll
          for (int f = 0; f < F; ++f) {
            const double other = __shfl_xor_sync(0xffffffffu, acc[f][g][1], 4);
            sts_f64(mine + (g * 4) * kColStride + f * 64, alpha * fma(sgn, other, acc[f][g][0]));
          }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (threadIdx.x < BN / 2) {
          double* gp = c00 + (long long)threadIdx.x * ldc;
          asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f64 [%0], [%1], %2;" ::"l"(gp),
                       "r"(stg + threadIdx.x * kColStride), "r"(BM * 8)
                       : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        stg_pending = true;
        continue;
      }
#pragma unroll
      for (int g = 0; g < G; ++g) {
        const int cl = wn * (BN / 4) + g * 8 + 2 * kk;  // product column (even) inside the CTA block
        const int cc = (col0 + cl) >> 1;               // complex column of C
        double* colp = p.C + (long long)cc * ldc;
#pragma unroll
        for (int f = 0; f < F; ++f) {
          const double other = __shfl_xor_sync(0xffffffffu, acc[f][g][1], 4);  // all lanes take part
          const int rl = wm * (BM / 2) + f * 8 + r;
          const int rr = row0 + rl;
          if (col0 + cl >= n_lim || rr >= m_lim)
            continue;
          const int ci = (c.trow + rl) >> 1, cj = (c.tcol + cl) >> 1;  // complex position inside the tile / rectangle
          if (!full_block && ci < cj)
            continue;
          double v = alpha * fma(sgn, other, acc[f][g][0]);
          if (beta != 0.0)
            v = fma(beta, colp[rr], v);
          if (c.lower && ci == cj && (rr & 1))
            v = 0.0;  // HERK: the diagonal stays real
          colp[rr] = v;
        }
      }
      continue;
    }
    if (full_block && row0 + BM <= m_lim && col0 + BN <= n_lim) {
      double* c00 = p.C + (long long)col0 * ldc + row0;
      if (beta == 1.0 && (reinterpret_cast<uintptr_t>(c00) & 15) == 0 && (ldc & 1) == 0) {
        // C += alpha*acc without reading C into the SM: half a block at a time is staged in shared memory and added
        // to global memory by bulk reduce copies, one per column; the adds execute in L2 while this CTA goes on
        // with the next block. One rounding per element, exactly like fma(1, c, v).
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          if (stg_pending && threadIdx.x < BN / 2)
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
          __syncthreads();  // staging buffer free
          if ((wn >> 1) == h) {
            const uint32_t mine = stg + ((wn & 1) * (BN / 4) + 2 * kk) * kColStride + (wm * (BM / 2) + r) * 8;
#pragma unroll
            for (int g = 0; g < G; ++g)
#pragma unroll
              for (int e = 0; e < 2; ++e)
#pragma unroll
                for (int f = 0; f < F; ++f)
                  sts_f64(mine + (g * 8 + e) * kColStride + f * 64, alpha * acc[f][g][e]);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          }
          __syncthreads();
          if (threadIdx.x < BN / 2) {
            double* gp = c00 + (long long)(h * (BN / 2) + threadIdx.x) * ldc;
            asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f64 [%0], [%1], %2;" ::"l"(gp),
                         "r"(stg + threadIdx.x * kColStride), "r"(BM * 8)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
          stg_pending = true;
        }
        continue;
      }
      // interior block: no predicates; the loads of column group g+1 are in flight while group g is stored
      double* cb = p.C + (long long)(col0 + wn * (BN / 4) + 2 * kk) * ldc + (row0 + wm * (BM / 2) + r);
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
      continue;
    }
#pragma unroll
    for (int g = 0; g < G; ++g) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int cl = wn * (BN / 4) + g * 8 + 2 * kk + e;  // column inside the CTA block
        const int cc = col0 + cl;
        if (cc >= n_lim)
          continue;
        double* colp = p.C + (long long)cc * ldc;
#pragma unroll
        for (int f = 0; f < F; ++f) {
          const int rl = wm * (BM / 2) + f * 8 + r;
          const int rr = row0 + rl;
          if (rr >= m_lim)
            continue;
          if (!full_block && (c.trow + rl < c.tcol + cl))
            continue;
          double v = alpha * acc[f][g][e];
          if (beta != 0.0)
            v = fma(beta, colp[rr], v);
          colp[rr] = v;
        }
      }
    }
  }
  // the staging buffer must outlive the bulk copies that read it
  if (stg_pending && threadIdx.x < BN / 2)
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
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
                     int kind, int box_rows, long long tile_cols) {
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
    // tile layout (contiguous tiles of ld rows x tile_cols columns, leading dimension ld): rows = tiles * ld
    const cuuint64_t tr = (cuuint64_t)ld, tc = (cuuint64_t)(tile_cols > 0 ? tile_cols : ld);
    cuuint64_t dims[4] = {16, tc, tr / 16, (cuuint64_t)(rows / ld)};
    cuuint64_t strides[3] = {tr * 8, 128, tr * tc * 8};
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
int sm_count_impl() {
  static int n = [] {
    int dev = 0, v = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0)
      v = 148;
    return v;
  }();
  return n;
}

template <int BM, int BN, int STAGES, int MINB, bool CPLX>
cudaError_t launch_shape(cudaStream_t stream, const CUtensorMap& map_a, const CUtensorMap& map_b, GemmParams p,
                         unsigned nx, unsigned ny, unsigned max_ctas) {
  constexpr int smem = gemm_smem_bytes(BM, BN, STAGES);
  static_assert(MINB * (smem + 1024) <= 228 * 1024, "MINB CTAs must fit the SM's shared memory");
  static cudaError_t attr = cudaFuncSetAttribute(gemm_nt_dmma_kernel<BM, BN, STAGES, MINB, CPLX>,
                                                 cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (attr != cudaSuccess)
    return attr;
  p.ntiles = (int)(nx * ny);
  p.rect_nx = (int)nx;
  // One block per CTA by default: a persistent grid would keep every SM busy for the whole launch and starve the
  // high-priority stream (diagonal tile, panel solve, look-ahead column), which gets an SM only when a CTA retires.
  // DLAFB_GEMM_PERSISTENT=1 makes CTA b loop over blocks b, b + grid, ... (kernel microbenchmarks).
  static const bool persistent = [] {
    const char* e = getenv("DLAFB_GEMM_PERSISTENT");
    return e && e[0] == '1';
  }();
  unsigned grid = (unsigned)p.ntiles;
  if (persistent)
    grid = std::min<unsigned>(grid, (unsigned)sm_count_impl() * MINB);
  if (max_ctas > 0)
    grid = std::min<unsigned>(grid, max_ctas);  // persistent on purpose: the caller keeps SM slots free
  if (p.list_scale < 1)
    p.list_scale = 1;
  gemm_nt_dmma_kernel<BM, BN, STAGES, MINB, CPLX><<<grid, kGemmThreads, smem, stream>>>(map_a, map_b, p);
  ++g_launches;
  return cudaGetLastError();
}
}  // namespace

int gemm_sm_count() {
  return sm_count_impl();
}

cudaError_t launch_gemm_nt_dmma(cudaStream_t stream, const CUtensorMap& map_a, const CUtensorMap& map_b,
                                const GemmParams& p, unsigned nblocks_x, unsigned nblocks_y, GemmShape shape,
                                bool cplx, unsigned max_ctas) {
  if (nblocks_x == 0 || nblocks_y == 0)
    return cudaSuccess;
#define DLAFB_SHAPE(BM, BN, ST, MB)                                                                  \
  return cplx ? launch_shape<BM, BN, ST, MB, true>(stream, map_a, map_b, p, nblocks_x, nblocks_y, max_ctas)    \
              : launch_shape<BM, BN, ST, MB, false>(stream, map_a, map_b, p, nblocks_x, nblocks_y, max_ctas)
  switch (shape) {
    case GemmShape::k128x128:
      DLAFB_SHAPE(128, 128, 4, 1);
    case GemmShape::k64x64:
      DLAFB_SHAPE(64, 64, 4, 2);
    case GemmShape::k32x128:
      DLAFB_SHAPE(32, 128, 4, 2);
    case GemmShape::k64x128:
      DLAFB_SHAPE(64, 128, 3, 2);
  }
#undef DLAFB_SHAPE
  return cudaErrorInvalidValue;
}

}  // namespace dlafb
