// Panel solve as ONE kernel (north_star: "Panel TRSM: one kernel"): B (m x nk) <- B * L^-H for the Lower factorization
// (reference call site: trsmPanelTile, include/dlaf/factorization/cholesky/impl.h:55-67 -> one cublasDtrsm per tile,
// include/dlaf/blas/tile.h:458-470), optionally also written in tile layout into the row panel that travels to the
// other ranks (fuses the pack_tiles launch that followed the solve).
//
// Round 1 ran the solve as 2*nk/128 - 1 dependent launches of the contraction kernel (7 at nb = 512): on a GPU whose
// SMs are full of bulk-update CTAs every one of them waits for slots to free up, which is what made the solve take
// 0.77 ms inside the 8-GPU run against 0.09 ms alone. The rows of B are independent, so here a
// CTA owns R rows, keeps them in shared memory for the whole substitution and streams the blocks of L and of the
// inverted diagonal blocks W_j = inv(L_jj) (left by potrf_tile) from L2:
//   for j = 0 .. nk/128 - 1:   B_j -= sum_{i<j} X_i * L_ji^T   (left-looking: K = 128 i, one accumulation)
//                              X_j  = B_j * W_j^T              (the 128-block "solve" is a product, as in round 1)
// Every product is R x 128 x 128 on the FP64 tensor pipe (mma.sync.m8n8k4.f64): 8 warps, warp w owns columns
// 16 w .. 16 w + 15 of the block; the 128 x 16 K-slices of the streamed operand are double-buffered with cp.async.
// R = 16: 116 KB of shared memory and <= 128 registers, so a CTA fits on an SM beside one resident bulk-update CTA
// (108 KB) instead of waiting for a whole SM; R = 32 for tall panels (half the L2 traffic per row).
// Measured alone (profiles/r02/chain_probe_*.json): m = 1024: 73.6 -> 54.2 us, m = 4096: 92.5 -> 69.8 us; about 60 % of
// the DMMA rate per CTA. A 4-stage ring of 8-column slices was slower (68.9 / 86.4 us: twice the barriers, and the
// loop is bound by the tensor pipe and the barriers, not by the L2 latency of the streamed operand).
// Bound: latency / L2 (it is on the critical path); algorithmic HBM bytes 8 * nk^2 / 2 + 2 * 8 * m * nk.
#include <cuda_runtime.h>

#include <cstdlib>

#include "gemm_dmma.cuh"
#include "tile_ops.cuh"

namespace dlafb {
namespace {

constexpr int kTfThreads = 256;
constexpr int kTfBlk = 128;          // column block = inner block of the diagonal tile
constexpr int kTfSlice = 16;         // K-slice of the streamed operand
constexpr int kTfWld = kTfBlk + 4;   // row stride of a staged slice: (k * 132 + n) hits 16 distinct bank pairs

struct TrsmFusedParams {
  double* B;
  long long ldb;
  int m, nk;            // nk = q * 128
  const double* L;      // nk x nk lower factor of the diagonal tile
  long long ldl;
  const double* W;      // inverted 128 x 128 diagonal blocks, block j at W + j * 128 * 128, ld = 128
  double* packed;       // optional: the solved panel in tile layout (row i -> slot i / tile), else nullptr
  int tile;
};

__device__ __forceinline__ void tf_dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}
__device__ __forceinline__ void tf_cp_async16(void* smem, const void* gmem) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void tf_cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tf_cp_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

template <int R>
constexpr size_t trsm_fused_smem(int nk) {
  return ((size_t)nk * (R + 4) + 2 * (size_t)kTfSlice * kTfWld) * sizeof(double);
}

template <int R>
__global__ void __launch_bounds__(kTfThreads, R == 16 ? 2 : 1) trsm_fused_kernel(const TrsmFusedParams p) {
  constexpr int BLD = R + 4;  // row stride of Bs (doubles): k * BLD + row hits 16 distinct bank pairs per half-warp
  constexpr int F = R / 8;    // 8-row fragments per warp
  constexpr int G = 2;        // 8-column fragments per warp (16 of the block's 128 columns)
  extern __shared__ __align__(16) unsigned char smraw[];
  double* Bs = reinterpret_cast<double*>(smraw);  // [nk][BLD]: this CTA's rows of B, column-major
  double* Ws = Bs + (size_t)p.nk * BLD;           // [2][kTfSlice][kTfWld]: staged K-slices of the streamed operand

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int r = lane >> 2, kk = lane & 3;
  const int row0 = blockIdx.x * R;
  const int nrows = min(R, p.m - row0);
  const int q = p.nk / kTfBlk;

  // ---- this CTA's rows of B -> shared memory (rows beyond m: zero) ----
  for (int i = tid; i < p.nk * R; i += kTfThreads) {
    const int c = i / R, rr = i - c * R;
    Bs[c * BLD + rr] = rr < nrows ? p.B[(long long)c * p.ldb + row0 + rr] : 0.0;
  }

  // products in the order (1) B_j -= X_i L_ji^T for i < j, (2) X_j = B_j W_j^T; slice g = product * 8 + K-slice
  const int nslices = (q * (q + 1) / 2) * (kTfBlk / kTfSlice);
  auto decode = [&](int prod, int& j, int& i) {
    j = 0;
    while (prod >= j + 1) {
      prod -= j + 1;
      ++j;
    }
    i = prod;  // i < j: update with block i; i == j: the diagonal block
  };
  // thread t stages n = 2 (t & 63) .. + 1 for the four k = (t >> 6) + 4 h of a slice
  const int sn = 2 * (tid & 63), sk = tid >> 6;
  auto issue = [&](int g) {
    int j, i;
    decode(g >> 3, j, i);
    const int k0 = (g & 7) * kTfSlice;
    const double* src;
    long long ld;
    if (i < j) {
      src = p.L + (long long)j * kTfBlk + ((long long)i * kTfBlk + k0) * p.ldl;
      ld = p.ldl;
    }
    else {
      src = p.W + (size_t)j * kTfBlk * kTfBlk + (size_t)k0 * kTfBlk;
      ld = kTfBlk;
    }
    double* dst = Ws + (size_t)(g & 1) * kTfSlice * kTfWld;
#pragma unroll
    for (int h = 0; h < 4; ++h) {
      const int k = sk + 4 * h;
      tf_cp_async16(dst + k * kTfWld + sn, src + (long long)k * ld + sn);
    }
    tf_cp_commit();
  };

  double acc[F][G][2];
#pragma unroll
  for (int f = 0; f < F; ++f)
#pragma unroll
    for (int gg = 0; gg < G; ++gg)
      acc[f][gg][0] = acc[f][gg][1] = 0.0;

  issue(0);
  for (int g = 0; g < nslices; ++g) {
    tf_cp_wait_all();
    __syncthreads();  // slice g has landed for everyone; everyone is done with slice g - 1 (and with Bs writes)
    if (g + 1 < nslices)
      issue(g + 1);
    int j, i;
    decode(g >> 3, j, i);
    const int s = g & 7;
    const double* a_base = Bs + ((size_t)i * kTfBlk + s * kTfSlice + kk) * BLD + r;
    const double* b_base = Ws + (size_t)(g & 1) * kTfSlice * kTfWld + kk * kTfWld + 16 * warp + r;
#pragma unroll
    for (int ks = 0; ks < kTfSlice / 4; ++ks) {
      double a[F], b[G];
#pragma unroll
      for (int f = 0; f < F; ++f)
        a[f] = a_base[ks * 4 * BLD + 8 * f];
#pragma unroll
      for (int gg = 0; gg < G; ++gg)
        b[gg] = b_base[ks * 4 * kTfWld + 8 * gg];
#pragma unroll
      for (int f = 0; f < F; ++f)
#pragma unroll
        for (int gg = 0; gg < G; ++gg)
          tf_dmma(acc[f][gg][0], acc[f][gg][1], a[f], b[gg]);
    }
    if (s != 7)
      continue;
    // ---- end of a product ----
    double* cj = Bs + ((size_t)j * kTfBlk + 16 * warp + 2 * kk) * BLD + r;  // this lane's first element of block j
    if (i < j - 1)
      continue;  // more update products for block j follow: keep accumulating
    if (i == j - 1) {
      // B_j -= acc on this warp's own columns (nobody else touches them; the next loop-top barrier publishes them)
#pragma unroll
      for (int f = 0; f < F; ++f)
#pragma unroll
        for (int gg = 0; gg < G; ++gg)
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            cj[(8 * gg + e) * BLD + 8 * f] -= acc[f][gg][e];
            acc[f][gg][e] = 0.0;
          }
    }
    else {
      // X_j = acc replaces B_j: every warp must have finished reading B_j first
      __syncthreads();
#pragma unroll
      for (int f = 0; f < F; ++f)
#pragma unroll
        for (int gg = 0; gg < G; ++gg)
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            cj[(8 * gg + e) * BLD + 8 * f] = acc[f][gg][e];
            acc[f][gg][e] = 0.0;
          }
    }
  }
  __syncthreads();

  // ---- the solved rows back to B, and into the packed row panel ----
  for (int i = tid; i < p.nk * R; i += kTfThreads) {
    const int c = i / R, rr = i - c * R;
    if (rr < nrows) {
      const double v = Bs[c * BLD + rr];
      const int grow = row0 + rr;
      p.B[(long long)c * p.ldb + grow] = v;
      if (p.packed != nullptr)
        p.packed[(size_t)(grow / p.tile) * p.tile * p.tile + (size_t)c * p.tile + grow % p.tile] = v;
    }
  }
}

template <int R>
cudaError_t launch_trsm_fused(cudaStream_t stream, const TrsmFusedParams& p) {
  const size_t smem = trsm_fused_smem<R>(p.nk);
  if (smem > 227 * 1024)
    return cudaErrorInvalidValue;
  // the largest size ever requested is kept as the kernel's limit (the attribute is per kernel, not per launch)
  static size_t granted = 0;
  if (smem > granted) {
    const cudaError_t e =
        cudaFuncSetAttribute(trsm_fused_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess)
      return e;
    granted = smem;
  }
  trsm_fused_kernel<R><<<(unsigned)ceil_div(p.m, R), kTfThreads, smem, stream>>>(p);
  ++g_launches;
  return cudaGetLastError();
}

}  // namespace

// Returns false when this solve is not the fused kernel's (type, block sizes, alignment, switched off): the caller
// then runs the multi-launch path. *err carries the launch status when it returns true.
bool trsm_panel_fused_d(cudaStream_t stream, int nk, const double* L, long long ldl, const double* W, int ib, double* B,
                        long long ldb, int m, double* packed, int tile, cudaError_t* err) {
  static const int mode = [] {  // DLAFB_TRSM_FUSED=0 off, 16 / 32 force the row-block size
    const char* e = getenv("DLAFB_TRSM_FUSED");
    return e ? atoi(e) : 1;
  }();
  // Tall panels stay on the multi-launch path: a CTA of the fused kernel streams all of L and W for its rows (4x the L2
  // traffic per row of the 128-row contraction blocks; alone m = 16384 takes 310 us against 222 us), and the early
  // steps, where panels are tall, hide the solve behind a long bulk update anyway. Measured with 16-row CTAs:
  // no limit: 1 GPU n = 32768 348.0 ms (345.1 without), 8 GPUs n = 65536 384.7 ms (389.1 with the first default, limit
  // 8192 and 32-row CTAs above 4736 rows; 387.6 ms on the multi-launch path); limit 20480: 1 GPU 346.0 ms
  // (profiles/r02/session3_summary.txt, profiles/r02/n8/). DLAFB_TRSM_FUSED_MAX_M overrides.
  static const int max_m = [] {
    const char* e = getenv("DLAFB_TRSM_FUSED_MAX_M");
    return e ? atoi(e) : 20480;
  }();
  if (mode == 0 || ib != kTfBlk || nk <= 0 || nk % kTfBlk != 0 || m <= 0 || m > max_m)
    return false;
  if ((reinterpret_cast<uintptr_t>(L) & 15) || (reinterpret_cast<uintptr_t>(W) & 15) || (ldl & 1))
    return false;  // 16-byte cp.async of the streamed operand
  if (packed != nullptr && (tile <= 0 || nk > tile))
    return false;
  TrsmFusedParams p{B, ldb, m, nk, L, ldl, W, packed, tile};
  // tall panels: 32 rows per CTA (half the operand traffic per row, one CTA per SM); otherwise 16 rows per CTA, two
  // CTAs per SM, a CTA fits beside a resident bulk-update CTA
  const bool fits16 = trsm_fused_smem<16>(nk) <= 227 * 1024, fits32 = trsm_fused_smem<32>(nk) <= 227 * 1024;
  if (!fits16)
    return false;
  // 16-row CTAs by default: a CTA fits beside a resident bulk-update CTA and starts as soon as one retires (in situ on
  // 8 GPUs the solve took 0.36 ms instead of 0.76 ms at k = 60); DLAFB_TRSM_FUSED=32 selects 32-row CTAs
  const bool tall = mode == 32;
  *err = (tall && fits32) ? launch_trsm_fused<32>(stream, p) : launch_trsm_fused<16>(stream, p);
  return true;
}

}  // namespace dlafb
