// Included by potf2.cu inside its anonymous namespace (uses kPotf2Threads, STAMP, potf2_trailing, shfl).
//
// Round 2: the full 128 x 128 double block (the inner block of the 512 x 512 diagonal tile), factor + inverse, with
// DELAYED trailing updates. ncu of the generic kernel (profiles/r02/ncu_potf2_v1.json): 152 k cycles, 280 k warp
// instructions, issue slots 48 % busy, FP64 pipe 21 %: the rank-8 update of the whole trailing matrix after every
// panel spends as many instructions on the read-modify-write of S as on the FMAs (K = 8), and the panel issues the
// next pivot's rsqrt only after the whole column's updates. Here:
//   * columns are grouped in super-panels of 32. After each 8-column panel only the REST OF THE SUPER-PANEL (<= 24
//     columns) gets the rank-8 update (the scalar code of potf2.cu, column-limited);
//   * at the end of a super-panel everything to its right gets ONE rank-32 update on the FP64 tensor pipe
//     (mma.sync.m8n8k4.f64 on 8 x 8 fragments of S, operands from the super-panel buffer XS[l][row]): one
//     read-modify-write of S per 32 columns instead of four;
//   * in the panel the reciprocal square root of the NEXT pivot is issued as soon as that pivot is known (a
//     branch-free MUFU.RSQ64H + the cubic Newton step ::rsqrt uses), ahead of the current column's updates.
// Same augmented-matrix formulation as the generic kernel: E = inv(L)^H grows in the strict upper triangle of S,
// W = E^H. Row r of XS holds, for the big update, L(r, c0 + l) when r >= c0 + 32 and E(r, c0 + l) when r < c0 + 32
// (zero where row r of E has not started yet).
// Measured (profiles/r02/potf2_d128_timing.json): 75.3 -> 65.2 us for the block with the scalar rank-8 update and one
// DMMA fragment at a time, 58.0 us with both updates on p128_dmma_update; the 512 tile 358.7 -> 292 - 304 us.
__device__ __forceinline__ double rsqrt_nobranch(double d) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
  const double e = fma(-(y * y), d, 1.0);
  const double p = fma(e, 0.375, 0.5);
  return fma(p, y * e, y);
}

__device__ __forceinline__ void dmma884_acc(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

constexpr int kP128 = 128;           // block order handled by the kernel below
// Strides chosen for the DMMA fragment accesses (lane = 4 r + kk; a half-warp is r = 0..3, kk = 0..3; 16 bank pairs of
// 8 bytes): operands X[(4 i + kk) * XLD + 8 f + r] need kk * XLD + r distinct mod 16 -> XLD = 4 mod 16; accumulators
// S[(8 fa + r) + (8 fb + 2 kk + e) * LDS] need 2 kk * LDS + r distinct mod 16 -> LDS = 2 mod 4. (ncu of the first
// version, LDS = 137 and XLD = 136: 13.4 k bank conflicts in 50 k shared-memory wavefronts,
// profiles/r02/ncu_potf2_d128.json.) Row walks (lanes along a column) are conflict-free for any stride; only the final
// read of E along rows pays a 2-way conflict (1 k wavefronts, once).
constexpr int kP128Lds = kP128 + 2;
constexpr int kP128Xld = kP128 + 4;
constexpr int kP128Sp = 32;          // super-panel width
constexpr int kP128Pld = kP128 + 4;  // Xp row stride: k * 132 + row hits 16 distinct bank pairs per half-warp
constexpr size_t kP128Smem =
    ((size_t)kP128Lds * (kP128 + 1) + (size_t)8 * kP128Pld + (size_t)kP128Sp * kP128Xld + kP128) * sizeof(double) + 16;

// S(:, 8 cb0 .. 8 fb_end) -= X X^T on the FP64 tensor pipe, X = the K columns just factored ([k][row], row stride XLD):
// E rows (fragment rows fa < cb0) take every column, L rows only the fragments on or below the diagonal. Warp w owns the
// 8-row fragment fa = w: its A operand is loaded once and reused for all its column fragments, four of which are in
// flight at a time (the first version did one fragment at a time: 8 dependent DMMAs, 1.3 k cycles per fragment).
template <int K, int XLD>
__device__ __forceinline__ void p128_dmma_update(double* __restrict__ S, const double* __restrict__ X, int cb0,
                                                 int fb_end, int warp, int lane) {
  constexpr int LDS = kP128Lds;
  const int r = lane >> 2, kk = lane & 3;
  const int fa = warp;
  const int fb_hi = fa < cb0 ? fb_end : min(fb_end, fa + 1);  // exclusive
  if (cb0 >= fb_hi)
    return;  // warp-uniform
  double xa[K / 4];
#pragma unroll
  for (int i = 0; i < K / 4; ++i)
    xa[i] = -X[(4 * i + kk) * XLD + 8 * fa + r];
  for (int fb0 = cb0; fb0 < fb_hi; fb0 += 4) {
    double acc[4][2];
    double* cp[4];
    const double* xb[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int fb = min(fb0 + u, fb_hi - 1);  // beyond the end: recompute the last fragment, do not store it
      cp[u] = S + (8 * fa + r) + (8 * fb + 2 * kk) * LDS;
      xb[u] = X + kk * XLD + 8 * fb + r;
      acc[u][0] = cp[u][0];
      acc[u][1] = cp[u][LDS];
    }
#pragma unroll
    for (int i = 0; i < K / 4; ++i)
#pragma unroll
      for (int u = 0; u < 4; ++u)
        dmma884_acc(acc[u][0], acc[u][1], xa[i], xb[u][4 * i * XLD]);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int fb = fb0 + u;
      // the diagonal fragment of the L part keeps its strict upper triangle (E values live there)
      const bool diag = fa == fb;
      if (fb < fb_hi && (!diag || r >= 2 * kk))
        cp[u][0] = acc[u][0];
      if (fb < fb_hi && (!diag || r >= 2 * kk + 1))
        cp[u][LDS] = acc[u][1];
    }
  }
}

// One 8-column panel (columns jb .. jb+7) for this thread's row: see the generic kernel for the roles.
// Leaves the finished panel row in x[], the inverse pivots in dinv, the first failing column (1-based) in fail.
__device__ __forceinline__ void p128_panel(const double* __restrict__ Sp, int jb, int tid, int row, bool roleA,
                                           bool act, double (&x)[8], double* __restrict__ dinv, int& fail) {
  constexpr int BS = 8, LDS = kP128Lds;
  const int t = row - jb;
  const int r = tid & (BS - 1);
  double Dr[BS], dg[BS];
#pragma unroll
  for (int c = 0; c < BS; ++c) {
    const double dv = Sp[(jb + r) + c * LDS];
    const double gv = Sp[(jb + c) + c * LDS];
    const double xv = Sp[row + c * LDS];
    Dr[c] = (r >= c) ? dv : 0.0;
    dg[c] = gv;
    const bool keep = roleA ? (act && (t >= BS || c <= t)) : (act && t < 0);
    x[c] = keep ? xv : ((!roleA && act && t == c) ? 1.0 : 0.0);
  }
  if (jb == 32) {  // fine-grained timing of one panel (DLAFB_TIMING builds only)
    STAMPT(52, 0);
  }
  double lraw[BS];
#pragma unroll
  for (int c = 1; c < BS; ++c)
    lraw[c] = shfl(Dr[0], (tid & 24) + c);  // element (c, 0) from the lane that keeps row c
  fail = (!(dg[0] > 0.0) && fail == 0) ? jb + 1 : fail;
  double is = rsqrt_nobranch(dg[0]);
#pragma unroll
  for (int j = 0; j < BS; ++j) {
    double is_next = 0.0, lcn = 0.0;
    if (j + 1 < BS) {
      // the next pivot first: its reciprocal square root is in flight during this column's updates
      lcn = lraw[j + 1] * is;
      dg[j + 1] = fma(-lcn, lcn, dg[j + 1]);
      fail = (!(dg[j + 1] > 0.0) && fail == 0) ? jb + j + 2 : fail;
      is_next = rsqrt_nobranch(dg[j + 1]);
    }
    if (tid == j)
      dinv[jb + j] = is;
    const double xs = x[j] * is;
    x[j] = xs;
    const double drj = Dr[j] * is;
#pragma unroll
    for (int c = j + 1; c < BS; ++c) {
      const double lc = (c == j + 1) ? lcn : lraw[c] * is;
      if (c > j + 1)
        dg[c] = fma(-lc, lc, dg[c]);
      Dr[c] = fma(-drj, lc, Dr[c]);
      x[c] = fma(-xs, lc, x[c]);
    }
    if (j + 1 < BS) {
#pragma unroll
      for (int c = j + 2; c < BS; ++c)
        lraw[c] = shfl(Dr[j + 1], (tid & 24) + c);  // element (c, j + 1), final after this column's update
    }
    is = is_next;
    if (jb == 32 && j == 0) {
      STAMPT(53, 0);
    }
    if (jb == 32 && j == 3) {
      STAMPT(54, 0);
    }
  }
  if (jb == 32) {
    STAMPT(55, 0);
    STAMPT(57, 128);  // first E-role warp
    STAMPT(58, 96);   // last A-role warp
  }
}

__global__ void __launch_bounds__(kPotf2Threads, 1)
    potf2_d128_kernel(double* __restrict__ a, long long lda, double* __restrict__ winv, long long ldw, int* info,
                      int info_base) {
  constexpr int N = kP128, LDS = kP128Lds, XLD = kP128Xld, BS = 8, SP = kP128Sp;
  extern __shared__ __align__(16) unsigned char smraw[];
  double* S = reinterpret_cast<double*>(smraw);   // lower: A -> L; strict upper: E = inv(L)^H
  double* Xp = S + (size_t)LDS * (N + 1);          // current panel, BS x N (rows < r0 from E, rows >= r0 from L)
  double* XS = Xp + BS * kP128Pld;                 // super-panel operand, SP x XLD
  double* dinv = XS + SP * XLD;                    // 1 / L_jj
  int* sfail = reinterpret_cast<int*>(dinv + N);

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int li = tid & (N - 1);
  const int lj0 = tid / N;
  constexpr int lstep = kPotf2Threads / N;

  for (int j = lj0; j < N; j += lstep)
    S[li + j * LDS] = (li >= j) ? a[li + j * lda] : 0.0;
  if (tid == 0)
    *sfail = 0;
  STAMP(0);
  __syncthreads();
  STAMP(1);

  const bool roleA = tid < N;
  const int row = roleA ? tid : tid - N;
  int fail = 0;
  for (int c0 = 0; c0 < N; c0 += SP) {
    for (int jb = c0; jb < c0 + SP; jb += BS) {
      const int r0 = jb + BS;
      if (tid < 2 * N) {
        double* Sp = S + jb * LDS;
        const int t = row - jb;
        const bool act = roleA ? (row >= jb) : (row < r0);
        double x[BS];
        p128_panel(Sp, jb, tid, row, roleA, act, x, dinv, fail);
        if (tid == 0)
          *sfail = fail;
        STAMP(4 + 3 * (jb / BS));
        // write the panel back: S (final values), Xp (rank-8 update inside the super-panel), XS (rank-32 update)
        const int l0 = jb - c0;
#pragma unroll
        for (int c = 0; c < BS; ++c) {
          const bool ps = act && (roleA ? (t >= BS || c <= t) : (row < jb + c));
          if (ps)
            Sp[row + c * LDS] = x[c];
          if (roleA ? row >= r0 : row < r0)
            Xp[c * kP128Pld + row] = x[c];
          if (roleA) {
            if (row >= r0)
              XS[(l0 + c) * XLD + row] = (row >= c0 + SP) ? x[c] : 0.0;  // rows of E that have not started: zero
          }
          else if (row < r0)
            XS[(l0 + c) * XLD + row] = x[c];
        }
        if (jb == 32) {
          STAMPT(56, 0);
          STAMPT(59, 128);
        }
      }
      __syncthreads();
      fail = *sfail;
      if (fail)
        break;  // uniform
      STAMP(5 + 3 * (jb / BS));
      if (r0 < c0 + SP) {  // rank-8 update of the rest of the super-panel
        p128_dmma_update<BS, kP128Pld>(S, Xp, r0 / 8, (c0 + SP) / 8, warp, lane);
        __syncthreads();
        STAMP(6 + 3 * (jb / BS));
      }
    }
    if (fail)
      break;
    const int cb0 = (c0 + SP) / 8;  // first fragment column right of the super-panel
    if (cb0 < N / 8) {
      // rank-32 update of everything right of the super-panel
      p128_dmma_update<SP, XLD>(S, XS, cb0, N / 8, warp, lane);
      __syncthreads();
      STAMP(6 + 3 * ((c0 + SP) / BS - 1));
    }
  }
  STAMP(60);
  if (fail) {
    if (tid == 0 && info != nullptr)
      atomicCAS(info, 0, info_base + fail);
    return;
  }
  for (int j = lj0; j <= li; j += lstep)
    a[li + j * lda] = S[li + j * LDS];
  STAMP(61);
  for (int c = lj0; c < N; c += lstep) {
    double v = 0.0;  // W(li, c) = conj(E(c, li))
    if (c < li)
      v = S[c + li * LDS];
    else if (c == li)
      v = dinv[li];
    winv[li + c * ldw] = v;
  }
  STAMP(62);
}
