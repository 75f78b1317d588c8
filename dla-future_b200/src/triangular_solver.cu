// Device executor and C ABI of the local triangular solver (algorithm: triangular_solver.h).
#include "triangular_solver.h"

#include <cuda_runtime.h>

#include <cstdlib>
#include <new>

#include "dlafb_solver.h"
#include "op_trace.h"
#include "tile_ops.cuh"

namespace dlafb {
namespace {

// Operations go to the current lane's stream in program order; without look-ahead there is one lane, the caller's
// stream.
template <class T>
struct DeviceOps {
  cudaStream_t s;
  cudaStream_t lanes[2];
  cudaEvent_t* events;
  using R = typename Tr<T>::Real;
  void lane(int l) { s = lanes[l]; }
  int signal(int ev) { return (int)cudaEventRecord(events[ev], s); }
  int wait(int ev) { return (int)cudaStreamWaitEvent(s, events[ev], 0); }
  int gemm_nt(int M, int N, int K, R alpha, const T* A, long long lda, const T* B, long long ldb, R beta, T* C,
              long long ldc) {
    DLAFB_TRACE_OP("gemm_nt M=%d N=%d K=%d alpha=%g beta=%g A=%p lda=%lld B=%p ldb=%lld C=%p ldc=%lld", M, N, K, (double)alpha,
                   (double)beta, (const void*)A, lda, (const void*)B, ldb, (void*)C, ldc);
    return op_trace_finish(s, (int)dlafb::gemm_nt<T>(s, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, 0), "gemm_nt");
  }
  int transpose(int m, int n, const T* in, long long ldin, T* out, long long ldout, int conj, char tri) {
    DLAFB_TRACE_OP("transpose m=%d n=%d conj=%d tri=%c in=%p ldin=%lld out=%p ldout=%lld", m, n, conj, tri, (const void*)in,
                   ldin, (void*)out, ldout);
    return op_trace_finish(s, (int)dlafb::transpose<T>(s, m, n, in, ldin, out, ldout, conj, tri), "transpose");
  }
  int lacpy(int m, int n, const T* a, long long lda, T* b, long long ldb) {
    DLAFB_TRACE_OP("lacpy G m=%d n=%d a=%p lda=%lld b=%p ldb=%lld", m, n, (const void*)a, lda, (void*)b, ldb);
    return op_trace_finish(s, (int)dlafb::lacpy<T>(s, 'G', m, n, a, lda, b, ldb), "lacpy");
  }
  int tri_inverse(int upper, int unit, int n, int nb, const T* t, long long ldt, T* w) {
    TriInvParams p{n, nb, ldt, upper, unit};
    DLAFB_TRACE_OP("tri_inverse upper=%d unit=%d n=%d nb=%d t=%p ldt=%lld w=%p", upper, unit, n, nb, (const void*)t, ldt,
                   (void*)w);
    return op_trace_finish(s, (int)launch_tri_inverse<T>(s, p, t, w), "tri_inverse");
  }
  int scale(int m, int n, T* b, long long ldb, T alpha, int conj) {
    DLAFB_TRACE_OP("scale m=%d n=%d conj=%d b=%p ldb=%lld", m, n, conj, (void*)b, ldb);
    return op_trace_finish(s, (int)launch_scale<T>(s, m, n, b, ldb, alpha, conj), "scale");
  }
};

// Look-ahead on two streams (DLAFB_TRSOLVE_LOOKAHEAD=0/1 overrides).
constexpr int kDefaultLookahead = 0;

struct Plan {
  char type;
  TriSolveDesc d;
  size_t elem;
  void* tcopy = nullptr;
  void* winv = nullptr;
  void* panel = nullptr;
  size_t bytes = 0;
  // look-ahead: lane 0 = high-priority stream (block solves), lane 1 = bulk updates
  int lookahead = 0;
  cudaStream_t hp = nullptr, np = nullptr;
  cudaEvent_t ev[kTriSolveEvents] = {};
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  ~Plan() {
    if (hp)
      cudaStreamSynchronize(hp);
    if (np)
      cudaStreamSynchronize(np);
    for (cudaEvent_t e : ev)
      if (e)
        cudaEventDestroy(e);
    if (ev_fork)
      cudaEventDestroy(ev_fork);
    if (ev_join)
      cudaEventDestroy(ev_join);
    if (hp)
      cudaStreamDestroy(hp);
    if (np)
      cudaStreamDestroy(np);
    cudaFree(tcopy);
    cudaFree(winv);
    cudaFree(panel);
  }
};

template <class T>
int run_typed(Plan& p, const void* alpha, const void* a, long long lda, void* b, long long ldb, cudaStream_t s) {
  TriSolveWork<T> w{static_cast<T*>(p.tcopy), tri_solve_ldt(p.d), static_cast<T*>(p.winv), static_cast<T*>(p.panel),
                    tri_solve_ldp(p.d)};
  if (!p.lookahead) {
    DeviceOps<T> ops{s, {s, s}, p.ev};
    return triangular_solve<T>(ops, p.d, *static_cast<const T*>(alpha), static_cast<const T*>(a), lda,
                               static_cast<T*>(b), ldb, w);
  }
  w.lookahead = 1;
  w.panel_stride = tri_solve_panel_elems(p.d);
  // fork from the caller's stream, join back into it
  cudaError_t ce;
  if ((ce = cudaEventRecord(p.ev_fork, s)) != cudaSuccess || (ce = cudaStreamWaitEvent(p.hp, p.ev_fork, 0)) != cudaSuccess)
    return (int)ce;
  DeviceOps<T> ops{p.hp, {p.hp, p.np}, p.ev};
  const int e = triangular_solve<T>(ops, p.d, *static_cast<const T*>(alpha), static_cast<const T*>(a), lda,
                                    static_cast<T*>(b), ldb, w);
  // join even after an error so that the caller's stream stays ordered after whatever was enqueued
  cudaEventRecord(p.ev_join, p.np);
  cudaStreamWaitEvent(p.hp, p.ev_join, 0);
  cudaEventRecord(p.ev_join, p.hp);
  cudaStreamWaitEvent(s, p.ev_join, 0);
  return e;
}

bool one_of(char c, const char* set) {
  for (; *set; ++set)
    if (*set == c)
      return true;
  return false;
}

}  // namespace
}  // namespace dlafb

using namespace dlafb;

extern "C" {

int dlafb_triangular_solver_plan_create(dlafb_trsolve_plan_t* out, char type, char side, char uplo, char op, char diag,
                                        int64_t m, int64_t n, int64_t nb) {
  if (!out || !one_of(type, "sdcz") || !one_of(side, "LR") || !one_of(uplo, "LU") || !one_of(op, "NTC") ||
      !one_of(diag, "NU") || m < 0 || n < 0 || nb < 1 || m > (1LL << 30) || n > (1LL << 30) || nb > (1 << 14))
    return -1;
  Plan* p = new (std::nothrow) Plan;
  if (!p)
    return (int)cudaErrorMemoryAllocation;
  p->type = type;
  p->d = TriSolveDesc{side, uplo, op, diag, (int)m, (int)n, (int)nb};
  p->elem = type == 's' ? 4 : type == 'z' ? 16 : 8;
  const size_t bt = tri_solve_tcopy_elems(p->d) * p->elem;
  const size_t bw = tri_solve_winv_elems(p->d) * p->elem;
  const char* la = getenv("DLAFB_TRSOLVE_LOOKAHEAD");
  p->lookahead = la ? (la[0] != '0') : kDefaultLookahead;
  const size_t bp = tri_solve_panel_elems(p->d) * p->elem * (p->lookahead ? 2 : 1);
  cudaError_t e = cudaSuccess;
  if (m > 0 && n > 0 && p->lookahead) {
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    e = cudaStreamCreateWithPriority(&p->hp, cudaStreamNonBlocking, hi);
    if (e == cudaSuccess)
      e = cudaStreamCreateWithPriority(&p->np, cudaStreamNonBlocking, lo);
    for (int i = 0; i < kTriSolveEvents && e == cudaSuccess; ++i)
      e = cudaEventCreateWithFlags(&p->ev[i], cudaEventDisableTiming);
    if (e == cudaSuccess)
      e = cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming);
    if (e == cudaSuccess)
      e = cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming);
  }
  if (m > 0 && n > 0) {
    if (bt && e == cudaSuccess)
      e = cudaMalloc(&p->tcopy, bt);
    if (e == cudaSuccess)
      e = cudaMalloc(&p->winv, bw);
    if (e == cudaSuccess)
      e = cudaMalloc(&p->panel, bp);
    p->bytes = bt + bw + bp;
  }
  if (e != cudaSuccess) {
    delete p;
    return (int)e;
  }
  *out = reinterpret_cast<dlafb_trsolve_plan_t>(p);
  return 0;
}

int dlafb_triangular_solver_plan_run(dlafb_trsolve_plan_t plan, const void* alpha, const void* a, int64_t lda, void* b,
                                     int64_t ldb, void* stream) {
  Plan* p = reinterpret_cast<Plan*>(plan);
  if (!p || !alpha)
    return -1;
  if (p->d.m == 0 || p->d.n == 0)
    return 0;
  const int64_t order = tri_solve_order(p->d);
  if (!a || !b || lda < order || ldb < p->d.m)
    return -1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (p->type) {
    case 's': return run_typed<float>(*p, alpha, a, lda, b, ldb, s);
    case 'd': return run_typed<double>(*p, alpha, a, lda, b, ldb, s);
    case 'c': return run_typed<float2>(*p, alpha, a, lda, b, ldb, s);
    default: return run_typed<double2>(*p, alpha, a, lda, b, ldb, s);
  }
}

size_t dlafb_triangular_solver_plan_workspace_bytes(dlafb_trsolve_plan_t plan) {
  return plan ? reinterpret_cast<Plan*>(plan)->bytes : 0;
}

int dlafb_triangular_solver_plan_destroy(dlafb_trsolve_plan_t plan) {
  delete reinterpret_cast<Plan*>(plan);
  return 0;
}

}  // extern "C"
