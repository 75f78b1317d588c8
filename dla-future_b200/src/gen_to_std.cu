// Device executor and C ABI of the generalized-to-standard reduction on one GPU (algorithm: gen_to_std.h).
#include "gen_to_std.h"

#include <cuda_runtime.h>

#include <cstdlib>
#include <new>

#include "dist_device_ops.h"
#include "dlafb_gen_to_std.h"
#include "gen_to_std_dist.h"
#include "op_trace.h"
#include "tile_ops.cuh"

namespace dlafb {
namespace {

// Operations go to the current lane's stream in program order; without look-ahead there is one lane, the caller's
// stream.
template <class T>
struct G2sDeviceOps {
  cudaStream_t s;
  cudaStream_t lanes[2];
  cudaEvent_t* events;
  using R = typename Tr<T>::Real;
  void lane(int l) { s = lanes[l]; }
  int signal(int ev) { return (int)cudaEventRecord(events[ev], s); }
  int wait(int ev) { return (int)cudaStreamWaitEvent(s, events[ev], 0); }
  int gemm_nt(int M, int N, int K, R alpha, const T* A, long long lda, const T* B, long long ldb, R beta, T* C,
              long long ldc, int lower) {
    DLAFB_TRACE_OP("gemm_nt M=%d N=%d K=%d alpha=%g beta=%g lower=%d A=%p lda=%lld B=%p ldb=%lld C=%p ldc=%lld", M, N, K,
                   (double)alpha, (double)beta, lower, (const void*)A, lda, (const void*)B, ldb, (void*)C, ldc);
    return op_trace_finish(s, (int)dlafb::gemm_nt<T>(s, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, lower), "gemm_nt");
  }
  int transpose(int m, int n, const T* in, long long ldin, T* out, long long ldout, int conj, char tri) {
    DLAFB_TRACE_OP("transpose m=%d n=%d conj=%d tri=%c in=%p ldin=%lld out=%p ldout=%lld", m, n, conj, tri, (const void*)in,
                   ldin, (void*)out, ldout);
    return op_trace_finish(s, (int)dlafb::transpose<T>(s, m, n, in, ldin, out, ldout, conj, tri), "transpose");
  }
  int lacpy(char uplo, int m, int n, const T* a, long long lda, T* b, long long ldb) {
    DLAFB_TRACE_OP("lacpy %c m=%d n=%d a=%p lda=%lld b=%p ldb=%lld", uplo, m, n, (const void*)a, lda, (void*)b, ldb);
    return op_trace_finish(s, (int)dlafb::lacpy<T>(s, uplo, m, n, a, lda, b, ldb), "lacpy");
  }
  int tri_inverse(int upper, int unit, int n, int nb, const T* t, long long ldt, T* w) {
    TriInvParams p{n, nb, ldt, upper, unit};
    DLAFB_TRACE_OP("tri_inverse upper=%d unit=%d n=%d nb=%d t=%p ldt=%lld w=%p", upper, unit, n, nb, (const void*)t, ldt,
                   (void*)w);
    return op_trace_finish(s, (int)launch_tri_inverse<T>(s, p, t, w), "tri_inverse");
  }
};

constexpr int kG2sEvents = 5;

struct G2sPlan {
  char type, uplo;
  int n, nb;
  size_t elem;
  void* base = nullptr;  // one allocation, parts at 256-byte aligned offsets
  size_t off_winv = 0, off_akk = 0, off_tmp = 0, off_pa = 0, off_pb = 0, off_acopy = 0, off_lcopy = 0;
  size_t bytes = 0;
  // look-ahead (DLAFB_G2S_LOOKAHEAD=1, read when the plan is created): lane 0 = high-priority stream, lane 1 = bulk
  int lookahead = 0;
  cudaStream_t hp = nullptr, np = nullptr;
  cudaEvent_t ev[kG2sEvents] = {};
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  ~G2sPlan() {
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
    cudaFree(base);
  }
};

size_t align_up(size_t x) { return (x + 255) & ~size_t(255); }

template <class T>
int g2s_run_typed(G2sPlan& p, void* a, long long lda, const void* f, long long ldf, cudaStream_t s) {
  char* b = static_cast<char*>(p.base);
  GenToStdWork<T> w{};
  w.winv = reinterpret_cast<T*>(b + p.off_winv);
  w.akk = reinterpret_cast<T*>(b + p.off_akk);
  w.tmp = reinterpret_cast<T*>(b + p.off_tmp);
  w.ldk = g2s_even(p.nb);
  w.pa = reinterpret_cast<T*>(b + p.off_pa);
  w.pb = reinterpret_cast<T*>(b + p.off_pb);
  w.ldp = g2s_ldp(p.n);
  w.acopy = reinterpret_cast<T*>(b + p.off_acopy);
  w.lcopy = reinterpret_cast<T*>(b + p.off_lcopy);
  w.ldc = g2s_ldp(p.n);
  if (!p.lookahead) {
    G2sDeviceOps<T> ops{s, {s, s}, p.ev};
    return gen_to_std<T>(ops, p.uplo, p.n, p.nb, static_cast<T*>(a), lda, static_cast<const T*>(f), ldf, w);
  }
  w.lookahead = 1;
  w.panel_stride = g2s_panel_elems(p.n, p.nb);
  // fork from the caller's stream, join back into it
  cudaError_t ce;
  if ((ce = cudaEventRecord(p.ev_fork, s)) != cudaSuccess || (ce = cudaStreamWaitEvent(p.hp, p.ev_fork, 0)) != cudaSuccess)
    return (int)ce;
  G2sDeviceOps<T> ops{p.hp, {p.hp, p.np}, p.ev};
  const int e = gen_to_std<T>(ops, p.uplo, p.n, p.nb, static_cast<T*>(a), lda, static_cast<const T*>(f), ldf, w);
  // join even after an error so that the caller's stream stays ordered after whatever was enqueued
  cudaEventRecord(p.ev_join, p.np);
  cudaStreamWaitEvent(p.hp, p.ev_join, 0);
  cudaEventRecord(p.ev_join, p.hp);
  cudaStreamWaitEvent(s, p.ev_join, 0);
  return e;
}

struct G2sDistPlan {
  char type;
  G2sDistDesc d;
  dlafb_grid* grid;
  size_t elem;
  void* base = nullptr;
  size_t off[7] = {};
  size_t bytes = 0;
  ~G2sDistPlan() { cudaFree(base); }
};

template <class T>
int g2s_dist_run_typed(G2sDistPlan& p, void* a, long long lda, const void* f, long long ldf, cudaStream_t s) {
  char* b = static_cast<char*>(p.base);
  G2sDistWork<T> w{reinterpret_cast<T*>(b + p.off[0]), reinterpret_cast<T*>(b + p.off[1]),
                   reinterpret_cast<T*>(b + p.off[2]), reinterpret_cast<T*>(b + p.off[3]),
                   reinterpret_cast<T*>(b + p.off[4]), reinterpret_cast<T*>(b + p.off[5]),
                   reinterpret_cast<T*>(b + p.off[6])};
  DistDeviceOps<T> ops{s, p.grid};
  return gen_to_std_dist<T>(ops, p.d, static_cast<T*>(a), lda, static_cast<const T*>(f), ldf, w);
}

}  // namespace
}  // namespace dlafb

using namespace dlafb;

extern "C" {

int dlafb_gen_to_std_plan_create(dlafb_gen_to_std_plan_t* out, char type, char uplo, int64_t n, int64_t nb) {
  if (!out || (type != 's' && type != 'd' && type != 'c' && type != 'z') || (uplo != 'L' && uplo != 'U') || n < 0 ||
      nb < 1 || n > (1LL << 30) || nb > (1 << 14))
    return -1;
  G2sPlan* p = new (std::nothrow) G2sPlan;
  if (!p)
    return (int)cudaErrorMemoryAllocation;
  p->type = type;
  p->uplo = uplo;
  p->n = (int)n;
  p->nb = (int)nb;
  p->elem = type == 's' ? 4 : type == 'z' ? 16 : 8;
  const char* la = getenv("DLAFB_G2S_LOOKAHEAD");
  p->lookahead = (la && la[0] == '1') ? 1 : 0;
  if (n > 0 && p->lookahead) {
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    cudaError_t e = cudaStreamCreateWithPriority(&p->hp, cudaStreamNonBlocking, hi);
    if (e == cudaSuccess)
      e = cudaStreamCreateWithPriority(&p->np, cudaStreamNonBlocking, lo);
    for (int i = 0; i < kG2sEvents && e == cudaSuccess; ++i)
      e = cudaEventCreateWithFlags(&p->ev[i], cudaEventDisableTiming);
    if (e == cudaSuccess)
      e = cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming);
    if (e == cudaSuccess)
      e = cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming);
    if (e != cudaSuccess) {
      delete p;
      return (int)e;
    }
  }
  if (n > 0) {
    size_t off = 0;
    auto part = [&](size_t elems) {
      const size_t at = off;
      off = align_up(off + elems * p->elem);
      return at;
    };
    const size_t copies = p->lookahead ? 2 : 1;  // the panels are double-buffered with look-ahead
    p->off_winv = part(g2s_winv_elems(p->n, p->nb));
    p->off_akk = part(g2s_tile_elems(p->nb));
    p->off_tmp = part(g2s_tile_elems(p->nb));
    p->off_pa = part(copies * g2s_panel_elems(p->n, p->nb));
    p->off_pb = part(copies * g2s_panel_elems(p->n, p->nb));
    p->off_acopy = part(g2s_copy_elems(p->n, uplo));
    p->off_lcopy = part(g2s_copy_elems(p->n, uplo));
    p->bytes = off;
    const cudaError_t e = cudaMalloc(&p->base, p->bytes);
    if (e != cudaSuccess) {
      delete p;
      return (int)e;
    }
  }
  *out = reinterpret_cast<dlafb_gen_to_std_plan_t>(p);
  return 0;
}

int dlafb_gen_to_std_plan_run(dlafb_gen_to_std_plan_t plan, void* a, int64_t lda, const void* factor, int64_t ldf,
                              void* stream) {
  G2sPlan* p = reinterpret_cast<G2sPlan*>(plan);
  if (!p)
    return -1;
  if (p->n == 0)
    return 0;
  if (!a || !factor || lda < p->n || ldf < p->n)
    return -1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (p->type) {
    case 's': return g2s_run_typed<float>(*p, a, lda, factor, ldf, s);
    case 'd': return g2s_run_typed<double>(*p, a, lda, factor, ldf, s);
    case 'c': return g2s_run_typed<float2>(*p, a, lda, factor, ldf, s);
    default: return g2s_run_typed<double2>(*p, a, lda, factor, ldf, s);
  }
}

size_t dlafb_gen_to_std_plan_workspace_bytes(dlafb_gen_to_std_plan_t plan) {
  return plan ? reinterpret_cast<G2sPlan*>(plan)->bytes : 0;
}

int dlafb_gen_to_std_plan_destroy(dlafb_gen_to_std_plan_t plan) {
  delete reinterpret_cast<G2sPlan*>(plan);
  return 0;
}

int dlafb_gen_to_std_dist_plan_create(dlafb_gen_to_std_dist_plan_t* out, dlafb_grid_t grid, char type, char uplo,
                                      int64_t n, int64_t nb, int src_r, int src_c) {
  if (!out || !grid || (type != 's' && type != 'd' && type != 'c' && type != 'z') || (uplo != 'L' && uplo != 'U') ||
      n < 0 || nb < 1 || n > (1LL << 30) || nb > (1 << 14) || src_r < 0 || src_r >= grid->P || src_c < 0 ||
      src_c >= grid->Q)
    return -1;
  G2sDistPlan* p = new (std::nothrow) G2sDistPlan;
  if (!p)
    return (int)cudaErrorMemoryAllocation;
  p->type = type;
  p->grid = grid;
  p->d = G2sDistDesc{uplo, (int)n, (int)nb, grid->P, grid->Q, grid->pr, grid->pc, src_r, src_c};
  p->elem = type == 's' ? 4 : type == 'z' ? 16 : 8;
  if (n > 0) {
    const G2sDistSizes sz = g2s_dist_sizes(p->d);
    const size_t parts[7] = {sz.winv, sz.dk, sz.tmp, sz.xr, sz.ytiles, sz.acopy, sz.lcopy};
    size_t at = 0;
    for (int i = 0; i < 7; ++i) {
      p->off[i] = at;
      at = align_up(at + (parts[i] > 0 ? parts[i] : 1) * p->elem);
    }
    p->bytes = at;
    const cudaError_t e = cudaMalloc(&p->base, p->bytes);
    if (e != cudaSuccess) {
      delete p;
      return (int)e;
    }
  }
  *out = reinterpret_cast<dlafb_gen_to_std_dist_plan_t>(p);
  return 0;
}

int dlafb_gen_to_std_dist_plan_local_size(dlafb_gen_to_std_dist_plan_t plan, int64_t* rows, int64_t* cols) {
  G2sDistPlan* p = reinterpret_cast<G2sDistPlan*>(plan);
  if (!p)
    return -1;
  if (rows) *rows = Dist1D{p->d.n, p->d.nb, p->d.P, p->d.pr, p->d.src_r}.local_size();
  if (cols) *cols = Dist1D{p->d.n, p->d.nb, p->d.Q, p->d.pc, p->d.src_c}.local_size();
  return 0;
}

int dlafb_gen_to_std_dist_plan_run(dlafb_gen_to_std_dist_plan_t plan, void* a, int64_t lda, const void* factor,
                                   int64_t ldf, void* stream) {
  G2sDistPlan* p = reinterpret_cast<G2sDistPlan*>(plan);
  if (!p)
    return -1;
  if (p->d.n == 0)
    return 0;
  const int64_t rows = Dist1D{p->d.n, p->d.nb, p->d.P, p->d.pr, p->d.src_r}.local_size();
  const int64_t cols = Dist1D{p->d.n, p->d.nb, p->d.Q, p->d.pc, p->d.src_c}.local_size();
  if (rows > 0 && cols > 0 && (!a || !factor || lda < rows || ldf < rows))
    return -1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (p->type) {
    case 's': return g2s_dist_run_typed<float>(*p, a, lda, factor, ldf, s);
    case 'd': return g2s_dist_run_typed<double>(*p, a, lda, factor, ldf, s);
    case 'c': return g2s_dist_run_typed<float2>(*p, a, lda, factor, ldf, s);
    default: return g2s_dist_run_typed<double2>(*p, a, lda, factor, ldf, s);
  }
}

size_t dlafb_gen_to_std_dist_plan_workspace_bytes(dlafb_gen_to_std_dist_plan_t plan) {
  return plan ? reinterpret_cast<G2sDistPlan*>(plan)->bytes : 0;
}

int dlafb_gen_to_std_dist_plan_destroy(dlafb_gen_to_std_dist_plan_t plan) {
  delete reinterpret_cast<G2sDistPlan*>(plan);
  return 0;
}

}  // extern "C"
