// Device executor and C ABI of the distributed triangular solver (algorithm: triangular_solver_dist.h).
#include "triangular_solver_dist.h"

#include <cuda_runtime.h>
#include <nccl.h>

#include <new>

#include "dist_device_ops.h"
#include "dlafb_solver.h"

namespace dlafb {
namespace {

struct DistPlan {
  char type;
  TriDistDesc d;
  dlafb_grid* grid;
  size_t elem;
  void* base = nullptr;
  size_t off[6] = {};
  size_t bytes = 0;
  ~DistPlan() { cudaFree(base); }
};

size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

template <class T>
int dist_run_typed(DistPlan& p, const void* alpha, const void* a, long long lda, void* b, long long ldb, cudaStream_t s) {
  char* base = static_cast<char*>(p.base);
  TriDistWork<T> w{reinterpret_cast<T*>(base + p.off[0]), reinterpret_cast<T*>(base + p.off[1]),
                   reinterpret_cast<T*>(base + p.off[2]), reinterpret_cast<T*>(base + p.off[3]),
                   reinterpret_cast<T*>(base + p.off[4]), reinterpret_cast<T*>(base + p.off[5])};
  DistDeviceOps<T> ops{s, p.grid};
  return triangular_solve_dist<T>(ops, p.d, *static_cast<const T*>(alpha), static_cast<const T*>(a), lda,
                                  static_cast<T*>(b), ldb, w);
}

bool in_set(char c, const char* set) {
  for (; *set; ++set)
    if (*set == c)
      return true;
  return false;
}

}  // namespace
}  // namespace dlafb

using namespace dlafb;

extern "C" {

int dlafb_triangular_solver_dist_plan_create(dlafb_trsolve_dist_plan_t* out, dlafb_grid_t grid, char type, char side,
                                             char uplo, char op, char diag, int64_t m, int64_t n, int64_t mb, int64_t nb,
                                             int src_r, int src_c) {
  if (!out || !grid || !in_set(type, "sdcz") || !in_set(side, "LR") || !in_set(uplo, "LU") || !in_set(op, "NTC") ||
      !in_set(diag, "NU") || m < 0 || n < 0 || mb < 1 || nb < 1 || m > (1LL << 30) || n > (1LL << 30) ||
      mb > (1 << 14) || nb > (1 << 14) || src_r < 0 || src_r >= grid->P || src_c < 0 || src_c >= grid->Q)
    return -1;
  DistPlan* p = new (std::nothrow) DistPlan;
  if (!p)
    return (int)cudaErrorMemoryAllocation;
  p->type = type;
  p->grid = grid;
  p->d = TriDistDesc{side, uplo, op, diag, (int)m, (int)n, (int)mb, (int)nb, grid->P, grid->Q, grid->pr, grid->pc,
                     src_r, src_c};
  p->elem = type == 's' ? 4 : type == 'z' ? 16 : 8;
  if (m > 0 && n > 0) {
    const TriDistSizes sz = tri_dist_sizes(p->d);
    const size_t parts[6] = {sz.winv, sz.pan, sz.pan2, sz.xb, sz.wkh, sz.bt};
    size_t at = 0;
    for (int i = 0; i < 6; ++i) {
      p->off[i] = at;
      at = align256(at + (parts[i] > 0 ? parts[i] : 1) * p->elem);
    }
    p->bytes = at;
    const cudaError_t e = cudaMalloc(&p->base, p->bytes);
    if (e != cudaSuccess) {
      delete p;
      return (int)e;
    }
  }
  *out = reinterpret_cast<dlafb_trsolve_dist_plan_t>(p);
  return 0;
}

int dlafb_triangular_solver_dist_plan_local_sizes(dlafb_trsolve_dist_plan_t plan, int64_t* a_rows, int64_t* a_cols,
                                                  int64_t* b_rows, int64_t* b_cols) {
  DistPlan* p = reinterpret_cast<DistPlan*>(plan);
  if (!p)
    return -1;
  const TriDistGeom g = tri_dist_geom(p->d);
  if (a_rows) *a_rows = g.larows;
  if (a_cols) *a_cols = g.lacols;
  if (b_rows) *b_rows = g.lrows;
  if (b_cols) *b_cols = g.lcols;
  return 0;
}

int dlafb_triangular_solver_dist_plan_run(dlafb_trsolve_dist_plan_t plan, const void* alpha, const void* a, int64_t lda,
                                          void* b, int64_t ldb, void* stream) {
  DistPlan* p = reinterpret_cast<DistPlan*>(plan);
  if (!p || !alpha)
    return -1;
  if (p->d.m == 0 || p->d.n == 0)
    return 0;
  const TriDistGeom g = tri_dist_geom(p->d);
  if ((g.larows > 0 && g.lacols > 0 && (!a || lda < g.larows)) || (g.lrows > 0 && g.lcols > 0 && (!b || ldb < g.lrows)))
    return -1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (p->type) {
    case 's': return dist_run_typed<float>(*p, alpha, a, lda, b, ldb, s);
    case 'd': return dist_run_typed<double>(*p, alpha, a, lda, b, ldb, s);
    case 'c': return dist_run_typed<float2>(*p, alpha, a, lda, b, ldb, s);
    default: return dist_run_typed<double2>(*p, alpha, a, lda, b, ldb, s);
  }
}

size_t dlafb_triangular_solver_dist_plan_workspace_bytes(dlafb_trsolve_dist_plan_t plan) {
  return plan ? reinterpret_cast<DistPlan*>(plan)->bytes : 0;
}

int dlafb_triangular_solver_dist_plan_destroy(dlafb_trsolve_dist_plan_t plan) {
  delete reinterpret_cast<DistPlan*>(plan);
  return 0;
}

}  // extern "C"
