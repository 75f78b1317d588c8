// Device plan and C ABI of the matrix redistribution (algorithm and index algebra: redistribute.h).
#include "redistribute.h"

#include <cuda_runtime.h>
#include <nccl.h>

#include <new>

#include "dlafb_engine.h"
#include "dlafb_matrix.h"
#include "grid.h"
#include "tile_ops.cuh"

namespace dlafb {
namespace {

struct RedistPlan {
  char type;
  size_t elem;
  RedistDesc d;
  dlafb_grid* grid;
  // the rectangle lists depend on the caller's leading dimensions: rebuilt (and uploaded) when those change
  long long ld_src = -1, ld_dst = -1;
  RedistLists lists;
  CopyRect* dev_rects = nullptr;  // pack | self | unpack
  size_t dev_rects_cap = 0;
  void* sendbuf = nullptr;
  void* recvbuf = nullptr;
  size_t send_cap = 0, recv_cap = 0;
  ~RedistPlan() {
    cudaFree(dev_rects);
    cudaFree(sendbuf);
    cudaFree(recvbuf);
  }
};

inline int nccl_rc(ncclResult_t r) { return r == ncclSuccess ? 0 : 1000 + (int)r; }

template <class T>
struct RedistDeviceOps {
  cudaStream_t s;
  RedistPlan* p;
  int copy_rects(int which, const T* src, T* dst) {
    const RedistLists& L = p->lists;
    const size_t off = which == 0 ? 0 : which == 1 ? L.pack.size() : L.pack.size() + L.self.size();
    const size_t n = which == 0 ? L.pack.size() : which == 1 ? L.self.size() : L.unpack.size();
    return (int)launch_copy_rects<T>(s, (int)n, p->dev_rects + off, src, dst);
  }
  int alltoallv(const T* sendbuf, T* recvbuf) {
    const RedistLists& L = p->lists;
    const int world = p->d.P * p->d.Q;
    if (world == 1)
      return 0;
    // bytes travel as ncclInt8 so that one code path serves the four element types
    int e = nccl_rc(ncclGroupStart());
    for (int r = 0; r < world && e == 0; ++r) {
      if (L.send_cnt[r] > 0)
        e = nccl_rc(ncclSend(reinterpret_cast<const char*>(sendbuf + L.send_off[r]), (size_t)L.send_cnt[r] * sizeof(T),
                             ncclInt8, r, p->grid->world, s));
      if (e == 0 && L.recv_cnt[r] > 0)
        e = nccl_rc(ncclRecv(reinterpret_cast<char*>(recvbuf + L.recv_off[r]), (size_t)L.recv_cnt[r] * sizeof(T),
                             ncclInt8, r, p->grid->world, s));
    }
    const int e2 = nccl_rc(ncclGroupEnd());
    return e != 0 ? e : e2;
  }
};

// (Re)builds the lists for these leading dimensions and makes sure the device buffers are large enough.
// Synchronises the stream before it frees or overwrites anything a previous run may still be using.
int prepare(RedistPlan& p, long long ld_src, long long ld_dst, cudaStream_t s) {
  if (p.ld_src == ld_src && p.ld_dst == ld_dst)
    return 0;
  cudaError_t e = cudaStreamSynchronize(s);
  if (e != cudaSuccess)
    return (int)e;
  p.lists = redist_build(p.d, ld_src, ld_dst);
  const RedistLists& L = p.lists;
  const size_t nrect = L.pack.size() + L.self.size() + L.unpack.size();
  if (nrect > p.dev_rects_cap) {
    cudaFree(p.dev_rects);
    p.dev_rects = nullptr;
    if ((e = cudaMalloc(&p.dev_rects, nrect * sizeof(CopyRect))) != cudaSuccess)
      return (int)e;
    p.dev_rects_cap = nrect;
  }
  CopyRect* at = p.dev_rects;
  for (const std::vector<CopyRect>* v : {&L.pack, &L.self, &L.unpack}) {
    if (!v->empty() &&
        (e = cudaMemcpy(at, v->data(), v->size() * sizeof(CopyRect), cudaMemcpyHostToDevice)) != cudaSuccess)
      return (int)e;
    at += v->size();
  }
  const size_t sb = (size_t)L.send_total * p.elem, rb = (size_t)L.recv_total * p.elem;
  if (sb > p.send_cap) {
    cudaFree(p.sendbuf);
    p.sendbuf = nullptr;
    if ((e = cudaMalloc(&p.sendbuf, sb)) != cudaSuccess)
      return (int)e;
    p.send_cap = sb;
  }
  if (rb > p.recv_cap) {
    cudaFree(p.recvbuf);
    p.recvbuf = nullptr;
    if ((e = cudaMalloc(&p.recvbuf, rb)) != cudaSuccess)
      return (int)e;
    p.recv_cap = rb;
  }
  // The lists were uploaded with cudaMemcpy from pageable memory: it is ordered on the legacy stream only and may return
  // before its DMA has finished, while the copy kernel runs on the caller's (possibly non-blocking) stream. Without
  // this the kernel could read a list that had not arrived: rectangles of 0 x 0 elements, nothing copied
  // (order-dependent failure of dlaf_p?potrf with DLAFB_POTRF_REDISTRIBUTE=1, profiles/r02/flaky_upload.md).
  if ((e = cudaDeviceSynchronize()) != cudaSuccess)
    return (int)e;
  p.ld_src = ld_src;
  p.ld_dst = ld_dst;
  return 0;
}

template <class T>
int run_typed(RedistPlan& p, const void* src, void* dst, cudaStream_t s) {
  RedistDeviceOps<T> ops{s, &p};
  return redistribute<T>(ops, p.lists, static_cast<const T*>(src), static_cast<T*>(dst), static_cast<T*>(p.sendbuf),
                         static_cast<T*>(p.recvbuf));
}

// max |a_ij| over this rank's part of the (global) uplo region: for 'L' the tiles strictly below the global diagonal in
// full and the diagonal tiles' lower triangles - per local column tile one launch for the rows below and one for a
// local diagonal tile (auxiliary/norm/mc.h:35-95 walks the same tiles one task per tile).
template <class T>
int local_max_abs(cudaStream_t s, char uplo, int64_t m, int64_t n, int64_t mb, int64_t nb, const dlafb_grid* g, int src_r,
                  int src_c, const void* a, int64_t lda, double* dev_out) {
  const Dist1D R{m, mb, g->P, g->pr, src_r}, C{n, nb, g->Q, g->pc, src_c};
  const int64_t lrows = R.local_size(), lcols = C.local_size();
  const T* A = static_cast<const T*>(a);
  if (lrows <= 0 || lcols <= 0)
    return 0;
  if (uplo == 'G')
    return (int)max_abs<T>(s, 'G', (int)lrows, (int)lcols, A, lda, dev_out);
  for (int64_t jl = 0; jl < C.nloc_tiles(); ++jl) {
    const int64_t j = C.global_of(jl);
    const int wj = (int)C.tile_size(j);
    int64_t below = R.next_local(j) * mb;
    const T* cj = A + jl * nb * lda;
    cudaError_t e;
    if (R.owner(j) == g->pr && below < lrows) {
      const int hj = (int)R.tile_size(j);
      if ((e = max_abs<T>(s, 'L', hj, wj, cj + below, lda, dev_out)) != cudaSuccess)
        return (int)e;
      below += mb;
    }
    if (below < lrows && (e = max_abs<T>(s, 'G', (int)(lrows - below), wj, cj + below, lda, dev_out)) != cudaSuccess)
      return (int)e;
  }
  return 0;
}

}  // namespace
}  // namespace dlafb

using namespace dlafb;

extern "C" {

int dlafb_redistribute_plan_create(dlafb_redistribute_plan_t* out, dlafb_grid_t grid, char type, int64_t m, int64_t n,
                                   int64_t mb_src, int64_t nb_src, int src_r_src, int src_c_src, int64_t mb_dst,
                                   int64_t nb_dst, int src_r_dst, int src_c_dst) {
  if (!out || !grid || (type != 's' && type != 'd' && type != 'c' && type != 'z') || m > (1LL << 30) || n > (1LL << 30) ||
      mb_src > (1 << 20) || nb_src > (1 << 20) || mb_dst > (1 << 20) || nb_dst > (1 << 20))
    return -1;
  RedistDesc d{m, n, (int)mb_src, (int)nb_src, src_r_src, src_c_src, (int)mb_dst, (int)nb_dst, src_r_dst, src_c_dst,
               grid->P, grid->Q, grid->pr, grid->pc, grid->ordering};
  if (!redist_valid(d) || src_r_src < 0 || src_r_src >= grid->P || src_r_dst < 0 || src_r_dst >= grid->P ||
      src_c_src < 0 || src_c_src >= grid->Q || src_c_dst < 0 || src_c_dst >= grid->Q)
    return -1;
  RedistPlan* p = new (std::nothrow) RedistPlan;
  if (!p)
    return (int)cudaErrorMemoryAllocation;
  p->type = type;
  p->elem = type == 's' ? 4 : type == 'z' ? 16 : 8;
  p->d = d;
  p->grid = grid;
  *out = reinterpret_cast<dlafb_redistribute_plan_t>(p);
  return 0;
}

int dlafb_redistribute_plan_local_sizes(dlafb_redistribute_plan_t plan, int64_t* src_rows, int64_t* src_cols,
                                        int64_t* dst_rows, int64_t* dst_cols) {
  RedistPlan* p = reinterpret_cast<RedistPlan*>(plan);
  if (!p)
    return -1;
  const RedistDesc& d = p->d;
  if (src_rows) *src_rows = Dist1D{d.m, d.mb_s, d.P, d.pr, d.src_r_s}.local_size();
  if (src_cols) *src_cols = Dist1D{d.n, d.nb_s, d.Q, d.pc, d.src_c_s}.local_size();
  if (dst_rows) *dst_rows = Dist1D{d.m, d.mb_d, d.P, d.pr, d.src_r_d}.local_size();
  if (dst_cols) *dst_cols = Dist1D{d.n, d.nb_d, d.Q, d.pc, d.src_c_d}.local_size();
  return 0;
}

int dlafb_redistribute_plan_run(dlafb_redistribute_plan_t plan, const void* src, int64_t ld_src, void* dst,
                                int64_t ld_dst, void* stream) {
  RedistPlan* p = reinterpret_cast<RedistPlan*>(plan);
  if (!p)
    return -1;
  int64_t sr, sc, dr, dc;
  dlafb_redistribute_plan_local_sizes(plan, &sr, &sc, &dr, &dc);
  if ((sr > 0 && sc > 0 && (!src || ld_src < sr)) || (dr > 0 && dc > 0 && (!dst || ld_dst < dr)))
    return -1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int e = prepare(*p, ld_src > 0 ? ld_src : 1, ld_dst > 0 ? ld_dst : 1, s);
  if (e != 0)
    return e;
  switch (p->type) {
    case 's': return run_typed<float>(*p, src, dst, s);
    case 'd': return run_typed<double>(*p, src, dst, s);
    case 'c': return run_typed<float2>(*p, src, dst, s);
    default: return run_typed<double2>(*p, src, dst, s);
  }
}

int dlafb_redistribute_plan_destroy(dlafb_redistribute_plan_t plan) {
  delete reinterpret_cast<RedistPlan*>(plan);
  return 0;
}

int dlafb_max_norm(dlafb_grid_t grid, char type, char uplo, int64_t m, int64_t n, int64_t mb, int64_t nb, int src_r,
                   int src_c, const void* a, int64_t lda, void* stream, double* result) {
  if (!grid || !result || (type != 's' && type != 'd' && type != 'c' && type != 'z') || (uplo != 'L' && uplo != 'G') ||
      m < 0 || n < 0 || mb < 1 || nb < 1 || src_r < 0 || src_r >= grid->P || src_c < 0 || src_c >= grid->Q ||
      (uplo == 'L' && (m != n || mb != nb)))
    return -1;
  *result = 0.0;  // LAPACK: the norm of an empty matrix is 0 (auxiliary/norm.h:41-43)
  if (m == 0 || n == 0)
    return 0;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  double* dev = nullptr;
  cudaError_t ce = cudaMalloc(&dev, sizeof(double));
  if (ce != cudaSuccess)
    return (int)ce;
  cudaMemsetAsync(dev, 0, sizeof(double), s);
  int e;
  switch (type) {
    case 's': e = local_max_abs<float>(s, uplo, m, n, mb, nb, grid, src_r, src_c, a, lda, dev); break;
    case 'd': e = local_max_abs<double>(s, uplo, m, n, mb, nb, grid, src_r, src_c, a, lda, dev); break;
    case 'c': e = local_max_abs<float2>(s, uplo, m, n, mb, nb, grid, src_r, src_c, a, lda, dev); break;
    default: e = local_max_abs<double2>(s, uplo, m, n, mb, nb, grid, src_r, src_c, a, lda, dev); break;
  }
  double local = 0.0;
  if (e == 0)
    e = (int)cudaMemcpyAsync(&local, dev, sizeof(double), cudaMemcpyDeviceToHost, s);
  if (e == 0)
    e = (int)cudaStreamSynchronize(s);
  cudaFree(dev);
  if (e != 0)
    return e;
  if ((e = dlafb_grid_allreduce_max(grid, &local)) != 0)
    return e;
  *result = local;
  return 0;
}

}  // extern "C"
