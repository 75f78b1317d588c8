// C API of the Cholesky path: dlaf_initialize / dlaf_finalize, the grid registry, descriptor conversion and the
// dlaf_cholesky_factorization_{s,d,c,z} / dlaf_p{s,d,c,z}potrf entry points.
// Reference: src/c_api/init.cpp:21-64, src/c_api/grid.cpp:26-92, src/c_api/utils.cpp:25-69,
// src/c_api/factorization/cholesky.h:32-77, src/c_api/factorization/cholesky.cpp:16-56.
// As in the reference the caller's host block is mirrored to the device, factorized there and copied back
// (matrix::MatrixMirror, include/dlaf/matrix/matrix_mirror.h:110-177); device pointers are used in place.
#include <cuda_runtime.h>
#include <nccl.h>

#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <tuple>
#include <unordered_map>

#include "../../include/dlafb_engine.h"
#include "../../include/dlafb_matrix.h"
// MPI_Comm: with -DDLAFB_WITH_MPI the application's real <mpi.h> (the reference's contract: dlaf_create_grid takes a
// real communicator, include/dlaf_c/grid.h:31, src/communication/communicator_grid.cpp:36-38); otherwise the
// single-node shim, whose communicator is {rank, size, NCCL world} set up by dlafb_mpi_init.
#ifdef DLAFB_WITH_MPI
#include <mpi.h>
#else
#include "../../include/dlafb_mpi_shim/mpi.h"
#endif
#include "grid.h"

#include <dlaf_c/factorization/cholesky.h>
#include <dlaf_c/init.h>
// dlaf_c/grid.h includes <mpi.h>; the shim above already defined MPI_Comm
DLAF_EXTERN_C int dlaf_create_grid(MPI_Comm comm, int nprow, int npcol, char order) DLAF_NOEXCEPT;
DLAF_EXTERN_C void dlaf_free_grid(int context) DLAF_NOEXCEPT;
DLAF_EXTERN_C char grid_ordering(MPI_Comm comm, int nprow, int npcol, int myprow, int mypcol) DLAF_NOEXCEPT;
DLAF_EXTERN_C void dlaf_create_grid_from_blacs_info(int blacs_ctxt, MPI_Comm comm, int nprow, int npcol,
                                                    char order) DLAF_NOEXCEPT;

#ifndef DLAFB_WITH_MPI
struct dlafb_mpi_comm {
  int rank = 0, size = 1;
  ncclComm_t world = nullptr;
};
#endif

namespace {

bool g_initialized = false;
#ifndef DLAFB_WITH_MPI
dlafb_mpi_comm* g_world = nullptr;
#endif

[[noreturn]] void die(const char* what, int code);

// ---- what the grid code needs from a communicator: rank, size, an NCCL communicator over the same ranks, and a
//      logical-and over the ranks (grid_ordering) ----
#ifdef DLAFB_WITH_MPI
int comm_rank(MPI_Comm c) {
  int r = 0;
  if (MPI_Comm_rank(c, &r) != MPI_SUCCESS)
    die("MPI_Comm_rank failed", -1);
  return r;
}
int comm_size(MPI_Comm c) {
  int n = 1;
  if (MPI_Comm_size(c, &n) != MPI_SUCCESS)
    die("MPI_Comm_size failed", -1);
  return n;
}
// One NCCL communicator per grid over the ranks of `c`: rank 0 creates the id, MPI_Bcast hands it out. The grid owns it.
ncclComm_t comm_nccl(MPI_Comm c, bool* owned) {
  ncclUniqueId id;
  std::memset(&id, 0, sizeof(id));
  if (comm_rank(c) == 0 && ncclGetUniqueId(&id) != ncclSuccess)
    die("ncclGetUniqueId failed", -1);
  if (MPI_Bcast(&id, (int)sizeof(id), MPI_BYTE, 0, c) != MPI_SUCCESS)
    die("MPI_Bcast of the NCCL id failed", -1);
  ncclComm_t w = nullptr;
  const ncclResult_t r = ncclCommInitRank(&w, comm_size(c), id, comm_rank(c));
  if (r != ncclSuccess)
    die("ncclCommInitRank failed", 1000 + (int)r);
  *owned = true;
  return w;
}
void comm_land2(MPI_Comm c, int flags[2]) {
  int out[2] = {0, 0};
  if (MPI_Allreduce(flags, out, 2, MPI_INT, MPI_MIN, c) != MPI_SUCCESS)
    die("MPI_Allreduce failed", -1);
  flags[0] = out[0];
  flags[1] = out[1];
}
#else
int comm_rank(MPI_Comm c) { return c ? c->rank : 0; }
int comm_size(MPI_Comm c) { return c ? c->size : 1; }
ncclComm_t comm_nccl(MPI_Comm c, bool* owned) {
  *owned = false;  // the shim's world communicator outlives the grids
  return c ? c->world : nullptr;
}
void comm_land2(MPI_Comm c, int flags[2]);  // below: all-reduce on the shim's NCCL communicator
#endif
std::unordered_map<int, dlafb_grid_t> g_grids;

struct PlanKey {
  int ctx;
  char type, uplo;
  int n, nb, isrc, jsrc;
  bool operator<(const PlanKey& o) const {
    return std::tie(ctx, type, uplo, n, nb, isrc, jsrc) < std::tie(o.ctx, o.type, o.uplo, o.n, o.nb, o.isrc, o.jsrc);
  }
};
struct PlanEntry {
  dlafb_chol_plan_t plan = nullptr;
  void* dev = nullptr;  // device mirror of the local block
  size_t dev_bytes = 0;
  int64_t ldd = 0;
  cudaStream_t stream = nullptr;
  unsigned long long last_use = 0;
};
std::map<PlanKey, PlanEntry> g_plans;
unsigned long long g_plan_clock = 0;

// A plan owns streams, panel workspaces and (host-pointer callers) an n^2-byte device mirror: a caller sweeping over
// sizes must not accumulate them. At most DLAFB_PLAN_CACHE (default 4) plans stay alive, least recently used first
// out. Every rank of a grid issues the same sequence of factorizations, so all ranks evict the same plans.
size_t plan_cache_limit() {
  static const size_t v = [] {
    const char* e = getenv("DLAFB_PLAN_CACHE");
    const long long x = e ? atoll(e) : 4;
    return (size_t)(x < 1 ? 1 : x);
  }();
  return v;
}

[[noreturn]] void die(const char* what, int code) {
  std::fprintf(stderr, "[dlaf_b200] fatal: %s (error %d)\n", what, code);
  std::abort();  // the reference's DLAF_ASSERT / *_CHECK_ERROR convention: print and terminate
}

void destroy_entry(PlanEntry& pe) {
  dlafb_cholesky_plan_destroy(pe.plan);
  cudaFree(pe.dev);
  if (pe.stream)
    cudaStreamDestroy(pe.stream);
}

void evict_lru_plans(const PlanKey& keep) {
  while (g_plans.size() > plan_cache_limit()) {
    auto victim = g_plans.end();
    for (auto it = g_plans.begin(); it != g_plans.end(); ++it) {
      const bool is_kept = !(it->first < keep) && !(keep < it->first);
      if (!is_kept && (victim == g_plans.end() || it->second.last_use < victim->second.last_use))
        victim = it;
    }
    if (victim == g_plans.end())
      return;
    destroy_entry(victim->second);
    g_plans.erase(victim);
  }
}

void drop_redist_of(int ctx);

void drop_plans_of(int ctx) {
  drop_redist_of(ctx);
  for (auto it = g_plans.begin(); it != g_plans.end();) {
    if (ctx == INT_MIN || it->first.ctx == ctx) {
      destroy_entry(it->second);
      it = g_plans.erase(it);
    }
    else
      ++it;
  }
}

size_t elem_size(char t) { return t == 's' ? 4 : (t == 'z' ? 16 : 8); }

// Callers that come from ScaLAPACK bring block sizes of 32 - 128; the factorization wants 512 (the contraction's K and
// the number of dependent steps). Such a matrix can be redistributed on the device to the largest block size <= 512
// that is a multiple of both the caller's block size and 16, factorized there and redistributed back
// (matrix::copy(src, dst, grid), include/dlaf/matrix/copy.h:118-231 of the reference, is the same tool; the reference's
// C API does not use it for this call). Measured at n = 16384 on one GPU (profiles/r02/small_blocks.txt):
//   device pointers  nb = 64: 125.8 ms plain, 49.4 ms redistributed;  nb = 128: 49.8 / 49.5 ms
//   host pointers    nb = 64: 136.4 / 125.6 ms;  nb = 128: 63.5 / 125.7 ms  (the whole block crosses PCIe both ways,
//                    unoverlapped, where the plain path moves the triangle only and overlaps it with the factorization)
// so: DLAFB_POTRF_REDISTRIBUTE unset -> device pointers with nb < 128 only; = 1 -> every call with nb < 256; = 0 -> never.
int redistribution_block(int n, int nb_in, bool on_device) {
  const char* e = getenv("DLAFB_POTRF_REDISTRIBUTE");
  const bool forced = e && e[0] == '1';
  const bool never = e && e[0] == '0';
  if (never || n < 2048 || nb_in >= 256)
    return 0;
  if (!forced && !(on_device && nb_in < 128))
    return 0;
  int l = nb_in;
  while (l % 16 != 0)
    l += nb_in;  // lcm(nb_in, 16)
  return l <= 512 ? (512 / l) * l : 0;
}

#define CAPI_CUDA(call, what)            \
  do {                                   \
    const cudaError_t ce_ = (call);      \
    if (ce_ != cudaSuccess)              \
      die(what, (int)ce_);               \
  } while (0)

// What the redistributed path keeps between calls (round 2: creating the three plans, two device blocks and a stream on
// every call cost ~85 ms at n = 16384, more than the factorization): at most two entries, least recently used first out,
// dropped with their grid.
struct RedistEntry {
  dlafb_redistribute_plan_t fwd = nullptr, bwd = nullptr;
  dlafb_chol_plan_t plan = nullptr;
  cudaStream_t stream = nullptr;
  void* in_dev = nullptr;   // device image of the caller's block (host-pointer callers)
  void* work = nullptr;     // the block at the factorization's block size
  size_t in_bytes = 0, work_bytes = 0;
  unsigned long long last_use = 0;
};
std::map<PlanKey, RedistEntry> g_redist;

void destroy_redist(RedistEntry& r) {
  dlafb_redistribute_plan_destroy(r.fwd);
  dlafb_redistribute_plan_destroy(r.bwd);
  dlafb_cholesky_plan_destroy(r.plan);
  cudaFree(r.in_dev);
  cudaFree(r.work);
  if (r.stream)
    cudaStreamDestroy(r.stream);
}

void drop_redist_of(int ctx) {
  for (auto it = g_redist.begin(); it != g_redist.end();) {
    if (ctx == INT_MIN || it->first.ctx == ctx) {
      destroy_redist(it->second);
      it = g_redist.erase(it);
    }
    else
      ++it;
  }
}

int cholesky_redistributed(int ctx, dlafb_grid_t grid, char type, char uplo, void* a, int n, int nb_in, int isrc,
                           int jsrc, int64_t ld, int nb_to) {
  const size_t es = elem_size(type);
  const int64_t rows_in = dlafb_dist_local_size(n, nb_in, grid->P, grid->pr, isrc);
  const int64_t cols_in = dlafb_dist_local_size(n, nb_in, grid->Q, grid->pc, jsrc);
  const int64_t rows_w = dlafb_dist_local_size(n, nb_to, grid->P, grid->pr, 0);
  const int64_t cols_w = dlafb_dist_local_size(n, nb_to, grid->Q, grid->pc, 0);
  if (rows_in > 0 && ld < rows_in)
    die("leading dimension smaller than the local number of rows", (int)ld);
  const PlanKey key{ctx, type, uplo, n, nb_in, isrc, jsrc};
  RedistEntry& r = g_redist[key];
  r.last_use = ++g_plan_clock;
  int e;
  if (!r.plan) {
    while (g_redist.size() > 2) {  // every rank makes the same calls: the same entries go
      auto victim = g_redist.end();
      for (auto it = g_redist.begin(); it != g_redist.end(); ++it)
        if (&it->second != &r && (victim == g_redist.end() || it->second.last_use < victim->second.last_use))
          victim = it;
      if (victim == g_redist.end())
        break;
      destroy_redist(victim->second);
      g_redist.erase(victim);
    }
    CAPI_CUDA(cudaStreamCreateWithFlags(&r.stream, cudaStreamNonBlocking), "stream creation failed");
    if ((e = dlafb_redistribute_plan_create(&r.fwd, grid, type, n, n, nb_in, nb_in, isrc, jsrc, nb_to, nb_to, 0, 0)) != 0 ||
        (e = dlafb_redistribute_plan_create(&r.bwd, grid, type, n, n, nb_to, nb_to, 0, 0, nb_in, nb_in, isrc, jsrc)) != 0 ||
        (e = dlafb_cholesky_plan_create(&r.plan, grid, type, uplo, n, nb_to, 0, 0)) != 0)
      die("plan creation failed", e);
  }
  cudaStream_t s = r.stream;
  cudaPointerAttributes attr{};
  const bool on_device = (a != nullptr) && cudaPointerGetAttributes(&attr, a) == cudaSuccess &&
                         (attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged);
  cudaGetLastError();
  void* in_dev = a;
  int64_t ld_in = ld;
  const bool have_block = rows_in > 0 && cols_in > 0;
  if (!on_device) {
    ld_in = ((rows_in > 0 ? rows_in : 1) + 63) / 64 * 64;
    const size_t need = (size_t)ld_in * (size_t)(cols_in > 0 ? cols_in : 1) * es;
    if (r.in_bytes < need) {
      cudaFree(r.in_dev);
      CAPI_CUDA(cudaMalloc(&r.in_dev, need), "device allocation failed");
      r.in_bytes = need;
    }
    in_dev = r.in_dev;
    if (have_block)
      CAPI_CUDA(cudaMemcpy2DAsync(in_dev, ld_in * es, a, ld * es, rows_in * es, cols_in, cudaMemcpyHostToDevice, s),
                "upload failed");
  }
  const int64_t ldw = ((rows_w > 0 ? rows_w : 1) + 63) / 64 * 64;
  const size_t wneed = (size_t)ldw * (size_t)(cols_w > 0 ? cols_w : 1) * es;
  if (r.work_bytes < wneed) {
    cudaFree(r.work);
    CAPI_CUDA(cudaMalloc(&r.work, wneed), "device allocation failed");
    r.work_bytes = wneed;
  }
  void* work = r.work;
  if ((e = dlafb_redistribute_plan_run(r.fwd, in_dev, ld_in, work, ldw, s)) != 0)
    die("redistribution failed", e);
  if ((e = dlafb_cholesky_plan_run(r.plan, work, ldw, s)) != 0)
    die("factorization failed to launch", e);
  int info = 0;
  if ((e = dlafb_cholesky_plan_info(r.plan, &info)) != 0)
    die("factorization failed", e);
  if ((e = dlafb_redistribute_plan_run(r.bwd, work, ldw, in_dev, ld_in, s)) != 0)
    die("redistribution failed", e);
  if (!on_device && have_block)
    CAPI_CUDA(cudaMemcpy2DAsync(a, ld * es, in_dev, ld_in * es, rows_in * es, cols_in, cudaMemcpyDeviceToHost, s),
              "download failed");
  CAPI_CUDA(cudaStreamSynchronize(s), "stream synchronisation failed");
  return info;
}

int cholesky(int ctx, char type, char uplo, void* a, const DLAF_descriptor& d) {
  if (!g_initialized)
    die("dlaf_cholesky_factorization called before dlaf_initialize", -1);
  if (d.m != d.n || d.mb != d.nb)
    die("cholesky_factorization needs a square matrix with square blocks", -1);  // factorization/cholesky.h:69-70
  // Sub-matrix sub(A) = A(i : i+n, j : j+n) of the distributed matrix the descriptor describes. The reference stops at
  // i = j = 0 (src/c_api/factorization/cholesky.h:37-38, src/c_api/utils.cpp:25-33); ScaLAPACK's p?potrf accepts
  // tile-aligned offsets (IROFF = ICOFF = 0) and so does this build: a tile-aligned sub-matrix of a 2D block-cyclic
  // matrix is itself 2D block-cyclic, with the source rank advanced by the tiles skipped, and on every rank its local
  // block starts after the local tiles that precede the offset.
  if (d.i < 0 || d.j < 0 || d.i % d.mb != 0 || d.j % d.nb != 0)
    die("sub-matrix offsets must be non-negative multiples of the block size", d.i % d.mb != 0 ? d.i : d.j);
  auto git = g_grids.find(ctx);
  if (git == g_grids.end())
    die("no grid registered for this context", ctx);  // utils.cpp:55-69 terminates as well
  dlafb_grid_t grid = git->second;
  if (d.isrc < 0 || d.isrc >= grid->P || d.jsrc < 0 || d.jsrc >= grid->Q)
    die("source rank outside the grid", d.isrc);
  const int i0 = d.i / d.mb, j0 = d.j / d.nb;
  const int sub_isrc = (d.isrc + i0) % grid->P, sub_jsrc = (d.jsrc + j0) % grid->Q;
  if (i0 != 0 || j0 != 0) {
    const int64_t skip_r = dlafb_dist_next_local_tile_from_global(i0, grid->P, grid->pr, d.isrc) * d.mb;
    const int64_t skip_c = dlafb_dist_next_local_tile_from_global(j0, grid->Q, grid->pc, d.jsrc) * d.nb;
    a = static_cast<char*>(a) + ((size_t)skip_r + (size_t)skip_c * (size_t)d.ld) * elem_size(type);
  }
  const PlanKey key{ctx, type, uplo == 'u' ? 'U' : (uplo == 'l' ? 'L' : uplo), d.n, d.nb, sub_isrc, sub_jsrc};
  if (key.uplo != 'L' && key.uplo != 'U')
    die("uplo must be 'L' or 'U'", -1);
  bool dev_ptr = false;
  {
    cudaPointerAttributes pa{};
    dev_ptr = (a != nullptr) && cudaPointerGetAttributes(&pa, a) == cudaSuccess &&
              (pa.type == cudaMemoryTypeDevice || pa.type == cudaMemoryTypeManaged);
    cudaGetLastError();
  }
  if (const int nb_to = redistribution_block(d.n, d.nb, dev_ptr))
    return cholesky_redistributed(ctx, grid, type, key.uplo, a, d.n, d.nb, sub_isrc, sub_jsrc, d.ld, nb_to);
  PlanEntry& pe = g_plans[key];
  pe.last_use = ++g_plan_clock;
  int e;
  if (!pe.plan) {
    evict_lru_plans(key);
    if ((e = dlafb_cholesky_plan_create(&pe.plan, grid, type, key.uplo, d.n, d.nb, sub_isrc, sub_jsrc)) != 0)
      die("plan creation failed", e);
    if (cudaStreamCreateWithFlags(&pe.stream, cudaStreamNonBlocking) != cudaSuccess)
      die("stream creation failed", -1);
  }
  int64_t rows = 0, cols = 0;
  dlafb_cholesky_plan_local_size(pe.plan, &rows, &cols);
  if (rows > 0 && d.ld < rows)
    die("leading dimension smaller than the local number of rows", d.ld);
  const size_t es = elem_size(type);
  cudaPointerAttributes attr{};
  const bool on_device = (a != nullptr) && cudaPointerGetAttributes(&attr, a) == cudaSuccess &&
                         (attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged);
  cudaGetLastError();
  int info = 0;
  if (on_device) {
    if ((e = dlafb_cholesky_plan_run(pe.plan, a, d.ld, pe.stream)) != 0)
      die("factorization failed to launch", e);
  }
  else {
    const int64_t ldd = ((rows > 0 ? rows : 1) + 63) / 64 * 64;  // default device layout: matrix.h:87-89
    const size_t need = (size_t)ldd * (size_t)(cols > 0 ? cols : 1) * es + 1024;
    if (pe.dev_bytes < need) {
      cudaFree(pe.dev);
      if (cudaMalloc(&pe.dev, need) != cudaSuccess)
        die("device allocation of the mirror failed", -1);
      pe.dev_bytes = need;
    }
    pe.ldd = ldd;
    // only the referenced triangle travels, and (uplo = L) finished column panels overlap the remaining steps
    if ((e = dlafb_cholesky_plan_run_host(pe.plan, a, d.ld, pe.dev, ldd, pe.stream)) != 0)
      die("factorization failed to launch", e);
  }
  if ((e = dlafb_cholesky_plan_info(pe.plan, &info)) != 0)
    die("factorization failed", e);
  if (cudaStreamSynchronize(pe.stream) != cudaSuccess)
    die("stream synchronisation failed", -1);
  return info;
}

void pxpotrf(char type, char uplo, int n, void* a, int ia, int ja, const int desca[9], int* info) {
  if (desca[0] != 1)
    die("only dense ScaLAPACK descriptors (desca[0] == 1) are supported", desca[0]);
  if (ia < 1 || ja < 1)
    die("ia and ja are 1-based", ia < 1 ? ia : ja);
  const DLAF_descriptor d = make_dlaf_descriptor(n, n, ia, ja, desca);
  const int r = cholesky(desca[1], type, uplo, a, d);
  if (info)
    *info = r;
}

#ifndef DLAFB_WITH_MPI
void comm_land2(MPI_Comm c, int flags[2]) {
  int* dev = nullptr;
  CAPI_CUDA(cudaMalloc(&dev, 2 * sizeof(int)), "device allocation failed");
  CAPI_CUDA(cudaMemcpy(dev, flags, 2 * sizeof(int), cudaMemcpyHostToDevice), "upload failed");
  if (ncclAllReduce(dev, dev, 2, ncclInt, ncclMin, c->world, nullptr) != ncclSuccess)
    die("grid_ordering: all-reduce failed", -1);
  CAPI_CUDA(cudaStreamSynchronize(nullptr), "stream synchronisation failed");
  CAPI_CUDA(cudaMemcpy(flags, dev, 2 * sizeof(int), cudaMemcpyDeviceToHost), "download failed");
  cudaFree(dev);
}
#endif

}  // namespace

extern "C" {

int dlafb_capi_cached_plans(void) { return (int)(g_plans.size() + g_redist.size()); }

#ifndef DLAFB_WITH_MPI
// ---- MPI shim ----
MPI_Comm dlafb_mpi_comm_world(void) { return g_world; }

int dlafb_mpi_init(int rank, int size, const void* id128) {
  if (g_world)
    return 0;
  dlafb_mpi_comm* w = new dlafb_mpi_comm();
  w->rank = rank;
  w->size = size;
  if (size > 1) {
    if (!id128)
      return -1;
    ncclUniqueId id;
    std::memcpy(&id, id128, sizeof(id));
    const ncclResult_t r = ncclCommInitRank(&w->world, size, id, rank);
    if (r != ncclSuccess) {
      delete w;
      return 1000 + (int)r;
    }
  }
  g_world = w;
  return 0;
}

int dlafb_mpi_finalize(void) {
  if (g_world) {
    if (g_world->world)
      ncclCommDestroy(g_world->world);
    delete g_world;
    g_world = nullptr;
  }
  return 0;
}

int MPI_Comm_rank(MPI_Comm comm, int* rank) {
  *rank = comm ? comm->rank : 0;
  return MPI_SUCCESS;
}
int MPI_Comm_size(MPI_Comm comm, int* size) {
  *size = comm ? comm->size : 1;
  return MPI_SUCCESS;
}
#endif  // !DLAFB_WITH_MPI

// ---- init ----
void dlaf_initialize(int, const char**, int argc_dlaf, const char** argv_dlaf) noexcept {
  if (g_initialized)
    return;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    die("no CUDA device: this library has no CPU fallback", -1);
  for (int i = 0; i < argc_dlaf; ++i)
    if (argv_dlaf && argv_dlaf[i] && std::strcmp(argv_dlaf[i], "--dlaf:print-config") == 0) {
      int dev = 0;
      cudaGetDevice(&dev);
      cudaDeviceProp p;
      cudaGetDeviceProperties(&p, dev);
      std::printf("DLA-Future (B200 build) configuration: device %d %s, %d SMs; look-ahead 1; panel workspaces 2\n",
                  dev, p.name, p.multiProcessorCount);
    }
  g_initialized = true;
}

void dlaf_finalize(void) noexcept {
  drop_plans_of(INT_MIN);
  for (auto& kv : g_grids)
    dlafb_grid_destroy(kv.second);
  g_grids.clear();
  g_initialized = false;
}

// ---- grids ----
static dlafb_grid_t make_grid(MPI_Comm comm, int nprow, int npcol, char order) {
  dlafb_grid_t g = nullptr;
  int e;
  const int size = comm_size(comm);
  if (size != nprow * npcol)
    die("grid size does not match the communicator size", size);
  if (size == 1)
    e = dlafb_grid_create_local(&g);
  else {
    bool owned = false;
    ncclComm_t world = comm_nccl(comm, &owned);
    e = dlafb_grid_create_from_nccl(&g, world, owned ? 1 : 0, comm_rank(comm), size, nprow, npcol, order);
  }
  if (e != 0)
    die("grid creation failed", e);
  return g;
}

// used by the C++ API (include/dlaf/communication/communicator_grid.h)
int dlafb_grid_create_from_comm(dlafb_grid_t* grid, MPI_Comm comm, int nprow, int npcol, char order) {
  if (!grid)
    return -1;
  *grid = make_grid(comm, nprow, npcol, order);
  return 0;
}

int dlaf_create_grid(MPI_Comm comm, int nprow, int npcol, char order) noexcept {
  // contexts count down from INT_MAX so that they never collide with BLACS contexts (src/c_api/grid.cpp:28-40)
  int ctx = INT_MAX - (int)g_grids.size();
  while (g_grids.count(ctx))
    --ctx;
#ifdef DLAFB_WITH_MPI
  if (MPI_Barrier(comm) != MPI_SUCCESS)  // src/c_api/grid.cpp:35
    die("MPI_Barrier failed", -1);
#endif
  g_grids[ctx] = make_grid(comm, nprow, npcol, order);
  return ctx;
}

void dlaf_create_grid_from_blacs_info(int blacs_ctxt, MPI_Comm comm, int nprow, int npcol, char order) noexcept {
  if (g_grids.count(blacs_ctxt)) {
    drop_plans_of(blacs_ctxt);
    dlafb_grid_destroy(g_grids[blacs_ctxt]);
  }
  g_grids[blacs_ctxt] = make_grid(comm, nprow, npcol, order);
}

void dlaf_free_grid(int context) noexcept {
  auto it = g_grids.find(context);
  if (it == g_grids.end())
    return;
  drop_plans_of(context);
  dlafb_grid_destroy(it->second);
  g_grids.erase(it);
}

char grid_ordering(MPI_Comm comm, int nprow, int npcol, int myprow, int mypcol) noexcept {
  // src/c_api/grid.cpp:45-70: both tests are all-reduced (logical and) over the communicator, because rank 0 and every
  // rank on the "diagonal" of the two numberings satisfy both; column-major wins when both hold everywhere.
  const int rank = comm_rank(comm);
  int flags[2] = {rank == myprow * npcol + mypcol ? 1 : 0, rank == mypcol * nprow + myprow ? 1 : 0};
  if (comm_size(comm) > 1)
    comm_land2(comm, flags);
  if (!flags[0] && !flags[1])
    die("grid_ordering: grid layout must be row major or column major", rank);
  return flags[1] ? 'C' : 'R';
}

struct DLAF_descriptor make_dlaf_descriptor(const int m, const int n, const int i, const int j,
                                            const int desc[9]) noexcept {
  // ScaLAPACK's 1-based (ia, ja), stored 0-based like the reference does (src/c_api/utils.cpp:25-33). The reference
  // asserts i == j == 1; this build goes on to accept tile-aligned sub-matrices (checked where they are used).
  if (i < 1 || j < 1)
    die("make_dlaf_descriptor: i and j are 1-based (ScaLAPACK ia, ja)", i < 1 ? i : j);
  struct DLAF_descriptor d = {m, n, desc[4], desc[5], desc[6], desc[7], i - 1, j - 1, desc[8]};
  return d;
}

// ---- factorization ----
int dlaf_cholesky_factorization_s(const int ctx, const char uplo, float* a, const struct DLAF_descriptor d) noexcept {
  return cholesky(ctx, 's', uplo, a, d);
}
int dlaf_cholesky_factorization_d(const int ctx, const char uplo, double* a, const struct DLAF_descriptor d) noexcept {
  return cholesky(ctx, 'd', uplo, a, d);
}
int dlaf_cholesky_factorization_c(const int ctx, const char uplo, dlaf_complex_c* a,
                                  const struct DLAF_descriptor d) noexcept {
  return cholesky(ctx, 'c', uplo, a, d);
}
int dlaf_cholesky_factorization_z(const int ctx, const char uplo, dlaf_complex_z* a,
                                  const struct DLAF_descriptor d) noexcept {
  return cholesky(ctx, 'z', uplo, a, d);
}
void dlaf_pspotrf(const char uplo, const int n, float* a, const int ia, const int ja, const int desca[9],
                  int* info) noexcept {
  pxpotrf('s', uplo, n, a, ia, ja, desca, info);
}
void dlaf_pdpotrf(const char uplo, const int n, double* a, const int ia, const int ja, const int desca[9],
                  int* info) noexcept {
  pxpotrf('d', uplo, n, a, ia, ja, desca, info);
}
void dlaf_pcpotrf(const char uplo, const int n, dlaf_complex_c* a, const int ia, const int ja, const int desca[9],
                  int* info) noexcept {
  pxpotrf('c', uplo, n, a, ia, ja, desca, info);
}
void dlaf_pzpotrf(const char uplo, const int n, dlaf_complex_z* a, const int ia, const int ja, const int desca[9],
                  int* info) noexcept {
  pxpotrf('z', uplo, n, a, ia, ja, desca, info);
}

}  // extern "C"
