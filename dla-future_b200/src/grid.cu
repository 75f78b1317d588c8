// Process grid over NCCL, index-algebra and schedule exports (see include/dlafb_engine.h).
#include <cstdlib>
#include "grid.h"

#include <cstring>

#include "../../include/dlafb_engine.h"
#include "common.cuh"
#include "schedule.h"

using namespace dlafb;

namespace {
inline int nccl_err(ncclResult_t r) { return r == ncclSuccess ? 0 : 1000 + (int)r; }
}  // namespace

extern "C" {

int dlafb_nccl_unique_id(void* out128) {
  ncclUniqueId id;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
  const ncclResult_t r = ncclGetUniqueId(&id);
  if (r != ncclSuccess)
    return nccl_err(r);
  std::memcpy(out128, &id, sizeof(id));
  return 0;
}

// Builds the grid on top of an existing world communicator (owned or borrowed).
int dlafb_grid_create_from_nccl(dlafb_grid_t* out, ncclComm_t world, int owns_world, int world_rank, int world_size,
                                int P, int Q, char ordering) {
  dlafb_grid* g = new dlafb_grid();
  g->P = P;
  g->Q = Q;
  g->world_rank = world_rank;
  g->world_size = world_size;
  g->world = world;
  g->owns_world = owns_world != 0;
  g->ordering = (ordering == 'C' || ordering == 'c') ? 'C' : 'R';
  // common::computeCoords (include/dlaf/common/index2d.h): row-major rank = row * Q + col
  if (g->ordering == 'R') {
    g->pr = world_rank / Q;
    g->pc = world_rank % Q;
  }
  else {
    g->pr = world_rank % P;
    g->pc = world_rank / P;
  }
  ncclResult_t r = ncclCommSplit(g->world, g->pr, g->pc, &g->row, nullptr);  // my grid row, ordered by column
  if (r == ncclSuccess)
    r = ncclCommSplit(g->world, g->pc, g->pr, &g->col, nullptr);  // my grid column, ordered by row
  cudaError_t e = cudaSuccess;
  if (r == ncclSuccess)
    e = cudaStreamCreateWithFlags(&g->stream, cudaStreamNonBlocking);
  if (r == ncclSuccess && e == cudaSuccess)
    e = cudaMalloc(&g->scratch, 64);
  if (r != ncclSuccess || e != cudaSuccess) {
    g->owns_world = false;
    dlafb_grid_destroy(g);
    return r != ncclSuccess ? nccl_err(r) : (int)e;
  }
  *out = g;
  return 0;
}

int dlafb_grid_create(dlafb_grid_t* out, const void* unique_id128, int world_rank, int world_size, int P, int Q,
                      char ordering) {
  if (!out || P < 1 || Q < 1 || world_size != P * Q || world_rank < 0 || world_rank >= world_size)
    return -1;
  if (world_size == 1)
    return dlafb_grid_create_local(out);
  if (!unique_id128)
    return -1;
  ncclUniqueId id;
  std::memcpy(&id, unique_id128, sizeof(id));
  ncclComm_t world = nullptr;
  const ncclResult_t r = ncclCommInitRank(&world, world_size, id, world_rank);
  if (r != ncclSuccess)
    return nccl_err(r);
  const int e = dlafb_grid_create_from_nccl(out, world, 1, world_rank, world_size, P, Q, ordering);
  if (e != 0)
    ncclCommDestroy(world);
  return e;
}

int dlafb_grid_create_local(dlafb_grid_t* out) {
  if (!out)
    return -1;
  dlafb_grid* g = new dlafb_grid();
  cudaError_t e = cudaStreamCreateWithFlags(&g->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess)
    e = cudaMalloc(&g->scratch, 64);
  if (e != cudaSuccess) {
    delete g;
    return (int)e;
  }
  *out = g;
  return 0;
}

int dlafb_grid_destroy(dlafb_grid_t g) {
  if (!g)
    return 0;
  dlafb_cholesky_plan_cache_drop(g);  // cached plans hold this grid's communicators
  if (g->row)
    ncclCommDestroy(g->row);
  if (g->col)
    ncclCommDestroy(g->col);
  if (g->world && g->owns_world)
    ncclCommDestroy(g->world);
  if (g->stream)
    cudaStreamDestroy(g->stream);
  if (g->scratch)
    cudaFree(g->scratch);
  delete g;
  return 0;
}

int dlafb_grid_info(dlafb_grid_t g, int* P, int* Q, int* pr, int* pc, int* world_rank, char* ordering) {
  if (!g)
    return -1;
  if (P) *P = g->P;
  if (Q) *Q = g->Q;
  if (pr) *pr = g->pr;
  if (pc) *pc = g->pc;
  if (world_rank) *world_rank = g->world_rank;
  if (ordering) *ordering = g->ordering;
  return 0;
}

int dlafb_grid_allreduce_max(dlafb_grid_t g, double* value) {
  if (!g || !value)
    return -1;
  if (g->world_size == 1)
    return (int)cudaDeviceSynchronize();
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess)
    return (int)e;
  e = cudaMemcpyAsync(g->scratch, value, sizeof(double), cudaMemcpyHostToDevice, g->stream);
  if (e != cudaSuccess)
    return (int)e;
  const ncclResult_t r = ncclAllReduce(g->scratch, g->scratch, 1, ncclDouble, ncclMax, g->world, g->stream);
  if (r != ncclSuccess)
    return nccl_err(r);
  e = cudaMemcpyAsync(value, g->scratch, sizeof(double), cudaMemcpyDeviceToHost, g->stream);
  if (e != cudaSuccess)
    return (int)e;
  return (int)cudaStreamSynchronize(g->stream);
}

int dlafb_grid_barrier(dlafb_grid_t g) {
  double v = 0.0;
  return dlafb_grid_allreduce_max(g, &v);
}

// ---- index algebra ----
int64_t dlafb_dist_nr_tiles(int64_t n, int64_t nb) { return Dist1D{n, nb, 1, 0, 0}.nt(); }
int dlafb_dist_rank_global_tile(int64_t g, int P, int src) { return Dist1D{0, 1, P, 0, src}.owner(g); }
int64_t dlafb_dist_local_tile_from_global(int64_t g, int P) { return g / P; }
int64_t dlafb_dist_next_local_tile_from_global(int64_t g, int P, int rank, int src) {
  return Dist1D{0, 1, P, rank, src}.next_local(g);
}
int64_t dlafb_dist_global_tile_from_local(int64_t l, int P, int rank, int src) {
  return Dist1D{0, 1, P, rank, src}.global_of(l);
}
int64_t dlafb_dist_local_nr_tiles(int64_t n, int64_t nb, int P, int rank, int src) {
  return Dist1D{n, nb, P, rank, src}.nloc_tiles();
}
int64_t dlafb_dist_local_size(int64_t n, int64_t nb, int P, int rank, int src) {
  return Dist1D{n, nb, P, rank, src}.local_size();
}

int dlafb_schedule_step(int64_t n, int64_t nb, int P, int Q, int pr, int pc, int src_r, int src_c, int64_t k,
                        dlafb_step_info* info, int* colb, int* hp, int* np) {
  if (n < 0 || nb < 1 || P < 1 || Q < 1 || !info)
    return -1;
  const Schedule s = make_schedule(n, nb, P, Q, pr, pc, src_r, src_c);
  if (k < 0 || k >= (int64_t)s.steps.size())
    return -1;
  const Step& st = s.steps[k];
  info->k = st.k;
  info->own_diag = st.own_diag;
  info->in_col = st.in_col;
  info->diag_root = st.diag_root;
  info->row_root = st.row_root;
  info->lk_row = st.lk_row;
  info->lk_col = st.lk_col;
  info->il0 = st.il0;
  info->jl0 = st.jl0;
  info->il_end = st.il_end;
  info->jl_end = st.jl_end;
  info->n_colb = (int)st.colb.size();
  info->n_hp = (int)st.hp.size();
  info->n_np = (int)st.np.size();
  if (colb)
    for (size_t i = 0; i < st.colb.size(); ++i) {
      colb[4 * i + 0] = st.colb[i].jl;
      colb[4 * i + 1] = (int)st.colb[i].g;
      colb[4 * i + 2] = st.colb[i].root;
      colb[4 * i + 3] = st.colb[i].src_il;
    }
  auto dump = [](const std::vector<UpdateTile>& v, int* out) {
    if (out)
      for (size_t i = 0; i < v.size(); ++i) {
        out[3 * i + 0] = v[i].row;
        out[3 * i + 1] = v[i].col;
        out[3 * i + 2] = v[i].diag;
      }
  };
  dump(st.hp, hp);
  dump(st.np, np);
  return 0;
}

uint64_t dlafb_launch_count(void) { return g_launches.load(); }

}  // extern "C"

// The engine drives up to 14 streams per rank (two priority streams, bulk, diagonal, copy-in groups, copy-out, peer
// copies). With the default of 8 hardware work queues several of them alias and serialise (measured: the
// host-mirrored run 394 ms instead of 365 ms at n = 32768, profiles/r01/e2e_probe.txt). The variable is read when the
// CUDA context is created, so it is set when the library is loaded, unless the user already chose a value.
namespace {
__attribute__((constructor)) void dlafb_more_hw_queues() {
  setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);
}
}  // namespace

