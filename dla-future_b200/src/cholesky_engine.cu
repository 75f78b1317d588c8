// Look-ahead scheduler of the distributed right-looking Cholesky over CUDA streams and NCCL.
//
// Role in the reference: Cholesky<Backend::GPU, Device::GPU, T>::call_L / call_U, local and distributed
// (include/dlaf/factorization/cholesky/impl.h:150-452), together with the machinery it leans on there: pika
// senders + per-tile async_rw_mutex for ordering, a CUDA stream pool with two priorities, RoundRobin<Panel>
// workspaces (impl.h:218-221) and per-tile MPI_Ibcast with host staging. Here instead, per step k:
//
//   hp stream (high priority): POTRF(k,k) -> [bcast diag tile + its inverted blocks down the grid column]
//        -> panel solve of column k -> [pack into the row panel, bcast along the row, bcasts of the
//        transposed panel down the column] -> trailing update of global column k+1 only (look-ahead)
//   np stream: trailing update of everything right of column k+1, one launch
//
// so POTRF/solve/broadcast of step k+1 overlap the bulk update of step k. Two panel buffers alternate
// (the reference's n_workspaces = 2). Events express the only cross-stream dependencies:
//   np(k) waits for the panels of step k; the hp update of step k waits for np(k-1).
// The Upper variant runs the Lower algorithm on the conjugate-transposed local block (rows <-> columns of the
// process grid swapped), with one HBM-speed transpose before and after.
#include <cuda.h>
#include <cuda_runtime.h>
#include <nccl.h>

#include <algorithm>
#include <cmath>
#include <type_traits>
#include <cstdlib>
#include <vector>

#include "../../include/dlafb_engine.h"
#include "gemm_dmma.cuh"
#include "gemm_simt.cuh"
#include "grid.h"
#include "schedule.h"
#include "tile_ops.cuh"

using namespace dlafb;

struct dlafb_chol_plan {
  virtual ~dlafb_chol_plan() {}
  virtual int run(void* a, int64_t lda, cudaStream_t stream) = 0;
  virtual int run_host(void* host, int64_t ldh, void* dev, int64_t ldd, cudaStream_t stream) = 0;
  virtual int info(int* out) = 0;
  virtual void local_size(int64_t* r, int64_t* c) = 0;
  virtual int profile(double* ms, double* flops, int* launches) = 0;
  virtual int timeline(double* out, long long cap) = 0;
  virtual int exchange_mode() const = 0;
  virtual int residual(void* a0, int64_t lda, const void* l, int64_t ldl, cudaStream_t stream, double* out2) = 0;
  bool profiling = false;
};

namespace {

// launches with fewer 128 x 128 blocks than this use the 64 x 64 shape (148 SMs)
constexpr size_t kSmallShapeBelow = 120;

// One-thread kernel that publishes a step counter in a peer GPU's memory, ordered after everything that precedes it
// on its stream (the copies of a panel): the peer's stream waits on it with a stream memory operation, no SM.
__global__ void flag_write_kernel(volatile uint32_t* flag, uint32_t value) {
  __threadfence_system();
  *flag = value;
  __threadfence_system();
}

using StreamWaitValue32Fn = CUresult (*)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
StreamWaitValue32Fn stream_wait_value32() {
  static StreamWaitValue32Fn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      f = nullptr;
    return reinterpret_cast<StreamWaitValue32Fn>(f);
  }();
  return fn;
}

// bulk launches with fewer 128 x 128 blocks than this leave SM slots to the high-priority stream (see update()).
// Experiment, off by default: one persistent CTA per SM runs the update at ~55 % of the two-CTA rate and the
// look-ahead update then only gets the other half slots: n=16384 on 1 GPU 49.1 -> 51.7 ms with 3000.
size_t reserve_below() {
  static const size_t v = [] {
    const char* e = getenv("DLAFB_RESERVE_BELOW");
    return e ? (size_t)atoll(e) : (size_t)0;  // off: measured slower (profiles/r01/reserve_sweep.txt)
  }();
  return v;
}

// shape of the bulk launches (DLAFB_BULK_SHAPE=0: one 128 x 128 CTA per SM, 3: two 64 x 128 CTAs per SM)
GemmShape bulk_shape() {
  static const GemmShape s = [] {
    const char* e = getenv("DLAFB_BULK_SHAPE");
    return (e && e[0] == '0') ? GemmShape::k128x128 : (e && e[0] == '3') ? GemmShape::k64x128 : kDefaultBulkShape;
  }();
  return s;
}

#define CU(x)                    \
  do {                           \
    cudaError_t e_ = (x);        \
    if (e_ != cudaSuccess)       \
      return (int)e_;            \
  } while (0)
#define PP(x)                    \
  do {                           \
    const int p_ = (x);          \
    if (p_ != 0)                 \
      return p_;                 \
  } while (0)
#define NC(x)                    \
  do {                           \
    ncclResult_t r_ = (x);       \
    if (r_ != ncclSuccess)       \
      return 1000 + (int)r_;     \
  } while (0)

template <class T>
struct NcclType;
template <>
struct NcclType<float> {
  static constexpr ncclDataType_t t = ncclFloat;
  static constexpr int mult = 1;
};
template <>
struct NcclType<double> {
  static constexpr ncclDataType_t t = ncclDouble;
  static constexpr int mult = 1;
};
template <>
struct NcclType<float2> {
  static constexpr ncclDataType_t t = ncclFloat;
  static constexpr int mult = 2;
};
template <>
struct NcclType<double2> {
  static constexpr ncclDataType_t t = ncclDouble;
  static constexpr int mult = 2;
};

inline int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

template <class T>
struct Plan final : dlafb_chol_plan {
  using R = typename Tr<T>::Real;

  // configuration (L orientation: for uplo = 'U' rows and columns of the grid are swapped)
  bool upper = false;
  int src_r = 0, src_c = 0;
  int64_t n = 0, nb = 1;
  int P = 1, Q = 1, pr = 0, pc = 0;
  ncclComm_t rowc = nullptr, colc = nullptr, worldc = nullptr;
  // residual checker (lazy): inclusive schedule, its lists, panel buffers of a single-rank plan, results
  Schedule rsched;
  int2* rlists = nullptr;
  std::vector<size_t> r_off;
  double* r_out = nullptr;
  bool r_ready = false;
  Schedule sched;
  int64_t nt = 0, mloc = 0, nloc = 0;  // local rows / cols (L orientation)
  bool dist = false;                   // more than one rank
  bool fast = false;                   // TMA + DMMA path (double with nb % 128 == 0, complex<double> with nb % 64 == 0)
  bool cplx = false;                   // T is complex<double>: the DMMA kernel works on the real views
  int vs = 1;                          // doubles per element along a column of the view (2 for complex<double>)
  int ib = 64, nblk = 1;

  cudaStream_t hp = nullptr, np = nullptr;
  // diagonal tile k+1 is factored on its own stream as soon as ITS look-ahead update is done, next to the look-ahead
  // update of the rest of column k+1 (both are on the critical path of step k+1)
  cudaStream_t dp = nullptr;
  cudaEvent_t ev_ladiag = nullptr, ev_potrf = nullptr;
  bool diag_ahead = false;  // ev_ladiag of the previous step covers this step's diagonal tile
  // host-mirrored runs: finished column panels leave for the host on `dh` while later steps compute
  cudaStream_t dh = nullptr;
  cudaEvent_t ev_col = nullptr, ev_dh_end = nullptr;
  T* sink = nullptr;
  int64_t sink_ld = 0;
  // ... and (single rank) the host block ARRIVES in kUpGroups groups of tile columns (growing in size) on `uh`
  // while the first steps already run: during the first `up_steps` steps the bulk update is one launch per group on
  // the group's own stream, which waits for that group's upload only.
  static constexpr int kUpGroups = 8;
  cudaStream_t uh = nullptr, gs[kUpGroups] = {};
  cudaEvent_t ev_up[kUpGroups] = {}, ev_grp[2][kUpGroups] = {};  // ev_grp[step parity][group]
  std::vector<int> grp_of_col;                 // LOCAL tile column -> upload group
  std::vector<size_t> split_off;               // [step * (kUpGroups + 1) + g]: range in lists_split
  int2* lists_split = nullptr;
  int64_t up_steps = 0;                        // steps with a split bulk update (0: no overlap possible)
  bool up_on = false;                          // this run uploads while it computes
  cudaEvent_t ev_in = nullptr, ev_panel[2] = {nullptr, nullptr}, ev_np[2] = {nullptr, nullptr}, ev_hp_end = nullptr,
              ev_np_end = nullptr;
  T* D = nullptr;  // [inverted diagonal blocks (nblk*ib*ib) | diagonal tile (nb*nb, ld = nb)]
  T* Dk[2] = {nullptr, nullptr};  // D of even / odd steps (the same buffer unless peers write into it)
  size_t d_elems = 0;

  // ---- panel exchange over NVLink peer memory instead of NCCL (one node, one process per GPU) ----
  // Every rank owns one arena [flags | D0 D1 | Pr0 Pr1 | Pc0 Pc1] with the same layout everywhere, exported with CUDA
  // IPC to the ranks of its grid row and column. The owner of a panel COPIES it into the peers' arenas with the copy
  // engines (side streams, no SM) and then raises a step counter there; the receivers' streams wait for the counter
  // with cuStreamWaitValue32. Receivers acknowledge "step s consumed" the same way before a buffer is reused.
  // (ncclBroadcast kept resident, spinning kernels on every receiver: 6 % of the SM time at 8 GPUs.)
  struct P2P {
    bool on = false;
    uint8_t* base = nullptr;
    size_t bytes = 0, off_D[2] = {0, 0}, off_Pr[2] = {0, 0}, off_Pc[2] = {0, 0};
    std::vector<uint8_t*> row_peer, col_peer;   // arena of the rank in grid column q / grid row r (self: base)
    std::vector<cudaStream_t> cs_row, cs_col;   // one copy stream per peer
    std::vector<cudaEvent_t> ev_row_sent[2], ev_col_sent[2], ev_tail;
    cudaEvent_t ev_src = nullptr, ev_la = nullptr;
    uint64_t epoch = 0;  // factorizations run so far: the step counters never restart
  } pp;
  uint32_t* flag_at(uint8_t* arena, int idx) const { return reinterpret_cast<uint32_t*>(arena) + idx; }
  int f_sig_d(int b) const { return b; }
  int f_sig_row(int b) const { return 2 + b; }
  int f_sig_col(int b, int r) const { return 4 + b * P + r; }
  int f_done_row(int q) const { return 4 + 2 * P + q; }
  int f_done_col(int r) const { return 4 + 2 * P + Q + r; }
  T* Pr[2] = {nullptr, nullptr};  // row panel: one nb x nb slot per local row tile
  T* Pc[2] = {nullptr, nullptr};  // transposed panel: one slot per local column tile
  CUtensorMap map_pr[2], map_pc[2];      // boxes of 128 rows (throughput shape)
  CUtensorMap map_pr_s[2], map_pc_s[2];  // boxes of 64 rows (latency shape, few tiles left)
  int2* lists = nullptr;
  std::vector<size_t> hp_off, np_off;
  int* info_dev = nullptr;
  int* info_red = nullptr;  // grid-wide reduction of info
  T* tbuf = nullptr;  // transposed local block (Upper) or aligned mirror
  int64_t ldt = 0;
  // profiling of the bulk (np stream) update launches: one timed event pair per step
  std::vector<cudaEvent_t> pev;
  // ... and of the high-priority stream: per step [start, diagonal tile done, panel solved + packed, row broadcast
  // done, column broadcasts done, look-ahead column updated], plus the start of the run
  std::vector<cudaEvent_t> hev;
  cudaEvent_t ev_t0 = nullptr;
  std::vector<double> np_flops;
  std::vector<char> np_timed;

  ~Plan() override {
    cudaDeviceSynchronize();
    if (hp) cudaStreamDestroy(hp);
    if (np) cudaStreamDestroy(np);
    if (dp) cudaStreamDestroy(dp);
    if (ev_ladiag) cudaEventDestroy(ev_ladiag);
    if (ev_potrf) cudaEventDestroy(ev_potrf);
    if (dh) cudaStreamDestroy(dh);
    if (uh) cudaStreamDestroy(uh);
    for (int g = 0; g < kUpGroups; ++g) {
      if (gs[g]) cudaStreamDestroy(gs[g]);
      if (ev_up[g]) cudaEventDestroy(ev_up[g]);
      if (ev_grp[0][g]) cudaEventDestroy(ev_grp[0][g]);
      if (ev_grp[1][g]) cudaEventDestroy(ev_grp[1][g]);
    }
    cudaFree(lists_split);
    if (ev_col) cudaEventDestroy(ev_col);
    if (ev_dh_end) cudaEventDestroy(ev_dh_end);
    for (cudaEvent_t e : {ev_in, ev_panel[0], ev_panel[1], ev_np[0], ev_np[1], ev_hp_end, ev_np_end})
      if (e) cudaEventDestroy(e);
    if (!pp.base)
      cudaFree(D);
    for (uint8_t* q : pp.row_peer)
      if (q && q != pp.base) cudaIpcCloseMemHandle(q);
    for (uint8_t* q : pp.col_peer)
      if (q && q != pp.base) cudaIpcCloseMemHandle(q);
    for (cudaStream_t q : pp.cs_row)
      if (q) cudaStreamDestroy(q);
    for (cudaStream_t q : pp.cs_col)
      if (q) cudaStreamDestroy(q);
    for (int b = 0; b < 2; ++b) {
      for (cudaEvent_t e : pp.ev_row_sent[b]) cudaEventDestroy(e);
      for (cudaEvent_t e : pp.ev_col_sent[b]) cudaEventDestroy(e);
    }
    for (cudaEvent_t e : pp.ev_tail) cudaEventDestroy(e);
    if (pp.ev_src) cudaEventDestroy(pp.ev_src);
    if (pp.ev_la) cudaEventDestroy(pp.ev_la);
    if (!pp.base)
      for (int b = 0; b < 2; ++b) {
        cudaFree(Pr[b]);
        if (Pc[b] != Pr[b])  // the residual checker of a single-rank plan aliases them
          cudaFree(Pc[b]);
      }
    cudaFree(pp.base);
    cudaFree(lists);
    cudaFree(rlists);
    cudaFree(r_out);
    cudaFree(info_dev);
    cudaFree(info_red);
    cudaFree(tbuf);
    for (cudaEvent_t e : pev)
      cudaEventDestroy(e);
    for (cudaEvent_t e : hev)
      cudaEventDestroy(e);
    if (ev_t0) cudaEventDestroy(ev_t0);
  }

  // per step 8 times in ms since the start of the last profiled run: the 6 hp marks, then begin/end of the bulk update
  // (-1 where a mark was not recorded)
  int timeline(double* out, long long cap) override {
    CU(cudaStreamSynchronize(hp));
    CU(cudaStreamSynchronize(np));
    if (hev.empty() || cap < nt * 8)
      return 1;
    for (int64_t k = 0; k < nt; ++k) {
      for (int i = 0; i < 8; ++i)
        out[k * 8 + i] = -1.0;
      if (k == nt - 1)
        continue;
      float x = 0;
      for (int i = 0; i < 6; ++i) {
        CU(cudaEventElapsedTime(&x, ev_t0, hev[6 * k + i]));
        out[k * 8 + i] = x;
      }
      if (np_timed[k])
        for (int i = 0; i < 2; ++i) {
          CU(cudaEventElapsedTime(&x, ev_t0, pev[2 * k + i]));
          out[k * 8 + 6 + i] = x;
        }
    }
    return 0;
  }

  // algorithmic flops of one update tile (real flops; complex multiply-add = 8)
  double tile_flops(const UpdateTile& t) const {
    const double r = (double)std::min<int64_t>(nb, mloc - t.row), c = (double)std::min<int64_t>(nb, nloc - t.col);
    const double cm = Tr<T>::is_complex ? 4.0 : 1.0;
    return cm * (t.diag ? r * (r + 1) * (double)nb : 2.0 * r * c * (double)nb);
  }

  int profile(double* ms, double* flops, int* launches) override {
    CU(cudaStreamSynchronize(hp));
    CU(cudaStreamSynchronize(np));
    double t = 0, f = 0;
    int l = 0;
    for (int64_t k = 0; k < nt && 2 * k + 1 < (int64_t)pev.size(); ++k)
      if (np_timed[k]) {
        float x = 0;
        CU(cudaEventElapsedTime(&x, pev[2 * k], pev[2 * k + 1]));
        t += x;
        f += np_flops[k];
        ++l;
      }
    *ms = t;
    *flops = f;
    *launches = l;
    return 0;
  }

  void local_size(int64_t* r, int64_t* c) override {
    *r = upper ? nloc : mloc;
    *c = upper ? mloc : nloc;
  }

  // Collective over the grid. Falls back to NCCL (pp.on = false) unless EVERY rank could map its peers.
  int p2p_setup(dlafb_grid* g) {
    const char* env = getenv("DLAFB_P2P");
    int ok = (env && env[0] == '0') ? 0 : 1;
    if (ok && (stream_wait_value32() == nullptr || nt <= 0))
      ok = 0;
    // uniform layout: sizes from the largest local tile counts of the grid
    const size_t slotb = (size_t)nb * nb * sizeof(T);
    const size_t rt_max = (size_t)((nt + P - 1) / P), ct_max = (size_t)((nt + Q - 1) / Q);
    auto up = [](size_t x) { return (x + 1023) & ~(size_t)1023; };
    size_t off = up(sizeof(uint32_t) * (size_t)(4 + 3 * P + Q));
    for (int b = 0; b < 2; ++b) {
      pp.off_D[b] = off;
      off += up(d_elems * sizeof(T));
    }
    for (int b = 0; b < 2; ++b) {
      pp.off_Pr[b] = off;
      off += up((rt_max + 1) * slotb);
    }
    for (int b = 0; b < 2; ++b) {
      pp.off_Pc[b] = off;
      off += up((ct_max + 1) * slotb);
    }
    pp.bytes = off;
    struct Info {
      cudaIpcMemHandle_t h;
      int pr, pc, ok;
    } mine{};
    mine.pr = pr;
    mine.pc = pc;
    if (ok) {
      if (cudaMalloc(&pp.base, pp.bytes) != cudaSuccess || cudaMemset(pp.base, 0, pp.bytes) != cudaSuccess ||
          cudaIpcGetMemHandle(&mine.h, pp.base) != cudaSuccess) {
        cudaGetLastError();
        ok = 0;
      }
    }
    mine.ok = ok;
    // exchange the handles (bytes through the world communicator)
    const int W = g->world_size;
    std::vector<Info> all((size_t)W);
    uint8_t* xbuf = nullptr;
    CU(cudaMalloc(&xbuf, sizeof(Info) * (size_t)(W + 1)));
    // (the plain cudaMemcpy / cudaMemset calls of this file are ordered on the legacy stream only, and a host-to-device
    // copy from pageable memory may still be in flight when it returns: nothing on the plan's non-blocking streams is
    // ordered after them until the cudaDeviceSynchronize at the end of init())
    CU(cudaDeviceSynchronize());  // the memset of the arena above
    CU(cudaMemcpyAsync(xbuf, &mine, sizeof(Info), cudaMemcpyHostToDevice, g->stream));
    NC(ncclAllGather(xbuf, xbuf + sizeof(Info), sizeof(Info), ncclUint8, g->world, g->stream));
    CU(cudaStreamSynchronize(g->stream));
    CU(cudaMemcpy(all.data(), xbuf + sizeof(Info), sizeof(Info) * (size_t)W, cudaMemcpyDeviceToHost));
    for (const Info& i : all)
      ok = ok && i.ok;
    pp.row_peer.assign((size_t)Q, nullptr);
    pp.col_peer.assign((size_t)P, nullptr);
    if (ok) {
      for (const Info& i : all) {
        const bool same_row = i.pr == pr, same_col = i.pc == pc;
        if (!same_row && !same_col)
          continue;
        uint8_t* ptr = pp.base;
        if (!(same_row && same_col)) {
          void* v = nullptr;
          if (cudaIpcOpenMemHandle(&v, i.h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            cudaGetLastError();
            ok = 0;
            break;
          }
          ptr = static_cast<uint8_t*>(v);
        }
        if (same_row)
          pp.row_peer[(size_t)i.pc] = ptr;
        if (same_col)
          pp.col_peer[(size_t)i.pr] = ptr;
      }
    }
    // second agreement: every rank opened every handle it needs
    int* flag = reinterpret_cast<int*>(xbuf);
    CU(cudaMemcpyAsync(flag, &ok, sizeof(int), cudaMemcpyHostToDevice, g->stream));
    NC(ncclAllReduce(flag, flag, 1, ncclInt, ncclMin, g->world, g->stream));
    CU(cudaStreamSynchronize(g->stream));
    CU(cudaMemcpy(&ok, flag, sizeof(int), cudaMemcpyDeviceToHost));
    CU(cudaFree(xbuf));
    if (!ok) {
      for (uint8_t*& q : pp.row_peer) {
        if (q && q != pp.base) cudaIpcCloseMemHandle(q);
        q = nullptr;
      }
      for (uint8_t*& q : pp.col_peer) {
        if (q && q != pp.base) cudaIpcCloseMemHandle(q);
        q = nullptr;
      }
      cudaFree(pp.base);
      pp.base = nullptr;
      cudaGetLastError();
      return 0;  // NCCL path
    }
    int lo, hi;
    CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    pp.cs_row.assign((size_t)Q, nullptr);
    pp.cs_col.assign((size_t)P, nullptr);
    for (int q = 0; q < Q; ++q)
      if (q != pc)
        CU(cudaStreamCreateWithPriority(&pp.cs_row[(size_t)q], cudaStreamNonBlocking, hi));
    for (int r = 0; r < P; ++r)
      if (r != pr)
        CU(cudaStreamCreateWithPriority(&pp.cs_col[(size_t)r], cudaStreamNonBlocking, hi));
    for (int b = 0; b < 2; ++b) {
      pp.ev_row_sent[b].resize((size_t)Q);
      pp.ev_col_sent[b].resize((size_t)P);
      for (cudaEvent_t& e : pp.ev_row_sent[b])
        CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      for (cudaEvent_t& e : pp.ev_col_sent[b])
        CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    pp.ev_tail.resize((size_t)(P + Q));
    for (cudaEvent_t& e : pp.ev_tail)
      CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&pp.ev_src, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&pp.ev_la, cudaEventDisableTiming));
    for (int b = 0; b < 2; ++b) {
      Dk[b] = reinterpret_cast<T*>(pp.base + pp.off_D[b]);
      Pr[b] = reinterpret_cast<T*>(pp.base + pp.off_Pr[b]);
      Pc[b] = reinterpret_cast<T*>(pp.base + pp.off_Pc[b]);
    }
    D = Dk[0];
    pp.on = true;
    return 0;
  }

  int exchange_mode() const override { return !dist ? 0 : (pp.on ? 2 : 1); }

  int wait_flag(cudaStream_t s, int idx, uint32_t value) {
    const CUresult r = stream_wait_value32()(s, reinterpret_cast<CUdeviceptr>(flag_at(pp.base, idx)), value,
                                             CU_STREAM_WAIT_VALUE_GEQ);
    return r == CUDA_SUCCESS ? 0 : 1000 + (int)r;
  }
  int raise_flag(cudaStream_t s, uint8_t* arena, int idx, uint32_t value) {
    flag_write_kernel<<<1, 1, 0, s>>>(flag_at(arena, idx), value);
    ++g_launches;
    return cudaGetLastError() == cudaSuccess ? 0 : 1;
  }

  int init(dlafb_grid* g, char uplo, int64_t n_, int64_t nb_, int sr, int sc) {
    upper = (uplo == 'U' || uplo == 'u');
    n = n_;
    nb = nb_;
    if (!upper) {
      P = g->P, Q = g->Q, pr = g->pr, pc = g->pc, rowc = g->row, colc = g->col;
    }
    else {
      P = g->Q, Q = g->P, pr = g->pc, pc = g->pr, rowc = g->col, colc = g->row;
      const int t = sr;
      sr = sc;
      sc = t;
    }
    dist = P * Q > 1;
    worldc = g->world;
    src_r = sr;
    src_c = sc;
    sched = make_schedule(n, nb, P, Q, pr, pc, sr, sc);
    nt = sched.rows.nt();
    mloc = sched.rows.local_size();
    nloc = sched.cols.local_size();
    // DMMA path: double, and complex<double> through its real view (2*nb rows per tile must be a multiple of 128)
    cplx = std::is_same<T, double2>::value;
    vs = cplx ? 2 : 1;
    fast = (std::is_same<T, double>::value && (nb % 128 == 0)) || (cplx && (nb % 64 == 0));
    ib = (fast && !cplx) ? 128 : 64;
    nblk = (int)((nb + ib - 1) / ib);

    int lo, hi;
    CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CU(cudaStreamCreateWithPriority(&hp, cudaStreamNonBlocking, hi));
    CU(cudaStreamCreateWithPriority(&np, cudaStreamNonBlocking, lo));
    CU(cudaStreamCreateWithPriority(&dp, cudaStreamNonBlocking, hi));
    CU(cudaEventCreateWithFlags(&ev_ladiag, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&ev_potrf, cudaEventDisableTiming));
    CU(cudaStreamCreateWithFlags(&dh, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&ev_col, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&ev_dh_end, cudaEventDisableTiming));
    for (cudaEvent_t* e : {&ev_in, &ev_panel[0], &ev_panel[1], &ev_np[0], &ev_np[1], &ev_hp_end, &ev_np_end})
      CU(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    CU(cudaMalloc(&info_dev, sizeof(int)));
    CU(cudaMemset(info_dev, 0, sizeof(int)));
    d_elems = (size_t)nblk * ib * ib + (size_t)nb * nb;
    if (dist) {
      const int e = p2p_setup(g);
      if (e != 0)
        return e;
    }
    if (!pp.on) {
      CU(cudaMalloc(&D, d_elems * sizeof(T)));
      CU(cudaMemset(D, 0, d_elems * sizeof(T)));
      Dk[0] = Dk[1] = D;
    }
    if (dist) {
      const size_t rt = (size_t)sched.rows.nloc_tiles(), ct = (size_t)sched.cols.nloc_tiles();
      for (int b = 0; b < 2; ++b) {
        if (!pp.on) {
          CU(cudaMalloc(&Pr[b], (rt + 1) * nb * nb * sizeof(T)));
          CU(cudaMalloc(&Pc[b], (ct + 1) * nb * nb * sizeof(T)));
          CU(cudaMemset(Pr[b], 0, (rt + 1) * nb * nb * sizeof(T)));
          CU(cudaMemset(Pc[b], 0, (ct + 1) * nb * nb * sizeof(T)));
        }
        if (fast) {
          const long long tr = (long long)nb * vs;  // rows of a tile in the (real) view
          if (rt && make_operand_map(&map_pr[b], (const double*)Pr[b], (long long)rt * tr, nb, tr, 4, 128, nb) != 0)
            return -2;
          if (ct && make_operand_map(&map_pc[b], (const double*)Pc[b], (long long)ct * tr, nb, tr, 4, 128, nb) != 0)
            return -2;
          if (rt && make_operand_map(&map_pr_s[b], (const double*)Pr[b], (long long)rt * tr, nb, tr, 4, 64, nb) != 0)
            return -2;
          if (ct && make_operand_map(&map_pc_s[b], (const double*)Pc[b], (long long)ct * tr, nb, tr, 4, 64, nb) != 0)
            return -2;
        }
      }
    }
    // update lists of all steps, uploaded once
    std::vector<int2> all;
    hp_off.assign(nt + 1, 0);
    np_off.assign(nt + 1, 0);
    for (int64_t k = 0; k < nt; ++k) {
      const Step& st = sched.steps[k];
      hp_off[k] = all.size();
      for (const UpdateTile& t : st.hp)
        all.push_back(make_int2(t.row, t.col | (t.diag ? (int)0x80000000 : 0)));
      np_off[k] = all.size();
      for (const UpdateTile& t : st.np)
        all.push_back(make_int2(t.row, t.col | (t.diag ? (int)0x80000000 : 0)));
    }
    if (!all.empty()) {
      CU(cudaMalloc(&lists, all.size() * sizeof(int2)));
      CU(cudaMemcpy(lists, all.data(), all.size() * sizeof(int2), cudaMemcpyHostToDevice));
    }
    const int64_t ctl = sched.cols.nloc_tiles();
    if (!upper && nt >= 16 && ctl >= 4 && mloc > 0) {
      // upload groups of growing size over the LOCAL tile columns (single rank: local = global)
      grp_of_col.assign((size_t)ctl, 0);
      auto weight = [&](int64_t jl) {
        const int64_t gj = sched.cols.global_of(jl);
        return (double)(mloc - sched.rows.next_local(gj) * nb) * (double)sched.cols.tile_size(gj);
      };
      double total = 0, acc = 0;
      for (int64_t jl = 0; jl < ctl; ++jl)
        total += weight(jl);
      for (int64_t jl = 0; jl < ctl; ++jl) {
        // group g ends at the fraction ((g+1)/G)^2 of the bytes: small first groups start the factorization early
        grp_of_col[(size_t)jl] = total > 0 ? std::min<int>(kUpGroups - 1, (int)(std::sqrt(acc / total) * kUpGroups)) : 0;
        acc += weight(jl);
      }
      const char* ue = getenv("DLAFB_UP_STEPS");
      up_steps = std::min<int64_t>(ue ? atoll(ue) : (dist ? 16 : 10), nt - 2);
      std::vector<int2> sp;
      split_off.assign((size_t)up_steps * (kUpGroups + 1), 0);
      for (int64_t k = 0; k < up_steps; ++k)
        for (int g = 0; g <= kUpGroups; ++g) {
          split_off[(size_t)k * (kUpGroups + 1) + g] = sp.size();
          if (g < kUpGroups)
            for (const UpdateTile& t : sched.steps[k].np)
              if (grp_of_col[(size_t)(t.col / nb)] == g)
                sp.push_back(make_int2(t.row, t.col | (t.diag ? (int)0x80000000 : 0)));
        }
      if (!sp.empty()) {
        CU(cudaMalloc(&lists_split, sp.size() * sizeof(int2)));
        CU(cudaMemcpy(lists_split, sp.data(), sp.size() * sizeof(int2), cudaMemcpyHostToDevice));
      }
      CU(cudaStreamCreateWithFlags(&uh, cudaStreamNonBlocking));
      for (int g = 0; g < kUpGroups; ++g) {
        CU(cudaStreamCreateWithPriority(&gs[g], cudaStreamNonBlocking, lo));
        CU(cudaEventCreateWithFlags(&ev_up[g], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&ev_grp[0][g], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&ev_grp[1][g], cudaEventDisableTiming));
      }
    }
    np_flops.assign(nt + 1, 0.0);
    for (int64_t k = 0; k < nt; ++k)
      for (const UpdateTile& t : sched.steps[k].np)
        np_flops[k] += tile_flops(t);
    if (upper) {
      ldt = round_up(mloc > 0 ? mloc : 1, 64);
      CU(cudaMalloc(&tbuf, ((size_t)ldt * (nloc > 0 ? nloc : 1) + 64) * sizeof(T)));
    }
    // The uploads and memsets above went through the legacy stream; the plan works on non-blocking streams, which are
    // not ordered after it, and a cudaMemcpy from pageable host memory may return before its DMA has finished. Without
    // this a run right after plan creation could read update lists that had not arrived yet (seen as an
    // order-dependent failure of the redistribution lists, which had the same pattern: profiles/r02/flaky_upload.md).
    CU(cudaDeviceSynchronize());
    return 0;
  }

  // trailing update of the listed tiles: C(tile) -= rowpanel * colpanel^H
  // The 64 x 64 shape takes over when the listed tiles no longer fill the chip with 128 x 128 blocks: each CTA is a
  // serial chain of nb/16 slices, so 4x more, 4x shorter CTAs end the launch sooner (the look-ahead column and the
  // last steps are on the critical path).
  int update(cudaStream_t s, T* A, int64_t ld, const CUtensorMap* map_a_local, int nbox_local, int64_t k, int buf,
             size_t off, size_t count, bool bulk = false, const int2* list_base = nullptr, bool panels = false) {
    if (!list_base)
      list_base = lists;
    panels = panels || dist;  // operands from the packed panels Pr / Pc (always when distributed)
    if (count == 0)
      return 0;
    if (fast) {
      GemmParams p{};
      p.C = reinterpret_cast<double*>(A);
      p.ldc = ld * vs;
      p.m_total = (int)(mloc * vs);
      p.n_total = (int)(nloc * vs);
      p.K = (int)nb;
      p.alpha = -1.0;
      p.beta = 1.0;
      p.list = list_base + off;
      p.tile = (int)(nb * vs);
      p.list_scale = vs;
      const size_t big = count * (p.tile / 128) * (p.tile / 128);
      const GemmShape shape = big < kSmallShapeBelow ? GemmShape::k64x64 : bulk_shape();
      const int bm = shape_bm(shape), bn = shape_bn(shape);
      const unsigned ctas = (unsigned)(count * (p.tile / bm) * (p.tile / bn));
      // Bulk launch short enough that the dependent chain (diagonal tile -> panel solve -> exchange) decides the
      // step time: run it as one persistent CTA per SM on all SMs but one. Every high-priority kernel then finds a
      // free half SM at once, the single-CTA diagonal kernel a whole SM, instead of waiting for update CTAs to retire.
      const unsigned cap = (bulk && big < reserve_below()) ? (unsigned)std::max(1, gemm_sm_count() - 1) : 0u;
      if (panels) {
        p.nbox_a = p.nbox_b = 4;
        return (int)launch_gemm_nt_dmma(s, bm == 64 ? map_pr_s[buf] : map_pr[buf], bn == 64 ? map_pc_s[buf] : map_pc[buf],
                                        p, ctas, 1, shape, cplx, cap);
      }
      p.a_col0 = p.b_col0 = (int)(k * nb);
      p.nbox_a = p.nbox_b = nbox_local;
      return (int)launch_gemm_nt_dmma(s, map_a_local[bm == 64 ? 1 : 0], map_a_local[bn == 64 ? 1 : 0], p, ctas, 1, shape,
                                      cplx, cap);
    }
    SimtParams<T> p{};
    p.C = A;
    p.ldc = ld;
    p.m_total = (int)mloc;
    p.n_total = (int)nloc;
    p.K = (int)nb;
    p.alpha = R(-1);
    p.beta = R(1);
    p.list = list_base + off;
    p.sub = (int)((nb + kSimtTile - 1) / kSimtTile);
    p.tile = (int)nb;
    if (panels) {
      p.A = Pr[buf];
      p.B = Pc[buf];
      p.lda = p.ldb = nb;
      p.a_tiled = p.b_tiled = 1;
    }
    else {
      p.A = A + (size_t)k * nb * ld;
      p.B = p.A;
      p.lda = p.ldb = ld;
    }
    return (int)launch_gemm_simt<T>(s, p, (unsigned)(count * p.sub * p.sub), 1);
  }

  // column panel k is final on this rank (diagonal tile factored, rows below solved): send it to the host mirror
  int sink_column(const Step& st, int64_t k, const T* A, int64_t ld) {
    if (!sink || (dist && !st.in_col))
      return 0;
    const int64_t r0 = (st.own_diag ? (int64_t)st.lk_row : (int64_t)st.il0) * nb;
    const int64_t nrows = mloc - r0, ncols = sched.cols.tile_size(k);
    if (nrows <= 0 || ncols <= 0)
      return 0;
    const size_t c0 = (size_t)st.lk_col * nb;
    CU(cudaEventRecord(ev_col, hp));
    CU(cudaStreamWaitEvent(dh, ev_col, 0));
    CU(cudaMemcpy2DAsync(sink + r0 + c0 * sink_ld, (size_t)sink_ld * sizeof(T), A + r0 + c0 * ld, (size_t)ld * sizeof(T),
                         (size_t)nrows * sizeof(T), (size_t)ncols, cudaMemcpyDeviceToHost, dh));
    return 0;
  }

  // blocks of the referenced triangle of the local matrix in the caller's orientation, one per local tile column
  template <class F>
  int for_each_triangle_block(F&& f) const {
    const Dist1D& crow = upper ? sched.cols : sched.rows;  // caller's rows
    const Dist1D& ccol = upper ? sched.rows : sched.cols;  // caller's columns
    const int64_t lrows = crow.local_size(), lcols = ccol.local_size();
    for (int64_t jl = 0; jl < ccol.nloc_tiles(); ++jl) {
      const int64_t gj = ccol.global_of(jl);
      const int64_t c0 = jl * nb, nc = std::min<int64_t>(nb, lcols - c0);
      int64_t r0, nr;
      if (!upper) {
        r0 = crow.next_local(gj) * nb;
        nr = lrows - r0;
      }
      else {
        r0 = 0;
        nr = std::min<int64_t>(crow.next_local(gj + 1) * nb, lrows);
      }
      if (nr > 0 && nc > 0) {
        const int e = f(r0, nr, c0, nc);
        if (e != 0)
          return e;
      }
    }
    return 0;
  }

  int run_host(void* host_, int64_t ldh, void* dev_, int64_t ldd, cudaStream_t user) override {
    T* host = static_cast<T*>(host_);
    T* dev = static_cast<T*>(dev_);
    if (n <= 0)
      return run(dev, ldd, user);
    const bool misaligned = fast && ((reinterpret_cast<uintptr_t>(dev) & 15) || ((ldd * vs) & 1));
    const bool overlap = !upper && !misaligned;  // the other cases factor a transposed / realigned copy
    // single rank: the upload runs on its own stream, group after group, and the factorization starts with the
    // first group; otherwise the whole triangle is uploaded first
    const bool stream_in = overlap && up_steps > 0 && uh != nullptr;
    cudaStream_t us = stream_in ? uh : user;
    if (stream_in) {
      CU(cudaEventRecord(ev_in, user));
      CU(cudaStreamWaitEvent(uh, ev_in, 0));
    }
    int last_g = 0;
    int e = for_each_triangle_block([&](int64_t r0, int64_t nr, int64_t c0, int64_t nc) {
      if (stream_in) {
        const int g = grp_of_col[(size_t)(c0 / nb)];
        for (; last_g < g; ++last_g)
          CU(cudaEventRecord(ev_up[last_g], uh));
      }
      CU(cudaMemcpy2DAsync(dev + r0 + (size_t)c0 * ldd, (size_t)ldd * sizeof(T), host + r0 + (size_t)c0 * ldh,
                           (size_t)ldh * sizeof(T), (size_t)nr * sizeof(T), (size_t)nc, cudaMemcpyHostToDevice, us));
      return 0;
    });
    if (e != 0)
      return e;
    if (stream_in)
      for (; last_g < kUpGroups; ++last_g)
        CU(cudaEventRecord(ev_up[last_g], uh));
    up_on = stream_in;
    if (overlap) {
      sink = host;
      sink_ld = ldh;
    }
    e = run(dev, ldd, user);
    sink = nullptr;
    up_on = false;
    if (e != 0)
      return e;
    if (overlap) {
      CU(cudaEventRecord(ev_dh_end, dh));
      CU(cudaStreamWaitEvent(user, ev_dh_end, 0));
      return 0;
    }
    return for_each_triangle_block([&](int64_t r0, int64_t nr, int64_t c0, int64_t nc) {
      CU(cudaMemcpy2DAsync(host + r0 + (size_t)c0 * ldh, (size_t)ldh * sizeof(T), dev + r0 + (size_t)c0 * ldd,
                           (size_t)ldd * sizeof(T), (size_t)nr * sizeof(T), (size_t)nc, cudaMemcpyDeviceToHost, user));
      return 0;
    });
  }

  // Residual of a factorization, on the device and distributed like the matrix (the reference's checker,
  // miniapp_cholesky.cpp:243-442, gathers norms on the CPU with blocking MPI and costs more than the factorization):
  // R <- A0 - L*L^H by the same panel exchange and the same update kernels as the factorization (inclusive
  // schedule, no diagonal-tile or solve kernels), then max|.| over the referenced triangle and a max-reduction over
  // the grid. `r` holds this rank's block of the ORIGINAL matrix and is overwritten with the residual; `l` holds
  // the factor. out2 = {max|A0|, max|A0 - L*L^H|}. uplo = Lower only. Synchronises `user` at the end.
  int residual(void* r_, int64_t ldr, const void* l_, int64_t ldl, cudaStream_t user, double* out2) override {
    if (upper)
      return 2;
    T* Rm = static_cast<T*>(r_);
    const T* Lm = static_cast<const T*>(l_);
    out2[0] = out2[1] = 0.0;
    if (n <= 0)
      return 0;
    const size_t slot = (size_t)nb * nb;
    if (!r_ready) {
      rsched = make_schedule(n, nb, P, Q, pr, pc, src_r, src_c, true);
      std::vector<int2> all;
      r_off.assign((size_t)nt + 1, 0);
      for (int64_t k = 0; k < nt; ++k) {
        r_off[(size_t)k] = all.size();
        for (const UpdateTile& t : rsched.steps[k].np)
          all.push_back(make_int2(t.row, t.col | (t.diag ? (int)0x80000000 : 0)));
      }
      r_off[(size_t)nt] = all.size();
      if (!all.empty()) {
        CU(cudaMalloc(&rlists, all.size() * sizeof(int2)));
        CU(cudaMemcpy(rlists, all.data(), all.size() * sizeof(int2), cudaMemcpyHostToDevice));
      }
      CU(cudaMalloc(&r_out, 2 * sizeof(double)));
      if (!dist) {
        // a single-rank plan factors out of the matrix itself; the checker needs packed panels (the diagonal tile
        // enters with its strictly upper part zeroed)
        CU(cudaMalloc(&Pr[0], (size_t)(nt + 1) * slot * sizeof(T)));
        CU(cudaMemset(Pr[0], 0, (size_t)(nt + 1) * slot * sizeof(T)));
        Pc[0] = Pr[0];
        if (fast) {
          const long long tr = (long long)nb * vs;
          if (make_operand_map(&map_pr[0], (const double*)Pr[0], (long long)nt * tr, nb, tr, 4, 128, nb) != 0 ||
              make_operand_map(&map_pr_s[0], (const double*)Pr[0], (long long)nt * tr, nb, tr, 4, 64, nb) != 0)
            return -2;
          map_pc[0] = map_pr[0];
          map_pc_s[0] = map_pr_s[0];
        }
      }
      CU(cudaDeviceSynchronize());  // legacy-stream uploads / memsets above vs the non-blocking hp stream below
      r_ready = true;
    }
    const ncclDataType_t nt_ = NcclType<T>::t;
    const size_t mult = NcclType<T>::mult;
    CU(cudaEventRecord(ev_in, user));
    CU(cudaStreamWaitEvent(hp, ev_in, 0));
    CU(cudaMemsetAsync(r_out, 0, 2 * sizeof(double), hp));
    // max over this rank's part of the lower triangle: the diagonal tile's lower part + everything below it
    auto tri_max = [&](double* out) -> int {
      for (int64_t jl = 0; jl < rsched.cols.nloc_tiles(); ++jl) {
        const int64_t gj = rsched.cols.global_of(jl);
        const int w = (int)rsched.cols.tile_size(gj);
        int64_t il = rsched.rows.next_local(gj);
        const T* col = Rm + (size_t)jl * nb * ldr;
        if (il < rsched.rows.nloc_tiles() && rsched.rows.global_of(il) == gj) {
          CU(max_abs<T>(hp, 'L', w, w, col + (size_t)il * nb, ldr, out));
          ++il;
        }
        const int64_t r0 = il * nb;
        if (r0 < mloc)
          CU(max_abs<T>(hp, 'G', (int)(mloc - r0), w, col + r0, ldr, out));
      }
      return 0;
    };
    PP(tri_max(r_out));
    for (int64_t k = 0; k < nt; ++k) {
      const Step& st = rsched.steps[k];
      const int w = (int)rsched.rows.tile_size(k);
      const size_t row_cnt = (size_t)(st.il_end - st.il0) * slot;
      if (st.in_col && row_cnt > 0) {
        const int64_t r0 = (int64_t)st.il0 * nb;
        T* dst = Pr[0] + (size_t)st.il0 * slot;
        if (w < nb)
          CU(cudaMemsetAsync(dst, 0, row_cnt * sizeof(T), hp));  // ragged last tile column: zero-padded to nb
        CU(pack_tiles<T>(hp, (int)(mloc - r0), w, Lm + r0 + (size_t)st.lk_col * nb * ldl, ldl, dst, (int)nb));
        if (st.own_diag && w > 1)  // strictly upper part of the diagonal tile = upper triangle of T[0:w-1, 1:w]
          CU(laset<T>(hp, 'U', w - 1, w - 1, Tr<T>::zero(), Tr<T>::zero(), dst + nb, nb));
      }
      if (Q > 1 && row_cnt > 0) {
        T* p = Pr[0] + (size_t)st.il0 * slot;
        NC(ncclBroadcast(p, p, row_cnt * mult, nt_, st.row_root, rowc, hp));
      }
      if (P > 1 && !st.colb.empty()) {
        NC(ncclGroupStart());
        for (const ColBcast& cb : st.colb) {
          T* recv = Pc[0] + (size_t)cb.jl * slot;
          const T* send = (cb.root == pr) ? Pr[0] + (size_t)cb.src_il * slot : recv;
          NC(ncclBroadcast(send, recv, slot * mult, nt_, cb.root, colc, hp));
        }
        NC(ncclGroupEnd());
      }
      else if (dist) {
        for (const ColBcast& cb : st.colb)
          CU(cudaMemcpyAsync(Pc[0] + (size_t)cb.jl * slot, Pr[0] + (size_t)cb.src_il * slot, slot * sizeof(T),
                             cudaMemcpyDeviceToDevice, hp));
      }
      const int r = update(hp, Rm, ldr, nullptr, 1, k, 0, r_off[(size_t)k], r_off[(size_t)k + 1] - r_off[(size_t)k],
                           false, rlists, true);
      if (r != 0)
        return r;
    }
    PP(tri_max(r_out + 1));
    if (dist)
      NC(ncclAllReduce(r_out, r_out, 2, ncclDouble, ncclMax, worldc, hp));
    CU(cudaMemcpyAsync(out2, r_out, 2 * sizeof(double), cudaMemcpyDeviceToHost, hp));
    CU(cudaStreamSynchronize(hp));
    return 0;
  }

  int factor_lower(T* A, int64_t ld) {
    CUtensorMap map_local[2];  // [0] boxes of 128 rows, [1] of 64 rows
    int nbox_local = 1;
    if (fast && !dist && n > 0) {
      // rank-3 view reads whole 16-row groups: fine when the rows are a multiple of 16 or the columns are padded
      const long long vr = mloc * vs, vld = ld * vs;  // the (real) view
      const int kind = (vr % 16 == 0 || vld >= round_up(vr, 16)) ? 3 : 2;
      int r = make_operand_map(&map_local[0], (const double*)A, vr, nloc, vld, kind);
      if (r == 0)
        r = make_operand_map(&map_local[1], (const double*)A, vr, nloc, vld, kind, 64);
      nbox_local = kind == 3 ? 1 : 8;
      if (r != 0 && kind == 3) {
        r = make_operand_map(&map_local[0], (const double*)A, vr, nloc, vld, 2);
        map_local[1] = map_local[0];
        nbox_local = 8;
      }
      if (r != 0)
        return -2;
    }
    const ncclDataType_t nt_ = NcclType<T>::t;
    const size_t mult = NcclType<T>::mult;
    const size_t slot = (size_t)nb * nb;
    const size_t slotb = slot * sizeof(T);
    diag_ahead = false;

    for (int64_t k = 0; k < nt; ++k) {
      const Step& st = sched.steps[k];
      const int buf = (int)(k & 1);
      T* DW = Dk[buf];
      T* DL = DW + (size_t)nblk * ib * ib;
      // global step number: the counters in the peers' arenas never restart (pp.epoch = runs so far)
      const uint32_t gstep = (uint32_t)(pp.epoch * (uint64_t)nt + (uint64_t)k);
      if (pp.on) {
        // my copies of step k-2 out of Pr[buf] / Dk[buf] must have left before this step overwrites them
        for (int q = 0; q < Q; ++q)
          if (q != pc)
            CU(cudaStreamWaitEvent(hp, pp.ev_row_sent[buf][(size_t)q], 0));
        for (int r = 0; r < P; ++r)
          if (r != pr)
            CU(cudaStreamWaitEvent(hp, pp.ev_col_sent[buf][(size_t)r], 0));
      }
      const int nk = (int)sched.rows.tile_size(k);
      const bool last = (k == nt - 1);
      T* Akk = nullptr;
      const bool mark = profiling && !last;
      if (mark)
        CU(cudaEventRecord(hev[6 * k + 0], hp));
      // upload group of global tile column g on this rank, -1 when another grid column owns it
      auto col_group = [&](int64_t g) -> int {
        if (g >= nt || sched.cols.owner(g) != pc)
          return -1;
        return grp_of_col[(size_t)sched.cols.local_of(g)];
      };
      if (up_on) {
        // tile column k (and k+1 for the look-ahead update) must have arrived from the host
        const int g0 = col_group(k), g1 = last ? -1 : col_group(k + 1);
        if (g0 >= 0)
          CU(cudaStreamWaitEvent(hp, ev_up[g0], 0));
        if (g1 >= 0)
          CU(cudaStreamWaitEvent(hp, ev_up[g1], 0));
        // distributed: the split bulk update of step k-2 read the panel buffers this step is about to overwrite
        if (dist && k >= 2 && k <= up_steps + 1)
          for (int g = 0; g < kUpGroups; ++g)
            CU(cudaStreamWaitEvent(hp, ev_grp[k & 1][g], 0));
      }
      if (st.own_diag) {
        Akk = A + (size_t)st.lk_row * nb + (size_t)st.lk_col * nb * ld;
        if (diag_ahead) {
          // the tile got its last update first thing in the previous step's look-ahead: factor it on `dp` while `hp`
          // still updates the rest of this column
          CU(cudaStreamWaitEvent(dp, ev_ladiag, 0));
          CU(potrf_tile<T>(dp, 'L', nk, Akk, ld, DW, ib, info_dev, (int)(k * nb)));
          CU(cudaEventRecord(ev_potrf, dp));
          CU(cudaStreamWaitEvent(hp, ev_potrf, 0));
        }
        else {
          CU(potrf_tile<T>(hp, 'L', nk, Akk, ld, DW, ib, info_dev, (int)(k * nb)));
        }
      }
      diag_ahead = false;
      if (mark)
        CU(cudaEventRecord(hev[6 * k + 1], hp));
      if (last) {
        const int e = sink_column(st, k, A, ld);
        if (e != 0)
          return e;
      }
      if (!last) {
        if (!dist) {
          const int m_below = (int)(n - (k + 1) * nb);
          CU(trsm_panel<T>(hp, (int)nb, Akk, ld, DW, ib, A + (size_t)(k + 1) * nb + (size_t)k * nb * ld, ld, m_below));
          const int e = sink_column(st, k, A, ld);
          if (e != 0)
            return e;
        }
        else {
          if (st.in_col) {
            if (P > 1 && !pp.on) {
              if (st.own_diag)
                CU(lacpy<T>(hp, 'L', (int)nb, (int)nb, Akk, ld, DL, nb));
              NC(ncclBroadcast(DW, DW, d_elems * mult, nt_, st.diag_root, colc, hp));
            }
            else if (P > 1 && st.own_diag) {
              CU(lacpy<T>(hp, 'L', (int)nb, (int)nb, Akk, ld, DL, nb));
              CU(cudaEventRecord(pp.ev_src, hp));
              for (int r = 0; r < P; ++r) {
                if (r == pr)
                  continue;
                cudaStream_t cs = pp.cs_col[(size_t)r];
                CU(cudaStreamWaitEvent(cs, pp.ev_src, 0));
                if (gstep >= 2)
                  PP(wait_flag(cs, f_done_col(r), gstep - 1));
                CU(cudaMemcpyAsync(pp.col_peer[(size_t)r] + pp.off_D[buf], DW, d_elems * sizeof(T),
                                   cudaMemcpyDeviceToDevice, cs));
                PP(raise_flag(cs, pp.col_peer[(size_t)r], f_sig_d(buf), gstep + 1));
                CU(cudaEventRecord(pp.ev_col_sent[buf][(size_t)r], cs));
              }
            }
            else if (P > 1) {
              PP(wait_flag(hp, f_sig_d(buf), gstep + 1));
            }
            const T* Lkk = (P > 1) ? DL : Akk;
            const int64_t ldl = (P > 1) ? nb : ld;
            const int m_below = (int)(mloc - (int64_t)st.il0 * nb);
            if (m_below > 0) {
              T* B = A + (size_t)st.il0 * nb + (size_t)st.lk_col * nb * ld;
              // solve + copy into the row panel (tile layout): one kernel for double
              CU(trsm_panel_packed<T>(hp, (int)nb, Lkk, ldl, DW, ib, B, ld, m_below, Pr[buf] + (size_t)st.il0 * slot,
                                      (int)nb));
            }
            const int e = sink_column(st, k, A, ld);
            if (e != 0)
              return e;
          }
          if (mark)
            CU(cudaEventRecord(hev[6 * k + 2], hp));
          const size_t row_cnt = (size_t)(st.il_end - st.il0) * slot;
          if (Q > 1 && row_cnt > 0 && !pp.on) {
            T* p = Pr[buf] + (size_t)st.il0 * slot;
            NC(ncclBroadcast(p, p, row_cnt * mult, nt_, st.row_root, rowc, hp));
          }
          else if (Q > 1 && row_cnt > 0 && st.in_col) {
            const size_t o = (size_t)st.il0 * slotb;
            CU(cudaEventRecord(pp.ev_src, hp));
            for (int q = 0; q < Q; ++q) {
              if (q == pc)
                continue;
              cudaStream_t cs = pp.cs_row[(size_t)q];
              CU(cudaStreamWaitEvent(cs, pp.ev_src, 0));
              if (gstep >= 2)
                PP(wait_flag(cs, f_done_row(q), gstep - 1));
              CU(cudaMemcpyAsync(pp.row_peer[(size_t)q] + pp.off_Pr[buf] + o, reinterpret_cast<uint8_t*>(Pr[buf]) + o,
                                 row_cnt * sizeof(T), cudaMemcpyDeviceToDevice, cs));
              PP(raise_flag(cs, pp.row_peer[(size_t)q], f_sig_row(buf), gstep + 1));
              CU(cudaEventRecord(pp.ev_row_sent[buf][(size_t)q], cs));
            }
          }
          else if (Q > 1 && row_cnt > 0) {
            PP(wait_flag(hp, f_sig_row(buf), gstep + 1));
          }
          if (mark)
            CU(cudaEventRecord(hev[6 * k + 3], hp));
          if (!st.colb.empty()) {
            if (P > 1 && pp.on) {
              // every rank row forwards the tiles whose diagonal position it owns, straight into the peers' Pc
              bool sends = false;
              std::vector<char> from((size_t)P, 0);
              for (const ColBcast& cb : st.colb) {
                from[(size_t)cb.root] = 1;
                if (cb.root == pr) {
                  sends = true;
                  CU(cudaMemcpyAsync(Pc[buf] + (size_t)cb.jl * slot, Pr[buf] + (size_t)cb.src_il * slot, slotb,
                                     cudaMemcpyDeviceToDevice, hp));
                }
              }
              if (sends) {
                CU(cudaEventRecord(pp.ev_src, hp));
                for (int r = 0; r < P; ++r) {
                  if (r == pr)
                    continue;
                  cudaStream_t cs = pp.cs_col[(size_t)r];
                  CU(cudaStreamWaitEvent(cs, pp.ev_src, 0));
                  if (gstep >= 2)
                    PP(wait_flag(cs, f_done_col(r), gstep - 1));
                  for (const ColBcast& cb : st.colb)
                    if (cb.root == pr)
                      CU(cudaMemcpyAsync(pp.col_peer[(size_t)r] + pp.off_Pc[buf] + (size_t)cb.jl * slotb,
                                         Pr[buf] + (size_t)cb.src_il * slot, slotb, cudaMemcpyDeviceToDevice, cs));
                  PP(raise_flag(cs, pp.col_peer[(size_t)r], f_sig_col(buf, pr), gstep + 1));
                  CU(cudaEventRecord(pp.ev_col_sent[buf][(size_t)r], cs));
                }
              }
              for (int r = 0; r < P; ++r)
                if (r != pr && from[(size_t)r])
                  PP(wait_flag(hp, f_sig_col(buf, r), gstep + 1));
            }
            else if (P > 1) {
              NC(ncclGroupStart());
              for (const ColBcast& cb : st.colb) {
                T* recv = Pc[buf] + (size_t)cb.jl * slot;
                const T* send = (cb.root == pr) ? Pr[buf] + (size_t)cb.src_il * slot : recv;
                NC(ncclBroadcast(send, recv, slot * mult, nt_, cb.root, colc, hp));
              }
              NC(ncclGroupEnd());
            }
            else {
              for (const ColBcast& cb : st.colb)
                CU(cudaMemcpyAsync(Pc[buf] + (size_t)cb.jl * slot, Pr[buf] + (size_t)cb.src_il * slot,
                                   slot * sizeof(T), cudaMemcpyDeviceToDevice, hp));
            }
          }
        }
        if (mark) {
          if (!dist) {
            CU(cudaEventRecord(hev[6 * k + 2], hp));
            CU(cudaEventRecord(hev[6 * k + 3], hp));
          }
          CU(cudaEventRecord(hev[6 * k + 4], hp));
        }
        CU(cudaEventRecord(ev_panel[buf], hp));
        // look-ahead: global column k+1 on the high-priority stream, after np(k-1) which touched it last
        if (k >= 1 && up_on && k - 1 < up_steps) {
          const int g1 = col_group(k + 1);
          if (g1 >= 0)
            CU(cudaStreamWaitEvent(hp, ev_grp[(k - 1) & 1][g1], 0));  // split bulk update of step k-1
        }
        else if (k >= 1)
          CU(cudaStreamWaitEvent(hp, ev_np[(k - 1) & 1], 0));
        int r = 0;
        size_t la_off = hp_off[k], la_cnt = np_off[k] - hp_off[k];
        if (la_cnt > 0 && st.hp.front().diag) {
          // this rank owns A(k+1,k+1): its update goes first and releases the factorization of step k+1
          r = update(hp, A, ld, map_local, nbox_local, k, buf, la_off, 1);
          if (r != 0)
            return r;
          CU(cudaEventRecord(ev_ladiag, hp));
          diag_ahead = true;
          ++la_off;
          --la_cnt;
        }
        r = update(hp, A, ld, map_local, nbox_local, k, buf, la_off, la_cnt);
        if (r != 0)
          return r;
        if (mark)
          CU(cudaEventRecord(hev[6 * k + 5], hp));
        const bool split = up_on && k < up_steps;
        if (split) {
          // one launch per upload group on the group's stream: the groups that have arrived run ahead
          for (int g = 0; g < kUpGroups; ++g) {
            const size_t o = split_off[(size_t)k * (kUpGroups + 1) + g];
            const size_t c = split_off[(size_t)k * (kUpGroups + 1) + g + 1] - o;
            if (c > 0) {
              if (k == 0)
                CU(cudaStreamWaitEvent(gs[g], ev_in, 0));
              CU(cudaStreamWaitEvent(gs[g], ev_up[g], 0));
              CU(cudaStreamWaitEvent(gs[g], ev_panel[buf], 0));
              r = update(gs[g], A, ld, map_local, nbox_local, k, buf, o, c, true, lists_split);
              if (r != 0)
                return r;
            }
            CU(cudaEventRecord(ev_grp[buf][g], gs[g]));
          }
          np_timed[k] = 0;
        }
        else {
          if (up_on && k == up_steps)
            for (int g = 0; g < kUpGroups; ++g) {  // join: from here on one launch per step again
              CU(cudaStreamWaitEvent(np, ev_grp[0][g], 0));
              CU(cudaStreamWaitEvent(np, ev_grp[1][g], 0));
              CU(cudaStreamWaitEvent(np, ev_up[g], 0));
            }
          CU(cudaStreamWaitEvent(np, ev_panel[buf], 0));
          const size_t np_cnt = (k + 1 < nt ? hp_off[k + 1] : np_off[k]) - np_off[k];
          const bool timed = profiling && np_cnt > 0;
          if (timed)
            CU(cudaEventRecord(pev[2 * k], np));
          r = update(np, A, ld, map_local, nbox_local, k, buf, np_off[k], np_cnt, true);
          if (r != 0)
            return r;
          if (timed)
            CU(cudaEventRecord(pev[2 * k + 1], np));
          np_timed[k] = timed;
          CU(cudaEventRecord(ev_np[buf], np));
        }
        if (pp.on && split)
          for (int g = 0; g < kUpGroups; ++g)  // "step consumed" below is raised on np: after the group launches
            CU(cudaStreamWaitEvent(np, ev_grp[buf][g], 0));
      }
      if (pp.on) {
        // "step k consumed here": after the look-ahead update, the bulk update and my forwarding copies of this step
        CU(cudaEventRecord(pp.ev_la, hp));
        CU(cudaStreamWaitEvent(np, pp.ev_la, 0));
        for (int q = 0; q < Q; ++q)
          if (q != pc)
            CU(cudaStreamWaitEvent(np, pp.ev_row_sent[buf][(size_t)q], 0));
        for (int r = 0; r < P; ++r)
          if (r != pr)
            CU(cudaStreamWaitEvent(np, pp.ev_col_sent[buf][(size_t)r], 0));
        for (int q = 0; q < Q; ++q)
          if (q != pc)
            PP(raise_flag(np, pp.row_peer[(size_t)q], f_done_row(pc), gstep + 1));
        for (int r = 0; r < P; ++r)
          if (r != pr)
            PP(raise_flag(np, pp.col_peer[(size_t)r], f_done_col(pr), gstep + 1));
      }
    }
    if (pp.on) {
      // the side streams end before the run does
      size_t i = 0;
      for (cudaStream_t cs : pp.cs_row)
        if (cs) {
          CU(cudaEventRecord(pp.ev_tail[i], cs));
          CU(cudaStreamWaitEvent(hp, pp.ev_tail[i++], 0));
        }
      for (cudaStream_t cs : pp.cs_col)
        if (cs) {
          CU(cudaEventRecord(pp.ev_tail[i], cs));
          CU(cudaStreamWaitEvent(hp, pp.ev_tail[i++], 0));
        }
      ++pp.epoch;
    }
    return 0;
  }

  int run(void* a_, int64_t lda, cudaStream_t user) override {
    T* a = static_cast<T*>(a_);
    if (profiling && pev.empty()) {
      pev.resize(2 * nt);
      for (cudaEvent_t& e : pev)
        CU(cudaEventCreate(&e));
    }
    if (profiling && hev.empty()) {
      hev.resize(6 * nt);
      for (cudaEvent_t& e : hev)
        CU(cudaEventCreate(&e));
      CU(cudaEventCreate(&ev_t0));
    }
    np_timed.assign(nt + 1, 0);
    if (profiling)
      CU(cudaEventRecord(ev_t0, user));
    CU(cudaEventRecord(ev_in, user));
    CU(cudaStreamWaitEvent(hp, ev_in, 0));
    CU(cudaStreamWaitEvent(np, ev_in, 0));
    // a plan may be reused from another caller stream before the previous run has drained (cached plans of the C++
    // API): each of the two streams also waits for the other one's end of the previous run (no-ops the first time)
    CU(cudaStreamWaitEvent(hp, ev_np_end, 0));
    CU(cudaStreamWaitEvent(np, ev_hp_end, 0));
    CU(cudaMemsetAsync(info_dev, 0, sizeof(int), hp));
    int r = 0;
    if (n > 0) {
      if (!upper) {
        const bool misaligned = fast && ((reinterpret_cast<uintptr_t>(a) & 15) || ((lda * vs) & 1));
        if (misaligned) {
          if (!tbuf) {
            ldt = round_up(mloc > 0 ? mloc : 1, 64);
            CU(cudaMalloc(&tbuf, ((size_t)ldt * (nloc > 0 ? nloc : 1) + 64) * sizeof(T)));
          }
          CU(lacpy<T>(hp, 'G', (int)mloc, (int)nloc, a, lda, tbuf, ldt));
          r = factor_lower(tbuf, ldt);
          if (r == 0) {
            CU(cudaStreamWaitEvent(hp, ev_np[0], 0));
            CU(cudaStreamWaitEvent(hp, ev_np[1], 0));
            CU(lacpy<T>(hp, 'G', (int)mloc, (int)nloc, tbuf, ldt, a, lda));
          }
        }
        else {
          r = factor_lower(a, lda);
        }
      }
      else {
        // local block of A is nloc x mloc in the caller's orientation; its conjugate transpose is mloc x nloc
        CU(transpose<T>(hp, (int)nloc, (int)mloc, a, lda, tbuf, ldt, 1, 'G'));
        r = factor_lower(tbuf, ldt);
        if (r == 0) {
          CU(cudaStreamWaitEvent(hp, ev_np[0], 0));
          CU(cudaStreamWaitEvent(hp, ev_np[1], 0));
          CU(transpose<T>(hp, (int)mloc, (int)nloc, tbuf, ldt, a, lda, 1, 'G'));
        }
      }
    }
    CU(cudaEventRecord(ev_hp_end, hp));
    CU(cudaEventRecord(ev_np_end, np));
    CU(cudaStreamWaitEvent(user, ev_hp_end, 0));
    CU(cudaStreamWaitEvent(user, ev_np_end, 0));
    return r;
  }

  int info(int* out) override {
    CU(cudaStreamSynchronize(hp));
    CU(cudaStreamSynchronize(np));
    int v = 0;
    CU(cudaMemcpy(&v, info_dev, sizeof(int), cudaMemcpyDeviceToHost));
    if (dist) {
      // ScaLAPACK convention: the same info on every rank. The kernel sets it only where the failing diagonal tile
      // lives: take the smallest non-zero value over the grid (first failing column). Collective, like the run.
      if (!info_red)
        CU(cudaMalloc(&info_red, sizeof(int)));
      const int mine = v != 0 ? v : 0x7fffffff;
      CU(cudaMemcpyAsync(info_red, &mine, sizeof(int), cudaMemcpyHostToDevice, hp));
      NC(ncclAllReduce(info_red, info_red, 1, ncclInt, ncclMin, worldc, hp));
      CU(cudaMemcpyAsync(&v, info_red, sizeof(int), cudaMemcpyDeviceToHost, hp));
      CU(cudaStreamSynchronize(hp));
      if (v == 0x7fffffff)
        v = 0;
    }
    *out = v;
    return 0;
  }
};

template <class T>
int create(dlafb_chol_plan_t* out, dlafb_grid* g, char uplo, int64_t n, int64_t nb, int sr, int sc) {
  Plan<T>* p = new Plan<T>();
  const int r = p->init(g, uplo, n, nb, sr, sc);
  if (r != 0) {
    delete p;
    return r;
  }
  *out = p;
  return 0;
}

}  // namespace

extern "C" {

int dlafb_cholesky_plan_create(dlafb_chol_plan_t* plan, dlafb_grid_t grid, char type, char uplo, int64_t n,
                               int64_t nb, int src_r, int src_c) {
  if (!plan || !grid || n < 0 || nb < 1 || n > 0x7fffff00ll)
    return -1;
  if (src_r < 0 || src_r >= grid->P || src_c < 0 || src_c >= grid->Q)
    return -1;
  switch (type) {
    case 's': return create<float>(plan, grid, uplo, n, nb, src_r, src_c);
    case 'd': return create<double>(plan, grid, uplo, n, nb, src_r, src_c);
    case 'c': return create<float2>(plan, grid, uplo, n, nb, src_r, src_c);
    case 'z': return create<double2>(plan, grid, uplo, n, nb, src_r, src_c);
  }
  return -1;
}

int dlafb_cholesky_plan_run_host(dlafb_chol_plan_t plan, void* host, int64_t ldh, void* dev, int64_t ldd,
                                 void* stream) {
  if (!plan || !dev)
    return 1;
  if (!host) {
    // a rank whose local block is empty may pass NULL (ScaLAPACK callers do): it still takes part in the exchange
    int64_t r = 0, c = 0;
    plan->local_size(&r, &c);
    if (r > 0 && c > 0)
      return 1;
  }
  return plan->run_host(host, ldh, dev, ldd, static_cast<cudaStream_t>(stream));
}

int dlafb_cholesky_plan_run(dlafb_chol_plan_t plan, void* a, int64_t lda, void* stream) {
  if (!plan)
    return -1;
  return plan->run(a, lda, static_cast<cudaStream_t>(stream));
}

int dlafb_cholesky_plan_info(dlafb_chol_plan_t plan, int* info) {
  if (!plan || !info)
    return -1;
  return plan->info(info);
}

int dlafb_cholesky_plan_set_profiling(dlafb_chol_plan_t plan, int on) {
  if (!plan)
    return -1;
  plan->profiling = on != 0;
  return 0;
}

int dlafb_cholesky_plan_residual(dlafb_chol_plan_t plan, void* a0, int64_t lda, const void* l, int64_t ldl, void* stream,
                                 double* max_abs_a, double* max_abs_diff) {
  if (!plan || !a0 || !l || !max_abs_a || !max_abs_diff)
    return 1;
  double out[2] = {0.0, 0.0};
  const int r = plan->residual(a0, lda, l, ldl, static_cast<cudaStream_t>(stream), out);
  *max_abs_a = out[0];
  *max_abs_diff = out[1];
  return r;
}

int dlafb_cholesky_plan_exchange_mode(dlafb_chol_plan_t plan) {
  return plan ? plan->exchange_mode() : -1;
}

int dlafb_cholesky_plan_timeline(dlafb_chol_plan_t plan, double* out, long long cap) {
  if (!plan || !out)
    return 1;
  return plan->timeline(out, cap);
}

int dlafb_cholesky_plan_profile(dlafb_chol_plan_t plan, double* update_ms, double* update_flops, int* launches) {
  if (!plan || !update_ms || !update_flops || !launches)
    return -1;
  return plan->profile(update_ms, update_flops, launches);
}

int dlafb_cholesky_plan_destroy(dlafb_chol_plan_t plan) {
  delete plan;
  return 0;
}

// ---- plans kept alive between calls (the C++ front end: plan creation is a collective with allocations, IPC
//      mappings and list uploads; destruction synchronises the device). At most kPlanCacheLimit idle plans stay;
//      a plan that is in use (acquired, not yet released) is never destroyed. Every rank of a grid makes the same
//      sequence of calls, so all ranks keep and drop the same plans. ----
namespace {
struct CachedPlan {
  dlafb_grid_t grid;
  char type, uplo;
  int64_t n, nb;
  int sr, sc;
  dlafb_chol_plan_t plan;
  int refs;
  unsigned long long last_use;
};
std::vector<CachedPlan>& plan_cache() {
  static std::vector<CachedPlan>* v = new std::vector<CachedPlan>();  // never destroyed: no CUDA calls at exit
  return *v;
}
unsigned long long g_cache_clock = 0;
constexpr size_t kPlanCacheLimit = 4;
void cache_trim() {
  auto& v = plan_cache();
  while (v.size() > kPlanCacheLimit) {
    size_t lru = v.size();
    for (size_t i = 0; i < v.size(); ++i)
      if (v[i].refs == 0 && (lru == v.size() || v[i].last_use < v[lru].last_use))
        lru = i;
    if (lru == v.size())
      return;  // everything is in use
    delete v[lru].plan;
    v.erase(v.begin() + (long)lru);
  }
}
}  // namespace

int dlafb_cholesky_plan_acquire(dlafb_chol_plan_t* plan, dlafb_grid_t grid, char type, char uplo, int64_t n, int64_t nb,
                                int src_r, int src_c) {
  if (!plan || !grid)
    return -1;
  uplo = (uplo == 'u' || uplo == 'U') ? 'U' : 'L';
  auto& v = plan_cache();
  for (CachedPlan& c : v)
    if (c.grid == grid && c.type == type && c.uplo == uplo && c.n == n && c.nb == nb && c.sr == src_r &&
        c.sc == src_c) {
      ++c.refs;
      c.last_use = ++g_cache_clock;
      *plan = c.plan;
      return 0;
    }
  dlafb_chol_plan_t p = nullptr;
  const int e = dlafb_cholesky_plan_create(&p, grid, type, uplo, n, nb, src_r, src_c);
  if (e != 0)
    return e;
  v.push_back(CachedPlan{grid, type, uplo, n, nb, src_r, src_c, p, 1, ++g_cache_clock});
  cache_trim();
  *plan = p;
  return 0;
}

int dlafb_cholesky_plan_release(dlafb_chol_plan_t plan) {
  for (CachedPlan& c : plan_cache())
    if (c.plan == plan) {
      if (c.refs > 0)
        --c.refs;
      cache_trim();
      return 0;
    }
  return -1;
}

int dlafb_cholesky_plan_cache_drop(dlafb_grid_t grid) {
  auto& v = plan_cache();
  for (size_t i = 0; i < v.size();) {
    if (grid == nullptr || v[i].grid == grid) {
      delete v[i].plan;  // the grid goes: its plans cannot run any more
      v.erase(v.begin() + (long)i);
    }
    else
      ++i;
  }
  return 0;
}

int dlafb_cholesky_plan_cache_size(void) { return (int)plan_cache().size(); }

int dlafb_cholesky_plan_local_size(dlafb_chol_plan_t plan, int64_t* rows, int64_t* cols) {
  if (!plan || !rows || !cols)
    return -1;
  plan->local_size(rows, cols);
  return 0;
}

}  // extern "C"
