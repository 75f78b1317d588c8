// TEST INFRASTRUCTURE ONLY. CPU execution of the triangular solver for boxes without a GPU: the product's own
// algorithm template (dla-future_b200/src/triangular_solver.h) and its own kernel bodies
// (dla-future_b200/csrc/solver_kernels.cuh, run once per thread of the launch geometry) over an executor whose
// remaining operations (contraction, transpose, copy) are restated as plain loops with the device kernels'
// documented semantics. Built host-only by nvcc (no __global__ code here); never linked into the product.
//
// What this pins without a GPU: the reduction of the 24 side/uplo/op/diag cases to the two canonical solves, every
// pointer offset and leading dimension handed to the kernels, the tile-inverse recurrence (lower/upper/unit), the
// workspace sizes (workspaces are NaN-filled, so reading anything the algorithm did not write poisons the result),
// that no contraction is asked to run with C aliasing an operand, and - for the two-stream look-ahead variant - that
// the events order every pair of operations with overlapping operands (replay under adversarial interleavings plus
// a happens-before check).
#include <cmath>
#include <cstring>
#include <functional>
#include <limits>
#include <utility>
#include <vector>

#include <condition_variable>
#include <mutex>
#include <thread>

#include "gen_to_std.h"
#include "op_trace.h"
#include "redistribute.h"
#include "gen_to_std_dist.h"
#include "triangular_solver.h"
#include "triangular_solver_dist.h"

namespace {
using namespace dlafb;

template <class T>
T nan_value() {
  using R = typename Tr<T>::Real;
  const R q = std::numeric_limits<R>::quiet_NaN();
  if constexpr (Tr<T>::is_complex)
    return Tr<T>::make(q, q);
  else
    return q;
}

template <class T>
bool is_nan_value(T v) {
  if constexpr (Tr<T>::is_complex)
    return v.x != v.x && v.y != v.y;
  else
    return v != v;
}

// A workspace part as the device plan would allocate it (exactly `n` elements) followed by a guard zone: anything the
// algorithm writes past the end of a part shows up as a damaged guard (error 9300) instead of silent corruption.
template <class T>
struct Guarded {
  static constexpr size_t kGuard = 64;
  std::vector<T> v;
  size_t n;
  explicit Guarded(size_t elems) : v(elems + kGuard, nan_value<T>()), n(elems) {}
  T* data() { return v.data(); }
  bool intact() const {
    for (size_t i = n; i < v.size(); ++i)
      if (!is_nan_value<T>(v[i]))
        return false;
    return true;
  }
};

// do the element sets {p + i + j*ld : i < rows, j < cols} of two operands intersect?
template <class T>
bool overlap(const T* a, long long lda, int ra, int ca, const T* b, long long ldb, int rb, int cb) {
  for (int j = 0; j < ca; ++j) {
    const T* a0 = a + j * lda;
    const T* a1 = a0 + ra;
    for (int l = 0; l < cb; ++l) {
      const T* b0 = b + l * ldb;
      const T* b1 = b0 + rb;
      if (a0 < b1 && b0 < a1)
        return true;
    }
  }
  return false;
}

// One operand: the elements {p + (i + j*ld) * es : i < rows, j < cols}, in bytes.
struct Rect {
  const char* p;
  long long ld;
  int rows, cols;
  size_t es;
};

inline bool rects_overlap(const Rect& a, const Rect& b) {
  if (a.rows <= 0 || a.cols <= 0 || b.rows <= 0 || b.cols <= 0)
    return false;
  const char* a_end = a.p + ((long long)(a.cols - 1) * a.ld + a.rows) * (long long)a.es;
  const char* b_end = b.p + ((long long)(b.cols - 1) * b.ld + b.rows) * (long long)b.es;
  if (a_end <= b.p || b_end <= a.p)
    return false;
  for (int j = 0; j < a.cols; ++j) {
    const char* a0 = a.p + (long long)j * a.ld * (long long)a.es;
    const char* a1 = a0 + (long long)a.rows * (long long)a.es;
    for (int l = 0; l < b.cols; ++l) {
      const char* b0 = b.p + (long long)l * b.ld * (long long)b.es;
      const char* b1 = b0 + (long long)b.rows * (long long)b.es;
      if (a0 < b1 && b0 < a1)
        return true;
    }
  }
  return false;
}

// Executor on the host. Immediate mode: every operation runs when it is issued (one lane). Deferred mode: the
// operations are queued per lane, signal / wait become markers, and drain() replays the two queues preferring one
// lane whenever the events allow it (the other lane then runs as late as possible) while tracking, per operation,
// what it is ordered after on the other lane. races() then looks at every pair of operations the events leave
// unordered and reports overlapping write/write or read/write operands.
template <class T>
struct HostOps {
  using R = typename Tr<T>::Real;
  struct Op {
    int kind = 0;  // 0 operation, 1 signal, 2 wait
    int ev = 0, gen = 0;
    std::function<int()> fn;
    std::vector<Rect> reads, writes;
    int own = 0, known = 0;  // filled by drain(): index on its lane, highest index of the other lane ordered before it
  };
  bool deferred = false;
  int cur = 0;
  int hazards = 0;
  std::vector<Op> q[2];
  int sig_enq[kTriSolveEvents] = {};

  template <class P>
  static Rect rect(const P* p, long long ld, int rows, int cols) {
    return Rect{reinterpret_cast<const char*>(p), ld, rows, cols, sizeof(P)};
  }
  int issue(Op&& op) {
    if (!deferred)
      return op.fn();
    q[cur].push_back(std::move(op));
    return 0;
  }
  void lane(int l) { cur = l; }
  int signal(int ev) {
    if (!deferred)
      return 0;
    Op op;
    op.kind = 1;
    op.ev = ev;
    op.gen = ++sig_enq[ev];
    q[cur].push_back(std::move(op));
    return 0;
  }
  int wait(int ev) {
    if (!deferred)
      return 0;
    Op op;
    op.kind = 2;
    op.ev = ev;
    op.gen = sig_enq[ev];
    q[cur].push_back(std::move(op));
    return 0;
  }

  // Replays the queues; `prefer` = the lane that runs whenever it can. Returns the first error (9100: deadlock).
  int drain(int prefer) {
    size_t head[2] = {0, 0};
    int own[2] = {0, 0}, known[2] = {0, 0};
    int sig_done[kTriSolveEvents] = {};
    std::vector<std::pair<int, int>> sig_at[kTriSolveEvents];  // per generation: (lane, ops executed on it so far)
    for (auto& v : sig_at)
      v.push_back({0, 0});
    int first_error = 0;
    while (head[0] < q[0].size() || head[1] < q[1].size()) {
      bool progressed = false;
      for (int pass = 0; pass < 2 && !progressed; ++pass) {
        const int l = pass == 0 ? prefer : 1 - prefer;
        if (head[l] >= q[l].size())
          continue;
        Op& op = q[l][head[l]];
        if (op.kind == 2 && sig_done[op.ev] < op.gen)
          continue;  // blocked
        if (op.kind == 0) {
          op.own = ++own[l];
          op.known = known[l];
          const int e = op.fn();
          if (e != 0 && first_error == 0)
            first_error = e;
        }
        else if (op.kind == 1) {
          sig_done[op.ev] = op.gen;
          sig_at[op.ev].push_back({l, own[l]});
        }
        else if (op.gen > 0) {
          const auto& at = sig_at[op.ev][op.gen];
          if (at.first != l && at.second > known[l])
            known[l] = at.second;
        }
        ++head[l];
        progressed = true;
      }
      if (!progressed)
        return 9100;
    }
    return first_error;
  }

  int races() const {
    int n = 0;
    for (const Op& a : q[0]) {
      if (a.kind != 0)
        continue;
      for (const Op& b : q[1]) {
        if (b.kind != 0)
          continue;
        if (b.known >= a.own || a.known >= b.own)
          continue;  // ordered by an event
        bool clash = false;
        for (const Rect& wa : a.writes) {
          for (const Rect& wb : b.writes)
            clash = clash || rects_overlap(wa, wb);
          for (const Rect& rb : b.reads)
            clash = clash || rects_overlap(wa, rb);
        }
        for (const Rect& wb : b.writes)
          for (const Rect& ra : a.reads)
            clash = clash || rects_overlap(wb, ra);
        n += clash;
      }
    }
    return n;
  }

  // lower != 0: only the elements with row >= col are touched (gemm_dmma.cu / gemm_simt.cu RECT mode)
  int gemm_nt(int M, int N, int K, R alpha, const T* A, long long lda, const T* B, long long ldb, R beta, T* C,
              long long ldc, int lower = 0) {
    if (M <= 0 || N <= 0)
      return 0;
    // same line format as the device executors (src/op_trace.h) so that the two logs can be compared
    DLAFB_TRACE_OP("gemm_nt M=%d N=%d K=%d alpha=%g beta=%g lower=%d A=%p lda=%lld B=%p ldb=%lld C=%p ldc=%lld", M, N, K,
                   (double)alpha, (double)beta, lower, (const void*)A, lda, (const void*)B, ldb, (void*)C, ldc);
    Op op;
    op.reads = {rect(A, lda, M, K), rect(B, ldb, N, K)};
    if (beta != R(0))
      op.reads.push_back(rect(C, ldc, M, N));
    op.writes = {rect(C, ldc, M, N)};
    op.fn = [=]() -> int {
      if (overlap<T>(C, ldc, M, N, A, lda, M, K) || overlap<T>(C, ldc, M, N, B, ldb, N, K)) {
        ++this->hazards;
        return 9001;
      }
      for (int j = 0; j < N; ++j)
        for (int i = lower ? j : 0; i < M; ++i) {
          T acc = Tr<T>::zero();
          for (int k = 0; k < K; ++k)
            acc = Tr<T>::fma(A[i + k * lda], Tr<T>::conj(B[j + k * ldb]), acc);
          T v = Tr<T>::scale(acc, alpha);
          if (beta != R(0))
            v = Tr<T>::add(v, Tr<T>::scale(C[i + j * ldc], beta));
          if (Tr<T>::is_complex && lower && i == j)
            v = Tr<T>::from_real(Tr<T>::real(v));  // the kernels' lower mode keeps the diagonal real (herk semantics)
          C[i + j * ldc] = v;
        }
      return 0;
    };
    return issue(std::move(op));
  }
  // aux.cu transpose_kernel: out(j, i) = op(in(i, j)); tri 'L' writes only out rows >= out cols, 'U' rows <= cols
  int transpose(int m, int n, const T* in, long long ldin, T* out, long long ldout, int conj, char tri) {
    if (m <= 0 || n <= 0)
      return 0;
    DLAFB_TRACE_OP("transpose m=%d n=%d conj=%d tri=%c in=%p ldin=%lld out=%p ldout=%lld", m, n, conj, tri, (const void*)in,
                   ldin, (void*)out, ldout);
    Op op;
    op.reads = {rect(in, ldin, m, n)};
    op.writes = {rect(out, ldout, n, m)};
    op.fn = [=]() -> int {
      if (overlap<T>(out, ldout, n, m, in, ldin, m, n)) {
        ++this->hazards;
        return 9002;
      }
      for (int i = 0; i < m; ++i)
        for (int j = 0; j < n; ++j) {
          if ((tri == 'L' && j < i) || (tri == 'U' && j > i))
            continue;
          const T v = in[i + j * ldin];
          out[j + i * ldout] = conj ? Tr<T>::conj(v) : v;
        }
      return 0;
    };
    return issue(std::move(op));
  }
  int lacpy(int m, int n, const T* a, long long lda, T* b, long long ldb) {
    if (m <= 0 || n <= 0)
      return 0;
    Op op;
    op.reads = {rect(a, lda, m, n)};
    op.writes = {rect(b, ldb, m, n)};
    op.fn = [=]() -> int {
      for (int j = 0; j < n; ++j)
        for (int i = 0; i < m; ++i)
          b[i + j * ldb] = a[i + j * lda];
      return 0;
    };
    return issue(std::move(op));
  }
  // aux.cu lacpy_kernel with uplo: 'L' copies rows >= cols, 'U' rows <= cols
  int lacpy(char uplo, int m, int n, const T* a, long long lda, T* b, long long ldb) {
    if (m <= 0 || n <= 0)
      return 0;
    DLAFB_TRACE_OP("lacpy %c m=%d n=%d a=%p lda=%lld b=%p ldb=%lld", uplo, m, n, (const void*)a, lda, (void*)b, ldb);
    Op op;
    op.reads = {rect(a, lda, m, n)};
    op.writes = {rect(b, ldb, m, n)};
    op.fn = [=]() -> int {
      if (overlap<T>(b, ldb, m, n, a, lda, m, n)) {
        ++this->hazards;
        return 9003;
      }
      for (int j = 0; j < n; ++j)
        for (int i = 0; i < m; ++i) {
          if ((uplo == 'L' && i < j) || (uplo == 'U' && i > j))
            continue;
          b[i + j * ldb] = a[i + j * lda];
        }
      return 0;
    };
    return issue(std::move(op));
  }
  int tri_inverse(int upper, int unit, int n, int nb, const T* t, long long ldt, T* w) {
    DLAFB_TRACE_OP("tri_inverse upper=%d unit=%d n=%d nb=%d t=%p ldt=%lld w=%p", upper, unit, n, nb, (const void*)t, ldt,
                   (void*)w);
    Op op;
    op.reads = {rect(t, ldt, n, n)};
    op.writes = {rect(w, 1, tri_inverse_blocks(n, nb) * nb * nb, 1)};
    op.fn = [=]() -> int {
      TriInvParams p{n, nb, ldt, upper, unit};
      const int blocks = tri_inverse_blocks(n, nb);
      for (int b = 0; b < blocks; ++b)
        for (int th = 0; th < kTriInvThreads; ++th)
          tri_inverse_body<T>(p, t, w, b, th, kTriInvThreads);
      return 0;
    };
    return issue(std::move(op));
  }
  int scale(int m, int n, T* b, long long ldb, T alpha, int conj) {
    if (m <= 0 || n <= 0)
      return 0;
    DLAFB_TRACE_OP("scale m=%d n=%d conj=%d b=%p ldb=%lld", m, n, conj, (void*)b, ldb);
    Op op;
    op.reads = {rect(b, ldb, m, n)};
    op.writes = {rect(b, ldb, m, n)};
    op.fn = [=]() -> int {
      for (int by = 0; by < scale_blocks_y(n); ++by)
        for (int bx = 0; bx < scale_blocks_x(m); ++bx)
          for (int th = 0; th < kScaleRows; ++th)
            scale_body<T>(m, n, b, ldb, alpha, conj, bx, by, th);
      return 0;
    };
    return issue(std::move(op));
  }
};

// lookahead = 0: one lane, immediate. lookahead = 1: the two-lane variant, replayed with lane `prefer` running
// whenever it can; *races = number of unordered operation pairs with overlapping operands.
template <class T>
int solve(char side, char uplo, char op, char diag, int m, int n, int nb, const void* alpha, const void* a,
          long long lda, void* b, long long ldb, int lookahead = 0, int prefer = 0, int* races = nullptr) {
  TriSolveDesc d{side, uplo, op, diag, m, n, nb};
  std::vector<T> tcopy(tri_solve_tcopy_elems(d), nan_value<T>());
  std::vector<T> winv(tri_solve_winv_elems(d), nan_value<T>());
  std::vector<T> panel(tri_solve_panel_elems(d) * (lookahead ? 2 : 1), nan_value<T>());
  TriSolveWork<T> w{tcopy.data(), tri_solve_ldt(d), winv.data(), panel.data(), tri_solve_ldp(d)};
  HostOps<T> ops;
  if (lookahead) {
    w.lookahead = 1;
    w.panel_stride = tri_solve_panel_elems(d);
    ops.deferred = true;
  }
  int e = triangular_solve<T>(ops, d, *static_cast<const T*>(alpha), static_cast<const T*>(a), lda, static_cast<T*>(b),
                              ldb, w);
  if (lookahead) {
    if (e == 0)
      e = ops.drain(prefer);
    if (races)
      *races = ops.races();
  }
  return e;
}

template <class T>
int inverse(int upper, int unit, int n, int nb, const void* t, long long ldt, void* w) {
  HostOps<T> ops;
  return ops.tri_inverse(upper, unit, n, nb, static_cast<const T*>(t), ldt, static_cast<T*>(w));
}

// Generalized-to-standard reduction: the product's algorithm template over the host executor, NaN-filled workspaces.
// lookahead = 0: one lane, immediate. lookahead = 1: the two-lane variant, replayed with lane `prefer` running
// whenever it can; *races = number of unordered operation pairs with overlapping operands.
template <class T>
int gen_to_std_emu(char uplo, int n, int nb, void* a, long long lda, const void* f, long long ldf, int lookahead = 0,
                   int prefer = 0, int* races = nullptr) {
  const size_t copies = lookahead ? 2 : 1;
  Guarded<T> winv(g2s_winv_elems(n, nb)), akk(g2s_tile_elems(nb)), tmp(g2s_tile_elems(nb)),
      pa(copies * g2s_panel_elems(n, nb)), pb(copies * g2s_panel_elems(n, nb)), ac(g2s_copy_elems(n, uplo)),
      lc(g2s_copy_elems(n, uplo));
  GenToStdWork<T> w{winv.data(), akk.data(), tmp.data(), g2s_even(nb), pa.data(), pb.data(), g2s_ldp(n),
                    ac.data(), lc.data(), g2s_ldp(n)};
  HostOps<T> ops;
  if (lookahead) {
    w.lookahead = 1;
    w.panel_stride = g2s_panel_elems(n, nb);
    ops.deferred = true;
  }
  int e = gen_to_std<T>(ops, uplo, n, nb, static_cast<T*>(a), lda, static_cast<const T*>(f), ldf, w);
  if (lookahead) {
    if (e == 0)
      e = ops.drain(prefer);
    if (races)
      *races = ops.races();
  }
  if (e == 0 && !(winv.intact() && akk.intact() && tmp.intact() && pa.intact() && pb.intact() && ac.intact() && lc.intact()))
    return 9300;
  return e;
}

// ---- simulated grid: one host thread per rank, collectives through shared memory -------------------------------
// One communicator (a grid row or a grid column): a reusable barrier and the buffer pointers its members published.
struct SimComm {
  std::mutex mu;
  std::condition_variable cv;
  int size = 1, arrived = 0;
  unsigned long long gen = 0;
  std::vector<void*> ptr;
  std::vector<size_t> cnt;
  int mismatches = 0;
  void barrier() {
    std::unique_lock<std::mutex> lk(mu);
    const unsigned long long g = gen;
    if (++arrived == size) {
      arrived = 0;
      ++gen;
      cv.notify_all();
    }
    else {
      cv.wait(lk, [&] { return gen != g; });
    }
  }
};

// What the distributed templates need besides HostOps and the collectives: the strided tile inverse and zero-fill.
template <class T>
struct DistLocalOps : HostOps<T> {
  int tri_inverse_local(int upper, int unit, int n, int nb, const T* t, long long ldt, T* w, int g0, int gstep, int row0,
                        int row_step, int col0, int col_step, int count) {
    TriInvParams p{n, nb, ldt, upper, unit, g0, gstep, row0, row_step, col0, col_step, count};
    DLAFB_TRACE_OP("tri_inverse upper=%d unit=%d n=%d nb=%d t=%p ldt=%lld w=%p g0=%d gstep=%d count=%d", upper, unit, n, nb,
                   (const void*)t, ldt, (void*)w, g0, gstep, count);
    for (int b = 0; b < count; ++b)
      for (int th = 0; th < kTriInvThreads; ++th)
        tri_inverse_body<T>(p, t, w, b, th, kTriInvThreads);
    return 0;
  }
  int zero(int m, int n, T* a, long long lda) {
    DLAFB_TRACE_OP("zero m=%d n=%d a=%p lda=%lld", m, n, (void*)a, lda);
    for (int j = 0; j < n; ++j)
      for (int i = 0; i < m; ++i)
        a[i + j * lda] = Tr<T>::zero();
    return 0;
  }
  int group_begin() { return 0; }  // the emulated collectives complete one by one; grouping changes nothing here
  int group_end() { return 0; }
};

template <class T>
struct DistHostOps : DistLocalOps<T> {
  SimComm* row = nullptr;  // members ordered by grid column
  SimComm* col = nullptr;  // members ordered by grid row
  int me_row = 0, me_col = 0;  // my index inside them
  // every member must arrive with the same count: a mismatch is what NCCL would turn into a hang or corruption
  static int bcast(SimComm* c, int me, int root, T* p, size_t count) {
    c->ptr[me] = p;
    c->cnt[me] = count;
    c->barrier();
    int bad = 0;
    for (int i = 0; i < c->size; ++i)
      bad += c->cnt[i] != count;
    if (me != root && !bad)
      std::memcpy(p, c->ptr[root], count * sizeof(T));
    c->barrier();
    return bad ? 9200 : 0;
  }
  static int reduce(SimComm* c, int me, int root, T* p, size_t count) {
    c->ptr[me] = p;
    c->cnt[me] = count;
    c->barrier();
    int bad = 0;
    for (int i = 0; i < c->size; ++i)
      bad += c->cnt[i] != count;
    if (me == root && !bad)
      for (int i = 0; i < c->size; ++i)
        if (i != root) {
          const T* q = static_cast<const T*>(c->ptr[i]);
          for (size_t x = 0; x < count; ++x)
            p[x] = Tr<T>::add(p[x], q[x]);
        }
    c->barrier();
    // NCCL leaves the non-root buffers undefined: poison them so that nothing can depend on their content
    if (me != root && !bad)
      for (size_t x = 0; x < count; ++x)
        p[x] = nan_value<T>();
    return bad ? 9201 : 0;
  }
  int bcast_row(int root, T* p, size_t count) {
    DLAFB_TRACE_OP("bcast_row root=%d count=%zu p=%p", root, count, (void*)p);
    return bcast(row, me_row, root, p, count);
  }
  int bcast_col(int root, T* p, size_t count) {
    DLAFB_TRACE_OP("bcast_col root=%d count=%zu p=%p", root, count, (void*)p);
    return bcast(col, me_col, root, p, count);
  }
  int reduce_row(int root, T* p, size_t count) {
    DLAFB_TRACE_OP("reduce_row root=%d count=%zu p=%p", root, count, (void*)p);
    return reduce(row, me_row, root, p, count);
  }
  int reduce_col(int root, T* p, size_t count) {
    DLAFB_TRACE_OP("reduce_col root=%d count=%zu p=%p", root, count, (void*)p);
    return reduce(col, me_col, root, p, count);
  }
};

// All the ranks of a P x Q grid in one process; a[r], b[r] = local blocks of rank r = pr * Q + pc.
template <class T>
int solve_dist(int P, int Q, int src_r, int src_c, char side, char uplo, char op, char diag, int m, int n, int mb, int nb,
               const void* alpha, void* const* a, const long long* lda, void* const* b, const long long* ldb) {
  std::vector<SimComm> rows(P), cols(Q);
  for (auto& c : rows) {
    c.size = Q;
    c.ptr.assign(Q, nullptr);
    c.cnt.assign(Q, 0);
  }
  for (auto& c : cols) {
    c.size = P;
    c.ptr.assign(P, nullptr);
    c.cnt.assign(P, 0);
  }
  std::vector<int> err(P * Q, 0);
  std::vector<std::thread> th;
  for (int pr = 0; pr < P; ++pr)
    for (int pc = 0; pc < Q; ++pc)
      th.emplace_back([&, pr, pc] {
        const int r = pr * Q + pc;
        TriDistDesc d{side, uplo, op, diag, m, n, mb, nb, P, Q, pr, pc, src_r, src_c};
        const TriDistSizes sz = tri_dist_sizes(d);
        Guarded<T> winv(sz.winv), pan(sz.pan), pan2(sz.pan2), xb(sz.xb), wkh(sz.wkh), bt(sz.bt);
        TriDistWork<T> w{winv.data(), pan.data(), pan2.data(), xb.data(), wkh.data(), bt.data()};
        DistHostOps<T> ops;
        ops.row = &rows[pr];
        ops.col = &cols[pc];
        ops.me_row = pc;
        ops.me_col = pr;
        err[r] = triangular_solve_dist<T>(ops, d, *static_cast<const T*>(alpha), static_cast<const T*>(a[r]), lda[r],
                                          static_cast<T*>(b[r]), ldb[r], w);
        if (err[r] == 0 && ops.hazards)
          err[r] = 9001;
        if (err[r] == 0 && !(winv.intact() && pan.intact() && pan2.intact() && xb.intact() && wkh.intact() && bt.intact()))
          err[r] = 9300;
      });
  for (auto& t : th)
    t.join();
  for (int e : err)
    if (e)
      return e;
  return 0;
}

// Distributed generalized-to-standard reduction for all the ranks of a simulated grid.
template <class T>
int gen_to_std_dist_emu(int P, int Q, int src_r, int src_c, char uplo, int n, int nb, void* const* a,
                        const long long* lda, void* const* f, const long long* ldf) {
  std::vector<SimComm> rows(P), cols(Q);
  for (auto& c : rows) {
    c.size = Q;
    c.ptr.assign(Q, nullptr);
    c.cnt.assign(Q, 0);
  }
  for (auto& c : cols) {
    c.size = P;
    c.ptr.assign(P, nullptr);
    c.cnt.assign(P, 0);
  }
  std::vector<int> err(P * Q, 0);
  std::vector<std::thread> th;
  for (int pr = 0; pr < P; ++pr)
    for (int pc = 0; pc < Q; ++pc)
      th.emplace_back([&, pr, pc] {
        const int r = pr * Q + pc;
        G2sDistDesc d{uplo, n, nb, P, Q, pr, pc, src_r, src_c};
        const G2sDistSizes sz = g2s_dist_sizes(d);
        Guarded<T> winv(sz.winv), dk(sz.dk), tmp(sz.tmp), xr(sz.xr), yt(sz.ytiles), ac(sz.acopy), lc(sz.lcopy);
        G2sDistWork<T> w{winv.data(), dk.data(), tmp.data(), xr.data(), yt.data(), ac.data(), lc.data()};
        DistHostOps<T> ops;
        ops.row = &rows[pr];
        ops.col = &cols[pc];
        ops.me_row = pc;
        ops.me_col = pr;
        err[r] = gen_to_std_dist<T>(ops, d, static_cast<T*>(a[r]), lda[r], static_cast<const T*>(f[r]), ldf[r], w);
        if (err[r] == 0 && ops.hazards)
          err[r] = 9001;
        if (err[r] == 0 && !(winv.intact() && dk.intact() && tmp.intact() && xr.intact() && yt.intact() && ac.intact() &&
                             lc.intact()))
          err[r] = 9300;
      });
  for (auto& t : th)
    t.join();
  for (int e : err)
    if (e)
      return e;
  return 0;
}

// ---- redistribution for all the ranks of a simulated grid: the product's list builder and kernel body, the
// all-to-all through shared memory (every rank reads its segments straight out of the senders' buffers)
struct SimWorld {
  SimComm comm;
  std::vector<const void*> sendbuf;
  std::vector<const RedistLists*> lists;
};

template <class T>
struct RedistHostOps {
  const RedistLists* L;
  SimWorld* world;
  int me;
  int copy_rects(int which, const T* src, T* dst) {
    const std::vector<CopyRect>& v = which == 0 ? L->pack : which == 1 ? L->self : L->unpack;
    for (int b = 0; b < (int)v.size(); ++b)
      for (int th = 0; th < kCopyRectThreads; ++th)
        copy_rects_body<T>(v.data(), src, dst, b, th, kCopyRectThreads);
    return 0;
  }
  int alltoallv(const T* sendbuf, T* recvbuf) {
    world->sendbuf[me] = sendbuf;
    world->lists[me] = L;
    world->comm.barrier();
    int bad = 0;
    for (int r = 0; r < world->comm.size; ++r) {
      if (r == me)
        continue;
      const RedistLists* theirs = world->lists[r];
      if (theirs->send_cnt[me] != L->recv_cnt[r]) {  // NCCL would hang or corrupt on a size mismatch
        ++bad;
        continue;
      }
      if (L->recv_cnt[r] > 0)
        std::memcpy(recvbuf + L->recv_off[r], static_cast<const T*>(world->sendbuf[r]) + theirs->send_off[me],
                    (size_t)L->recv_cnt[r] * sizeof(T));
    }
    world->comm.barrier();
    return bad ? 9400 : 0;
  }
};

template <class T>
int redistribute_emu(int P, int Q, char ordering, long long m, long long n, int mb_s, int nb_s, int sr_s, int sc_s, int mb_d,
                     int nb_d, int sr_d, int sc_d, void* const* src, const long long* ld_src, void* const* dst,
                     const long long* ld_dst) {
  SimWorld world;
  world.comm.size = P * Q;
  world.sendbuf.assign(P * Q, nullptr);
  world.lists.assign(P * Q, nullptr);
  std::vector<int> err(P * Q, 0);
  std::vector<std::thread> th;
  for (int pr = 0; pr < P; ++pr)
    for (int pc = 0; pc < Q; ++pc)
      th.emplace_back([&, pr, pc] {
        RedistDesc d{m, n, mb_s, nb_s, sr_s, sc_s, mb_d, nb_d, sr_d, sc_d, P, Q, pr, pc, ordering};
        const int r = redist_linear_rank(d, pr, pc);
        if (!redist_valid(d)) {
          err[r] = 9401;
          return;
        }
        const RedistLists L = redist_build(d, ld_src[r], ld_dst[r]);
        Guarded<T> sb((size_t)L.send_total), rb((size_t)L.recv_total);
        RedistHostOps<T> ops{&L, &world, r};
        err[r] = redistribute<T>(ops, L, static_cast<const T*>(src[r]), static_cast<T*>(dst[r]), sb.data(), rb.data());
        if (err[r] == 0 && !(sb.intact() && rb.intact()))
          err[r] = 9300;
      });
  for (auto& t : th)
    t.join();
  for (int e : err)
    if (e)
      return e;
  return 0;
}

// ---- one rank of a REAL process group: the collectives are callbacks into the test harness (torch.distributed / gloo)
extern "C" {
typedef int (*emu_bcast_fn)(int axis, int root, void* ptr, long long bytes);  // axis 0: my grid row, 1: my grid column
typedef int (*emu_reduce_fn)(int axis, int root, void* ptr, long long count, int is_double);  // sum of real numbers
typedef int (*emu_alltoallv_fn)(const void* sendbuf, const long long* send_off, const long long* send_cnt, void* recvbuf,
                                const long long* recv_off, const long long* recv_cnt, int elem_bytes, int world);
}

template <class T>
struct CallbackOps : DistLocalOps<T> {
  emu_bcast_fn bc = nullptr;
  emu_reduce_fn rd = nullptr;
  static constexpr int kRealPerElem = Tr<T>::is_complex ? 2 : 1;
  static constexpr int kIsDouble = sizeof(typename Tr<T>::Real) == 8;
  int bcast_row(int root, T* p, size_t count) { return bc(0, root, p, (long long)(count * sizeof(T))); }
  int bcast_col(int root, T* p, size_t count) { return bc(1, root, p, (long long)(count * sizeof(T))); }
  int reduce_row(int root, T* p, size_t count) { return rd(0, root, p, (long long)count * kRealPerElem, kIsDouble); }
  int reduce_col(int root, T* p, size_t count) { return rd(1, root, p, (long long)count * kRealPerElem, kIsDouble); }
};

template <class T>
int solve_dist_rank(int P, int Q, int pr, int pc, int src_r, int src_c, char side, char uplo, char op, char diag, int m,
                    int n, int mb, int nb, const void* alpha, const void* a, long long lda, void* b, long long ldb,
                    emu_bcast_fn bc, emu_reduce_fn rd) {
  TriDistDesc d{side, uplo, op, diag, m, n, mb, nb, P, Q, pr, pc, src_r, src_c};
  const TriDistSizes sz = tri_dist_sizes(d);
  Guarded<T> winv(sz.winv), pan(sz.pan), pan2(sz.pan2), xb(sz.xb), wkh(sz.wkh), bt(sz.bt);
  TriDistWork<T> w{winv.data(), pan.data(), pan2.data(), xb.data(), wkh.data(), bt.data()};
  CallbackOps<T> ops;
  ops.bc = bc;
  ops.rd = rd;
  int e = triangular_solve_dist<T>(ops, d, *static_cast<const T*>(alpha), static_cast<const T*>(a), lda,
                                   static_cast<T*>(b), ldb, w);
  if (e == 0 && ops.hazards)
    e = 9001;
  if (e == 0 && !(winv.intact() && pan.intact() && pan2.intact() && xb.intact() && wkh.intact() && bt.intact()))
    e = 9300;
  return e;
}

template <class T>
int gen_to_std_dist_rank(int P, int Q, int pr, int pc, int src_r, int src_c, char uplo, int n, int nb, void* a,
                         long long lda, const void* f, long long ldf, emu_bcast_fn bc, emu_reduce_fn rd) {
  G2sDistDesc d{uplo, n, nb, P, Q, pr, pc, src_r, src_c};
  const G2sDistSizes sz = g2s_dist_sizes(d);
  Guarded<T> winv(sz.winv), dk(sz.dk), tmp(sz.tmp), xr(sz.xr), yt(sz.ytiles), ac(sz.acopy), lc(sz.lcopy);
  G2sDistWork<T> w{winv.data(), dk.data(), tmp.data(), xr.data(), yt.data(), ac.data(), lc.data()};
  CallbackOps<T> ops;
  ops.bc = bc;
  ops.rd = rd;
  int e = gen_to_std_dist<T>(ops, d, static_cast<T*>(a), lda, static_cast<const T*>(f), ldf, w);
  if (e == 0 && ops.hazards)
    e = 9001;
  if (e == 0 && !(winv.intact() && dk.intact() && tmp.intact() && xr.intact() && yt.intact() && ac.intact() && lc.intact()))
    e = 9300;
  return e;
}

template <class T>
struct RedistCallbackOps {
  const RedistLists* L;
  emu_alltoallv_fn a2a;
  int world;
  int copy_rects(int which, const T* src, T* dst) {
    const std::vector<CopyRect>& v = which == 0 ? L->pack : which == 1 ? L->self : L->unpack;
    for (int b = 0; b < (int)v.size(); ++b)
      for (int th = 0; th < kCopyRectThreads; ++th)
        copy_rects_body<T>(v.data(), src, dst, b, th, kCopyRectThreads);
    return 0;
  }
  int alltoallv(const T* sendbuf, T* recvbuf) {
    return a2a(sendbuf, L->send_off.data(), L->send_cnt.data(), recvbuf, L->recv_off.data(), L->recv_cnt.data(),
               (int)sizeof(T), world);
  }
};

template <class T>
int redistribute_rank(int P, int Q, int pr, int pc, char ordering, long long m, long long n, int mb_s, int nb_s, int sr_s,
                      int sc_s, int mb_d, int nb_d, int sr_d, int sc_d, const void* src, long long ld_src, void* dst,
                      long long ld_dst, emu_alltoallv_fn a2a) {
  RedistDesc d{m, n, mb_s, nb_s, sr_s, sc_s, mb_d, nb_d, sr_d, sc_d, P, Q, pr, pc, ordering};
  if (!redist_valid(d))
    return 9401;
  const RedistLists L = redist_build(d, ld_src, ld_dst);
  Guarded<T> sb((size_t)L.send_total), rb((size_t)L.recv_total);
  RedistCallbackOps<T> ops{&L, a2a, P * Q};
  int e = redistribute<T>(ops, L, static_cast<const T*>(src), static_cast<T*>(dst), sb.data(), rb.data());
  if (e == 0 && !(sb.intact() && rb.intact()))
    e = 9300;
  return e;
}

}  // namespace

#define EMU_TYPED(X, T)                                                                                             \
  int emu_triangular_solver_##X(char side, char uplo, char op, char diag, int m, int n, int nb, const void* alpha, \
                                const void* a, long long lda, void* b, long long ldb) {                             \
    return solve<T>(side, uplo, op, diag, m, n, nb, alpha, a, lda, b, ldb);                                         \
  }                                                                                                                 \
  int emu_triangular_solver_la_##X(char side, char uplo, char op, char diag, int m, int n, int nb,                  \
                                   const void* alpha, const void* a, long long lda, void* b, long long ldb,        \
                                   int prefer, int* races) {                                                        \
    return solve<T>(side, uplo, op, diag, m, n, nb, alpha, a, lda, b, ldb, 1, prefer, races);                       \
  }                                                                                                                 \
  int emu_tri_inverse_##X(int upper, int unit, int n, int nb, const void* t, long long ldt, void* w) {              \
    return inverse<T>(upper, unit, n, nb, t, ldt, w);                                                               \
  }                                                                                                                 \
  int emu_gen_to_std_##X(char uplo, int n, int nb, void* a, long long lda, const void* f, long long ldf) {          \
    return gen_to_std_emu<T>(uplo, n, nb, a, lda, f, ldf);                                                          \
  }                                                                                                                 \
  int emu_gen_to_std_la_##X(char uplo, int n, int nb, void* a, long long lda, const void* f, long long ldf,        \
                            int prefer, int* races) {                                                               \
    return gen_to_std_emu<T>(uplo, n, nb, a, lda, f, ldf, 1, prefer, races);                                        \
  }                                                                                                                 \
  int emu_triangular_solver_dist_rank_##X(int P, int Q, int pr, int pc, int src_r, int src_c, char side, char uplo, \
                                          char op, char diag, int m, int n, int mb, int nb, const void* alpha,      \
                                          const void* a, long long lda, void* b, long long ldb, emu_bcast_fn bc,     \
                                          emu_reduce_fn rd) {                                                        \
    return solve_dist_rank<T>(P, Q, pr, pc, src_r, src_c, side, uplo, op, diag, m, n, mb, nb, alpha, a, lda, b, ldb, \
                              bc, rd);                                                                              \
  }                                                                                                                 \
  int emu_gen_to_std_dist_rank_##X(int P, int Q, int pr, int pc, int src_r, int src_c, char uplo, int n, int nb,    \
                                   void* a, long long lda, const void* f, long long ldf, emu_bcast_fn bc,           \
                                   emu_reduce_fn rd) {                                                               \
    return gen_to_std_dist_rank<T>(P, Q, pr, pc, src_r, src_c, uplo, n, nb, a, lda, f, ldf, bc, rd);                \
  }                                                                                                                 \
  int emu_redistribute_rank_##X(int P, int Q, int pr, int pc, char ordering, long long m, long long n, int mb_s,    \
                                int nb_s, int sr_s, int sc_s, int mb_d, int nb_d, int sr_d, int sc_d,               \
                                const void* src, long long ld_src, void* dst, long long ld_dst,                      \
                                emu_alltoallv_fn a2a) {                                                              \
    return redistribute_rank<T>(P, Q, pr, pc, ordering, m, n, mb_s, nb_s, sr_s, sc_s, mb_d, nb_d, sr_d, sc_d, src,  \
                                ld_src, dst, ld_dst, a2a);                                                          \
  }                                                                                                                 \
  int emu_redistribute_##X(int P, int Q, char ordering, long long m, long long n, int mb_s, int nb_s, int sr_s,     \
                           int sc_s, int mb_d, int nb_d, int sr_d, int sc_d, void* const* src,                      \
                           const long long* ld_src, void* const* dst, const long long* ld_dst) {                    \
    return redistribute_emu<T>(P, Q, ordering, m, n, mb_s, nb_s, sr_s, sc_s, mb_d, nb_d, sr_d, sc_d, src, ld_src,   \
                               dst, ld_dst);                                                                        \
  }                                                                                                                 \
  int emu_gen_to_std_dist_##X(int P, int Q, int src_r, int src_c, char uplo, int n, int nb, void* const* a,         \
                              const long long* lda, void* const* f, const long long* ldf) {                          \
    return gen_to_std_dist_emu<T>(P, Q, src_r, src_c, uplo, n, nb, a, lda, f, ldf);                                  \
  }                                                                                                                 \
  int emu_triangular_solver_dist_##X(int P, int Q, int src_r, int src_c, char side, char uplo, char op, char diag,  \
                                     int m, int n, int mb, int nb, const void* alpha, void* const* a,               \
                                     const long long* lda, void* const* b, const long long* ldb) {                  \
    return solve_dist<T>(P, Q, src_r, src_c, side, uplo, op, diag, m, n, mb, nb, alpha, a, lda, b, ldb);            \
  }

extern "C" {
EMU_TYPED(s, float)
EMU_TYPED(d, double)
EMU_TYPED(c, float2)
EMU_TYPED(z, double2)
}
