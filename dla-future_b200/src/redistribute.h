// Redistribution of a distributed matrix to another block size / source rank on the same grid:
//   dst (mb_d x nb_d tiles) <- src (mb_s x nb_s tiles), same global size, same P x Q grid.
// Reference: matrix::copy(src, dst, grid), include/dlaf/matrix/copy.h:118-231 - it cuts both matrices into sub-tiles of
// the smaller block size (which must divide the larger one, :124-131) and posts one MPI send / receive per sub-tile.
// Why it is on this path: ScaLAPACK callers come with block sizes of 32-128, the factorization wants 512 (the
// contraction's K); redistribute - factorize - redistribute back costs 4 passes over the local block at HBM / NVLink
// speed against an O(n^3) gain.
//
// B200 design: an all-to-all with ONE message per pair of ranks instead of one per sub-tile.
//   pack    one launch of the rectangle-list kernel gathers, for every peer, the sub-tiles it will own into one
//           contiguous segment of the send buffer (sub-tiles that stay on this rank go straight into dst)
//   exchange  grouped ncclSend / ncclRecv of the segments over NVLink (all pairs at once)
//   unpack  one launch scatters the received segments into the local block of dst
// Both sides enumerate the sub-tiles of a (sender, receiver) pair in the same global column-major order, so no
// per-message metadata travels. The rectangle lists are host index algebra (built once per plan, shared with the CPU
// emulation) and live on the device.
#pragma once
#include <cstddef>
#include <vector>

#include "copy_rects.cuh"
#include "schedule.h"

namespace dlafb {

struct RedistDesc {
  long long m, n;
  int mb_s, nb_s, src_r_s, src_c_s;  // source: tile size and rank owning the first tile
  int mb_d, nb_d, src_r_d, src_c_d;  // destination
  int P, Q, pr, pc;
  char ordering;  // 'R' / 'C': linear rank of (pr, pc) in the full communicator
};

struct RedistLists {
  std::vector<CopyRect> pack;    // local src -> send buffer
  std::vector<CopyRect> self;    // local src -> local dst (sub-tiles that do not move)
  std::vector<CopyRect> unpack;  // receive buffer -> local dst
  // per linear rank of the full communicator: element offset and count of its segment (0 for this rank itself)
  std::vector<long long> send_off, send_cnt, recv_off, recv_cnt;
  long long send_total = 0, recv_total = 0;
  long long ls_rows = 0, ls_cols = 0, ld_rows = 0, ld_cols = 0;  // local sizes of src and dst
};

inline bool redist_valid(const RedistDesc& d) {
  if (d.m < 0 || d.n < 0 || d.mb_s < 1 || d.nb_s < 1 || d.mb_d < 1 || d.nb_d < 1 || d.P < 1 || d.Q < 1)
    return false;
  const int mb = d.mb_s < d.mb_d ? d.mb_s : d.mb_d, nb = d.nb_s < d.nb_d ? d.nb_s : d.nb_d;
  return d.mb_s % mb == 0 && d.mb_d % mb == 0 && d.nb_s % nb == 0 && d.nb_d % nb == 0;  // copy.h:124-131
}

inline int redist_linear_rank(const RedistDesc& d, int pr, int pc) {
  return d.ordering == 'C' ? pc * d.P + pr : pr * d.Q + pc;
}

// One axis: for every sub-tile i (of size sb) the owner and local element offset in a distribution with tiles of tb.
struct RedistAxis {
  std::vector<int> owner;
  std::vector<long long> local_off;
  std::vector<int> size;
};

inline RedistAxis redist_axis(long long n, int sb, int tb, int nprocs, int src) {
  RedistAxis a;
  const long long ns = n > 0 ? (n + sb - 1) / sb : 0;
  a.owner.resize(ns);
  a.local_off.resize(ns);
  a.size.resize(ns);
  for (long long i = 0; i < ns; ++i) {
    const long long first = i * sb;
    const long long tile = first / tb;
    a.owner[i] = (int)((tile + src) % nprocs);
    a.local_off[i] = (tile / nprocs) * tb + (first - tile * tb);
    a.size[i] = (int)(n - first < sb ? n - first : sb);
  }
  return a;
}

// ld_src / ld_dst: leading dimensions of the caller's local blocks.
inline RedistLists redist_build(const RedistDesc& d, long long ld_src, long long ld_dst) {
  RedistLists L;
  const int world = d.P * d.Q;
  L.send_off.assign(world, 0);
  L.send_cnt.assign(world, 0);
  L.recv_off.assign(world, 0);
  L.recv_cnt.assign(world, 0);
  L.ls_rows = Dist1D{d.m, d.mb_s, d.P, d.pr, d.src_r_s}.local_size();
  L.ls_cols = Dist1D{d.n, d.nb_s, d.Q, d.pc, d.src_c_s}.local_size();
  L.ld_rows = Dist1D{d.m, d.mb_d, d.P, d.pr, d.src_r_d}.local_size();
  L.ld_cols = Dist1D{d.n, d.nb_d, d.Q, d.pc, d.src_c_d}.local_size();
  if (d.m <= 0 || d.n <= 0)
    return L;
  const int mb = d.mb_s < d.mb_d ? d.mb_s : d.mb_d, nb = d.nb_s < d.nb_d ? d.nb_s : d.nb_d;
  const RedistAxis rs = redist_axis(d.m, mb, d.mb_s, d.P, d.src_r_s), rd = redist_axis(d.m, mb, d.mb_d, d.P, d.src_r_d);
  const RedistAxis cs = redist_axis(d.n, nb, d.nb_s, d.Q, d.src_c_s), cd = redist_axis(d.n, nb, d.nb_d, d.Q, d.src_c_d);
  const long long nI = (long long)rs.owner.size(), nJ = (long long)cs.owner.size();
  // counts first (segment offsets), then the rectangles, both in global column-major order of the sub-tiles
  for (int pass = 0; pass < 2; ++pass) {
    std::vector<long long> send_pos(world, 0), recv_pos(world, 0);
    if (pass == 1) {
      long long s = 0, r = 0;
      for (int p = 0; p < world; ++p) {
        L.send_off[p] = s;
        s += L.send_cnt[p];
        L.recv_off[p] = r;
        r += L.recv_cnt[p];
      }
      L.send_total = s;
      L.recv_total = r;
    }
    for (long long J = 0; J < nJ; ++J) {
      const bool col_src_mine = cs.owner[J] == d.pc, col_dst_mine = cd.owner[J] == d.pc;
      if (!col_src_mine && !col_dst_mine)
        continue;
      for (long long I = 0; I < nI; ++I) {
        const bool src_mine = col_src_mine && rs.owner[I] == d.pr;
        const bool dst_mine = col_dst_mine && rd.owner[I] == d.pr;
        if (!src_mine && !dst_mine)
          continue;
        const int rows = rs.size[I], cols = cs.size[J];
        const long long elems = (long long)rows * cols;
        const long long soff = rs.local_off[I] + cs.local_off[J] * ld_src;
        const long long doff = rd.local_off[I] + cd.local_off[J] * ld_dst;
        if (src_mine && dst_mine) {
          if (pass == 1)
            L.self.push_back(CopyRect{soff, doff, ld_src, ld_dst, rows, cols});
          continue;
        }
        if (src_mine) {
          const int to = redist_linear_rank(d, rd.owner[I], cd.owner[J]);
          if (pass == 0)
            L.send_cnt[to] += elems;
          else {
            L.pack.push_back(CopyRect{soff, L.send_off[to] + send_pos[to], ld_src, (long long)rows, rows, cols});
            send_pos[to] += elems;
          }
        }
        if (dst_mine) {
          const int from = redist_linear_rank(d, rs.owner[I], cs.owner[J]);
          if (pass == 0)
            L.recv_cnt[from] += elems;
          else {
            L.unpack.push_back(CopyRect{L.recv_off[from] + recv_pos[from], doff, (long long)rows, ld_dst, rows, cols});
            recv_pos[from] += elems;
          }
        }
      }
    }
  }
  return L;
}

// Ops interface:
//   copy_rects(which, src, dst)     which = 0 pack, 1 self, 2 unpack: run that rectangle list
//   alltoallv(sendbuf, recvbuf)     exchange the segments of the plan's lists over the full communicator
template <class T, class Ops>
int redistribute(Ops& ops, const RedistLists& L, const T* src, T* dst, T* sendbuf, T* recvbuf) {
  int e;
  if (!L.pack.empty() && (e = ops.copy_rects(0, src, sendbuf)) != 0)
    return e;
  if (!L.self.empty() && (e = ops.copy_rects(1, src, dst)) != 0)
    return e;
  if ((e = ops.alltoallv(sendbuf, recvbuf)) != 0)
    return e;
  if (!L.unpack.empty() && (e = ops.copy_rects(2, recvbuf, dst)) != 0)
    return e;
  return 0;
}

}  // namespace dlafb
