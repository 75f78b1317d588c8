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
  // Round 2: the m x n matrix may be a SUB-MATRIX of a larger distributed matrix on either side: it starts at global
  // element (off_r, off_c) of a parent with the tile size and source rank above, and the local pointer is the base of
  // the parent's local block (ScaLAPACK's sub(A) = A(ia : ia + m, ja : ja + n) with any ia, ja; the reference's
  // SubDistributionSpec, include/dlaf/matrix/distribution.h:60-63). With an offset the tile boundaries of the two sides
  // need not nest, so the pieces that travel are the intersections of the two tilings.
  long long off_r_s = 0, off_c_s = 0, off_r_d = 0, off_c_d = 0;
  bool has_offset() const { return off_r_s != 0 || off_c_s != 0 || off_r_d != 0 || off_c_d != 0; }
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
  if (d.off_r_s < 0 || d.off_c_s < 0 || d.off_r_d < 0 || d.off_c_d < 0)
    return false;
  if (d.has_offset())
    return true;  // sub-matrices: any pair of tilings (the pieces are their intersections)
  const int mb = d.mb_s < d.mb_d ? d.mb_s : d.mb_d, nb = d.nb_s < d.nb_d ? d.nb_s : d.nb_d;
  return d.mb_s % mb == 0 && d.mb_d % mb == 0 && d.nb_s % nb == 0 && d.nb_d % nb == 0;  // copy.h:124-131
}

inline int redist_linear_rank(const RedistDesc& d, int pr, int pc) {
  return d.ordering == 'C' ? pc * d.P + pr : pr * d.Q + pc;
}

// One axis: the pieces [first, first + size) of 0 .. n between consecutive tile boundaries of EITHER side, each with its
// owner and local element offset on the source side and on the destination side. Without offsets and with nesting
// tile sizes these are the sub-tiles of the smaller tile size (copy.h:133-216).
struct RedistAxis {
  std::vector<int> size;
  std::vector<int> owner_s, owner_d;
  std::vector<long long> loff_s, loff_d;
};

inline RedistAxis redist_axis(long long n, int tb_s, long long off_s, int src_s, int tb_d, long long off_d, int src_d,
                              int nprocs) {
  RedistAxis a;
  for (long long first = 0; first < n;) {
    const long long gs = off_s + first, gd = off_d + first;  // positions in the two parents
    const long long to_s = tb_s - gs % tb_s, to_d = tb_d - gd % tb_d;  // distance to the next tile boundary
    long long size = to_s < to_d ? to_s : to_d;
    if (size > n - first)
      size = n - first;
    const long long ts = gs / tb_s, td = gd / tb_d;
    a.size.push_back((int)size);
    a.owner_s.push_back((int)((ts + src_s) % nprocs));
    a.loff_s.push_back((ts / nprocs) * tb_s + gs % tb_s);
    a.owner_d.push_back((int)((td + src_d) % nprocs));
    a.loff_d.push_back((td / nprocs) * tb_d + gd % tb_d);
    first += size;
  }
  return a;
}

// Local element range [lo, hi) of one side of an axis on rank `rank` (hi = 0: nothing local). Contiguous: the local
// elements between the first and the last one of the (sub-)matrix all belong to it.
inline void redist_local_range(const std::vector<int>& owner, const std::vector<long long>& loff,
                               const std::vector<int>& size, int rank, long long* lo, long long* hi) {
  *lo = 0;
  *hi = 0;
  bool any = false;
  for (size_t i = 0; i < owner.size(); ++i)
    if (owner[i] == rank) {
      if (!any || loff[i] < *lo)
        *lo = loff[i];
      if (!any || loff[i] + size[i] > *hi)
        *hi = loff[i] + size[i];
      any = true;
    }
}

// ld_src / ld_dst: leading dimensions of the caller's local blocks.
inline RedistLists redist_build(const RedistDesc& d, long long ld_src, long long ld_dst) {
  RedistLists L;
  const int world = d.P * d.Q;
  L.send_off.assign(world, 0);
  L.send_cnt.assign(world, 0);
  L.recv_off.assign(world, 0);
  L.recv_cnt.assign(world, 0);
  const RedistAxis ra = redist_axis(d.m, d.mb_s, d.off_r_s, d.src_r_s, d.mb_d, d.off_r_d, d.src_r_d, d.P);
  const RedistAxis ca = redist_axis(d.n, d.nb_s, d.off_c_s, d.src_c_s, d.nb_d, d.off_c_d, d.src_c_d, d.Q);
  // local extents the caller's blocks must cover (without offsets: the local sizes)
  long long lo;
  redist_local_range(ra.owner_s, ra.loff_s, ra.size, d.pr, &lo, &L.ls_rows);
  redist_local_range(ca.owner_s, ca.loff_s, ca.size, d.pc, &lo, &L.ls_cols);
  redist_local_range(ra.owner_d, ra.loff_d, ra.size, d.pr, &lo, &L.ld_rows);
  redist_local_range(ca.owner_d, ca.loff_d, ca.size, d.pc, &lo, &L.ld_cols);
  if (d.m <= 0 || d.n <= 0)
    return L;
  const long long nI = (long long)ra.size.size(), nJ = (long long)ca.size.size();
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
      const bool col_src_mine = ca.owner_s[J] == d.pc, col_dst_mine = ca.owner_d[J] == d.pc;
      if (!col_src_mine && !col_dst_mine)
        continue;
      for (long long I = 0; I < nI; ++I) {
        const bool src_mine = col_src_mine && ra.owner_s[I] == d.pr;
        const bool dst_mine = col_dst_mine && ra.owner_d[I] == d.pr;
        if (!src_mine && !dst_mine)
          continue;
        const int rows = ra.size[I], cols = ca.size[J];
        const long long elems = (long long)rows * cols;
        const long long soff = ra.loff_s[I] + ca.loff_s[J] * ld_src;
        const long long doff = ra.loff_d[I] + ca.loff_d[J] * ld_dst;
        if (src_mine && dst_mine) {
          if (pass == 1)
            L.self.push_back(CopyRect{soff, doff, ld_src, ld_dst, rows, cols});
          continue;
        }
        if (src_mine) {
          const int to = redist_linear_rank(d, ra.owner_d[I], ca.owner_d[J]);
          if (pass == 0)
            L.send_cnt[to] += elems;
          else {
            L.pack.push_back(CopyRect{soff, L.send_off[to] + send_pos[to], ld_src, (long long)rows, rows, cols});
            send_pos[to] += elems;
          }
        }
        if (dst_mine) {
          const int from = redist_linear_rank(d, ra.owner_s[I], ca.owner_s[J]);
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
