// TEST INFRASTRUCTURE: replays known-answer rows (written by tests/test_cxx_matrix_api.py from the fixtures in
// tests/golden/) against the C++ API classes matrix::LayoutInfo and matrix::Distribution of include/dlaf/.
// Input (stdin), one record per line:
//   L rows cols brows bcols ld row_off col_off min_mem
//   C rows cols brows bcols ld row_off col_off equal          (compared with LayoutInfo({25,25},{5,5},50,8,1000))
//   D rows cols brows bcols pr pc P Q sr sc gt_r gt_c lt_r lt_c ls_r ls_c
// Prints the number of checks and of failures; exit code = failures != 0.
#include <cstdio>
#include <iostream>
#include <sstream>
#include <string>

#include "dlaf/common/range2d.h"
#include "dlaf/common/round_robin.h"
#include "dlaf/matrix/distribution.h"
#include "dlaf/matrix/layout_info.h"

using namespace dlaf;

static long checks = 0, failures = 0;
#define EXPECT(cond, line)                                             \
  do {                                                                 \
    ++checks;                                                          \
    if (!(cond)) {                                                     \
      ++failures;                                                      \
      std::fprintf(stderr, "FAILED %s on: %s\n", #cond, line.c_str()); \
    }                                                                  \
  } while (0)

// host-only pieces that need no table: iterate_range2d (column-major, end exclusive, empty ranges), RoundRobin
static void check_small_utilities() {
  const std::string where = "utilities";
  std::string seq;
  for (auto ij : common::iterate_range2d(LocalTileSize(3, 2)))
    seq += std::to_string(ij.row()) + std::to_string(ij.col()) + " ";
  EXPECT(seq == "00 10 20 01 11 21 ", where);
  seq.clear();
  for (auto ij : common::iterate_range2d(LocalTileIndex(1, 1), LocalTileIndex(3, 3)))
    seq += std::to_string(ij.row()) + std::to_string(ij.col()) + " ";
  EXPECT(seq == "11 21 12 22 ", where);
  int n = 0;
  for (auto ij : common::iterate_range2d(LocalTileSize(0, 5))) {
    (void)ij;
    ++n;
  }
  for (auto ij : common::iterate_range2d(LocalTileSize(4, 0))) {
    (void)ij;
    ++n;
  }
  EXPECT(n == 0, where);
  common::RoundRobin<int> rr(3, 7);
  EXPECT(rr.size() == 3 && rr.currentResource() == 7, where);
  rr.nextResource() = 1;
  rr.nextResource() = 2;
  EXPECT(rr.nextResource() == 7 && rr.nextResource() == 1 && rr.nextResource() == 2, where);
}

int main() {
  check_small_utilities();
  std::string line;
  while (std::getline(std::cin, line)) {
    if (line.empty())
      continue;
    std::istringstream in(line);
    char kind;
    in >> kind;
    if (kind == 'L' || kind == 'C') {
      SizeType r, c, br, bc, ld, ro, co, last;
      in >> r >> c >> br >> bc >> ld >> ro >> co >> last;
      matrix::LayoutInfo lay(LocalElementSize(r, c), TileElementSize(br, bc), ld, ro, co);
      if (kind == 'C') {
        matrix::LayoutInfo ref(LocalElementSize(25, 25), TileElementSize(5, 5), 50, 8, 1000);
        EXPECT((lay == ref) == (last != 0), line);
        EXPECT((lay != ref) == (last == 0), line);
        continue;
      }
      EXPECT(lay.size() == LocalElementSize(r, c), line);
      EXPECT(lay.blockSize() == TileElementSize(br, bc), line);
      EXPECT(lay.ldTile() == ld, line);
      EXPECT(lay.nrTiles() == LocalTileSize((r + br - 1) / br, (c + bc - 1) / bc), line);
      EXPECT(lay.minMemSize() == last, line);
      for (SizeType j = 0; j < lay.nrTiles().cols(); ++j) {
        const SizeType jb = std::min(bc, c - j * bc);
        for (SizeType i = 0; i < lay.nrTiles().rows(); ++i) {
          const SizeType ib = std::min(br, r - i * br);
          const LocalTileIndex idx(i, j);
          EXPECT(lay.tileOffset(idx) == i * ro + j * co, line);
          EXPECT(lay.tileSize(idx) == TileElementSize(ib, jb), line);
          EXPECT(lay.minTileMemSize(idx) == ib + ld * (jb - 1), line);
          EXPECT(lay.minTileMemSize(TileElementSize(ib, jb)) == ib + ld * (jb - 1), line);
        }
      }
    }
    else if (kind == 'D') {
      SizeType r, c, br, bc, gtr, gtc, ltr, ltc, lsr, lsc;
      int pr, pc, P, Q, sr, sc;
      in >> r >> c >> br >> bc >> pr >> pc >> P >> Q >> sr >> sc >> gtr >> gtc >> ltr >> ltc >> lsr >> lsc;
      matrix::Distribution d(GlobalElementSize(r, c), TileElementSize(br, bc), comm::Size2D(P, Q), comm::Index2D(pr, pc),
                             comm::Index2D(sr, sc));
      EXPECT(d.size() == GlobalElementSize(r, c), line);
      EXPECT(d.block_size() == TileElementSize(br, bc), line);
      EXPECT(d.rank_index() == comm::Index2D(pr, pc), line);
      EXPECT(d.grid_size() == comm::Size2D(P, Q), line);
      EXPECT(d.source_rank_index() == comm::Index2D(sr, sc), line);
      EXPECT(d.nr_tiles() == GlobalTileSize(gtr, gtc), line);
      EXPECT(d.local_nr_tiles() == LocalTileSize(ltr, ltc), line);
      EXPECT(d.local_size() == LocalElementSize(lsr, lsc), line);
      // round trips over the local tiles: local -> global -> owner / local
      for (SizeType il = 0; il < ltr; ++il) {
        const SizeType g = d.global_tile_from_local_tile<Coord::Row>(il);
        EXPECT(d.rank_global_tile<Coord::Row>(g) == pr, line);
        EXPECT(d.local_tile_from_global_tile<Coord::Row>(g) == il, line);
        EXPECT(d.next_local_tile_from_global_tile<Coord::Row>(g) == il, line);
      }
      for (SizeType jl = 0; jl < ltc; ++jl) {
        const SizeType g = d.global_tile_from_local_tile<Coord::Col>(jl);
        EXPECT(d.rank_global_tile<Coord::Col>(g) == pc, line);
        EXPECT(d.local_tile_from_global_tile<Coord::Col>(g) == jl, line);
      }
      // layouts built from the distribution
      const matrix::LayoutInfo cm = matrix::colMajorLayout(d, std::max<SizeType>(1, lsr) + 3);
      EXPECT(cm.size() == d.local_size(), line);
      EXPECT(cm.tileOffset(LocalTileIndex(1, 2)) == br + 2 * bc * (std::max<SizeType>(1, lsr) + 3), line);
      const matrix::LayoutInfo tl = matrix::tileLayout(d);
      EXPECT(tl.ldTile() == br, line);
      EXPECT(tl.tileOffset(LocalTileIndex(1, 0)) == br * bc, line);
    }
  }
  std::printf("checks=%ld failures=%ld\n", checks, failures);
  return failures != 0;
}
