// Host-side index algebra and per-step schedule of the distributed right-looking Cholesky (Lower form).
// Pure C++ (no CUDA): the same code drives the GPU engine and is exported through the C ABI so that the CPU
// test-suite can replay it with the oracle's tile arithmetic over a gloo process group.
//
// Follows the reference's
//   2D block-cyclic maps     include/dlaf/matrix/util_distribution.h:85-175, src/matrix/distribution.cpp:151-162
//   loop structure           include/dlaf/factorization/cholesky/impl.h:191-313
//   panel broadcast routing  include/dlaf/communication/broadcast_panel.h:123-213
#pragma once
#include <algorithm>
#include <cstdint>
#include <vector>

namespace dlafb {

struct Dist1D {
  int64_t n = 0, nb = 1;
  int P = 1, rank = 0, src = 0;
  int64_t nt() const { return n > 0 ? (n + nb - 1) / nb : 0; }
  int off() const { return (rank + P - src) % P; }                    // position of this rank counted from src
  int owner(int64_t g) const { return (int)((g + src) % P); }         // rank_global_tile
  int64_t local_of(int64_t g) const { return g / P; }                 // local_tile_from_global_tile (on its owner)
  int64_t next_local(int64_t g) const {                               // next_local_tile_from_global_tile
    return g / P + ((off() < (int)(g % P)) ? 1 : 0);
  }
  int64_t nloc_tiles() const { return next_local(nt()); }
  int64_t global_of(int64_t l) const { return l * P + off(); }        // global_tile_from_local_tile
  int64_t tile_size(int64_t g) const {
    const int64_t r = n - g * nb;
    return r < nb ? (r > 0 ? r : 0) : nb;
  }
  int64_t local_size() const {
    const int64_t lt = nloc_tiles();
    if (lt == 0)
      return 0;
    return (lt - 1) * nb + tile_size(global_of(lt - 1));
  }
};

struct ColBcast {
  int jl;      // slot of the transposed (column) panel = local column tile
  int64_t g;   // global tile index
  int root;    // rank row that owns the diagonal position g: root inside the column communicator
  int src_il;  // on the root: slot of its row panel holding the tile
};

struct UpdateTile {
  int row, col;  // first local element row / column of the tile
  int diag;      // tile lies on the global diagonal (HERK: lower part only, both operands from the row panel)
};

struct Step {
  int64_t k = 0;
  bool own_diag = false;   // this rank factors A(k,k)
  bool in_col = false;     // this rank's grid column owns global column k (takes part in the panel solve)
  int diag_root = 0;       // rank row owning A(k,k): root of the diagonal-tile broadcast in the column communicator
  int row_root = 0;        // rank column owning column k: root of the row-panel broadcast in the row communicator
  int lk_row = -1;         // local row tile of k on the owner row (valid when this rank's row owns k)
  int lk_col = -1;         // local column tile of k (valid when in_col)
  int il0 = 0, jl0 = 0;    // first local row / column tile with global index > k
  int il_end = 0, jl_end = 0;
  std::vector<ColBcast> colb;
  std::vector<UpdateTile> hp, np;  // trailing tiles of global column k+1 (look-ahead) and the rest, in row bands
};

struct Schedule {
  Dist1D rows, cols;
  std::vector<Step> steps;
};

// inclusive = false: the factorization (step k touches the tiles with global column > k).
// inclusive = true: the residual A - L*L^H (step k subtracts L(:,k)*L(:,k)^H from every tile with column >= k, the
// panel starts at the diagonal tile, there is no look-ahead list and the last step is a step like the others).
inline Schedule make_schedule(int64_t n, int64_t nb, int P, int Q, int pr, int pc, int sr, int sc,
                              bool inclusive = false) {
  Schedule s;
  s.rows = Dist1D{n, nb, P, pr, sr};
  s.cols = Dist1D{n, nb, Q, pc, sc};
  const Dist1D& R = s.rows;
  const Dist1D& C = s.cols;
  const int64_t nt = R.nt();
  s.steps.resize(nt);
  for (int64_t k = 0; k < nt; ++k) {
    Step& st = s.steps[k];
    st.k = k;
    st.diag_root = R.owner(k);
    st.row_root = C.owner(k);
    st.in_col = (st.row_root == pc);
    st.own_diag = st.in_col && (st.diag_root == pr);
    if (st.diag_root == pr)
      st.lk_row = (int)R.local_of(k);
    if (st.in_col)
      st.lk_col = (int)C.local_of(k);
    st.il0 = (int)R.next_local(inclusive ? k : k + 1);
    st.jl0 = (int)C.next_local(inclusive ? k : k + 1);
    st.il_end = (int)R.nloc_tiles();
    st.jl_end = (int)C.nloc_tiles();
    if (k == nt - 1 && !inclusive)
      continue;
    for (int jl = st.jl0; jl < st.jl_end; ++jl) {
      const int64_t g = C.global_of(jl);
      if (g == nt - 1)
        continue;  // the last tile has nothing below it: not needed in the transposed panel (:186-191)
      st.colb.push_back(ColBcast{jl, g, R.owner(g), (int)R.local_of(g)});
    }
    // Bands of kBandTiles local tile rows, column after column inside a band: the band's tiles of the row panel stay
    // in L2 while the column-panel tiles stream through once per band. (Plain column-major order re-read the whole
    // row panel from HBM for every tile column: 7.9 GB read instead of 4.2 GB for the first update of n = 32768,
    // profiles/r01/ncu_update_kernel.txt.)
    const int kBandTiles = (int)std::max<int64_t>(1, 4096 / nb);  // 16 MB of FP64 row panel at nb = 512
    for (int ib0 = st.il0; ib0 < st.il_end; ib0 += kBandTiles) {
      const int ib1 = std::min(ib0 + kBandTiles, st.il_end);
      for (int jl = st.jl0; jl < st.jl_end; ++jl) {
        const int64_t gj = C.global_of(jl);
        for (int il = std::max((int)R.next_local(gj), ib0); il < ib1; ++il) {
          const int64_t gi = R.global_of(il);
          UpdateTile t{(int)(il * nb), (int)(jl * nb), gi == gj ? 1 : 0};
          (gj == k + 1 && !inclusive ? st.hp : st.np).push_back(t);
        }
      }
    }
  }
  return s;
}

}  // namespace dlafb
