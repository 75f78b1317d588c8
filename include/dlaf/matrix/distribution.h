// matrix::Distribution: 2D block-cyclic index algebra of the reference (include/dlaf/matrix/distribution.h:111-733,
// include/dlaf/matrix/util_distribution.h:27-200) for the one-tile-per-block, offset-free case the Cholesky path
// uses (factorization/cholesky.h:69-71 asserts exactly that). Index maps call the engine's C ABI so that the C++
// API, the scheduler and the tests share one implementation.
#pragma once
#include "dlaf/common/index2d.h"
#include "dlafb_engine.h"

namespace dlaf::matrix {

class Distribution {
public:
  Distribution() noexcept = default;

  /// Local (non distributed) matrix.
  Distribution(const LocalElementSize& size, const TileElementSize& block_size)
      : Distribution(GlobalElementSize(size.rows(), size.cols()), block_size, comm::Size2D(1, 1), comm::Index2D(0, 0),
                     comm::Index2D(0, 0)) {}

  /// Distributed matrix: @p grid_size ranks, this rank at @p rank_index, first tile on @p source_rank_index.
  Distribution(const GlobalElementSize& size, const TileElementSize& block_size, const comm::Size2D& grid_size,
               const comm::Index2D& rank_index, const comm::Index2D& source_rank_index)
      : size_(size), block_(block_size), grid_(grid_size), rank_(rank_index), src_(source_rank_index) {
    DLAF_ASSERT(size.rows() >= 0 && size.cols() >= 0, size);
    DLAF_ASSERT(block_size.rows() >= 1 && block_size.cols() >= 1, block_size);
    DLAF_ASSERT(grid_size.rows() >= 1 && grid_size.cols() >= 1, grid_size);
    DLAF_ASSERT(rank_index.row() >= 0 && rank_index.row() < grid_size.rows(), rank_index);
    DLAF_ASSERT(rank_index.col() >= 0 && rank_index.col() < grid_size.cols(), rank_index);
    DLAF_ASSERT(source_rank_index.row() >= 0 && source_rank_index.row() < grid_size.rows(), source_rank_index);
    DLAF_ASSERT(source_rank_index.col() >= 0 && source_rank_index.col() < grid_size.cols(), source_rank_index);
  }

  const GlobalElementSize& size() const noexcept { return size_; }
  const TileElementSize& block_size() const noexcept { return block_; }
  const TileElementSize& tile_size() const noexcept { return block_; }
  // camelCase aliases kept by the reference for a while
  const TileElementSize& blockSize() const noexcept { return block_; }
  const comm::Size2D& grid_size() const noexcept { return grid_; }
  const comm::Size2D& commGridSize() const noexcept { return grid_; }
  const comm::Index2D& rank_index() const noexcept { return rank_; }
  const comm::Index2D& rankIndex() const noexcept { return rank_; }
  const comm::Index2D& source_rank_index() const noexcept { return src_; }
  const comm::Index2D& sourceRankIndex() const noexcept { return src_; }

  GlobalTileSize nr_tiles() const noexcept {
    return {dlafb_dist_nr_tiles(size_.rows(), block_.rows()), dlafb_dist_nr_tiles(size_.cols(), block_.cols())};
  }
  GlobalTileSize nrTiles() const noexcept { return nr_tiles(); }
  LocalTileSize local_nr_tiles() const noexcept {
    return {dlafb_dist_local_nr_tiles(size_.rows(), block_.rows(), grid_.rows(), rank_.row(), src_.row()),
            dlafb_dist_local_nr_tiles(size_.cols(), block_.cols(), grid_.cols(), rank_.col(), src_.col())};
  }
  LocalTileSize localNrTiles() const noexcept { return local_nr_tiles(); }
  LocalElementSize local_size() const noexcept {
    return {dlafb_dist_local_size(size_.rows(), block_.rows(), grid_.rows(), rank_.row(), src_.row()),
            dlafb_dist_local_size(size_.cols(), block_.cols(), grid_.cols(), rank_.col(), src_.col())};
  }
  LocalElementSize localSize() const noexcept { return local_size(); }

  template <Coord rc>
  int rank_global_tile(SizeType g) const noexcept {
    return dlafb_dist_rank_global_tile(g, grid_.get<rc>(), src_.get<rc>());
  }
  comm::Index2D rank_global_tile(const GlobalTileIndex& g) const noexcept {
    return {rank_global_tile<Coord::Row>(g.row()), rank_global_tile<Coord::Col>(g.col())};
  }
  comm::Index2D rankGlobalTile(const GlobalTileIndex& g) const noexcept { return rank_global_tile(g); }

  template <Coord rc>
  SizeType local_tile_from_global_tile(SizeType g) const noexcept {
    return rank_global_tile<rc>(g) == rank_.get<rc>() ? dlafb_dist_local_tile_from_global(g, grid_.get<rc>()) : -1;
  }
  LocalTileIndex local_tile_index(const GlobalTileIndex& g) const noexcept {
    return {local_tile_from_global_tile<Coord::Row>(g.row()), local_tile_from_global_tile<Coord::Col>(g.col())};
  }
  LocalTileIndex localTileIndex(const GlobalTileIndex& g) const noexcept { return local_tile_index(g); }

  template <Coord rc>
  SizeType next_local_tile_from_global_tile(SizeType g) const noexcept {
    return dlafb_dist_next_local_tile_from_global(g, grid_.get<rc>(), rank_.get<rc>(), src_.get<rc>());
  }
  template <Coord rc>
  SizeType global_tile_from_local_tile(SizeType l) const noexcept {
    return dlafb_dist_global_tile_from_local(l, grid_.get<rc>(), rank_.get<rc>(), src_.get<rc>());
  }
  GlobalTileIndex global_tile_index(const LocalTileIndex& l) const noexcept {
    return {global_tile_from_local_tile<Coord::Row>(l.row()), global_tile_from_local_tile<Coord::Col>(l.col())};
  }
  GlobalTileIndex globalTileIndex(const LocalTileIndex& l) const noexcept { return global_tile_index(l); }

  /// Size of the global tile @p g (edge tiles are smaller).
  TileElementSize tile_size_of(const GlobalTileIndex& g) const noexcept {
    auto sz = [](SizeType n, SizeType nb, SizeType i) {
      const SizeType r = n - i * nb;
      return r < nb ? (r > 0 ? r : 0) : nb;
    };
    return {sz(size_.rows(), block_.rows(), g.row()), sz(size_.cols(), block_.cols(), g.col())};
  }
  TileElementSize tileSize(const GlobalTileIndex& g) const noexcept { return tile_size_of(g); }

  bool operator==(const Distribution& o) const noexcept {
    return size_ == o.size_ && block_ == o.block_ && grid_ == o.grid_ && rank_ == o.rank_ && src_ == o.src_;
  }

private:
  GlobalElementSize size_{0, 0};
  TileElementSize block_{1, 1};
  comm::Size2D grid_{1, 1};
  comm::Index2D rank_{0, 0};
  comm::Index2D src_{0, 0};
};

}  // namespace dlaf::matrix
