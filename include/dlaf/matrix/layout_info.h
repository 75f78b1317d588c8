// matrix::LayoutInfo: where tile (i, j) of the local block lives in memory (reference:
// include/dlaf/matrix/layout_info.h:26-161, src/matrix/layout_info.cpp:17-68).
#pragma once
#include "dlaf/matrix/distribution.h"

namespace dlaf::matrix {

class LayoutInfo {
public:
  LayoutInfo(const LocalElementSize& size, const TileElementSize& block_size, SizeType tile_ld, SizeType tile_offset_row,
             SizeType tile_offset_col)
      : size_(size), block_(block_size), ld_tile_(tile_ld), off_row_(tile_offset_row), off_col_(tile_offset_col) {
    DLAF_ASSERT(tile_ld >= 1, tile_ld);
  }

  /// Offset (in elements) of the first element of local tile @p index.
  SizeType tile_offset(const LocalTileIndex& index) const noexcept {
    return index.row() * off_row_ + index.col() * off_col_;
  }
  SizeType tileOffset(const LocalTileIndex& index) const noexcept { return tile_offset(index); }
  SizeType ld_tile() const noexcept { return ld_tile_; }
  SizeType ldTile() const noexcept { return ld_tile_; }
  const LocalElementSize& size() const noexcept { return size_; }
  const TileElementSize& block_size() const noexcept { return block_; }
  const TileElementSize& blockSize() const noexcept { return block_; }
  LocalTileSize nr_tiles() const noexcept {
    return {(size_.rows() + block_.rows() - 1) / block_.rows(), (size_.cols() + block_.cols() - 1) / block_.cols()};
  }
  LocalTileSize nrTiles() const noexcept { return nr_tiles(); }
  /// Size of local tile @p index (edge tiles are smaller; layout_info.h:96-103).
  TileElementSize tile_size_of(const LocalTileIndex& index) const noexcept {
    const SizeType r = size_.rows() - index.row() * block_.rows(), c = size_.cols() - index.col() * block_.cols();
    return {r < block_.rows() ? r : block_.rows(), c < block_.cols() ? c : block_.cols()};
  }
  TileElementSize tileSize(const LocalTileIndex& index) const noexcept { return tile_size_of(index); }
  /// Minimum number of elements a tile occupies (layout_info.h:78-94).
  SizeType min_tile_mem_size(const TileElementSize& tile_size) const noexcept {
    return tile_size.isEmpty() ? 0 : tile_size.rows() + ld_tile_ * (tile_size.cols() - 1);
  }
  SizeType minTileMemSize(const TileElementSize& tile_size) const noexcept { return min_tile_mem_size(tile_size); }
  SizeType minTileMemSize(const LocalTileIndex& index) const noexcept { return min_tile_mem_size(tile_size_of(index)); }
  bool operator==(const LayoutInfo& o) const noexcept {
    return size_ == o.size_ && block_ == o.block_ && ld_tile_ == o.ld_tile_ && off_row_ == o.off_row_ &&
           off_col_ == o.off_col_;
  }
  bool operator!=(const LayoutInfo& o) const noexcept { return !(*this == o); }
  bool is_col_major() const noexcept { return off_row_ == block_.rows() && off_col_ == block_.cols() * ld_tile_; }

  /// Minimum number of elements the memory has to hold (layout_info.cpp:40-60).
  SizeType min_mem_size() const noexcept {
    if (size_.isEmpty())
      return 0;
    const SizeType rt = (size_.rows() + block_.rows() - 1) / block_.rows();
    const SizeType ct = (size_.cols() + block_.cols() - 1) / block_.cols();
    const SizeType last_rows = size_.rows() - (rt - 1) * block_.rows();
    const SizeType last_cols = size_.cols() - (ct - 1) * block_.cols();
    return (rt - 1) * off_row_ + (ct - 1) * off_col_ + (last_cols - 1) * ld_tile_ + last_rows;
  }
  SizeType minMemSize() const noexcept { return min_mem_size(); }

private:
  LocalElementSize size_;
  TileElementSize block_;
  SizeType ld_tile_, off_row_, off_col_;
};

/// Column-major local block with leading dimension @p ld (layout_info.h:125-132).
inline LayoutInfo colMajorLayout(const LocalElementSize& size, const TileElementSize& block_size, SizeType ld) {
  DLAF_ASSERT(ld >= (size.rows() > 1 ? size.rows() : 1), ld);
  return LayoutInfo(size, block_size, ld, block_size.rows(), block_size.cols() * ld);
}
inline LayoutInfo colMajorLayout(const Distribution& distribution, SizeType ld) {
  return colMajorLayout(distribution.local_size(), distribution.block_size(), ld);
}

/// Contiguous tiles (layout_info.h:135-159), the layout of the panel workspaces.
inline LayoutInfo tileLayout(const LocalElementSize& size, const TileElementSize& block_size) {
  const SizeType rt = (size.rows() + block_size.rows() - 1) / block_size.rows();
  const SizeType tile_elems = block_size.rows() * block_size.cols();
  return LayoutInfo(size, block_size, block_size.rows(), tile_elems, tile_elems * (rt > 1 ? rt : 1));
}
inline LayoutInfo tileLayout(const Distribution& distribution) {
  return tileLayout(distribution.local_size(), distribution.block_size());
}

}  // namespace dlaf::matrix
