// common::iterate_range2d: column-major iteration over a 2D index range (reference: include/dlaf/common/range2d.h;
// the reference's own loops over tiles use it, e.g. auxiliary/norm/mc.h:62, matrix/copy.h:199).
//   for (auto ij : iterate_range2d(LocalTileSize(3, 2)))            -> (0,0) (1,0) (2,0) (0,1) (1,1) (2,1)
//   for (auto ij : iterate_range2d(LocalTileIndex(1, 1), LocalTileIndex(3, 2)))   begin inclusive, end exclusive
#pragma once
#include "dlaf/common/index2d.h"

namespace dlaf::common {

template <class IndexT, class Tag>
class IterableRange2D {
public:
  using Index = Index2D<IndexT, Tag>;
  class iterator {
  public:
    iterator(IndexT row_begin, IndexT row_end, IndexT row, IndexT col)
        : row_begin_(row_begin), row_end_(row_end), row_(row), col_(col) {}
    Index operator*() const noexcept { return Index(row_, col_); }
    iterator& operator++() noexcept {
      if (++row_ >= row_end_) {
        row_ = row_begin_;
        ++col_;
      }
      return *this;
    }
    bool operator!=(const iterator& o) const noexcept { return row_ != o.row_ || col_ != o.col_; }
    bool operator==(const iterator& o) const noexcept { return !(*this != o); }

  private:
    IndexT row_begin_, row_end_, row_, col_;
  };

  IterableRange2D(Index begin, Index end) : begin_(begin), end_(end) {}
  iterator begin() const noexcept {
    if (empty())
      return end();
    return iterator(begin_.row(), end_.row(), begin_.row(), begin_.col());
  }
  iterator end() const noexcept { return iterator(begin_.row(), end_.row(), begin_.row(), empty() ? begin_.col() : end_.col()); }
  bool empty() const noexcept { return begin_.row() >= end_.row() || begin_.col() >= end_.col(); }

private:
  Index begin_, end_;
};

/// [begin, end): rows fastest.
template <class IndexT, class Tag>
IterableRange2D<IndexT, Tag> iterate_range2d(Index2D<IndexT, Tag> begin, Index2D<IndexT, Tag> end) {
  return {begin, end};
}
/// [(0, 0), size)
template <class IndexT, class Tag>
IterableRange2D<IndexT, Tag> iterate_range2d(Size2D<IndexT, Tag> size) {
  return {Index2D<IndexT, Tag>(0, 0), Index2D<IndexT, Tag>(size.rows(), size.cols())};
}

}  // namespace dlaf::common
