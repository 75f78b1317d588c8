// 2D index / size strong types (reference: include/dlaf/common/index2d.h, include/dlaf/matrix/index.h,
// include/dlaf/communication/index.h). Only what the Cholesky path touches.
#pragma once
#include <array>
#include <ostream>

#include "dlaf/types.h"

namespace dlaf {

enum class Coord { Row, Col };
constexpr Coord orthogonal(Coord c) { return c == Coord::Row ? Coord::Col : Coord::Row; }

namespace common {
enum class Ordering { RowMajor, ColumnMajor };

template <class IndexT, class Tag>
class Index2D {
public:
  constexpr Index2D() noexcept : row_(-1), col_(-1) {}
  constexpr Index2D(IndexT row, IndexT col) noexcept : row_(row), col_(col) {}
  constexpr explicit Index2D(const std::array<IndexT, 2>& c) noexcept : row_(c[0]), col_(c[1]) {}
  constexpr IndexT row() const noexcept { return row_; }
  constexpr IndexT col() const noexcept { return col_; }
  template <Coord rc>
  constexpr IndexT get() const noexcept { return rc == Coord::Row ? row_ : col_; }
  constexpr IndexT get(Coord rc) const noexcept { return rc == Coord::Row ? row_ : col_; }
  constexpr bool isValid() const noexcept { return row_ >= 0 && col_ >= 0; }
  constexpr bool operator==(const Index2D& o) const noexcept { return row_ == o.row_ && col_ == o.col_; }
  constexpr bool operator!=(const Index2D& o) const noexcept { return !(*this == o); }
  friend std::ostream& operator<<(std::ostream& os, const Index2D& i) {
    return os << "(" << i.row_ << ", " << i.col_ << ")";
  }

protected:
  IndexT row_, col_;
};

template <class IndexT, class Tag>
class Size2D : public Index2D<IndexT, Tag> {
public:
  constexpr Size2D() noexcept : Index2D<IndexT, Tag>(0, 0) {}
  constexpr Size2D(IndexT rows, IndexT cols) noexcept : Index2D<IndexT, Tag>(rows, cols) {}
  constexpr IndexT rows() const noexcept { return this->row_; }
  constexpr IndexT cols() const noexcept { return this->col_; }
  constexpr bool isEmpty() const noexcept { return this->row_ == 0 || this->col_ == 0; }
  constexpr IndexT linear_size() const noexcept { return this->row_ * this->col_; }
};

// computeCoords: linear rank -> 2D coordinates (include/dlaf/common/index2d.h)
template <class Index2DT, class IndexT>
Index2DT computeCoords(Ordering ordering, IndexT index, const std::array<IndexT, 2>& dims) {
  return ordering == Ordering::RowMajor ? Index2DT(index / dims[1], index % dims[1])
                                        : Index2DT(index % dims[0], index / dims[0]);
}
}  // namespace common

namespace matrix {
struct GlobalElement_TAG;
struct LocalElement_TAG;
struct GlobalTile_TAG;
struct LocalTile_TAG;
struct TileElement_TAG;
}  // namespace matrix

using GlobalElementIndex = common::Index2D<SizeType, matrix::GlobalElement_TAG>;
using GlobalElementSize = common::Size2D<SizeType, matrix::GlobalElement_TAG>;
using LocalElementSize = common::Size2D<SizeType, matrix::LocalElement_TAG>;
using GlobalTileIndex = common::Index2D<SizeType, matrix::GlobalTile_TAG>;
using GlobalTileSize = common::Size2D<SizeType, matrix::GlobalTile_TAG>;
using LocalTileIndex = common::Index2D<SizeType, matrix::LocalTile_TAG>;
using LocalTileSize = common::Size2D<SizeType, matrix::LocalTile_TAG>;
using TileElementIndex = common::Index2D<SizeType, matrix::TileElement_TAG>;
using TileElementSize = common::Size2D<SizeType, matrix::TileElement_TAG>;

namespace comm {
using IndexT_MPI = int;
struct Comm_TAG;
using Index2D = common::Index2D<IndexT_MPI, Comm_TAG>;
using Size2D = common::Size2D<IndexT_MPI, Comm_TAG>;
}  // namespace comm

}  // namespace dlaf
