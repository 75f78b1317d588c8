// matrix::Matrix<T, Device>: the local block of a (possibly distributed) matrix plus its Distribution; constructor
// set and accessors follow the reference (include/dlaf/matrix/matrix.h:58-338, matrix_base.h). The reference hands
// out per-tile senders ordered by async_rw_mutex pipelines; here every Matrix owns a CUDA stream and operations on
// the matrix are ordered on it: algorithms return after enqueueing, waitLocalTiles() blocks until they are done.
#pragma once
#include <cuda_runtime.h>

#include <functional>
#include <memory>
#include <vector>
#include <utility>

#include "dlaf/communication/communicator_grid.h"
#include "dlaf/matrix/layout_info.h"

namespace dlaf {
namespace matrix {

/// Lightweight view of one tile (reference: matrix::Tile, include/dlaf/matrix/tile.h:203-460).
template <class T, Device D>
class Tile {
public:
  Tile(T* ptr, SizeType ld, TileElementSize size) : ptr_(ptr), ld_(ld), size_(size) {}
  T* ptr() const noexcept { return ptr_; }
  T* ptr(const TileElementIndex& i) const noexcept { return ptr_ + i.row() + i.col() * ld_; }
  T& operator()(const TileElementIndex& i) const noexcept { return *ptr(i); }  // Device::CPU only
  SizeType ld() const noexcept { return ld_; }
  const TileElementSize& size() const noexcept { return size_; }

private:
  T* ptr_;
  SizeType ld_;
  TileElementSize size_;
};

template <class T, Device D>
class Matrix {
public:
  using ElementType = T;

  /// Local matrix, memory allocated here (default layout: column-major, ld rounded up to 64, matrix.h:87-89).
  Matrix(const LocalElementSize& size, const TileElementSize& tile_size)
      : Matrix(Distribution(size, tile_size)) {}
  /// Distributed matrix on @p comm, first tile on rank (0, 0), memory allocated here.
  Matrix(const GlobalElementSize& size, const TileElementSize& tile_size, const comm::CommunicatorGrid& comm)
      : Matrix(Distribution(size, tile_size, comm.size(), comm.rank(), comm::Index2D(0, 0))) {}
  explicit Matrix(Distribution distribution)
      : Matrix(distribution, colMajorLayout(distribution, default_ld(distribution))) {}
  Matrix(Distribution distribution, const LayoutInfo& layout) : dist_(std::move(distribution)), layout_(layout) {
    DLAF_ASSERT(layout.is_col_major(), "only column-major local blocks are supported on this path");
    allocate();
  }
  /// Wraps user memory (not copied, not freed).
  Matrix(const LayoutInfo& layout, T* ptr) noexcept
      : Matrix(Distribution(layout.size(), layout.block_size()), layout, ptr) {}
  Matrix(Distribution distribution, const LayoutInfo& layout, T* ptr) noexcept
      : dist_(std::move(distribution)), layout_(layout), ptr_(ptr) {
    DLAF_ASSERT(layout.is_col_major(), "only column-major local blocks are supported on this path");
    cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking);
  }

  Matrix(const Matrix&) = delete;
  Matrix& operator=(const Matrix&) = delete;
  Matrix(Matrix&& o) noexcept : dist_(o.dist_), layout_(o.layout_) { swap(o); }
  Matrix& operator=(Matrix&& o) noexcept {
    swap(o);
    return *this;
  }
  ~Matrix() {
    if (stream_) {
      cudaStreamSynchronize(stream_);
      keep_alive_.clear();
      cudaStreamDestroy(stream_);
    }
    if (owned_) {
      if (D == Device::GPU)
        cudaFree(owned_);
      else
        cudaFreeHost(owned_);
    }
  }

  const Distribution& distribution() const noexcept { return dist_; }
  const GlobalElementSize& size() const noexcept { return dist_.size(); }
  const TileElementSize& block_size() const noexcept { return dist_.block_size(); }
  const TileElementSize& blockSize() const noexcept { return dist_.block_size(); }
  const TileElementSize& tile_size() const noexcept { return dist_.block_size(); }
  GlobalTileSize nr_tiles() const noexcept { return dist_.nr_tiles(); }
  GlobalTileSize nrTiles() const noexcept { return dist_.nr_tiles(); }
  const comm::Index2D& rank_index() const noexcept { return dist_.rank_index(); }
  const comm::Index2D& rankIndex() const noexcept { return dist_.rank_index(); }
  const comm::Size2D& grid_size() const noexcept { return dist_.grid_size(); }
  const comm::Size2D& commGridSize() const noexcept { return dist_.grid_size(); }

  /// Local block: base pointer and leading dimension.
  T* ptr() const noexcept { return ptr_; }
  SizeType ld() const noexcept { return layout_.ld_tile(); }
  const LayoutInfo& layout() const noexcept { return layout_; }
  /// Stream that orders every operation on this matrix.
  cudaStream_t stream() const noexcept { return stream_; }

  /// Tile access (reference: read / readwrite return senders of tiles). The handle is valid immediately; its data
  /// is ordered after the work already enqueued on stream(): call waitLocalTiles() before touching it on the host.
  Tile<const T, D> read(const LocalTileIndex& index) const noexcept {
    return {ptr_ + layout_.tile_offset(index), ld(), dist_.tile_size_of(dist_.global_tile_index(index))};
  }
  Tile<T, D> readwrite(const LocalTileIndex& index) noexcept {
    return {ptr_ + layout_.tile_offset(index), ld(), dist_.tile_size_of(dist_.global_tile_index(index))};
  }
  Tile<const T, D> read(const GlobalTileIndex& index) const noexcept { return read(dist_.local_tile_index(index)); }
  Tile<T, D> readwrite(const GlobalTileIndex& index) noexcept { return readwrite(dist_.local_tile_index(index)); }

  /// Blocks until every operation enqueued on the local tiles has completed (matrix.h:322-338).
  void waitLocalTiles() noexcept {
    cudaStreamSynchronize(stream_);
    // completion hooks of the algorithms enqueued since the last wait (e.g. the factorization's info check), in
    // order; then the resources they parked are released - nothing on the stream can still be using them
    auto hooks = std::move(on_wait_);
    on_wait_.clear();
    for (auto& f : hooks)
      f();  // the hooks may use the parked resources (the factorization's plan): release those afterwards
    keep_alive_.clear();
  }

  /// Internal: algorithms park resources that must outlive the enqueued work, and a completion hook. Several
  /// algorithms may be enqueued on the same matrix before it is waited for: everything is kept until then.
  void keep_alive(std::shared_ptr<void> p, std::function<void()> on_wait = nullptr) {
    if (p)
      keep_alive_.push_back(std::move(p));
    if (on_wait)
      on_wait_.push_back(std::move(on_wait));
  }

private:
  static SizeType default_ld(const Distribution& d) {
    const SizeType r = d.local_size().rows();
    return r > 0 ? (r + 63) / 64 * 64 : 1;
  }
  void allocate() {
    const LocalElementSize ls = dist_.local_size();
    const size_t elems = (size_t)layout_.ld_tile() * (size_t)(ls.cols() > 0 ? ls.cols() : 1) + 64;
    cudaError_t e = (D == Device::GPU) ? cudaMalloc(&owned_, elems * sizeof(T)) : cudaMallocHost(&owned_, elems * sizeof(T));
    DLAF_ASSERT(e == cudaSuccess, e);
    ptr_ = static_cast<T*>(owned_);
    cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking);
  }
  void swap(Matrix& o) noexcept {
    std::swap(dist_, o.dist_);
    std::swap(layout_, o.layout_);
    std::swap(ptr_, o.ptr_);
    std::swap(owned_, o.owned_);
    std::swap(stream_, o.stream_);
    std::swap(keep_alive_, o.keep_alive_);
    std::swap(on_wait_, o.on_wait_);
  }

  Distribution dist_;
  LayoutInfo layout_;
  T* ptr_ = nullptr;
  void* owned_ = nullptr;
  cudaStream_t stream_ = nullptr;
  std::vector<std::shared_ptr<void>> keep_alive_;
  std::vector<std::function<void()>> on_wait_;
};

/// Copies the local block between two matrices with the same distribution, on the destination's stream
/// (reference: matrix::copy, include/dlaf/matrix/copy.h:38-80; per-tile cudaMemcpy2DAsync there, one 2D copy here).
template <class T, Device Ds, Device Dd>
void copy(Matrix<T, Ds>& source, Matrix<T, Dd>& dest) {
  DLAF_ASSERT(source.distribution() == dest.distribution(), "copy: distributions differ");
  const LocalElementSize ls = source.distribution().local_size();
  if (ls.isEmpty())
    return;
  source.waitLocalTiles();
  const cudaMemcpyKind kind = Ds == Device::CPU ? (Dd == Device::CPU ? cudaMemcpyHostToHost : cudaMemcpyHostToDevice)
                                                : (Dd == Device::CPU ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice);
  cudaError_t e = cudaMemcpy2DAsync(dest.ptr(), dest.ld() * sizeof(T), source.ptr(), source.ld() * sizeof(T),
                                    ls.rows() * sizeof(T), ls.cols(), kind, dest.stream());
  DLAF_ASSERT(e == cudaSuccess, e);
}

/// Mirror of a matrix on another device: copy-in on construction, copy-back on destruction (reference:
/// matrix::MatrixMirror, include/dlaf/matrix/matrix_mirror.h:32-177).
template <class T, Device Target, Device Source>
class MatrixMirror {
public:
  explicit MatrixMirror(Matrix<T, Source>& source) : source_(source), target_(source.distribution()) {
    copy(source_, target_);
  }
  ~MatrixMirror() { copyTargetToSource(); }
  Matrix<T, Target>& get() noexcept { return target_; }
  void copySourceToTarget() { copy(source_, target_); }
  void copyTargetToSource() {
    copy(target_, source_);
    source_.waitLocalTiles();
  }

private:
  Matrix<T, Source>& source_;
  Matrix<T, Target> target_;
};

}  // namespace matrix

using matrix::Matrix;

}  // namespace dlaf
