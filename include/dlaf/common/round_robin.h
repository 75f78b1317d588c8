// common::RoundRobin<T>: a fixed pool of resources handed out cyclically (reference:
// include/dlaf/common/round_robin.h:21-59; the factorization keeps its two panel workspaces in one,
// cholesky/impl.h:218-221 - in this build those live inside the engine, the template is kept for callers' code).
#pragma once
#include <cstddef>
#include <vector>

#include "dlaf/types.h"

namespace dlaf::common {

template <class T>
class RoundRobin {
public:
  RoundRobin() = default;
  /// @p n resources, each constructed from (a copy of) @p args.
  template <class... Args>
  RoundRobin(std::size_t n, Args... args) {
    pool_.reserve(n);
    for (std::size_t i = 0; i < n; ++i)
      pool_.emplace_back(args...);
  }
  RoundRobin(RoundRobin&&) = default;
  RoundRobin& operator=(RoundRobin&&) = default;
  RoundRobin(const RoundRobin&) = delete;
  RoundRobin& operator=(const RoundRobin&) = delete;

  T& currentResource() {
    DLAF_ASSERT(current_ < pool_.size(), current_);
    return pool_[current_];
  }
  /// Advances to the next resource (wrapping around) and returns it.
  T& nextResource() {
    DLAF_ASSERT(!pool_.empty(), pool_.size());
    current_ = (current_ + 1) % pool_.size();
    return pool_[current_];
  }
  std::size_t size() const noexcept { return pool_.size(); }

private:
  std::size_t current_ = 0;
  std::vector<T> pool_;
};

}  // namespace dlaf::common
