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
  std::vector<T> items_;
  std::size_t at_ = 0;  // position of the resource handed out last

  void step() noexcept { at_ = at_ + 1 == items_.size() ? 0 : at_ + 1; }

public:
  RoundRobin() = default;
  /// @p count resources, every one constructed from the same arguments.
  template <class... CtorArgs>
  explicit RoundRobin(std::size_t count, const CtorArgs&... ctor_args) {
    items_.reserve(count);
    while (items_.size() < count)
      items_.emplace_back(ctor_args...);
  }
  RoundRobin(const RoundRobin&) = delete;
  RoundRobin& operator=(const RoundRobin&) = delete;
  RoundRobin(RoundRobin&&) = default;
  RoundRobin& operator=(RoundRobin&&) = default;

  std::size_t size() const noexcept { return items_.size(); }

  /// The resource handed out last (the first one before any call of nextResource).
  T& currentResource() {
    DLAF_ASSERT(!items_.empty(), items_.size());
    return items_[at_];
  }
  /// Moves on to the following resource, wrapping around after the last, and returns it.
  T& nextResource() {
    DLAF_ASSERT(!items_.empty(), items_.size());
    step();
    return items_[at_];
  }
};

}  // namespace dlaf::common
