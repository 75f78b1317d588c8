// common::Timer<clock>: seconds since construction (reference: include/dlaf/common/timer.h; the miniapps time the
// factorization with it between two barriers, miniapp_cholesky.cpp:141-153).
#pragma once
#include <chrono>

namespace dlaf::common {

template <class clock = std::chrono::high_resolution_clock>
class Timer {
public:
  Timer() : start_(clock::now()) {}
  double elapsed() const { return std::chrono::duration<double>(clock::now() - start_).count(); }

private:
  std::chrono::time_point<clock> start_;
};

}  // namespace dlaf::common
