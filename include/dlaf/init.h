// dlaf::initialize / finalize / ScopedInitializer (reference: include/dlaf/init.h, src/init.cpp:306-360).
// There is no task runtime to start: initialization checks for a CUDA device and records the configuration.
#pragma once
#include <dlaf_c/init.h>

#include "dlaf/tune.h"

namespace dlaf {

struct configuration {
  // knobs of the B200 build (the reference's stream-pool / Umpire knobs have no counterpart here)
  int lookahead = 1;          // steps of panel look-ahead
  int panel_workspaces = 2;   // reference: n_workspaces = 2, cholesky/impl.h:218-221
  bool print_config = false;
};

inline void initialize(int argc, const char* const argv[], const configuration& = {}) {
  dlaf_initialize(0, nullptr, argc, const_cast<const char**>(argv));
}
inline void initialize(const configuration& cfg = {}) {
  const char* args[] = {"--dlaf:print-config"};
  dlaf_initialize(0, nullptr, cfg.print_config ? 1 : 0, cfg.print_config ? args : nullptr);
}
inline void finalize() { dlaf_finalize(); }

struct [[nodiscard]] ScopedInitializer {
  ScopedInitializer(int argc, const char* const argv[], const configuration& cfg = {}) { initialize(argc, argv, cfg); }
  explicit ScopedInitializer(const configuration& cfg = {}) { initialize(cfg); }
  ~ScopedInitializer() { finalize(); }
  ScopedInitializer(const ScopedInitializer&) = delete;
  ScopedInitializer& operator=(const ScopedInitializer&) = delete;
};

}  // namespace dlaf
