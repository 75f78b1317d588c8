// dlaf::TuneParameters / getTuneParameters(): the reference's run-time knobs (include/dlaf/tune.h, src/tune.cpp) are
// about pika pools, MPI pipelines and eigensolver block sizes - none of which exist here. This build's knobs are the
// ones its kernels and schedulers actually read; each is an environment variable of the library (so that C, Fortran and
// Python callers reach them too) and a field here. Fields are initialised from the environment; apply() exports them
// again. Knobs marked [process] are read once per process by the library (set them before the first call), the others
// when a plan is created.
#pragma once
#include <cstdlib>
#include <string>

namespace dlaf {

struct TuneParameters {
  /// DLAFB_P2P: panels travel by copies into the peers' memory over NVLink (1, default when all peers can be mapped) or
  /// by ncclBroadcast (0). Read at plan creation.
  int peer_memory_exchange = env_int("DLAFB_P2P", 1);
  /// DLAFB_BULK_SHAPE [process]: CTA block of the bulk FP64 update, 3 = 64x128 with two CTAs per SM (default), 0 = 128x128.
  int bulk_update_shape = env_int("DLAFB_BULK_SHAPE", 3);
  /// DLAFB_TRSOLVE_LOOKAHEAD: two-stream look-ahead variant of the one-GPU triangular solver. Read at plan creation.
  int triangular_solver_lookahead = env_int("DLAFB_TRSOLVE_LOOKAHEAD", 0);
  /// DLAFB_G2S_LOOKAHEAD: two-stream look-ahead variant of the one-GPU generalized-to-standard reduction. Read at plan
  /// creation.
  int gen_to_std_lookahead = env_int("DLAFB_G2S_LOOKAHEAD", 0);
  /// DLAFB_SGEMM_WIDE [process]: float contractions on the 128x128 kernel (default since it was timed: 38 vs 27 TFLOP/s
  /// for the n = 32768 factorization); 0 selects the 64x64 kernel.
  int sgemm_wide = env_int("DLAFB_SGEMM_WIDE", 1);
  /// DLAFB_POTRF_REDISTRIBUTE: dlaf_p?potrf redistributes matrices with small block sizes to <= 512, factorizes and
  /// redistributes back. -1 (unset): device pointers with nb < 128 only; 1: every call with nb < 256; 0: never. Read at
  /// every call.
  int potrf_redistribute = env_int("DLAFB_POTRF_REDISTRIBUTE", -1);

  /// Exports the fields to the environment variables the library reads.
  void apply() const {
    set("DLAFB_P2P", peer_memory_exchange);
    set("DLAFB_BULK_SHAPE", bulk_update_shape);
    set("DLAFB_TRSOLVE_LOOKAHEAD", triangular_solver_lookahead);
    set("DLAFB_G2S_LOOKAHEAD", gen_to_std_lookahead);
    set("DLAFB_SGEMM_WIDE", sgemm_wide);
    set("DLAFB_POTRF_REDISTRIBUTE", potrf_redistribute);
  }

private:
  static int env_int(const char* name, int fallback) {
    const char* v = std::getenv(name);
    return v && *v ? std::atoi(v) : fallback;
  }
  static void set(const char* name, int value) { setenv(name, std::to_string(value).c_str(), 1); }
};

/// The process-wide parameters (reference: getTuneParameters(), include/dlaf/tune.h). Modify, then call apply().
inline TuneParameters& getTuneParameters() {
  static TuneParameters params;
  return params;
}

}  // namespace dlaf
