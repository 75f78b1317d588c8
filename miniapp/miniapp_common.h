// Shared by miniapp_triangular_solver and miniapp_gen_to_std: option parsing in the reference's spelling
// (miniapp/include/dlaf/miniapp/options.h), device-side fill. (Rank bootstrap: dlaf/communication/init.h.)
#pragma once
#include <unistd.h>

#include <complex>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include <dlaf/init.h>
#include <dlaf/matrix/matrix.h>
#include <dlafb_kernels.h>

namespace miniapp {

struct CommonOptions {
  int grid_rows = 1, grid_cols = 1;
  int64_t nruns = 1, nwarmups = 1;  // options.h:252-255
  char type = 'd';
  std::string check = "none";
  bool csv = false, local = false;
};

// Returns the value of --name (either "--name value" or "--name=value") or nullptr; advances i when it consumes two.
inline const char* option_value(int argc, char** argv, int& i, const char* name) {
  const std::string a = argv[i], n = std::string("--") + name;
  if (a == n && i + 1 < argc)
    return argv[++i];
  if (a.rfind(n + "=", 0) == 0)
    return argv[i] + n.size() + 1;
  return nullptr;
}

inline bool parse_common(int argc, char** argv, int& i, CommonOptions& o, const char* app) {
  const char* v;
  const std::string a = argv[i];
  if ((v = option_value(argc, argv, i, "grid-rows"))) o.grid_rows = std::atoi(v);
  else if ((v = option_value(argc, argv, i, "grid-cols"))) o.grid_cols = std::atoi(v);
  else if ((v = option_value(argc, argv, i, "nruns"))) o.nruns = std::atoll(v);
  else if ((v = option_value(argc, argv, i, "nwarmups"))) o.nwarmups = std::atoll(v);
  else if ((v = option_value(argc, argv, i, "type"))) o.type = v[0];
  else if ((v = option_value(argc, argv, i, "check-result"))) o.check = v;
  else if ((v = option_value(argc, argv, i, "backend"))) {
    if (std::string(v) == "mc") {
      std::fprintf(stderr, "%s: this build has no MC (CPU) backend\n", app);
      std::exit(2);
    }
  }
  else if (a == "--csv") o.csv = true;
  else if (a == "--local") o.local = true;
  else return false;
  return true;
}

#define MINIAPP_GEN(T, CT, X)                                                                                   \
  inline int k_gen(void* s, int64_t n, int nb, int r, int c, int rf, int rs, int cf, int cs, T* a, int64_t ld,  \
                   uint64_t seed) {                                                                             \
    return dlafb_gen_hpd_##X(s, n, nb, r, c, rf, rs, cf, cs, (CT*)a, ld, seed);                                 \
  }
MINIAPP_GEN(float, float, s)
MINIAPP_GEN(double, double, d)
MINIAPP_GEN(std::complex<float>, void, c)
MINIAPP_GEN(std::complex<double>, void, z)
#undef MINIAPP_GEN

// Counter-based fill on the device (timing input): |off-diagonal| <= 1, diagonal = u + 2n, Hermitian by construction -
// a well-conditioned triangular factor when one triangle is read, a Hermitian matrix when mirrored.
template <class T>
void generate(dlaf::Matrix<T, dlaf::Device::GPU>& m, uint64_t seed) {
  using dlaf::Coord;
  const auto& d = m.distribution();
  const auto ls = d.local_size();
  const int64_t n = d.size().rows() > d.size().cols() ? d.size().rows() : d.size().cols();
  k_gen(m.stream(), n, (int)d.block_size().rows(), (int)ls.rows(), (int)ls.cols(),
        (int)d.template global_tile_from_local_tile<Coord::Row>(0), d.grid_size().rows(),
        (int)d.template global_tile_from_local_tile<Coord::Col>(0), d.grid_size().cols(), m.ptr(), m.ld(), seed);
}

inline const char* type_name(char t) { return t == 's' ? "S" : t == 'd' ? "D" : t == 'c' ? "C" : "Z"; }

}  // namespace miniapp
