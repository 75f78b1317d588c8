// Basic types of the C++ API; names follow the reference (include/dlaf/types.h).
#pragma once
#include <complex>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

namespace dlaf {

using SizeType = std::ptrdiff_t;

enum class Device { CPU, GPU, Default = GPU };
enum class Backend { MC, GPU, Default = GPU };

template <Backend B>
struct DefaultDevice {
  static constexpr Device value = (B == Backend::GPU) ? Device::GPU : Device::CPU;
};
template <Backend B>
inline constexpr Device DefaultDevice_v = DefaultDevice<B>::value;

// flop weights of one add / mul (include/dlaf/types.h:115-133)
template <class T>
struct TypeInfo {
  using BaseType = T;
  static constexpr int ops_add = 1;
  static constexpr int ops_mul = 1;
  static constexpr bool is_complex = false;
};
template <class T>
struct TypeInfo<std::complex<T>> {
  using BaseType = T;
  static constexpr int ops_add = 2;
  static constexpr int ops_mul = 6;
  static constexpr bool is_complex = true;
};
template <class T>
using BaseType = typename TypeInfo<T>::BaseType;

// total_ops<T>(adds, muls): include/dlaf/types.h:159-162
template <class T>
constexpr double total_ops(const double add, const double mul) {
  return TypeInfo<T>::ops_add * add + TypeInfo<T>::ops_mul * mul;
}

namespace internal {
template <class T>
struct TypeCode;
template <>
struct TypeCode<float> { static constexpr char value = 's'; };
template <>
struct TypeCode<double> { static constexpr char value = 'd'; };
template <>
struct TypeCode<std::complex<float>> { static constexpr char value = 'c'; };
template <>
struct TypeCode<std::complex<double>> { static constexpr char value = 'z'; };
}  // namespace internal

// DLAF_ASSERT convention: print and terminate (include/dlaf/common/assert.h)
#define DLAF_ASSERT(cond, ...)                                                              \
  do {                                                                                      \
    if (!(cond)) {                                                                          \
      std::fprintf(stderr, "[ERROR] %s:%d: assertion failed: %s\n", __FILE__, __LINE__, #cond); \
      std::abort();                                                                         \
    }                                                                                       \
  } while (0)

}  // namespace dlaf

namespace blas {
enum class Uplo : char { Upper = 'U', Lower = 'L', General = 'G' };
enum class Side : char { Left = 'L', Right = 'R' };
enum class Op : char { NoTrans = 'N', Trans = 'T', ConjTrans = 'C' };
enum class Diag : char { NonUnit = 'N', Unit = 'U' };
}
