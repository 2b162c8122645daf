// TEST INFRASTRUCTURE ONLY.  Stand-in for eigenmath/scalar_utils.h (x_edr_eigenmath 1.0.0).
#ifndef ORACLE_SHIMS_EIGENMATH_SCALAR_UTILS_H_
#define ORACLE_SHIMS_EIGENMATH_SCALAR_UTILS_H_
#include <cmath>
namespace eigenmath {
template <typename T> constexpr T Square(T x) { return x * x; }
// radians * 180 / pi evaluated in double and rounded to T (same restatement as oracle/pcseg_oracle.cc)
template <typename T> constexpr T Degrees(T radians) { return static_cast<T>(static_cast<double>(radians) * (180.0 / M_PI)); }
template <typename T> constexpr T Radians(T degrees) { return degrees * static_cast<T>(M_PI) / static_cast<T>(180); }
}  // namespace eigenmath
#endif
