// TEST INFRASTRUCTURE ONLY (oracle). Minimal host-only stand-in for the parts of Kokkos that the reference's
// header-only hot-path code (mundy_core/throw_assert.hpp, mundy_math, mundy_geom, mundy_math/convex.hpp) touches, so that
// those headers can be compiled UNMODIFIED, from where they lie under /root/reference, into oracle/_ref/.
// Surface listed in SURVEY.md Appendix B. parallel_for / parallel_reduce are OpenMP loops (the loop structure the
// Kokkos OpenMP backend uses for a RangePolicy) when compiled with -fopenmp, serial otherwise.
#ifndef MB200_ORACLE_KOKKOS_SHIM_CORE_HPP_
#define MB200_ORACLE_KOKKOS_SHIM_CORE_HPP_

#include <algorithm>
#include <array>
#include <bit>
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <memory>
#include <numbers>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

#define KOKKOS_FUNCTION
#define KOKKOS_INLINE_FUNCTION inline
#define KOKKOS_FORCEINLINE_FUNCTION inline
#define KOKKOS_DEFAULTED_FUNCTION
#define KOKKOS_LAMBDA [=]
#define KOKKOS_CLASS_LAMBDA [ =, *this ]
#define KOKKOS_IF_ON_HOST(x) x
#define KOKKOS_IF_ON_DEVICE(x)
#define KOKKOS_IMPL_CTOR_DEFAULT_ARG
#define KOKKOS_VERSION 40301

namespace Kokkos {

using std::abs;
using std::ceil;
using std::copysign;
using std::cos;
using std::exp;
using std::fabs;
using std::floor;
using std::fmod;
using std::isfinite;
using std::isnan;
using std::log;
using std::pow;
using std::round;
using std::sin;
using std::sqrt;
using std::tan;
using std::acos;
using std::asin;
using std::atan2;
using std::pair;

template <class T>
constexpr const T& min(const T& a, const T& b) {
  return (b < a) ? b : a;
}
template <class T>
constexpr const T& max(const T& a, const T& b) {
  return (a < b) ? b : a;
}
// Mixed-type forms (the reference calls Kokkos::max(0.0, float_val) in one residual policy)
template <class T, class U>
  requires(!std::is_same_v<T, U>)
constexpr std::common_type_t<T, U> min(const T& a, const U& b) {
  using C = std::common_type_t<T, U>;
  return (static_cast<C>(b) < static_cast<C>(a)) ? static_cast<C>(b) : static_cast<C>(a);
}
template <class T, class U>
  requires(!std::is_same_v<T, U>)
constexpr std::common_type_t<T, U> max(const T& a, const U& b) {
  using C = std::common_type_t<T, U>;
  return (static_cast<C>(a) < static_cast<C>(b)) ? static_cast<C>(b) : static_cast<C>(a);
}

[[noreturn]] inline void abort(const char* msg) {
  std::fprintf(stderr, "Kokkos::abort: %s\n", msg);
  std::abort();
}

template <class To, class From>
constexpr To bit_cast(const From& f) {
  return std::bit_cast<To>(f);
}

namespace numbers {
template <class T>
inline constexpr T pi_v = std::numbers::pi_v<T>;
}

namespace Experimental {
template <class T>
inline constexpr T infinity_v = std::numeric_limits<T>::infinity();
template <class T>
inline constexpr T finite_min_v = std::numeric_limits<T>::lowest();
template <class T>
inline constexpr T finite_max_v = std::numeric_limits<T>::max();
template <class T>
inline constexpr T epsilon_v = std::numeric_limits<T>::epsilon();
template <class T>
inline constexpr T norm_min_v = std::numeric_limits<T>::min();
}  // namespace Experimental

struct HostSpace {};
struct Serial {
  using memory_space = HostSpace;
  using execution_space = Serial;
  void fence() const {}
};
using DefaultExecutionSpace = Serial;
using DefaultHostExecutionSpace = Serial;
using OpenMP = Serial;
inline void fence() {}
struct LayoutRight {};
struct LayoutLeft {};

template <class T, size_t N, class Proxy = void>
struct Array {
  T m_data[N > 0 ? N : 1];
  using value_type = T;
  constexpr T& operator[](size_t i) { return m_data[i]; }
  constexpr const T& operator[](size_t i) const { return m_data[i]; }
  static constexpr size_t size() { return N; }
  constexpr T* data() { return m_data; }
  constexpr const T* data() const { return m_data; }
};

// Serial "atomics" (only instantiated by accessor helpers the hot path never calls concurrently)
#define MB200_SHIM_ATOMIC(name, op)                          \
  template <class T, class U>                                \
  inline void atomic_##name(T* p, const U& v) {              \
    *p = *p op v;                                            \
  }                                                          \
  template <class T, class U>                                \
  inline T atomic_fetch_##name(T* p, const U& v) {           \
    T old = *p;                                              \
    *p = *p op v;                                            \
    return old;                                              \
  }                                                          \
  template <class T, class U>                                \
  inline T atomic_##name##_fetch(T* p, const U& v) {         \
    *p = *p op v;                                            \
    return *p;                                               \
  }
MB200_SHIM_ATOMIC(add, +)
MB200_SHIM_ATOMIC(sub, -)
MB200_SHIM_ATOMIC(mul, *)
MB200_SHIM_ATOMIC(div, /)
#undef MB200_SHIM_ATOMIC
template <class T>
inline T atomic_load(const T* p) {
  return *p;
}
template <class T, class U>
inline void atomic_store(T* p, const U& v) {
  *p = v;
}

}  // namespace Kokkos

#include "Kokkos_View_shim.hpp"
#include "Kokkos_Team_shim.hpp"

#endif  // MB200_ORACLE_KOKKOS_SHIM_CORE_HPP_
