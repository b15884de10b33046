// TEST INFRASTRUCTURE ONLY (oracle). Host View / RangePolicy / parallel_* / KokkosBlas::gemv stand-ins used by
// mundy_math/convex.hpp's KokkosBackend (reference: mundy/math/src/mundy_math/convex.hpp:141-285).
#ifndef MB200_ORACLE_KOKKOS_SHIM_VIEW_HPP_
#define MB200_ORACLE_KOKKOS_SHIM_VIEW_HPP_

namespace Kokkos {

struct WithoutInitializing_t {};
inline constexpr WithoutInitializing_t WithoutInitializing{};
struct ViewAllocLabel {
  std::string label;
};
template <class... Args>
inline ViewAllocLabel view_alloc(Args&&...) {
  return {};
}

namespace shim_impl {
template <class T>
struct view_traits;
template <class T>
struct view_traits<T*> {
  using value_type = T;
  static constexpr int rank = 1;
};
template <class T>
struct view_traits<T**> {
  using value_type = T;
  static constexpr int rank = 2;
};
}  // namespace shim_impl

/// Row-major (LayoutRight) shared-ownership host view of rank 1 or 2.
template <class DataType, class... Props>
class View {
 public:
  using traits = shim_impl::view_traits<DataType>;
  using value_type = typename traits::value_type;
  using non_const_value_type = std::remove_const_t<value_type>;
  using memory_space = HostSpace;
  using execution_space = Serial;
  using array_layout = LayoutRight;
  using HostMirror = View;
  static constexpr int rank = traits::rank;

  View() = default;
  explicit View(const std::string&, size_t n0 = 0, size_t n1 = 1) : n0_(n0), n1_(rank == 2 ? n1 : 1) { alloc(); }
  explicit View(const ViewAllocLabel&, size_t n0 = 0, size_t n1 = 1) : n0_(n0), n1_(rank == 2 ? n1 : 1) { alloc(); }
  View(value_type* ptr, size_t n0, size_t n1 = 1) : n0_(n0), n1_(rank == 2 ? n1 : 1), raw_(ptr) {}

  size_t extent(int d) const { return d == 0 ? n0_ : (d == 1 ? n1_ : 1); }
  size_t size() const { return n0_ * n1_; }
  size_t span() const { return size(); }
  value_type* data() const { return raw_; }
  value_type& operator()(size_t i) const { return raw_[i]; }
  value_type& operator()(size_t i, size_t j) const { return raw_[i * n1_ + j]; }
  value_type& operator[](size_t i) const { return raw_[i]; }
  std::string label() const { return ""; }

 private:
  void alloc() {
    store_ = std::shared_ptr<non_const_value_type[]>(new non_const_value_type[n0_ * n1_ > 0 ? n0_ * n1_ : 1]());
    raw_ = store_.get();
  }
  size_t n0_{0}, n1_{1};
  std::shared_ptr<non_const_value_type[]> store_;
  value_type* raw_{nullptr};
};

template <class T>
struct is_view : std::false_type {};
template <class D, class... P>
struct is_view<View<D, P...>> : std::true_type {};
template <class D, class... P>
struct is_view<const View<D, P...>> : std::true_type {};
template <class T>
inline constexpr bool is_view_v = is_view<std::remove_reference_t<T>>::value;

template <class V>
inline V create_mirror_view(const V& v) {
  return v;
}
template <class V>
inline V create_mirror_view_and_copy(HostSpace, const V& v) {
  return v;
}

template <class D1, class... P1, class D2, class... P2>
inline void deep_copy(const View<D1, P1...>& dst, const View<D2, P2...>& src) {
  if (dst.data() == src.data()) return;
  for (size_t i = 0; i < dst.size(); ++i) dst.data()[i] = src.data()[i];
}
template <class D1, class... P1, class S>
  requires std::is_arithmetic_v<S>
inline void deep_copy(const View<D1, P1...>& dst, const S& value) {
  for (size_t i = 0; i < dst.size(); ++i) dst.data()[i] = value;
}

template <class... Props>
struct RangePolicy {
  RangePolicy(size_t b, size_t e) : begin(b), end(e) {}
  template <class Exec>
  RangePolicy(const Exec&, size_t b, size_t e) : begin(b), end(e) {}
  size_t begin, end;
};

template <class T>
struct Max {
  using value_type = T;
  explicit Max(T& r) : ref(r) {}
  T& ref;
};
template <class T>
struct Sum {
  using value_type = T;
  explicit Sum(T& r) : ref(r) {}
  T& ref;
};

template <class Policy, class F>
inline void parallel_for(const std::string&, const Policy& p, const F& f) {
  const long b = static_cast<long>(p.begin), e = static_cast<long>(p.end);
#pragma omp parallel for schedule(static)
  for (long i = b; i < e; ++i) f(static_cast<int>(i));
}
template <class Policy, class F>
inline void parallel_for(const Policy& p, const F& f) {
  parallel_for("", p, f);
}

// Sum-to-scalar form
template <class Policy, class F, class T>
  requires std::is_arithmetic_v<T>
inline void parallel_reduce(const std::string&, const Policy& p, const F& f, T& result) {
  const long b = static_cast<long>(p.begin), e = static_cast<long>(p.end);
  T acc = 0;
#pragma omp parallel for schedule(static) reduction(+ : acc)
  for (long i = b; i < e; ++i) f(static_cast<int>(i), acc);
  result = acc;
}
// Max-reducer form
template <class Policy, class F, class T>
inline void parallel_reduce(const std::string&, const Policy& p, const F& f, Max<T> red) {
  const long b = static_cast<long>(p.begin), e = static_cast<long>(p.end);
  T acc = std::numeric_limits<T>::lowest();
#pragma omp parallel for schedule(static) reduction(max : acc)
  for (long i = b; i < e; ++i) f(static_cast<int>(i), acc);
  red.ref = acc;
}

// Three-Sum-reducer form (scrap/lcp_spheres/KokkosLCP.cpp:536-556). Serial and in index order: the only caller is the
// driver-LCP pin (oracle/ref_lcp.cpp), which wants the Kokkos Serial backend's order.
template <class Policy, class F, class T>
inline void parallel_reduce(const std::string&, const Policy& p, const F& f, Sum<T> r0, Sum<T> r1, Sum<T> r2) {
  T a0 = 0, a1 = 0, a2 = 0;
  for (long i = static_cast<long>(p.begin); i < static_cast<long>(p.end); ++i) f(static_cast<int>(i), a0, a1, a2);
  r0.ref = a0, r1.ref = a1, r2.ref = a2;
}

}  // namespace Kokkos

namespace KokkosBlas {
/// y = alpha * op(A) x + beta * y ; only trans == "N" is used by the reference (convex.hpp:173,183).
template <class Exec, class AView, class XView, class YView, class S1, class S2>
inline void gemv(const Exec&, const char* trans, S1 alpha, const AView& A, const XView& x, S2 beta, const YView& y) {
  (void)trans;
  const long m = static_cast<long>(A.extent(0)), n = static_cast<long>(A.extent(1));
#pragma omp parallel for schedule(static)
  for (long i = 0; i < m; ++i) {
    typename YView::non_const_value_type acc = 0;
    for (long j = 0; j < n; ++j) acc += A(i, j) * x(j);
    y(i) = (beta == S2(0)) ? alpha * acc : alpha * acc + beta * y(i);
  }
}
}  // namespace KokkosBlas

#endif  // MB200_ORACLE_KOKKOS_SHIM_VIEW_HPP_
