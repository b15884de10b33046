// TEST INFRASTRUCTURE ONLY (oracle). Host stand-ins for the hierarchical-parallelism pieces of Kokkos that the
// reference's mobility kernels use (scrap/parameter_interface/alens/src/mundy_alens/periphery/Periphery.hpp:489-676:
// TeamPolicy panels over target points, ThreadVectorRange reduction over source points with a custom reducer,
// single(PerTeam) + atomic_add into the target velocities). Everything runs on one thread in index order, which is the
// order of the Kokkos Serial backend: the sum over sources of one target is taken for s = 0, 1, 2, ...
#ifndef MB200_ORACLE_KOKKOS_SHIM_TEAM_HPP_
#define MB200_ORACLE_KOKKOS_SHIM_TEAM_HPP_

#include <initializer_list>

namespace Kokkos {

struct MemoryUnmanaged {};
struct AUTO_t {};
inline constexpr AUTO_t AUTO{};

struct ShimTeamMember {
  int league_rank_{0}, league_size_{1};
  int league_rank() const { return league_rank_; }
  int league_size() const { return league_size_; }
  int team_rank() const { return 0; }
  int team_size() const { return 1; }
  void team_barrier() const {}
};

template <class... Props>
struct TeamPolicy {
  using member_type = ShimTeamMember;
  TeamPolicy(int league, AUTO_t) : league_size(league) {}
  TeamPolicy(int league, int) : league_size(league) {}
  TeamPolicy(int league, int, int) : league_size(league) {}
  template <class Exec>
  TeamPolicy(const Exec&, int league, AUTO_t) : league_size(league) {}
  int league_size;
};

struct ShimTeamThreadRange {
  long begin, end;
};
struct ShimThreadVectorRange {
  long begin, end;
};
template <class I1, class I2>
inline ShimTeamThreadRange TeamThreadRange(const ShimTeamMember&, I1 b, I2 e) {
  return {static_cast<long>(b), static_cast<long>(e)};
}
template <class I>
inline ShimTeamThreadRange TeamThreadRange(const ShimTeamMember&, I n) {
  return {0, static_cast<long>(n)};
}
template <class I>
inline ShimThreadVectorRange ThreadVectorRange(const ShimTeamMember&, I n) {
  return {0, static_cast<long>(n)};
}
template <class I1, class I2>
inline ShimThreadVectorRange ThreadVectorRange(const ShimTeamMember&, I1 b, I2 e) {
  return {static_cast<long>(b), static_cast<long>(e)};
}
struct ShimPerTeam {};
struct ShimPerThread {};
inline ShimPerTeam PerTeam(const ShimTeamMember&) { return {}; }
inline ShimPerThread PerThread(const ShimTeamMember&) { return {}; }

// League loop: teams run one after another (panels of target points are independent).
template <class... P, class F>
inline void parallel_for(const std::string&, const TeamPolicy<P...>& pol, const F& f) {
  for (int l = 0; l < pol.league_size; ++l) f(ShimTeamMember{l, pol.league_size});
}
template <class... P, class F>
inline void parallel_for(const TeamPolicy<P...>& pol, const F& f) {
  parallel_for(std::string(), pol, f);
}
template <class F>
inline void parallel_for(const ShimTeamThreadRange& r, const F& f) {
  for (long i = r.begin; i < r.end; ++i) f(static_cast<int>(i));
}
template <class F>
inline void parallel_for(const ShimThreadVectorRange& r, const F& f) {
  for (long i = r.begin; i < r.end; ++i) f(static_cast<int>(i));
}
// Custom-reducer form (reducer exposes init / join / reference, Kokkos "ReducerConcept"): the Serial backend
// initialises the referenced value and accumulates into it in index order.
template <class F, class Reducer>
  requires requires(const Reducer& r) { r.reference(); }
inline void parallel_reduce(const ShimThreadVectorRange& r, const F& f, const Reducer& red) {
  red.init(red.reference());
  for (long i = r.begin; i < r.end; ++i) f(static_cast<int>(i), red.reference());
}
template <class F, class T>
  requires std::is_arithmetic_v<T>
inline void parallel_reduce(const ShimThreadVectorRange& r, const F& f, T& result) {
  T acc = 0;
  for (long i = r.begin; i < r.end; ++i) f(static_cast<int>(i), acc);
  result = acc;
}
template <class F>
inline void single(const ShimPerTeam&, const F& f) {
  f();
}
template <class F>
inline void single(const ShimPerThread&, const F& f) {
  f();
}

template <unsigned N>
struct Rank {};
template <class... Props>
struct MDRangePolicy {
  template <class A, class B>
  MDRangePolicy(std::initializer_list<A> lo, std::initializer_list<B> hi) {
    int k = 0;
    for (auto v : lo) begin[k++] = static_cast<long>(v);
    k = 0;
    for (auto v : hi) end[k++] = static_cast<long>(v);
  }
  long begin[2]{0, 0}, end[2]{0, 0};
};
template <class... P, class F>
inline void parallel_for(const std::string&, const MDRangePolicy<P...>& p, const F& f) {
  for (long i = p.begin[0]; i < p.end[0]; ++i)
    for (long j = p.begin[1]; j < p.end[1]; ++j) f(static_cast<size_t>(i), static_cast<size_t>(j));
}

}  // namespace Kokkos

#endif  // MB200_ORACLE_KOKKOS_SHIM_TEAM_HPP_
