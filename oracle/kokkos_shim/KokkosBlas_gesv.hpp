// TEST INFRASTRUCTURE ONLY (oracle). Declaration-only stand-in: Periphery.hpp's dense inverse (not on the path) names
// KokkosBlas::gesv; nothing the oracle calls reaches it.
#pragma once
#include <stdexcept>
namespace KokkosBlas {
template <class A, class B, class P>
inline void gesv(const A&, const B&, const P&) {
  throw std::logic_error("KokkosBlas::gesv is not provided by the oracle shim");
}
}  // namespace KokkosBlas
