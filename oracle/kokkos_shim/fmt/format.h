// TEST INFRASTRUCTURE ONLY (oracle). Periphery.hpp includes <fmt/format.h> for error strings only.
#pragma once
#include <string>
namespace fmt {
template <class... A>
inline std::string format(const char* s, A&&...) {
  return s;
}
}  // namespace fmt
