// mundy_b200/core.hpp — common pieces of the C++ veneer over the C-ABI (include/mundy_b200.h).
//
// The veneer keeps the reference's call signatures (MundyRepo/MuNDy @ 797a7f48) on top of libmundy_b200.so. It is
// header-only, needs no CUDA headers, and never computes on the host: every function forwards to a device entry point.
// Errors follow the reference's convention (mundy/core/src/mundy_core/throw_assert.hpp:156-212: MUNDY_THROW_REQUIRE throws
// on the host): a non-zero status becomes a C++ exception carrying mb200_last_error_string().
#ifndef MUNDY_B200_CORE_HPP_
#define MUNDY_B200_CORE_HPP_

#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../mundy_b200.h"

namespace mundy_b200 {

struct Error : std::runtime_error {
  int status;
  Error(int s, const std::string& where)
      : std::runtime_error(where + " failed with status " + std::to_string(s) + ": " + mb200_last_error_string()), status(s) {}
};
inline void check(int status, const char* where) {
  if (status != MB200_OK) throw Error(status, where);
}

/// Owning device array (the analogue of a Kokkos::View<T*> in device memory for code above the veneer).
template <typename T>
class DeviceVector {
 public:
  DeviceVector() = default;
  explicit DeviceVector(std::size_t n) { resize(n); }
  DeviceVector(const std::vector<T>& host, mb200_stream stream = nullptr) {
    resize(host.size());
    copy_from(host, stream);
  }
  DeviceVector(const DeviceVector&) = delete;
  DeviceVector& operator=(const DeviceVector&) = delete;
  DeviceVector(DeviceVector&& o) noexcept : p_(o.p_), n_(o.n_) { o.p_ = nullptr, o.n_ = 0; }
  DeviceVector& operator=(DeviceVector&& o) noexcept {
    std::swap(p_, o.p_), std::swap(n_, o.n_);
    return *this;
  }
  ~DeviceVector() { mb200_device_free(p_); }

  void resize(std::size_t n) {
    if (n == n_) return;
    check(mb200_device_free(p_), "mb200_device_free");
    p_ = nullptr, n_ = 0;
    void* q = nullptr;
    check(mb200_device_alloc(&q, static_cast<int64_t>(n * sizeof(T))), "mb200_device_alloc");
    p_ = static_cast<T*>(q), n_ = n;
  }
  void copy_from(const std::vector<T>& host, mb200_stream stream = nullptr) {
    if (host.size() != n_) resize(host.size());
    check(mb200_copy_to_device(stream, p_, host.data(), static_cast<int64_t>(n_ * sizeof(T))), "mb200_copy_to_device");
    check(mb200_stream_synchronize(stream), "mb200_stream_synchronize");
  }
  std::vector<T> to_host(mb200_stream stream = nullptr) const {
    std::vector<T> h(n_);
    check(mb200_copy_to_host(stream, h.data(), p_, static_cast<int64_t>(n_ * sizeof(T))), "mb200_copy_to_host");
    check(mb200_stream_synchronize(stream), "mb200_stream_synchronize");
    return h;
  }
  T* data() { return p_; }
  const T* data() const { return p_; }
  std::size_t size() const { return n_; }
  std::size_t extent(int) const { return n_; }  // Kokkos::View spelling

 private:
  T* p_ = nullptr;
  std::size_t n_ = 0;
};

}  // namespace mundy_b200
#endif  // MUNDY_B200_CORE_HPP_
