// Library-wide plumbing: error string, launch counter, device check.
#include "common.cuh"

namespace mb200 {

static thread_local std::string t_error;
std::atomic<long long> g_launches{0};

void set_error(const std::string& msg) { t_error = msg; }

int require_device() {
  static std::atomic<int> ok{0};
  if (ok.load(std::memory_order_acquire)) return MB200_OK;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count <= 0) {
    (void)cudaGetLastError();
    set_error(std::string("no CUDA device available (") + (e != cudaSuccess ? cudaGetErrorString(e) : "count = 0") +
              "); mundy_b200 has no CPU fallback");
    return MB200_ERR_CUDA;
  }
  int dev = 0;
  MB200_CUDA(cudaGetDevice(&dev));
  // Keep stream-ordered scratch memory cached in the pool between calls (a step allocates the same sizes again).
  cudaMemPool_t pool;
  MB200_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
  uint64_t threshold = UINT64_MAX;
  MB200_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold));
  ok.store(1, std::memory_order_release);
  return MB200_OK;
}

}  // namespace mb200

extern "C" {

int mb200_abi_version(void) { return MB200_ABI_VERSION; }

const char* mb200_last_error_string(void) { return mb200::t_error.c_str(); }

int mb200_device_count(void) {
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess) {
    (void)cudaGetLastError();
    return 0;
  }
  return count;
}

int64_t mb200_kernel_launch_count(void) { return mb200::g_launches.load(std::memory_order_relaxed); }

}  // extern "C"
