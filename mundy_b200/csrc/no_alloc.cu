// mb200_narrow_phase_no_alloc: the narrow phase for callers that may not allocate (CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc). No counterpart in the reference.
//
// mb200_narrow_phase packs the caller's bodies into two arrays of 32-byte records (stream-ordered scratch from the
// cudaMallocAsync pool) before the pair kernel gathers them. Here those two arrays live in ONE file-scope device array
// of a fixed size: every GPU that loads the library holds it (4 MB) until exit, the entry point refuses more than
// MB200_NO_ALLOC_MAX_BODIES bodies, and nothing on this path calls cudaMalloc, cudaMallocAsync, cudaFree or a library
// routine that allocates. The kernels are the ones of narrow_phase.cu, unchanged.
//
// The array is shared by every call on a device and keeps the previous call's records. Nothing has to start at zero:
// a call writes records [0, n) of each half it uses before its pair kernel reads them (pairs are range-checked by the
// caller's contract, as in mb200_narrow_phase), so stale records are never read. Two calls that overlap on the GPU
// (different streams, or different host threads on different streams) race on the array; calls on one stream do not.
#include <mutex>

#include "internal.cuh"

namespace mb200 {
namespace {

constexpr int64_t kMaxBodies = MB200_NO_ALLOC_MAX_BODIES;
constexpr int kMaxDevices = 64;

// {b0[kMaxBodies], b1[kMaxBodies]}: 2 x 32 B x 65536 = 4 MB.
// In host code this name is only a handle for cudaGetSymbolAddress, never a device address.
static __device__ __align__(256) double g_body_records[2 * 4 * kMaxBodies];

// Device address of g_body_records on the CURRENT device, looked up once per device with that device current.
int body_records(D4** out) {
  static std::mutex mu;
  static D4* addr[kMaxDevices] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    set_error(std::string("no CUDA device available (") + cudaGetErrorString(e) + "); mundy_b200 has no CPU fallback");
    return MB200_ERR_CUDA;
  }
  if (dev < 0 || dev >= kMaxDevices) {
    set_error("mb200_narrow_phase_no_alloc: device ordinal " + std::to_string(dev) + " out of range");
    return MB200_ERR_CUDA;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    MB200_CUDA(cudaGetSymbolAddress(&p, g_body_records));
    if (!p) {
      set_error("mb200_narrow_phase_no_alloc: cudaGetSymbolAddress returned a null address");
      return MB200_ERR_CUDA;
    }
    addr[dev] = static_cast<D4*>(p);
  }
  *out = addr[dev];
  return MB200_OK;
}

}  // namespace
}  // namespace mb200

using namespace mb200;

extern "C" int mb200_narrow_phase_no_alloc(mb200_stream stream, const mb200_bodies* bodies, int64_t npairs,
                                           const int32_t* pairs, double* phi, double* normal, double* arm_i,
                                           double* arm_j) {
  MB200_REQUIRE(bodies && bodies->n >= 0 && npairs >= 0, "bad sizes");
  if (bodies->n > kMaxBodies) {
    set_error("invalid argument: mb200_narrow_phase_no_alloc holds at most " + std::to_string(kMaxBodies) +
              " bodies (MB200_NO_ALLOC_MAX_BODIES) in its static scratch; got " + std::to_string(bodies->n) +
              " (use mb200_narrow_phase)");
    return MB200_ERR_INVALID_ARGUMENT;
  }
  if (npairs == 0) return MB200_OK;
  MB200_REQUIRE(bodies->center && bodies->radius && pairs && phi && normal, "null pointer");
  MB200_REQUIRE((arm_i == nullptr) == (arm_j == nullptr), "arm_i and arm_j must both be given or both be NULL");
  const bool rods = bodies->kind != nullptr;
  if (rods) MB200_REQUIRE(bodies->quat && bodies->length, "kind given without quat/length");
  D4* b0 = nullptr;
  MB200_TRY(body_records(&b0));
  D4* b1 = rods ? b0 + kMaxBodies : nullptr;
  cudaStream_t st = as_stream(stream);
  MB200_TRY(pack_bodies(st, bodies, nullptr, b0, b1));
  return narrow_phase_records(st, npairs, pairs, b0, b1, phi, normal, arm_i, arm_j, make_periodic(nullptr));
}
