// A. geom::compute_aabb, batched (mundy/geom/src/mundy_geom/compute_aabb.hpp:72-143).
// HBM-bound: 80 / 120 / 104 algorithmic bytes per body (sphere / spherocylinder / segment). Each CTA streams tiles of
// 256 bodies: inputs and outputs cross HBM as contiguous double2 runs staged in shared memory (tile_io.cuh), the
// per-body arithmetic runs out of shared memory. Grid = a multiple of the 148 SMs, tiles grid-strided.
#include "tile_io.cuh"

namespace mb200 {
namespace {

constexpr int kCtasPerSm = 4;

__global__ void __launch_bounds__(kTile) aabb_sphere_kernel(int64_t n, const double* __restrict__ center,
                                                            const double* __restrict__ radius,
                                                            double* __restrict__ aabb) {
  __shared__ __align__(16) double s_c[3 * kTile];
  __shared__ __align__(16) double s_o[6 * kTile];
  const int64_t ntiles = (n + kTile - 1) / kTile;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = tile * kTile;
    const int cnt = static_cast<int>(min(static_cast<int64_t>(kTile), n - base));
    load_tile<3>(s_c, center, base, cnt);
    __syncthreads();
    const int t = threadIdx.x;
    if (t < cnt) {
      const double r = __ldg(radius + base + t);
      const double one_r = 1.0 * r;  // center - ones * radius (:76-79)
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const double c = s_c[3 * t + k];
        s_o[6 * t + k] = c - one_r;
        s_o[6 * t + 3 + k] = c + one_r;
      }
    }
    __syncthreads();
    store_tile<6>(aabb, s_o, base, cnt);
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kTile)
    aabb_spherocylinder_kernel(int64_t n, const double* __restrict__ center, const double* __restrict__ quat,
                               const double* __restrict__ radius, const double* __restrict__ length,
                               double* __restrict__ aabb) {
  __shared__ __align__(16) double s_c[3 * kTile];
  __shared__ __align__(16) double s_o[6 * kTile];
  const int64_t ntiles = (n + kTile - 1) / kTile;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = tile * kTile;
    const int cnt = static_cast<int>(min(static_cast<int64_t>(kTile), n - base));
    load_tile<3>(s_c, center, base, cnt);
    __syncthreads();
    const int t = threadIdx.x;
    if (t < cnt) {
      const Q4 q = ldq(quat, base + t);  // 32-byte item: two 16-byte loads per thread, contiguous across the warp
      const double r = __ldg(radius + base + t);
      const double len = __ldg(length + base + t);
      const V3 axis = qrot(q, V3{0.0, 0.0, 1.0});
      const V3 d = (0.5 * len) * axis;  // 0.5 * length * (orientation * z_axis)  (:115-117)
      const V3 c{s_c[3 * t], s_c[3 * t + 1], s_c[3 * t + 2]};
      const V3 lo = c - d, hi = c + d;
      s_o[6 * t + 0] = fmin(lo.x, hi.x) - r;
      s_o[6 * t + 1] = fmin(lo.y, hi.y) - r;
      s_o[6 * t + 2] = fmin(lo.z, hi.z) - r;
      s_o[6 * t + 3] = fmax(lo.x, hi.x) + r;
      s_o[6 * t + 4] = fmax(lo.y, hi.y) + r;
      s_o[6 * t + 5] = fmax(lo.z, hi.z) + r;
    }
    __syncthreads();
    store_tile<6>(aabb, s_o, base, cnt);
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kTile) aabb_segment_kernel(int64_t n, const double* __restrict__ start,
                                                             const double* __restrict__ end,
                                                             const double* __restrict__ radius,
                                                             double* __restrict__ aabb) {
  __shared__ __align__(16) double s_a[3 * kTile];
  __shared__ __align__(16) double s_b[3 * kTile];
  __shared__ __align__(16) double s_o[6 * kTile];
  const int64_t ntiles = (n + kTile - 1) / kTile;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = tile * kTile;
    const int cnt = static_cast<int>(min(static_cast<int64_t>(kTile), n - base));
    load_tile<3>(s_a, start, base, cnt);
    load_tile<3>(s_b, end, base, cnt);
    __syncthreads();
    const int t = threadIdx.x;
    if (t < cnt) {
      const double r = __ldg(radius + base + t);
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const double a = s_a[3 * t + k], b = s_b[3 * t + k];
        s_o[6 * t + k] = fmin(a, b) - r;
        s_o[6 * t + 3 + k] = fmax(a, b) + r;
      }
    }
    __syncthreads();
    store_tile<6>(aabb, s_o, base, cnt);
    __syncthreads();
  }
}

inline int aabb_grid(int64_t n) { return grid_for(n, kTile, kSMs * kCtasPerSm); }

}  // namespace
}  // namespace mb200

using namespace mb200;

extern "C" {

int mb200_compute_aabb_sphere(mb200_stream stream, int64_t n, const double* center, const double* radius,
                              double* aabb) {
  MB200_REQUIRE(n >= 0, "n < 0");
  if (n == 0) return MB200_OK;
  MB200_REQUIRE(center && radius && aabb, "null pointer");
  MB200_TRY(require_device());
  aabb_sphere_kernel<<<aabb_grid(n), kTile, 0, as_stream(stream)>>>(n, center, radius, aabb);
  MB200_LAUNCHED();
  return MB200_OK;
}

int mb200_compute_aabb_spherocylinder(mb200_stream stream, int64_t n, const double* center, const double* quat,
                                      const double* radius, const double* length, double* aabb) {
  MB200_REQUIRE(n >= 0, "n < 0");
  if (n == 0) return MB200_OK;
  MB200_REQUIRE(center && quat && radius && length && aabb, "null pointer");
  MB200_REQUIRE((reinterpret_cast<uintptr_t>(quat) & 15) == 0, "quat must be 16-byte aligned");
  MB200_TRY(require_device());
  aabb_spherocylinder_kernel<<<aabb_grid(n), kTile, 0, as_stream(stream)>>>(n, center, quat, radius, length, aabb);
  MB200_LAUNCHED();
  return MB200_OK;
}

int mb200_compute_aabb_segment(mb200_stream stream, int64_t n, const double* start, const double* end,
                               const double* radius, double* aabb) {
  MB200_REQUIRE(n >= 0, "n < 0");
  if (n == 0) return MB200_OK;
  MB200_REQUIRE(start && end && radius && aabb, "null pointer");
  MB200_TRY(require_device());
  aabb_segment_kernel<<<aabb_grid(n), kTile, 0, as_stream(stream)>>>(n, start, end, radius, aabb);
  MB200_LAUNCHED();
  return MB200_OK;
}

}  // extern "C"
