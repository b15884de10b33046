// E3. Euler update: x += dt U (scrap/lcp_spheres/KokkosLCP.cpp:1015) and, for spherocylinders,
// q <- math::rotate_quaternion(q, omega, dt) (mundy/math/src/mundy_math/Quaternion.hpp:1359-1385, Delong 2015 App. A).
// HBM-bound: 72 B/body (spheres), 160 B/body (rods).
#include "internal.cuh"

namespace mb200 {

namespace {

__global__ void __launch_bounds__(256) euler_pos_kernel(int64_t n3, double* __restrict__ x, const double* __restrict__ u,
                                                        double dt) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; k < n3; k += stride) x[k] += dt * u[k];
}

__global__ void __launch_bounds__(256) euler_quat_kernel(int64_t n, double* __restrict__ quat,
                                                         const int32_t* __restrict__ kind, const double* __restrict__ W,
                                                         double dt) {
  const int64_t b = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (b >= n) return;
  if (kind && kind[b] != MB200_SPHEROCYLINDER) return;
  const Q4 q = rotate_quaternion(ldq(quat, b), ld3(W, b), dt);
  double2* o = reinterpret_cast<double2*>(quat + 4 * b);
  o[0] = make_double2(q.w, q.x);
  o[1] = make_double2(q.y, q.z);
}

}  // namespace

int euler_update(cudaStream_t st, int64_t n, double* center, double* quat, const int32_t* kind, const double* U,
                 const double* W, double dt) {
  if (n == 0) return MB200_OK;
  euler_pos_kernel<<<grid_for(3 * n, 256, kSMs * 8), 256, 0, st>>>(3 * n, center, U, dt);
  MB200_LAUNCHED();
  if (quat && W) {
    euler_quat_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, quat, kind, W, dt);
    MB200_LAUNCHED();
  }
  return MB200_OK;
}

}  // namespace mb200

using namespace mb200;

extern "C" int mb200_euler_update(mb200_stream stream, int64_t n, double* center, double* quat, const int32_t* kind,
                                  const double* U, const double* W, double dt) {
  MB200_REQUIRE(n >= 0 && (n == 0 || (center && U)), "bad arguments");
  MB200_TRY(require_device());
  return euler_update(as_stream(stream), n, center, quat, kind, U, W, dt);
}
