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

// "Attractive ABC flow" forcing of the multi-step driver, F += ... (scrap/parameter_interface/NGPSpheresLCP.cpp:971-998).
__global__ void __launch_bounds__(256) abc_flow_kernel(int64_t n, const double* __restrict__ center,
                                                       double* __restrict__ F) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double a = 1.0, b = sqrt(2.0), c = sqrt(3.0), k = 2.0;
  const V3 p = ld3(center, i);
  const double rr = sqrt(p.x * p.x + p.y * p.y + p.z * p.z);
  const double inv_r = 1.0 / rr;
  V3 f = ld3(F, i);
  f.x += a * sin(p.z) + c * cos(p.y) - k * p.x * inv_r;
  f.y += b * sin(p.x) + a * cos(p.z) - k * p.y * inv_r;
  f.z += c * sin(p.y) + b * cos(p.x) - k * p.z * inv_r;
  st3(F, i, f);
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

extern "C" int mb200_abc_flow_force(mb200_stream stream, int64_t n, const double* center, double* F) {
  MB200_REQUIRE(n >= 0 && (n == 0 || (center && F)), "bad arguments");
  MB200_TRY(require_device());
  if (n == 0) return MB200_OK;
  abc_flow_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(n, center, F);
  MB200_LAUNCHED();
  return MB200_OK;
}
