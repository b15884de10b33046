// G. Non-local mobility operators behind `ApplyMobilityOp`.
//   pair kernels  scrap/parameter_interface/alens/src/mundy_alens/periphery/Periphery.hpp
//                   apply_stokes_kernel :686-747, apply_rpy_kernel :847-942, apply_rpyc_kernel :954-1085,
//                   apply_local_drag :1316-1354, quake_inv_sqrt :72-82, panel driver :636-676
//                 scrap/lcp_spheres/KokkosLCP.cpp RPYKernel :219-280, apply_rpy_kernel_neighbors :305-333,
//                   compute_the_mobility_problem :593-614
//   operator      scrap/parameter_interface/alens/tests/performance_tests/NgpHP1.cpp :677-701, :804-925
//   solve         NgpHP1.cpp resolve_collisions<ApplyMobilityOp> :1487-1645
//
// B200 form. All-pairs: one thread per target, sources packed into two 32-byte records ({x, y, z, a}, {fx, fy, fz, -})
// and staged through shared memory in tiles that every thread of the CTA reads as broadcasts; the kernel is fp64-issue
// bound (~60 flops + one rsqrt per pair), not HBM bound. Neighbour form: one thread per target walking its CSR row,
// two 256-bit gathers per source. In both forms the sum over the sources of one target is taken in ascending source /
// row order from zero and then added to the target velocity: the order of the reference's panel driver on the Kokkos
// Serial backend, so results are bit-identical to the CPU oracle (-fmad=false, IEEE sqrt and division).
#include <cmath>
#include <cstring>
#include <utility>

#include "internal.cuh"
#include "lcp.cuh"

namespace mb200 {
namespace {

constexpr double kDoubleZero = 1e-12;  // DOUBLE_ZERO (Periphery.hpp:67, KokkosLCP.cpp:34)
constexpr double kPi = 3.14159265358979323846;
constexpr int kTargets = 64;   // targets (threads) per CTA: 65 536 targets = 1024 CTAs, 6.9 per SM (128 left a 4-vs-3 imbalance)
constexpr int kTileSrc = 256;  // sources staged per round: 16 KB

// Periphery.hpp:72-82: bit trick + one Newton iteration; it is the reference's arithmetic, error ~1.7e-3 included.
__device__ inline double quake_inv_sqrt(double number) {
  double y = number;
  const double x2 = y * 0.5;
  long long i = __double_as_longlong(y);
  i = 0x5fe6eb50c7b537a9ll - (i >> 1);
  y = __longlong_as_double(i);
  y = y * (1.5 - (x2 * y * y));
  return y;
}

struct Acc {
  double x, y, z;
};

// One (target, source) contribution. p = {x, y, z, a} of the source, f = its force, (tx, ty, tz, b) the target.
template <int KIND>
__device__ inline void pair_contrib(double mu, double tx, double ty, double tz, double b, const D4& p, const D4& f,
                                    Acc& acc) {
  const double dx = tx - p.x, dy = ty - p.y, dz = tz - p.z;
  const double fx = f.x, fy = f.y, fz = f.z;
  const double r2 = dx * dx + dy * dy + dz * dz;
  if (KIND == MB200_KERNEL_STOKES) {
    const double scale_factor = 1.0 / (8.0 * kPi * mu);
    const double rinv = r2 < kDoubleZero ? 0.0 : quake_inv_sqrt(r2);
    const double rinv3 = rinv * rinv * rinv;
    const double f_dot_r = fx * dx + fy * dy + fz * dz;
    const double scale_factor_rinv3 = scale_factor * rinv3;
    acc.x += scale_factor_rinv3 * (r2 * fx + dx * f_dot_r);
    acc.y += scale_factor_rinv3 * (r2 * fy + dy * f_dot_r);
    acc.z += scale_factor_rinv3 * (r2 * fz + dz * f_dot_r);
  } else if (KIND == MB200_KERNEL_RPY || KIND == MB200_KERNEL_RPY_EXACT) {
    constexpr bool exact = KIND == MB200_KERNEL_RPY_EXACT;
    const double scale_factor = 1.0 / (8.0 * kPi * mu);
    const double a = p.w;
    constexpr double one_over_three = 1.0 / 3.0;
    constexpr double one_over_six = 1.0 / 6.0;
    const double a2_over_three = one_over_three * a * a;
    const double rinv = r2 < kDoubleZero ? 0.0 : (exact ? 1.0 / sqrt(r2) : quake_inv_sqrt(r2));
    const double rinv3 = rinv * rinv * rinv;
    const double rinv5 = rinv * rinv * rinv3;
    const double fdotr = fx * dx + fy * dy + fz * dz;
    const double three_fdotr_rinv5 = 3.0 * fdotr * rinv5;
    const double cx = fx * rinv3 - three_fdotr_rinv5 * dx;
    const double cy = fy * rinv3 - three_fdotr_rinv5 * dy;
    const double cz = fz * rinv3 - three_fdotr_rinv5 * dz;
    const double fdotr_rinv3 = fdotr * rinv3;
    const double v0 = scale_factor * (fx * rinv + dx * fdotr_rinv3 + a2_over_three * cx);
    const double v1 = scale_factor * (fy * rinv + dy * fdotr_rinv3 + a2_over_three * cy);
    const double v2 = scale_factor * (fz * rinv + dz * fdotr_rinv3 + a2_over_three * cz);
    const double lap0 = 2.0 * scale_factor * cx;
    const double lap1 = 2.0 * scale_factor * cy;
    const double lap2 = 2.0 * scale_factor * cz;
    const double lap_coeff = exact ? 0.5 * a2_over_three : one_over_six * b * b;
    acc.x += v0 + lap_coeff * lap0;
    acc.y += v1 + lap_coeff * lap1;
    acc.z += v2 + lap_coeff * lap2;
  } else {  // corrected RPY
    constexpr double one_over_eight = 1.0 / 8.0;
    constexpr double one_over_six = 1.0 / 6.0;
    constexpr double one_over_three = 1.0 / 3.0;
    constexpr double one_over_32 = 1.0 / 32.0;
    constexpr double inv_pi = 1.0 / kPi;
    const double inv_viscosity = 1.0 / mu;
    const double a = p.w;
    const double r = sqrt(r2);
    const double r3 = r * r2;
    const double rinv = r2 < kDoubleZero ? 0.0 : 1.0 / r;
    const double rinv2 = rinv * rinv;
    const double rinv3 = rinv2 * rinv;
    const double dx_hat = dx * rinv, dy_hat = dy * rinv, dz_hat = dz * rinv;
    const double f_dot_rhat = fx * dx_hat + fy * dy_hat + fz * dz_hat;
    if (a + b < r) {
      const double a2 = a * a, b2 = b * b;
      const double a2_plus_b2_rinv2 = (a2 + b2) * rinv2;
      const double scale_factor = one_over_eight * inv_pi * inv_viscosity * rinv;
      const double tmp1_scaled = scale_factor * (1. + a2_plus_b2_rinv2 * one_over_three);
      const double tmp2_scaled = scale_factor * (1. - a2_plus_b2_rinv2);
      acc.x += tmp1_scaled * fx + tmp2_scaled * (f_dot_rhat * dx_hat);
      acc.y += tmp1_scaled * fy + tmp2_scaled * (f_dot_rhat * dy_hat);
      acc.z += tmp1_scaled * fz + tmp2_scaled * (f_dot_rhat * dz_hat);
    } else if (fabs(a - b) < r && a > kDoubleZero && b > kDoubleZero) {
      const double a_plus_b = a + b;
      const double a_minus_b = a - b;
      const double a_minus_b2 = a_minus_b * a_minus_b;
      const double tmp3 = a_minus_b2 + 3 * r2;
      const double tmp4 = a_minus_b2 - r2;
      const double scale_factor = one_over_six * inv_pi * inv_viscosity / (a * b);
      const double tmp1_scaled = scale_factor * (16.0 * r3 * a_plus_b - tmp3 * tmp3) * one_over_32 * rinv3;
      const double tmp2_scaled = scale_factor * 3.0 * tmp4 * tmp4 * one_over_32 * rinv3;
      acc.x += tmp1_scaled * fx + tmp2_scaled * (f_dot_rhat * dx_hat);
      acc.y += tmp1_scaled * fy + tmp2_scaled * (f_dot_rhat * dy_hat);
      acc.z += tmp1_scaled * fz + tmp2_scaled * (f_dot_rhat * dz_hat);
    } else {
      if (r2 < kDoubleZero) return;  // self interaction
      const double max_a_b = fmax(a, b);
      const double scale_factor = one_over_six * inv_pi * inv_viscosity / max_a_b;
      acc.x += scale_factor * fx;
      acc.y += scale_factor * fy;
      acc.z += scale_factor * fz;
    }
  }
}

// Self term: Periphery.hpp:1344-1353 (scale / r) or, with the exact RPY kernel, KokkosLCP.cpp:603-605 1/(6 pi r mu).
__device__ inline double self_coeff(int kind, double r, double mu) {
  if (kind == MB200_KERNEL_RPY_EXACT) return 1.0 / (6.0 * kPi * r * mu);
  const double scale = 1.0 / (6.0 * kPi * mu);
  return scale / r;
}

// pos[3n], rad[n] (nullable), force[3n] -> {x, y, z, a}, {fx, fy, fz, 0}
__global__ void __launch_bounds__(256) pack_sources_kernel(int64_t n, const double* __restrict__ pos,
                                                           const double* __restrict__ rad,
                                                           const double* __restrict__ force, D4* __restrict__ s0,
                                                           D4* __restrict__ s1) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (s0) st256(s0 + i, D4{pos[3 * i], pos[3 * i + 1], pos[3 * i + 2], rad ? rad[i] : 0.0});
  if (s1) st256(s1 + i, D4{force[3 * i], force[3 * i + 1], force[3 * i + 2], 0.0});
}

// u_t += sum_s K(t, s) f_s over ALL sources, s ascending.
template <int KIND>
__global__ void __launch_bounds__(kTargets)
    all_pairs_kernel(int64_t nt, int64_t ns, double mu, const double* __restrict__ tpos,
                     const double* __restrict__ trad, const D4* __restrict__ s0, const D4* __restrict__ s1,
                     double* __restrict__ tvel) {
  __shared__ D4 sh0[kTileSrc];
  __shared__ D4 sh1[kTileSrc];
  const int64_t t = static_cast<int64_t>(blockIdx.x) * kTargets + threadIdx.x;
  const bool live = t < nt;
  const double tx = live ? tpos[3 * t] : 0.0, ty = live ? tpos[3 * t + 1] : 0.0, tz = live ? tpos[3 * t + 2] : 0.0;
  const double b = (live && trad) ? trad[t] : 0.0;
  Acc acc{0.0, 0.0, 0.0};
  for (int64_t base = 0; base < ns; base += kTileSrc) {
    const int cnt = static_cast<int>(min(static_cast<int64_t>(kTileSrc), ns - base));
    __syncthreads();
    for (int k = threadIdx.x; k < cnt; k += kTargets) {
      sh0[k] = ldg256(s0 + base + k);
      sh1[k] = ldg256(s1 + base + k);
    }
    __syncthreads();
    if (live) {
#pragma unroll 4
      for (int k = 0; k < cnt; ++k) pair_contrib<KIND>(mu, tx, ty, tz, b, sh0[k], sh1[k], acc);
    }
  }
  if (live) {
    tvel[3 * t] += acc.x;
    tvel[3 * t + 1] += acc.y;
    tvel[3 * t + 2] += acc.z;
  }
}

// u_t += sum over the CSR row of t, in row order.
template <int KIND>
__global__ void __launch_bounds__(128)
    neighbor_kernel(int64_t n, double mu, const int32_t* __restrict__ offsets, const int32_t* __restrict__ indices,
                    const D4* __restrict__ s0, const D4* __restrict__ s1, double* __restrict__ vel) {
  const int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const D4 me = ldg256(s0 + t);
  Acc acc{0.0, 0.0, 0.0};
  const int beg = offsets[t], end = offsets[t + 1];
  for (int k = beg; k < end; ++k) {
    const int s = __ldg(indices + k);
    pair_contrib<KIND>(mu, me.x, me.y, me.z, me.w, ldg256(s0 + s), ldg256(s1 + s), acc);
  }
  vel[3 * t] += acc.x;
  vel[3 * t + 1] += acc.y;
  vel[3 * t + 2] += acc.z;
}

__global__ void __launch_bounds__(256) local_drag_kernel(int64_t n, int kind, double mu, const double* __restrict__ rad,
                                                         const double* __restrict__ force, double* __restrict__ vel) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double c = self_coeff(kind, rad[i], mu);
  vel[3 * i] += c * force[3 * i];
  vel[3 * i + 1] += c * force[3 * i + 1];
  vel[3 * i + 2] += c * force[3 * i + 2];
}

template <template <int> class Launch, typename... A>
int dispatch_kind(int kind, A... args) {
  switch (kind) {
    case MB200_KERNEL_STOKES: return Launch<MB200_KERNEL_STOKES>::run(args...);
    case MB200_KERNEL_RPY: return Launch<MB200_KERNEL_RPY>::run(args...);
    case MB200_KERNEL_RPYC: return Launch<MB200_KERNEL_RPYC>::run(args...);
    case MB200_KERNEL_RPY_EXACT: return Launch<MB200_KERNEL_RPY_EXACT>::run(args...);
  }
  set_error("invalid argument: unknown mobility kernel kind");
  return MB200_ERR_INVALID_ARGUMENT;
}
template <int KIND>
struct LaunchAll {
  static int run(cudaStream_t st, int64_t nt, int64_t ns, double mu, const double* tpos, const double* trad,
                 const D4* s0, const D4* s1, double* tvel) {
    all_pairs_kernel<KIND><<<grid_for(nt, kTargets), kTargets, 0, st>>>(nt, ns, mu, tpos, trad, s0, s1, tvel);
    MB200_LAUNCHED();
    return MB200_OK;
  }
};
template <int KIND>
struct LaunchNbr {
  static int run(cudaStream_t st, int64_t n, double mu, const int32_t* offsets, const int32_t* indices, const D4* s0,
                 const D4* s1, double* vel) {
    neighbor_kernel<KIND><<<grid_for(n, 128), 128, 0, st>>>(n, mu, offsets, indices, s0, s1, vel);
    MB200_LAUNCHED();
    return MB200_OK;
  }
};

bool kind_ok(int kind) { return kind >= MB200_KERNEL_STOKES && kind <= MB200_KERNEL_RPY_EXACT; }

int check_mobility(const mb200_mobility_op* m) {
  MB200_REQUIRE(m && m->n >= 0 && m->n < (int64_t(1) << 31), "bad mobility op");
  MB200_REQUIRE(m->mode >= MB200_MOBILITY_LOCAL_DRAG && m->mode <= MB200_MOBILITY_ALL, "unknown mobility mode");
  MB200_REQUIRE(kind_ok(m->kernel), "unknown mobility kernel kind");
  MB200_REQUIRE(m->viscosity > 0.0, "viscosity must be positive");
  if (m->n > 0) {
    MB200_REQUIRE(m->radius, "mobility op: null radius");
    if (m->mode != MB200_MOBILITY_LOCAL_DRAG) MB200_REQUIRE(m->pos, "mobility op: null positions");
    if (m->mode == MB200_MOBILITY_NEIGHBORS)
      MB200_REQUIRE(m->nbr_offsets && m->nbr_indices, "mobility op: null neighbour CSR");
  }
  return MB200_OK;
}

// U = M F with sources already packed (s0 fixed per solve, s1 scratch refilled from F here).
int mobility_apply_packed(cudaStream_t st, const mb200_mobility_op* m, const D4* s0, D4* s1, const double* F,
                          double* U) {
  const int64_t n = m->n;
  MB200_CUDA(cudaMemsetAsync(U, 0, 3 * n * sizeof(double), st));
  if (m->mode != MB200_MOBILITY_LOCAL_DRAG) {
    pack_sources_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, nullptr, nullptr, F, nullptr, s1);
    MB200_LAUNCHED();
    if (m->mode == MB200_MOBILITY_NEIGHBORS)
      MB200_TRY(dispatch_kind<LaunchNbr>(m->kernel, st, n, m->viscosity, m->nbr_offsets, m->nbr_indices, s0,
                                         static_cast<const D4*>(s1), U));
    else
      MB200_TRY(dispatch_kind<LaunchAll>(m->kernel, st, n, n, m->viscosity, m->pos, m->radius, s0,
                                         static_cast<const D4*>(s1), U));
  }
  local_drag_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, m->kernel, m->viscosity, m->radius, F, U);
  MB200_LAUNCHED();
  return MB200_OK;
}

// ---- BBPGD with the mobility behind an operator (NgpHP1.cpp:1487-1645), spheres / force only ----------------------
constexpr int kRedThreads = 256;

// gamma = max(gamma_prev - alpha (phi0 + dt gdot), 0); gdot == NULL on the first iteration (the reference reads the
// zeroed signed_sep_dot there, :1533,:1571).
__global__ void __launch_bounds__(256) gamma_update_kernel(int64_t m, double alpha, double dt,
                                                           const double* __restrict__ gamma_prev,
                                                           const double* __restrict__ phi0,
                                                           const double* __restrict__ gdot, double* __restrict__ gamma) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= m) return;
  const double sep_new = phi0[c] + dt * (gdot ? gdot[c] : 0.0);
  gamma[c] = fmax(gamma_prev[c] - alpha * sep_new, 0.0);
}

__global__ void __launch_bounds__(128)
    force_kernel(int64_t n, const int64_t* __restrict__ ptr, const uint32_t* __restrict__ list,
                 const double* __restrict__ gamma, const double* __restrict__ normal, double* __restrict__ F) {
  const int64_t b = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (b >= n) return;
  V3 f, t;
  body_wrench<false>(b, ptr, list, gamma, normal, nullptr, nullptr, f, t);
  st3(F, b, f);
}

// phi0 += dt D^T U_ext (:1524-1531)
__global__ void __launch_bounds__(256) phi_correct_kernel(int64_t m, const int2* __restrict__ pairs,
                                                          const double* __restrict__ normal,
                                                          const double* __restrict__ U, double dt,
                                                          double* __restrict__ phi0) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= m) return;
  const int2 p = pairs[c];
  const V3 z{0, 0, 0};
  const double g = contact_rate<false>(ld3(normal, c), ld3(U, p.x), ld3(U, p.y), z, z, z, z);
  phi0[c] += dt * g;
}

// gdot = D^T U; per-CTA partials of the residual (max) and of the three Barzilai-Borwein sums.
__global__ void __launch_bounds__(kRedThreads)
    rate_reduce_kernel(int64_t m, const int2* __restrict__ pairs, const double* __restrict__ normal,
                       const double* __restrict__ U, const double* __restrict__ phi0, double dt,
                       const double* __restrict__ gamma, const double* __restrict__ gamma_prev,
                       const double* __restrict__ gdot_prev, double* __restrict__ gdot, double* __restrict__ partials) {
  __shared__ double sh[4][kRedThreads];
  double res = -1.0, xx = 0.0, xg = 0.0, gg = 0.0;
  for (int64_t c = static_cast<int64_t>(blockIdx.x) * kRedThreads + threadIdx.x; c < m;
       c += static_cast<int64_t>(gridDim.x) * kRedThreads) {
    const int2 p = pairs[c];
    const V3 z{0, 0, 0};
    const double g = contact_rate<false>(ld3(normal, c), ld3(U, p.x), ld3(U, p.y), z, z, z, z);
    gdot[c] = g;
    const double gam = gamma[c];
    const double sep_new = phi0[c] + dt * g;
    const double v = (gam < 1e-12) ? fabs(fmin(sep_new, 0.0)) : fabs(sep_new);
    res = fmax(res, v);
    if (gamma_prev) {
      const double xd = gam - gamma_prev[c];
      const double gd = dt * (g - gdot_prev[c]);
      xx += xd * xd, xg += xd * gd, gg += gd * gd;
    }
  }
  sh[0][threadIdx.x] = res, sh[1][threadIdx.x] = xx, sh[2][threadIdx.x] = xg, sh[3][threadIdx.x] = gg;
  __syncthreads();
  for (int w = kRedThreads / 2; w > 0; w >>= 1) {  // fixed tree: deterministic
    if (threadIdx.x < w) {
      sh[0][threadIdx.x] = fmax(sh[0][threadIdx.x], sh[0][threadIdx.x + w]);
      for (int k = 1; k < 4; ++k) sh[k][threadIdx.x] += sh[k][threadIdx.x + w];
    }
    __syncthreads();
  }
  if (threadIdx.x < 4) partials[4 * blockIdx.x + threadIdx.x] = sh[threadIdx.x][0];
}
__global__ void reduce_finish_kernel(int blocks, const double* __restrict__ partials, double* __restrict__ out) {
  const int k = threadIdx.x;  // 4 threads
  double v = (k == 0) ? -1.0 : 0.0;
  for (int b = 0; b < blocks; ++b) v = (k == 0) ? fmax(v, partials[4 * b]) : v + partials[4 * b + k];
  out[k] = v;
}
__global__ void __launch_bounds__(kRedThreads) max_speed_kernel(int64_t n, const double* __restrict__ U,
                                                               double* __restrict__ out) {
  __shared__ double sh[kRedThreads];
  double v = 0.0;
  for (int64_t b = threadIdx.x; b < n; b += kRedThreads) {
    const V3 u = ld3(U, b);
    v = fmax(v, sqrt(u.x * u.x + u.y * u.y + u.z * u.z));
  }
  sh[threadIdx.x] = v;
  __syncthreads();
  for (int w = kRedThreads / 2; w > 0; w >>= 1) {
    if (threadIdx.x < w) sh[threadIdx.x] = fmax(sh[threadIdx.x], sh[threadIdx.x + w]);
    __syncthreads();
  }
  if (threadIdx.x == 0) *out = sh[0];
}

}  // namespace
}  // namespace mb200

using namespace mb200;

extern "C" {

int mb200_mobility_kernel_apply(mb200_stream stream, int kernel, int64_t nt, int64_t ns, double viscosity,
                                const double* target_pos, const double* source_pos, const double* source_radii,
                                const double* target_radii, const double* source_force, double* target_vel) {
  MB200_REQUIRE(kind_ok(kernel), "unknown mobility kernel kind");
  MB200_REQUIRE(nt >= 0 && ns >= 0 && nt < (int64_t(1) << 31) && ns < (int64_t(1) << 31), "bad sizes");
  MB200_REQUIRE(viscosity > 0.0, "viscosity must be positive");
  if (nt == 0 || ns == 0) return MB200_OK;
  MB200_REQUIRE(target_pos && source_pos && source_force && target_vel, "null pointer");
  MB200_REQUIRE(kernel == MB200_KERNEL_STOKES || (source_radii && target_radii), "null radii");
  MB200_TRY(require_device());
  cudaStream_t st = as_stream(stream);
  Scratch<D4> s0, s1;
  MB200_TRY(s0.alloc(ns, st));
  MB200_TRY(s1.alloc(ns, st));
  pack_sources_kernel<<<grid_for(ns, 256), 256, 0, st>>>(ns, source_pos, source_radii, source_force, s0.p, s1.p);
  MB200_LAUNCHED();
  return dispatch_kind<LaunchAll>(kernel, st, nt, ns, viscosity, target_pos, target_radii,
                                  static_cast<const D4*>(s0.p), static_cast<const D4*>(s1.p), target_vel);
}

int mb200_mobility_kernel_apply_neighbors(mb200_stream stream, int kernel, int64_t n, double viscosity,
                                          const double* pos, const double* radii, const int32_t* offsets,
                                          const int32_t* indices, const double* force, double* vel) {
  MB200_REQUIRE(kind_ok(kernel), "unknown mobility kernel kind");
  MB200_REQUIRE(n >= 0 && n < (int64_t(1) << 31), "bad size");
  MB200_REQUIRE(viscosity > 0.0, "viscosity must be positive");
  if (n == 0) return MB200_OK;
  MB200_REQUIRE(pos && offsets && indices && force && vel, "null pointer");
  MB200_REQUIRE(kernel == MB200_KERNEL_STOKES || radii, "null radii");
  MB200_TRY(require_device());
  cudaStream_t st = as_stream(stream);
  Scratch<D4> s0, s1;
  MB200_TRY(s0.alloc(n, st));
  MB200_TRY(s1.alloc(n, st));
  pack_sources_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, pos, radii, force, s0.p, s1.p);
  MB200_LAUNCHED();
  return dispatch_kind<LaunchNbr>(kernel, st, n, viscosity, offsets, indices, static_cast<const D4*>(s0.p),
                                  static_cast<const D4*>(s1.p), vel);
}

int mb200_apply_local_drag(mb200_stream stream, int kernel, int64_t n, double viscosity, const double* radii,
                           const double* force, double* vel) {
  MB200_REQUIRE(kind_ok(kernel), "unknown mobility kernel kind");
  MB200_REQUIRE(n >= 0 && viscosity > 0.0, "bad size or viscosity");
  if (n == 0) return MB200_OK;
  MB200_REQUIRE(radii && force && vel, "null pointer");
  MB200_TRY(require_device());
  local_drag_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(n, kernel, viscosity, radii, force, vel);
  MB200_LAUNCHED();
  return MB200_OK;
}

int mb200_mobility_apply(mb200_stream stream, const mb200_mobility_op* mob, const double* F, double* U) {
  MB200_TRY(check_mobility(mob));
  if (mob->n == 0) return MB200_OK;
  MB200_REQUIRE(F && U, "null pointer");
  MB200_TRY(require_device());
  cudaStream_t st = as_stream(stream);
  Scratch<D4> s0, s1;
  if (mob->mode != MB200_MOBILITY_LOCAL_DRAG) {
    MB200_TRY(s0.alloc(mob->n, st));
    MB200_TRY(s1.alloc(mob->n, st));
    pack_sources_kernel<<<grid_for(mob->n, 256), 256, 0, st>>>(mob->n, mob->pos, mob->radius, nullptr, s0.p, nullptr);
    MB200_LAUNCHED();
  }
  return mobility_apply_packed(st, mob, s0.p, s1.p, F, U);
}

int mb200_resolve_collisions_mobility(mb200_stream stream, const mb200_contact_op* op, const mb200_mobility_op* mob,
                                      const mb200_lcp_params* params, const double* F_ext, double* U_ext_out,
                                      double* phi0, double* gamma, double* F, double* U,
                                      mb200_collision_result* result_h) {
  MB200_REQUIRE(op && params && op->n >= 0 && op->npairs >= 0, "bad contact op");
  MB200_REQUIRE(op->arm_i == nullptr && op->arm_j == nullptr, "the operator-form solve is force-only (spheres)");
  MB200_REQUIRE(op->n < (int64_t(1) << 31) && op->npairs < (int64_t(1) << 30), "contact op too large for int32 ids");
  MB200_TRY(check_mobility(mob));
  MB200_REQUIRE(mob->n == op->n, "mobility op and contact op disagree on the number of bodies");
  const int64_t n = op->n, m = op->npairs;
  MB200_REQUIRE(n == 0 || (F && U), "null F/U");
  MB200_REQUIRE(m == 0 || (op->pairs && op->normal && phi0 && gamma), "null pointer");
  MB200_REQUIRE(op->dt > 0.0, "dt must be positive");
  MB200_TRY(require_device());
  cudaStream_t st = as_stream(stream);
  const bool fixed = params->fixed_iters >= 0;
  const int it_cap = fixed ? params->fixed_iters : params->max_iters;
  const double dt = op->dt;
  const int2* pairs = reinterpret_cast<const int2*>(op->pairs);

  Scratch<D4> s0, s1;
  if (mob->mode != MB200_MOBILITY_LOCAL_DRAG && n > 0) {
    MB200_TRY(s0.alloc(n, st));
    MB200_TRY(s1.alloc(n, st));
    pack_sources_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, mob->pos, mob->radius, nullptr, s0.p, nullptr);
    MB200_LAUNCHED();
  }
  if (F_ext && n > 0) {  // U_ext = M F_ext ; phi0 += dt D^T U_ext
    Scratch<double> ue;
    double* uext = U_ext_out;
    if (!uext) {
      MB200_TRY(ue.alloc(3 * n, st));
      uext = ue.p;
    }
    MB200_TRY(mobility_apply_packed(st, mob, s0.p, s1.p, F_ext, uext));
    if (m > 0) {
      phi_correct_kernel<<<grid_for(m, 256), 256, 0, st>>>(m, pairs, op->normal, uext, dt, phi0);
      MB200_LAUNCHED();
    }
  }
  if (m == 0) {  // nothing to solve
    if (n > 0) {
      MB200_CUDA(cudaMemsetAsync(F, 0, 3 * n * sizeof(double), st));
      MB200_CUDA(cudaMemsetAsync(U, 0, 3 * n * sizeof(double), st));
    }
    MB200_CUDA(cudaStreamSynchronize(st));
    if (result_h) *result_h = mb200_collision_result{-1.0, 0, 1, 0.0};
    return MB200_OK;
  }

  Incidence inc;
  MB200_TRY(build_incidence(st, n, m, op->pairs, &inc));
  MB200_TRY(check_incidence(st, inc));
  const int blocks = grid_for(m, kRedThreads, kSMs * 4);
  Scratch<double> gam_b, gd_a, gd_b, partials, red;
  MB200_TRY(gam_b.alloc(m, st));
  MB200_TRY(gd_a.alloc(m, st));
  MB200_TRY(gd_b.alloc(m, st));
  MB200_TRY(partials.alloc(4 * blocks, st));
  MB200_TRY(red.alloc(5, st));
  double red_h[5];

  // (gamma) -> F -> U -> gdot, residual and (when prev != NULL) the BB sums, read back to the host
  auto evaluate = [&](const double* gam, const double* gam_prev, const double* gd_prev, double* gd) -> int {
    force_kernel<<<grid_for(n, 128), 128, 0, st>>>(n, inc.ptr.p, inc.list.p, gam, op->normal, F);
    MB200_LAUNCHED();
    MB200_TRY(mobility_apply_packed(st, mob, s0.p, s1.p, F, U));
    rate_reduce_kernel<<<blocks, kRedThreads, 0, st>>>(m, pairs, op->normal, U, phi0, dt, gam, gam_prev, gd_prev, gd,
                                                      partials.p);
    MB200_LAUNCHED();
    reduce_finish_kernel<<<1, 4, 0, st>>>(blocks, partials.p, red.p);
    MB200_LAUNCHED();
    MB200_CUDA(cudaMemcpyAsync(red_h, red.p, 4 * sizeof(double), cudaMemcpyDeviceToHost, st));
    MB200_CUDA(cudaStreamSynchronize(st));
    return MB200_OK;
  };

  // x_{k-1} lives in `prev`, x_k is written to `cur`; the caller's gamma is one of the two buffers.
  double* prev = gamma;
  double* cur = gam_b.p;
  double* gd_prev = gd_a.p;
  double* gd_cur = gd_b.p;
  MB200_TRY(evaluate(prev, nullptr, nullptr, gd_prev));
  double res = red_h[0];
  int it = 0;
  const double* latest = prev;
  if (fixed || !(res < params->tol)) {
    double alpha = 1.0 / res;
    while (it < it_cap) {
      ++it;
      gamma_update_kernel<<<grid_for(m, 256), 256, 0, st>>>(m, alpha, dt, prev, phi0, it == 1 ? nullptr : gd_prev, cur);
      MB200_LAUNCHED();
      MB200_TRY(evaluate(cur, prev, gd_prev, gd_cur));
      res = red_h[0];
      latest = cur;
      if (!fixed && res < params->tol) break;
      double a = (it % 2 == 0) ? red_h[1] : red_h[2];
      double b = (it % 2 == 0) ? red_h[2] : red_h[3];
      if (fabs(b) < 1e-12) b += 1e-12;
      alpha = a / b;
      std::swap(prev, cur);
      std::swap(gd_prev, gd_cur);
    }
  }
  if (latest != gamma) MB200_CUDA(cudaMemcpyAsync(gamma, latest, m * sizeof(double), cudaMemcpyDeviceToDevice, st));
  max_speed_kernel<<<1, kRedThreads, 0, st>>>(n, U, red.p + 4);
  MB200_LAUNCHED();
  MB200_CUDA(cudaMemcpyAsync(red_h + 4, red.p + 4, sizeof(double), cudaMemcpyDeviceToHost, st));
  MB200_CUDA(cudaStreamSynchronize(st));
  if (result_h) *result_h = mb200_collision_result{res, it, res < params->tol ? 1 : 0, red_h[4] * dt};
  return MB200_OK;
}

}  // extern "C"
