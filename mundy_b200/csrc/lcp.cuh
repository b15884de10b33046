// Device pieces of the matrix-free contact operator, shared by the standalone operator entry points (lcp_ops.cu)
// and the BBPGD solver (lcp_solve.cu). Operation order follows the reference kernels so that per-element results
// are bit-identical to the CPU oracle:
//   D   : scrap/lcp_spheres/KokkosLCP.cpp:566-590 ; 6-DOF scrap/UnitTestMundy.cpp:652-674 (n_J = -n_I)
//   M   : scrap/parameter_interface/NGPSpheresLCP.cpp:597-613 ; rods scrap/UnitTestMundy.cpp:704-753
//   D^T : scrap/lcp_spheres/KokkosLCP.cpp:624-646 ; 6-DOF scrap/UnitTestMundy.cpp:783-824
#pragma once
#include "common.cuh"

namespace mb200 {

struct Mob {
  double inv_para, inv_perp, inv_rot;
};
// Spheres: 1/(6 pi mu r) evaluated as coeff * inv_radius; rods: prototype's zeta_para = zeta_perp = 6 pi mu r,
// zeta_rot = 8 pi mu r^3 (scrap/UnitTestMundy.cpp:719-724).
__host__ __device__ inline Mob mobility_of(double r, double mu) {
  const double pi = 3.14159265358979323846;
  const double coeff = 1.0 / (6.0 * pi * mu);
  const double inv_r = 1.0 / r;
  Mob m;
  m.inv_para = coeff * inv_r;
  m.inv_perp = m.inv_para;
  m.inv_rot = 1.0 / (8.0 * pi * r * r * r * mu);
  return m;
}

// Sum of the incident contact impulses of one body, in ascending (contact, side) order.
template <bool SIX>
__device__ inline void body_wrench(int64_t b, const int64_t* __restrict__ ptr, const uint32_t* __restrict__ list,
                                   const double* __restrict__ gamma, const double* __restrict__ normal,
                                   const double* __restrict__ arm_i, const double* __restrict__ arm_j, V3& F, V3& T) {
  double f0 = 0.0, f1 = 0.0, f2 = 0.0, t0 = 0.0, t1 = 0.0, t2 = 0.0;
  const int64_t beg = ptr[b], end = ptr[b + 1];
  for (int64_t k = beg; k < end; ++k) {
    const uint32_t e = __ldg(list + k);
    const int64_t c = e >> 1;
    const bool is_j = e & 1u;
    const double g = gamma[c];
    const double nx = __ldg(normal + 3 * c), ny = __ldg(normal + 3 * c + 1), nz = __ldg(normal + 3 * c + 2);
    if (!is_j) {
      f0 += -g * nx, f1 += -g * ny, f2 += -g * nz;
      if (SIX) {  // torque_I += -gamma * (r_I x n_I)
        const double a0 = __ldg(arm_i + 3 * c), a1 = __ldg(arm_i + 3 * c + 1), a2 = __ldg(arm_i + 3 * c + 2);
        t0 += -g * (a1 * nz - a2 * ny);
        t1 += -g * (a2 * nx - a0 * nz);
        t2 += -g * (a0 * ny - a1 * nx);
      }
    } else {
      f0 += g * nx, f1 += g * ny, f2 += g * nz;
      if (SIX) {  // torque_J += -gamma * (r_J x n_J), n_J = -n_I
        const double a0 = __ldg(arm_j + 3 * c), a1 = __ldg(arm_j + 3 * c + 1), a2 = __ldg(arm_j + 3 * c + 2);
        t0 += -g * (a1 * (-nz) - a2 * (-ny));
        t1 += -g * (a2 * (-nx) - a0 * (-nz));
        t2 += -g * (a0 * (-ny) - a1 * (-nx));
      }
    }
  }
  F = V3{f0, f1, f2};
  T = V3{t0, t1, t2};
}

// U = M F, W = M_r T for one body whose axis a = q (x) z_hat is known. `rod` selects the anisotropic form.
__device__ inline void body_mobility_axis(const Mob& m, bool rod, const V3& a, const V3& F, const V3& T, V3& U, V3& W) {
  if (!rod) {
    U = V3{m.inv_para * F.x, m.inv_para * F.y, m.inv_para * F.z};
  } else {
    const double av[3] = {a.x, a.y, a.z};
    double u[3];
#pragma unroll
    for (int row = 0; row < 3; ++row) {
      double acc[3];
#pragma unroll
      for (int col = 0; col < 3; ++col) {
        const double qq = av[row] * av[col];
        const double imqq = (row == col ? 1.0 - qq : -qq);
        acc[col] = m.inv_para * qq + m.inv_perp * imqq;
      }
      u[row] = acc[0] * F.x + acc[1] * F.y + acc[2] * F.z;
    }
    U = V3{u[0], u[1], u[2]};
  }
  W = V3{m.inv_rot * T.x, m.inv_rot * T.y, m.inv_rot * T.z};
}
__device__ inline void body_mobility(const Mob& m, bool rod, const Q4& q, const V3& F, const V3& T, V3& U, V3& W) {
  const V3 a = rod ? qrot(q, V3{0.0, 0.0, 1.0}) : V3{0.0, 0.0, 1.0};
  body_mobility_axis(m, rod, a, F, T, U, W);
}

// gdot of one contact.
template <bool SIX>
__device__ inline double contact_rate(const V3& n, const V3& Ui, const V3& Uj, const V3& Wi, const V3& Wj, const V3& ai,
                                      const V3& aj) {
  if (!SIX) {
    return -n.x * (Ui.x - Uj.x) - n.y * (Ui.y - Uj.y) - n.z * (Ui.z - Uj.z);
  } else {
    const V3 vi = Ui + cross(Wi, ai);
    const V3 vj = Uj + cross(Wj, aj);
    return -(n.x * vi.x + n.y * vi.y + n.z * vi.z) - ((-n.x) * vj.x + (-n.y) * vj.y + (-n.z) * vj.z);
  }
}

}  // namespace mb200
