// TEST INFRASTRUCTURE ONLY. CPU restatement ("port") of the reference's collision hot path, used as the parity checker by
// tests/, by __graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference legs. Nothing under mundy_b200/ or
// include/ may import, link or call it.
//
// Parity status: the arithmetic leaves (AABB, distance, quaternion ops, math::convex PGD) are PINNED bit-for-bit against
// the reference's own headers compiled into oracle/_ref/libmundy_ref.so (tests/test_oracle_goldens.py::test_oracle_vs_live_reference_random and the stored outputs tests/golden/ref_outputs.npz) and against the
// reference's golden vectors (tests/test_oracle_goldens.py). The driver loop (BBPGD, D / D^T / M kernels) and the
// 6-DOF rod Jacobian are restated from scrap/ programs that cannot be linked here (STK/ArborX absent) and are checked
// through their analytical properties. The neighbour SEARCH restates a predicate whose implementation lives in
// un-vendored third-party code (stk::search::coarse_search, Trilinos 16.0.0; ArborX e026f82): "parity unpinned" for
// the candidate set versus STK/ArborX; pinned here to the fp64 predicate of SURVEY.md section 8(a) row B2.
//
// Plain C++17, no reference headers. Compile with -ffp-contract=off so that op order below IS the rounding order.
// Every function cites the reference file:line (paths relative to /root/reference) it follows.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace {

constexpr double kZeroTol = 1e-15;  // mundy/math/src/mundy_math/Tolerance.hpp:37-50  get_zero_tolerance<double>()

struct V3 {
  double x, y, z;
};
inline V3 operator-(const V3& a, const V3& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator+(const V3& a, const V3& b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator*(double s, const V3& a) { return {s * a.x, s * a.y, s * a.z}; }
inline V3 operator*(const V3& a, double s) { return {a.x * s, a.y * s, a.z * s}; }
// mundy/math/src/mundy_math/impl/VectorImpl.hpp:338-345: C++ right fold  a0*b0 + (a1*b1 + a2*b2)
inline double dot(const V3& a, const V3& b) { return a.x * b.x + (a.y * b.y + a.z * b.z); }
// mundy/math/src/mundy_math/Vector.hpp:1151-1156: std::sqrt(dot(v, v))
inline double norm(const V3& a) { return std::sqrt(dot(a, a)); }
// mundy/math/src/mundy_math/Vector3.hpp:93-106
inline V3 cross(const V3& a, const V3& b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
inline V3 ld3(const double* p) { return {p[0], p[1], p[2]}; }
inline void st3(double* p, const V3& v) {
  p[0] = v.x;
  p[1] = v.y;
  p[2] = v.z;
}

struct Q4 {
  double w, x, y, z;
};
// mundy/math/src/mundy_math/impl/QuaternionImpl.hpp:142-164 quat_quat_multiplication_impl
inline Q4 qmul(const Q4& a, const Q4& b) {
  Q4 r;
  r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
  r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
  r.y = a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x;
  r.z = a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w;
  return r;
}
// q (x) v = q * (0, v) * inverse(q), inverse = conjugate * (1 / |q|^2): no unit-norm assumption.
// mundy/math/src/mundy_math/impl/QuaternionImpl.hpp:191-202 ; Quaternion.hpp:1207-1227
inline V3 qrot(const Q4& q, const V3& v) {
  const Q4 vq{0.0, v.x, v.y, v.z};
  const double inv_n2 = 1.0 / (q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
  const Q4 qi{q.w * inv_n2, -q.x * inv_n2, -q.y * inv_n2, -q.z * inv_n2};
  const Q4 r = qmul(qmul(q, vq), qi);
  return {r.x, r.y, r.z};
}

// distance(Point, Point, sep)  mundy/geom/src/mundy_geom/distance/PointPoint.hpp:54-62
inline double dist_pp(const V3& p1, const V3& p2, V3& sep) {
  sep = p2 - p1;
  return norm(sep);
}

// distance(Point, LineSegment, cp, t, sep)  mundy/geom/src/mundy_geom/distance/PointLineSegment.hpp:128-172
// NB: t is returned UNCLAMPED in cases 3.1/3.2 while cp is clamped (:156-167).
inline double dist_pseg(const V3& x, const V3& p1, const V3& p2, V3& cp, double& t, V3& sep) {
  const V3 p21 = p2 - p1;
  const double num = dot(p21, x - p1);
  if ((num < kZeroTol) & (num > -kZeroTol)) {
    cp = p1;
    t = 0.0;
  } else {
    const double denom = dot(p21, p21);
    if (denom < kZeroTol) {
      cp = p1;
      t = 0.0;
    } else {
      t = num / denom;
      if (t < 0.0) {
        cp = p1;
      } else if (t > 1.0) {
        cp = p2;
      } else {
        cp = p1 + t * p21;
      }
    }
  }
  return dist_pp(x, cp, sep);
}

// distance(LineSegment, LineSegment, cp1, cp2, s, t, sep)
// mundy/geom/src/mundy_geom/distance/LineSegmentLineSegment.hpp:189-318
inline double dist_segseg(const V3& l0, const V3& l1, const V3& m0, const V3& m1, V3& cp1, V3& cp2, double& s, double& t,
                          V3& sep) {
  const V3 u = l1 - l0, v = m1 - m0, w = l0 - m0;
  const double a = dot(u, u), b = dot(u, v), c = dot(v, v), d = dot(u, w), e = dot(v, w);
  const double D = a * c - b * b;
  if (D < std::sqrt(kZeroTol)) {  // :215 collinear / degenerate
    V3 c1, c2, c3, c4, s1, s2, s3, s4;
    double t1, t2, t3, t4;
    const double d1 = dist_pseg(l0, m0, m1, c1, t1, s1);
    const double d2 = dist_pseg(l1, m0, m1, c2, t2, s2);
    const double d3 = dist_pseg(m0, l0, l1, c3, t3, s3);
    const double d4 = dist_pseg(m1, l0, l1, c4, t4, s4);
    const double dmin = std::min(std::min(d1, d2), std::min(d3, d4));
    if (dmin == d1) {
      s = 0.0, t = t1, cp1 = l0, cp2 = c1, sep = s1;
    } else if (dmin == d2) {
      s = 1.0, t = t2, cp1 = l1, cp2 = c2, sep = s2;
    } else if (dmin == d3) {
      s = t3, t = 0.0, cp1 = c3, cp2 = m0, sep = s3;
    } else {
      s = t4, t = 1.0, cp1 = c4, cp2 = m1, sep = s4;
    }
    return dmin;
  }
  double sN = b * e - c * d, tN = a * e - b * d, sD = D, tD = D;  // :272-275
  if (sN < 0.0) {
    sN = 0.0, tN = e, tD = c;
  } else if (sN > sD) {
    sN = sD, tN = e + b, tD = c;
  }
  if (tN < 0.0) {
    tN = 0.0;
    if (-d < 0.0) {
      sN = 0.0;
    } else if (-d > a) {
      sN = sD;
    } else {
      sN = -d, sD = a;
    }
  } else if (tN > tD) {
    tN = tD;
    if ((-d + b) < 0.0) {
      sN = 0.0;
    } else if ((-d + b) > a) {
      sN = sD;
    } else {
      sN = (-d + b), sD = a;
    }
  }
  s = (std::fabs(sN) < kZeroTol) ? 0.0 : sN / sD;  // :313-314
  t = (std::fabs(tN) < kZeroTol) ? 0.0 : tN / tD;
  cp1 = l0 + s * u;
  cp2 = m0 + t * v;
  return dist_pp(cp1, cp2, sep);
}

}  // namespace

extern "C" {

int orc_abi_version() { return 1; }

// ---------------------------------------------------------------------------------------------------------------
// A. compute_aabb   mundy/geom/src/mundy_geom/compute_aabb.hpp
// ---------------------------------------------------------------------------------------------------------------
// Sphere :72-80  min = c - 1*r, max = c + 1*r
void orc_aabb_sphere(long n, const double* pos, const double* r, double* out) {
  for (long i = 0; i < n; ++i) {
    for (int k = 0; k < 3; ++k) {
      out[6 * i + k] = pos[3 * i + k] - 1.0 * r[i];
      out[6 * i + 3 + k] = pos[3 * i + k] + 1.0 * r[i];
    }
  }
}

// Spherocylinder :106-127  d = 0.5 * L * (q (x) z_hat); corners of centre-line box -/+ r. quat = (w,x,y,z).
void orc_aabb_spherocylinder(long n, const double* pos, const double* quat, const double* r, const double* len,
                             double* out) {
  for (long i = 0; i < n; ++i) {
    const Q4 q{quat[4 * i], quat[4 * i + 1], quat[4 * i + 2], quat[4 * i + 3]};
    const V3 axis = qrot(q, V3{0.0, 0.0, 1.0});
    const V3 d = (0.5 * len[i]) * axis;  // static_cast<scalar_t>(0.5) * length * (orientation * z_axis)
    const V3 c = ld3(pos + 3 * i);
    const V3 lo = c - d, hi = c + d;
    out[6 * i + 0] = std::min(lo.x, hi.x) - r[i];
    out[6 * i + 1] = std::min(lo.y, hi.y) - r[i];
    out[6 * i + 2] = std::min(lo.z, hi.z) - r[i];
    out[6 * i + 3] = std::max(lo.x, hi.x) + r[i];
    out[6 * i + 4] = std::max(lo.y, hi.y) + r[i];
    out[6 * i + 5] = std::max(lo.z, hi.z) + r[i];
  }
}

// SpherocylinderSegment :130-143
void orc_aabb_segment(long n, const double* p0, const double* p1, const double* r, double* out) {
  for (long i = 0; i < n; ++i) {
    for (int k = 0; k < 3; ++k) {
      out[6 * i + k] = std::min(p0[3 * i + k], p1[3 * i + k]) - r[i];
      out[6 * i + 3 + k] = std::max(p0[3 * i + k], p1[3 * i + k]) + r[i];
    }
  }
}

void orc_quat_rotate(long n, const double* quat, const double* v, double* out) {
  for (long i = 0; i < n; ++i) {
    const Q4 q{quat[4 * i], quat[4 * i + 1], quat[4 * i + 2], quat[4 * i + 3]};
    st3(out + 3 * i, qrot(q, ld3(v + 3 * i)));
  }
}

// math::rotate_quaternion(q, omega, dt)  mundy/math/src/mundy_math/Quaternion.hpp:1359-1385 (Delong 2015 App. A)
void orc_rotate_quaternion(long n, double* quat, const double* omega, double dt) {
  for (long i = 0; i < n; ++i) {
    double* q = quat + 4 * i;
    const V3 om = ld3(omega + 3 * i);
    const double w = norm(om);
    if (w < kZeroTol) continue;
    const double winv = 1.0 / w;
    const double sw = std::sin(0.5 * w * dt);
    const double cw = std::cos(0.5 * w * dt);
    const double s = q[0];
    const V3 p{q[1], q[2], q[3]};
    // xyz = s*sw*omega*winv + cw*p + sw*winv*cross(omega,p)   (left-to-right products, then left-to-right sums)
    const V3 t1 = ((s * sw) * om) * winv;
    const V3 t2 = cw * p;
    const V3 t3 = (sw * winv) * cross(om, p);
    const V3 xyz = (t1 + t2) + t3;
    const double qw = s * cw - dot(om, p) * sw * winv;
    const double inv_norm = 1.0 / std::sqrt(qw * qw + xyz.x * xyz.x + xyz.y * xyz.y + xyz.z * xyz.z);
    q[0] = qw * inv_norm;
    q[1] = xyz.x * inv_norm;
    q[2] = xyz.y * inv_norm;
    q[3] = xyz.z * inv_norm;
  }
}

// ---------------------------------------------------------------------------------------------------------------
// C. geom::distance
// ---------------------------------------------------------------------------------------------------------------
void orc_segseg(long n, const double* l0, const double* l1, const double* m0, const double* m1, double* dist, double* s,
                double* t, double* cp1, double* cp2, double* sep) {
#pragma omp parallel for schedule(static)
  for (long i = 0; i < n; ++i) {
    V3 c1, c2, sp;
    double ss, tt;
    dist[i] = dist_segseg(ld3(l0 + 3 * i), ld3(l1 + 3 * i), ld3(m0 + 3 * i), ld3(m1 + 3 * i), c1, c2, ss, tt, sp);
    s[i] = ss;
    t[i] = tt;
    st3(cp1 + 3 * i, c1);
    st3(cp2 + 3 * i, c2);
    st3(sep + 3 * i, sp);
  }
}

void orc_point_seg(long n, const double* x, const double* p1, const double* p2, double* dist, double* t, double* cp,
                   double* sep) {
  for (long i = 0; i < n; ++i) {
    V3 c, sp;
    double tt;
    dist[i] = dist_pseg(ld3(x + 3 * i), ld3(p1 + 3 * i), ld3(p2 + 3 * i), c, tt, sp);
    t[i] = tt;
    st3(cp + 3 * i, c);
    st3(sep + 3 * i, sp);
  }
}

// distance(Sphere, Sphere, sep)  mundy/geom/src/mundy_geom/distance/SphereSphere.hpp:65-76
void orc_sphere_sphere(long n, const double* c1, const double* r1, const double* c2, const double* r2, double* dist,
                       double* sep) {
  for (long i = 0; i < n; ++i) {
    V3 sp;
    const double cc = dist_pp(ld3(c1 + 3 * i), ld3(c2 + 3 * i), sp);
    const double sd = cc - r1[i] - r2[i];
    const double f = sd / cc;
    sp = {sp.x * f, sp.y * f, sp.z * f};  // sep *= surface_distance / center_center_distance
    dist[i] = sd;
    st3(sep + 3 * i, sp);
  }
}

// distance(LineSegment, Sphere, cp, t, sep)  mundy/geom/src/mundy_geom/distance/LineSegmentSphere.hpp:86-99
void orc_seg_sphere(long n, const double* p1, const double* p2, const double* c, const double* r, double* dist, double* t,
                    double* cp, double* sep) {
  for (long i = 0; i < n; ++i) {
    V3 cpt, sp;
    double tt;
    const double lc = dist_pseg(ld3(c + 3 * i), ld3(p1 + 3 * i), ld3(p2 + 3 * i), cpt, tt, sp);
    const double sd = lc - r[i];
    const double f = sd / lc;
    sp = {sp.x * f, sp.y * f, sp.z * f};
    dist[i] = sd;
    t[i] = tt;
    st3(cp + 3 * i, cpt);
    st3(sep + 3 * i, sp);
  }
}

}  // extern "C"

#include "mundy_oracle_collision.inc"
#include "mundy_oracle_convex.inc"
#include "mundy_oracle_mobility.inc"
#include "mundy_oracle_primitives.inc"
