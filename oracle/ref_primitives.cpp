// TEST INFRASTRUCTURE ONLY. Third translation unit of oracle/_ref/libmundy_ref.so: the reference's own headers for the
// remaining primitives, compiled unmodified from /root/reference behind oracle/kokkos_shim/.
//
// Reference entry points exercised:
//   geom::compute_aabb(Point / LineSegment / Ellipsoid)   mundy/geom/src/mundy_geom/compute_aabb.hpp:47-104
//   geom::distance(Point, Sphere, sep)                    mundy/geom/src/mundy_geom/distance/PointSphere.hpp
//   geom::distance(Point, Line, cp, arch, sep)            mundy/geom/src/mundy_geom/distance/PointLine.hpp
//   geom::distance(Line, Line, cp1, cp2, a1, a2, sep)     mundy/geom/src/mundy_geom/distance/LineLine.hpp
//   geom::distance(Line, Sphere, cp, arch, sep)           mundy/geom/src/mundy_geom/distance/LineSphere.hpp
//   geom::distance(Point, Ellipsoid, cp, normal)          mundy/geom/src/mundy_geom/distance/PointEllipsoid.hpp
//   geom::distance(Ellipsoid, Ellipsoid, cp1, cp2, n1, n2) mundy/geom/src/mundy_geom/distance/EllipsoidEllipsoid.hpp
//   geom::distance(Circle3D, Circle3D, cp1, cp2, n1, n2)  mundy/geom/src/mundy_geom/distance/Circle3DCircle3D.hpp
//   (the last three run math::find_min_using_approximate_derivatives, mundy/math/src/mundy_math/minimize.hpp and
//   impl/minimize_impl.hpp: L-BFGS with memory 10, central differences, Wolfe line search, nine starts)
#include <Kokkos_Core.hpp>

#include <mundy_geom/compute_aabb.hpp>
#include <mundy_geom/distance.hpp>
#include <mundy_geom/primitives.hpp>
#include <mundy_math/Quaternion.hpp>
#include <mundy_math/Vector3.hpp>

namespace mg = mundy::geom;
namespace mm = mundy::math;
using P3 = mg::Point<double>;
using V3 = mm::Vector3<double>;
using Quat = mm::Quaternion<double>;

static inline P3 pt(const double* p) { return P3(p[0], p[1], p[2]); }
static inline Quat qt(const double* q) { return Quat(q[0], q[1], q[2], q[3]); }
template <class T>
static inline void put3(double* o, const T& v) {
  o[0] = v[0];
  o[1] = v[1];
  o[2] = v[2];
}
template <class AABB>
static inline void put_aabb(double* o, const AABB& b) {
  for (int k = 0; k < 6; ++k) o[k] = b[k];
}

extern "C" {

void ref_aabb_point(long n, const double* p, double* out) {
  for (long i = 0; i < n; ++i) put_aabb(out + 6 * i, mg::compute_aabb(pt(p + 3 * i)));
}

void ref_aabb_line_segment(long n, const double* p0, const double* p1, double* out) {
  for (long i = 0; i < n; ++i)
    put_aabb(out + 6 * i, mg::compute_aabb(mg::LineSegment<double>(pt(p0 + 3 * i), pt(p1 + 3 * i))));
}

void ref_aabb_ellipsoid(long n, const double* c, const double* quat, const double* radii, double* out) {
  for (long i = 0; i < n; ++i)
    put_aabb(out + 6 * i, mg::compute_aabb(mg::Ellipsoid<double>(pt(c + 3 * i), qt(quat + 4 * i), pt(radii + 3 * i))));
}

void ref_point_sphere(long n, const double* x, const double* c, const double* r, double* dist, double* sep) {
  for (long i = 0; i < n; ++i) {
    V3 sp;
    dist[i] = mg::distance(pt(x + 3 * i), mg::Sphere<double>(pt(c + 3 * i), r[i]), sp);
    put3(sep + 3 * i, sp);
  }
}

void ref_point_line(long n, const double* x, const double* lc, const double* ld, double* dist, double* cp,
                    double* arch) {
  for (long i = 0; i < n; ++i) {
    P3 c;
    V3 sp;
    double a;
    dist[i] = mg::distance(pt(x + 3 * i), mg::Line<double>(pt(lc + 3 * i), pt(ld + 3 * i)), c, a, sp);
    arch[i] = a;
    put3(cp + 3 * i, c);
  }
}

void ref_line_line(long n, const double* c1, const double* d1, const double* c2, const double* d2, double* dist,
                   double* cp1, double* cp2, double* a1, double* a2, double* sep) {
  for (long i = 0; i < n; ++i) {
    P3 p1(0.0, 0.0, 0.0), p2(0.0, 0.0, 0.0);
    V3 sp(0.0, 0.0, 0.0);
    double s = 0.0, t = 0.0;
    dist[i] = mg::distance(mg::Line<double>(pt(c1 + 3 * i), pt(d1 + 3 * i)),
                           mg::Line<double>(pt(c2 + 3 * i), pt(d2 + 3 * i)), p1, p2, s, t, sp);
    a1[i] = s;
    a2[i] = t;
    put3(cp1 + 3 * i, p1);
    put3(cp2 + 3 * i, p2);
    put3(sep + 3 * i, sp);
  }
}

void ref_line_sphere(long n, const double* lc, const double* ld, const double* c, const double* r, double* dist,
                     double* cp, double* arch) {
  for (long i = 0; i < n; ++i) {
    P3 p;
    V3 sp(0.0, 0.0, 0.0);
    double a;
    dist[i] = mg::distance(mg::Line<double>(pt(lc + 3 * i), pt(ld + 3 * i)), mg::Sphere<double>(pt(c + 3 * i), r[i]),
                           p, a, sp);
    arch[i] = a;
    put3(cp + 3 * i, p);
  }
}

void ref_point_ellipsoid(long n, const double* x, const double* c, const double* quat, const double* radii,
                         double* dist, double* cp, double* normal) {
#pragma omp parallel for schedule(dynamic, 16)
  for (long i = 0; i < n; ++i) {
    P3 p;
    V3 nn;
    const P3 xx = pt(x + 3 * i);
    const mg::Ellipsoid<double> e(pt(c + 3 * i), qt(quat + 4 * i), pt(radii + 3 * i));
    dist[i] = mg::distance(mg::SharedNormalSigned{}, xx, e, p, nn);
    put3(cp + 3 * i, p);
    put3(normal + 3 * i, nn);
  }
}

void ref_ellipsoid_ellipsoid(long n, const double* c1, const double* q1, const double* rad1, const double* c2,
                             const double* q2, const double* rad2, double* dist, double* cp1, double* cp2,
                             double* n1) {
#pragma omp parallel for schedule(dynamic, 16)
  for (long i = 0; i < n; ++i) {
    P3 p1, p2;
    V3 na, nb;
    const mg::Ellipsoid<double> e1(pt(c1 + 3 * i), qt(q1 + 4 * i), pt(rad1 + 3 * i));
    const mg::Ellipsoid<double> e2(pt(c2 + 3 * i), qt(q2 + 4 * i), pt(rad2 + 3 * i));
    dist[i] = mg::distance(mg::SharedNormalSigned{}, e1, e2, p1, p2, na, nb);
    put3(cp1 + 3 * i, p1);
    put3(cp2 + 3 * i, p2);
    put3(n1 + 3 * i, na);
  }
}

void ref_circle_circle(long n, const double* c1, const double* q1, const double* r1, const double* c2,
                       const double* q2, const double* r2, double* dist, double* cp1, double* cp2, double* n1) {
#pragma omp parallel for schedule(dynamic, 16)
  for (long i = 0; i < n; ++i) {
    P3 p1, p2;
    V3 na, nb;
    const mg::Circle3D<double> a(pt(c1 + 3 * i), qt(q1 + 4 * i), r1[i]);
    const mg::Circle3D<double> b(pt(c2 + 3 * i), qt(q2 + 4 * i), r2[i]);
    dist[i] = mg::distance(mg::Euclidean{}, a, b, p1, p2, na, nb);
    put3(cp1 + 3 * i, p1);
    put3(cp2 + 3 * i, p2);
    put3(n1 + 3 * i, na);
  }
}

}  // extern "C"
