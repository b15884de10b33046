// mundy_b200/geom.hpp — the geom:: seam of the reference on top of the C-ABI.
//
// Reference signatures kept (all under mundy/geom/src/mundy_geom/):
//   AABB<Scalar> compute_aabb(const Sphere&)                 compute_aabb.hpp:72-80
//   AABB<Scalar> compute_aabb(const Spherocylinder&)         compute_aabb.hpp:106-127   (axis = q (x) z_hat)
//   AABB<Scalar> compute_aabb(const SpherocylinderSegment&)  compute_aabb.hpp:130-143
//   Scalar distance([SharedNormalSigned,] Sphere, Sphere[, sep])                         distance/SphereSphere.hpp:54-76
//   Scalar distance([tag,] LineSegment, LineSegment, cp1, cp2, s, t, sep)                distance/LineSegmentLineSegment.hpp:189-318
//   Scalar distance([tag,] Point, LineSegment, cp, t, sep)                               distance/PointLineSegment.hpp:128-172
//   Scalar distance([tag,] LineSegment, Sphere, cp, t, sep)                              distance/LineSegmentSphere.hpp:86-99
//   PeriodicScaledMetric / PeriodicMetricX,Y,XY,YZ: sep, wrap; unwrap_points_to_ref(Point | Sphere | Spherocylinder |
//   LineSegment, metric, ref)                                                            periodicity.hpp:335-845,1333-1516
// Value types mirror primitives/{Point,Sphere,Spherocylinder,SpherocylinderSegment,LineSegment,AABB}.hpp accessors.
// The single-object overloads run on the device through a one-element batch (they exist so that code and tests written
// against the reference compile unchanged); bulk work should use the *_batch functions, which take device arrays in the
// reference's layouts and are asynchronous on `stream`.
#ifndef MUNDY_B200_GEOM_HPP_
#define MUNDY_B200_GEOM_HPP_

#include <array>

#include "core.hpp"

namespace mundy_b200 {
namespace geom {

struct Vector3 {
  double v[3]{0, 0, 0};
  Vector3() = default;
  Vector3(double x, double y, double z) : v{x, y, z} {}
  double& operator[](int i) { return v[i]; }
  const double& operator[](int i) const { return v[i]; }
};
using Point = Vector3;

struct Quaternion {  // (w, x, y, z), mundy/math/src/mundy_math/Quaternion.hpp
  double q[4]{1, 0, 0, 0};
  Quaternion() = default;
  Quaternion(double w, double x, double y, double z) : q{w, x, y, z} {}
  double w() const { return q[0]; }
  const double* vector() const { return q + 1; }
};

struct SharedNormalSigned {};  // distance/Types.hpp:30-39 (the default distance type)

class AABB {  // primitives/AABB.hpp: flat index 0..5 = min xyz, max xyz (:199-202)
 public:
  AABB() = default;
  AABB(const Point& lo, const Point& hi) : c_{lo[0], lo[1], lo[2], hi[0], hi[1], hi[2]} {}
  double operator[](int i) const { return c_[i]; }
  double& operator[](int i) { return c_[i]; }
  Point min_corner() const { return {c_[0], c_[1], c_[2]}; }
  Point max_corner() const { return {c_[3], c_[4], c_[5]}; }
  double x_min() const { return c_[0]; }
  double y_min() const { return c_[1]; }
  double z_min() const { return c_[2]; }
  double x_max() const { return c_[3]; }
  double y_max() const { return c_[4]; }
  double z_max() const { return c_[5]; }
  double* data() { return c_.data(); }

 private:
  std::array<double, 6> c_{0, 0, 0, 0, 0, 0};
};
/// Closed-interval overlap test, touching counts (primitives/AABB.hpp:419-431). Pure predicate on two values.
inline bool intersects(const AABB& a, const AABB& b) {
  for (int d = 0; d < 3; ++d)
    if (a[3 + d] < b[d] || b[3 + d] < a[d]) return false;
  return true;
}

class Sphere {
 public:
  Sphere(const Point& c, double r) : c_(c), r_(r) {}
  const Point& center() const { return c_; }
  double radius() const { return r_; }

 private:
  Point c_;
  double r_;
};
class Spherocylinder {
 public:
  Spherocylinder(const Point& c, const Quaternion& q, double r, double len) : c_(c), q_(q), r_(r), len_(len) {}
  const Point& center() const { return c_; }
  const Quaternion& orientation() const { return q_; }
  double radius() const { return r_; }
  double length() const { return len_; }

 private:
  Point c_;
  Quaternion q_;
  double r_, len_;
};
class LineSegment {
 public:
  LineSegment(const Point& a, const Point& b) : a_(a), b_(b) {}
  const Point& start() const { return a_; }
  const Point& end() const { return b_; }

 private:
  Point a_, b_;
};
class SpherocylinderSegment {
 public:
  SpherocylinderSegment(const Point& a, const Point& b, double r) : a_(a), b_(b), r_(r) {}
  const Point& start() const { return a_; }
  const Point& end() const { return b_; }
  double radius() const { return r_; }

 private:
  Point a_, b_;
  double r_;
};

class Line {  // primitives/Line.hpp: a point on the line and its direction (not assumed unit)
 public:
  Line(const Point& c, const Vector3& d) : c_(c), d_(d) {}
  const Point& center() const { return c_; }
  const Vector3& direction() const { return d_; }

 private:
  Point c_;
  Vector3 d_;
};
class Ellipsoid {  // primitives/Ellipsoid.hpp:130-143: centre, orientation, three body-frame semi-axes
 public:
  Ellipsoid(const Point& c, const Quaternion& q, const Vector3& radii) : c_(c), q_(q), r_(radii) {}
  Ellipsoid(const Point& c, const Quaternion& q, double r1, double r2, double r3) : c_(c), q_(q), r_(r1, r2, r3) {}
  const Point& center() const { return c_; }
  const Quaternion& orientation() const { return q_; }
  const Vector3& radii() const { return r_; }

 private:
  Point c_;
  Quaternion q_;
  Vector3 r_;
};
class Circle3D {  // primitives/Circle3D.hpp:98-110: the circle lies in the body-frame xy plane
 public:
  Circle3D(const Point& c, const Quaternion& q, double r) : c_(c), q_(q), r_(r) {}
  const Point& center() const { return c_; }
  const Quaternion& orientation() const { return q_; }
  double radius() const { return r_; }

 private:
  Point c_;
  Quaternion q_;
  double r_;
};
struct Euclidean {};  // distance/Types.hpp

// ---- batched entry points (device arrays, reference layouts) --------------------------------------------------------
inline void compute_aabb_sphere_batch(mb200_stream st, int64_t n, const double* center, const double* radius, double* aabb) {
  check(mb200_compute_aabb_sphere(st, n, center, radius, aabb), "mb200_compute_aabb_sphere");
}
inline void compute_aabb_spherocylinder_batch(mb200_stream st, int64_t n, const double* center, const double* quat,
                                              const double* radius, const double* length, double* aabb) {
  check(mb200_compute_aabb_spherocylinder(st, n, center, quat, radius, length, aabb), "mb200_compute_aabb_spherocylinder");
}
inline void compute_aabb_segment_batch(mb200_stream st, int64_t n, const double* start, const double* end,
                                       const double* radius, double* aabb) {
  check(mb200_compute_aabb_segment(st, n, start, end, radius, aabb), "mb200_compute_aabb_segment");
}

namespace detail {
inline DeviceVector<double> up(std::initializer_list<double> v) { return DeviceVector<double>(std::vector<double>(v)); }
inline DeviceVector<double> up3(const Vector3& p) { return up({p[0], p[1], p[2]}); }
inline Vector3 down3(const DeviceVector<double>& d) {
  const std::vector<double> h = d.to_host();
  return {h[0], h[1], h[2]};
}
inline AABB down_aabb(const DeviceVector<double>& d) {
  const std::vector<double> h = d.to_host();
  return AABB({h[0], h[1], h[2]}, {h[3], h[4], h[5]});
}
}  // namespace detail

// ---- the reference's single-object signatures ---------------------------------------------------------------------
inline AABB compute_aabb(const Sphere& s) {
  auto c = detail::up3(s.center()), r = detail::up({s.radius()});
  DeviceVector<double> out(6);
  compute_aabb_sphere_batch(nullptr, 1, c.data(), r.data(), out.data());
  return detail::down_aabb(out);
}
inline AABB compute_aabb(const Spherocylinder& s) {
  auto c = detail::up3(s.center()), r = detail::up({s.radius()}), len = detail::up({s.length()});
  const Quaternion& o = s.orientation();
  auto q = detail::up({o.q[0], o.q[1], o.q[2], o.q[3]});
  DeviceVector<double> out(6);
  compute_aabb_spherocylinder_batch(nullptr, 1, c.data(), q.data(), r.data(), len.data(), out.data());
  return detail::down_aabb(out);
}
inline AABB compute_aabb(const SpherocylinderSegment& s) {
  auto a = detail::up3(s.start()), b = detail::up3(s.end()), r = detail::up({s.radius()});
  DeviceVector<double> out(6);
  compute_aabb_segment_batch(nullptr, 1, a.data(), b.data(), r.data(), out.data());
  return detail::down_aabb(out);
}

inline double distance(const Sphere& a, const Sphere& b, Vector3& sep) {
  auto c1 = detail::up3(a.center()), c2 = detail::up3(b.center()), r1 = detail::up({a.radius()}), r2 = detail::up({b.radius()});
  DeviceVector<double> d(1), s(3);
  check(mb200_distance_sphere_sphere(nullptr, 1, c1.data(), r1.data(), c2.data(), r2.data(), d.data(), s.data()),
        "mb200_distance_sphere_sphere");
  sep = detail::down3(s);
  return d.to_host()[0];
}
inline double distance(const Sphere& a, const Sphere& b) {
  Vector3 sep;
  return distance(a, b, sep);
}
inline double distance(SharedNormalSigned, const Sphere& a, const Sphere& b) { return distance(a, b); }
inline double distance(SharedNormalSigned, const Sphere& a, const Sphere& b, Vector3& sep) { return distance(a, b, sep); }

inline double distance(const LineSegment& l, const LineSegment& m, Point& cp1, Point& cp2, double& s, double& t,
                       Vector3& sep) {
  auto l0 = detail::up3(l.start()), l1 = detail::up3(l.end()), m0 = detail::up3(m.start()), m1 = detail::up3(m.end());
  DeviceVector<double> d(1), ds(1), dt(1), c1(3), c2(3), sv(3);
  check(mb200_distance_segment_segment(nullptr, 1, l0.data(), l1.data(), m0.data(), m1.data(), d.data(), ds.data(),
                                       dt.data(), c1.data(), c2.data(), sv.data()),
        "mb200_distance_segment_segment");
  cp1 = detail::down3(c1), cp2 = detail::down3(c2), sep = detail::down3(sv);
  s = ds.to_host()[0], t = dt.to_host()[0];
  return d.to_host()[0];
}
inline double distance(const LineSegment& l, const LineSegment& m) {
  Point a, b;
  Vector3 sep;
  double s, t;
  return distance(l, m, a, b, s, t, sep);
}
inline double distance(SharedNormalSigned, const LineSegment& l, const LineSegment& m, Point& cp1, Point& cp2, double& s,
                       double& t, Vector3& sep) {
  return distance(l, m, cp1, cp2, s, t, sep);
}

inline double distance(const Point& x, const LineSegment& seg, Point& cp, double& t, Vector3& sep) {
  auto px = detail::up3(x), p1 = detail::up3(seg.start()), p2 = detail::up3(seg.end());
  DeviceVector<double> d(1), dt(1), c(3), sv(3);
  check(mb200_distance_point_segment(nullptr, 1, px.data(), p1.data(), p2.data(), d.data(), dt.data(), c.data(), sv.data()),
        "mb200_distance_point_segment");
  cp = detail::down3(c), sep = detail::down3(sv), t = dt.to_host()[0];
  return d.to_host()[0];
}
inline double distance(const LineSegment& seg, const Sphere& sph, Point& cp, double& t, Vector3& sep) {
  auto p1 = detail::up3(seg.start()), p2 = detail::up3(seg.end()), c = detail::up3(sph.center()), r = detail::up({sph.radius()});
  DeviceVector<double> d(1), dt(1), cpd(3), sv(3);
  check(mb200_distance_segment_sphere(nullptr, 1, p1.data(), p2.data(), c.data(), r.data(), d.data(), dt.data(),
                                      cpd.data(), sv.data()),
        "mb200_distance_segment_sphere");
  cp = detail::down3(cpd), sep = detail::down3(sv), t = dt.to_host()[0];
  return d.to_host()[0];
}

// ---- the remaining primitives ------------------------------------------------------------------
namespace detail {
inline DeviceVector<double> upq(const Quaternion& o) { return up({o.q[0], o.q[1], o.q[2], o.q[3]}); }
}  // namespace detail
inline AABB compute_aabb(const Point& p) {  // compute_aabb.hpp:47-54
  auto a = detail::up3(p);
  DeviceVector<double> out(6);
  check(mb200_compute_aabb_point(nullptr, 1, a.data(), out.data()), "mb200_compute_aabb_point");
  return detail::down_aabb(out);
}
inline AABB compute_aabb(const LineSegment& s) {  // compute_aabb.hpp:57-69
  auto a = detail::up3(s.start()), b = detail::up3(s.end());
  DeviceVector<double> out(6);
  check(mb200_compute_aabb_line_segment(nullptr, 1, a.data(), b.data(), out.data()), "mb200_compute_aabb_line_segment");
  return detail::down_aabb(out);
}
inline AABB compute_aabb(const Ellipsoid& e) {  // compute_aabb.hpp:83-103
  auto c = detail::up3(e.center()), q = detail::upq(e.orientation()), r = detail::up3(e.radii());
  DeviceVector<double> out(6);
  check(mb200_compute_aabb_ellipsoid(nullptr, 1, c.data(), q.data(), r.data(), out.data()), "mb200_compute_aabb_ellipsoid");
  return detail::down_aabb(out);
}
inline double distance(const Point& x, const Sphere& s, Vector3& sep) {  // distance/PointSphere.hpp:68-76
  auto px = detail::up3(x), c = detail::up3(s.center()), r = detail::up({s.radius()});
  DeviceVector<double> d(1), sv(3);
  check(mb200_distance_point_sphere(nullptr, 1, px.data(), c.data(), r.data(), d.data(), sv.data()),
        "mb200_distance_point_sphere");
  sep = detail::down3(sv);
  return d.to_host()[0];
}
/// distance/PointLine.hpp:92-103. As in the reference, sep is not written.
inline double distance(const Point& x, const Line& line, Point& closest_point, double& arch_length, Vector3& /*sep*/) {
  auto px = detail::up3(x), c = detail::up3(line.center()), dir = detail::up3(line.direction());
  DeviceVector<double> d(1), cp(3), a(1);
  check(mb200_distance_point_line(nullptr, 1, px.data(), c.data(), dir.data(), d.data(), cp.data(), a.data()),
        "mb200_distance_point_line");
  closest_point = detail::down3(cp), arch_length = a.to_host()[0];
  return d.to_host()[0];
}
/// distance/LineLine.hpp:131-157. Parallel lines: returns 0 with sep = 0; closest points and arch lengths are then 0
/// here (the reference leaves them untouched).
inline double distance(const Line& l1, const Line& l2, Point& cp1, Point& cp2, double& arch1, double& arch2, Vector3& sep) {
  auto c1 = detail::up3(l1.center()), d1 = detail::up3(l1.direction()), c2 = detail::up3(l2.center()),
       d2 = detail::up3(l2.direction());
  DeviceVector<double> d(1), p1(3), p2(3), a1(1), a2(1), sv(3);
  check(mb200_distance_line_line(nullptr, 1, c1.data(), d1.data(), c2.data(), d2.data(), d.data(), p1.data(), p2.data(),
                                 a1.data(), a2.data(), sv.data()),
        "mb200_distance_line_line");
  cp1 = detail::down3(p1), cp2 = detail::down3(p2), sep = detail::down3(sv);
  arch1 = a1.to_host()[0], arch2 = a2.to_host()[0];
  return d.to_host()[0];
}
/// distance/LineSphere.hpp:93-102 (sep inherits PointLine's behaviour: not written).
inline double distance(const Line& line, const Sphere& s, Point& closest_point, double& arch_length, Vector3& /*sep*/) {
  auto lc = detail::up3(line.center()), dir = detail::up3(line.direction()), c = detail::up3(s.center()),
       r = detail::up({s.radius()});
  DeviceVector<double> d(1), cp(3), a(1);
  check(mb200_distance_line_sphere(nullptr, 1, lc.data(), dir.data(), c.data(), r.data(), d.data(), cp.data(), a.data()),
        "mb200_distance_line_sphere");
  closest_point = detail::down3(cp), arch_length = a.to_host()[0];
  return d.to_host()[0];
}
/// distance/PointEllipsoid.hpp:97-135
inline double distance(SharedNormalSigned, const Point& x, const Ellipsoid& e, Point& closest_point, Vector3& ellipsoid_normal) {
  auto px = detail::up3(x), c = detail::up3(e.center()), q = detail::upq(e.orientation()), r = detail::up3(e.radii());
  DeviceVector<double> d(1), cp(3), nn(3);
  check(mb200_distance_point_ellipsoid(nullptr, 1, px.data(), c.data(), q.data(), r.data(), d.data(), cp.data(), nn.data()),
        "mb200_distance_point_ellipsoid");
  closest_point = detail::down3(cp), ellipsoid_normal = detail::down3(nn);
  return d.to_host()[0];
}
inline double distance(SharedNormalSigned t, const Point& x, const Ellipsoid& e) {
  Point cp;
  Vector3 n;
  return distance(t, x, e, cp, n);
}
inline double distance(const Point& x, const Ellipsoid& e) { return distance(SharedNormalSigned{}, x, e); }
/// distance/EllipsoidEllipsoid.hpp:109-151
inline double distance(SharedNormalSigned, const Ellipsoid& e1, const Ellipsoid& e2, Point& cp1, Point& cp2,
                       Vector3& shared_normal1, Vector3& shared_normal2) {
  auto c1 = detail::up3(e1.center()), q1 = detail::upq(e1.orientation()), r1 = detail::up3(e1.radii());
  auto c2 = detail::up3(e2.center()), q2 = detail::upq(e2.orientation()), r2 = detail::up3(e2.radii());
  DeviceVector<double> d(1), p1(3), p2(3), nn(3);
  check(mb200_distance_ellipsoid_ellipsoid(nullptr, 1, c1.data(), q1.data(), r1.data(), c2.data(), q2.data(), r2.data(),
                                           d.data(), p1.data(), p2.data(), nn.data()),
        "mb200_distance_ellipsoid_ellipsoid");
  cp1 = detail::down3(p1), cp2 = detail::down3(p2), shared_normal1 = detail::down3(nn);
  shared_normal2 = Vector3(-shared_normal1[0], -shared_normal1[1], -shared_normal1[2]);
  return d.to_host()[0];
}
inline double distance(SharedNormalSigned t, const Ellipsoid& e1, const Ellipsoid& e2) {
  Point a, b;
  Vector3 n1, n2;
  return distance(t, e1, e2, a, b, n1, n2);
}
inline double distance(const Ellipsoid& e1, const Ellipsoid& e2) { return distance(SharedNormalSigned{}, e1, e2); }
/// distance/Circle3DCircle3D.hpp:88-151
inline double distance(Euclidean, const Circle3D& k1, const Circle3D& k2, Point& cp1, Point& cp2, Vector3& shared_normal1,
                       Vector3& shared_normal2) {
  auto c1 = detail::up3(k1.center()), q1 = detail::upq(k1.orientation()), r1 = detail::up({k1.radius()});
  auto c2 = detail::up3(k2.center()), q2 = detail::upq(k2.orientation()), r2 = detail::up({k2.radius()});
  DeviceVector<double> d(1), p1(3), p2(3), nn(3);
  check(mb200_distance_circle3d_circle3d(nullptr, 1, c1.data(), q1.data(), r1.data(), c2.data(), q2.data(), r2.data(),
                                         d.data(), p1.data(), p2.data(), nn.data()),
        "mb200_distance_circle3d_circle3d");
  cp1 = detail::down3(p1), cp2 = detail::down3(p2), shared_normal1 = detail::down3(nn);
  shared_normal2 = Vector3(-shared_normal1[0], -shared_normal1[1], -shared_normal1[2]);
  return d.to_host()[0];
}
inline double distance(const Circle3D& k1, const Circle3D& k2) {
  Point a, b;
  Vector3 n1, n2;
  return distance(Euclidean{}, k1, k2, a, b, n1, n2);
}

// ---- periodic metrics (periodicity.hpp) ------------------------------------------------------------------------------
/// Orthorhombic periodic metric. All three axes periodic = geom::PeriodicScaledMetric (periodicity.hpp:743-845); the
/// factories below give the partially periodic PeriodicMetricX / Y / XY / YZ (:335-741). cell_size()[d] <= 0: free axis.
class PeriodicScaledMetric {
 public:
  explicit PeriodicScaledMetric(const Vector3& cell_size) : cell_{cell_size[0], cell_size[1], cell_size[2]} {}
  bool is_periodic(unsigned d) const { return cell_[d] > 0.0; }
  unsigned num_periodic_dimensions() const { return unsigned(is_periodic(0)) + is_periodic(1) + is_periodic(2); }
  const double* cell_size() const { return cell_; }
  /// Minimum-image separation vector from point1 to point2 (:814-818).
  Vector3 sep(const Point& point1, const Point& point2) const {
    auto a = detail::up3(point1), b = detail::up3(point2);
    DeviceVector<double> out(3);
    check(mb200_periodic_sep(nullptr, 1, cell_, a.data(), b.data(), out.data()), "mb200_periodic_sep");
    return detail::down3(out);
  }
  /// Wrap a point into the primary cell (:821-824).
  Point wrap(const Point& point) const {
    auto a = detail::up3(point);
    DeviceVector<double> out(3);
    check(mb200_periodic_wrap(nullptr, 1, cell_, a.data(), out.data()), "mb200_periodic_wrap");
    return detail::down3(out);
  }

 private:
  double cell_[3];
};
inline PeriodicScaledMetric periodic_scaled_metric_from_unit_cell(const Vector3& cell_size) {  // :870-874
  return PeriodicScaledMetric(cell_size);
}
inline PeriodicScaledMetric PeriodicMetricX(double wx) { return PeriodicScaledMetric({wx, 0.0, 0.0}); }
inline PeriodicScaledMetric PeriodicMetricY(double wy) { return PeriodicScaledMetric({0.0, wy, 0.0}); }
inline PeriodicScaledMetric PeriodicMetricXY(double wx, double wy) { return PeriodicScaledMetric({wx, wy, 0.0}); }
inline PeriodicScaledMetric PeriodicMetricYZ(double wy, double wz) { return PeriodicScaledMetric({0.0, wy, wz}); }

/// unwrap_points_to_ref (:1333-1339 points; :1487-1516 spheres / spherocylinders move their centre; :1368-1382 segments).
inline Point unwrap_points_to_ref(const Point& point, const PeriodicScaledMetric& metric, const Point& ref_point) {
  auto a = detail::up3(point), r = detail::up3(ref_point);
  DeviceVector<double> out(3);
  check(mb200_periodic_unwrap_points(nullptr, 1, metric.cell_size(), a.data(), r.data(), out.data()),
        "mb200_periodic_unwrap_points");
  return detail::down3(out);
}
inline Sphere unwrap_points_to_ref(const Sphere& s, const PeriodicScaledMetric& metric, const Point& ref_point) {
  return Sphere(unwrap_points_to_ref(s.center(), metric, ref_point), s.radius());
}
inline Spherocylinder unwrap_points_to_ref(const Spherocylinder& s, const PeriodicScaledMetric& metric,
                                           const Point& ref_point) {
  return Spherocylinder(unwrap_points_to_ref(s.center(), metric, ref_point), s.orientation(), s.radius(), s.length());
}
inline LineSegment unwrap_points_to_ref(const LineSegment& seg, const PeriodicScaledMetric& metric, const Point& ref_point) {
  auto a = detail::up3(seg.start()), b = detail::up3(seg.end()), r = detail::up3(ref_point);
  DeviceVector<double> o0(3), o1(3);
  check(mb200_periodic_unwrap_segments(nullptr, 1, metric.cell_size(), a.data(), b.data(), r.data(), o0.data(), o1.data()),
        "mb200_periodic_unwrap_segments");
  return LineSegment(detail::down3(o0), detail::down3(o1));
}

}  // namespace geom
}  // namespace mundy_b200
#endif  // MUNDY_B200_GEOM_HPP_
