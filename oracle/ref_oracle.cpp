// TEST INFRASTRUCTURE ONLY. Never linked into, loaded by, or called from the product (mundy_b200/, include/).
//
// oracle/_ref/libmundy_ref.so: the reference's OWN header-only hot-path code, compiled unmodified from where it lies
// under /root/reference (mundy_core, mundy_math, mundy_geom) against oracle/kokkos_shim/, behind a small C ABI so that
// tests can (1) pin the CPU restatement in oracle/mundy_oracle.cpp against it and (2) generate tests/golden/ vectors.
// It cannot be rebuilt on the GPU box (no /root/reference there); the built .so travels with the repo snapshot.
//
// Reference entry points exercised:
//   geom::compute_aabb          mundy/geom/src/mundy_geom/compute_aabb.hpp:47-143
//   geom::distance(...)         mundy/geom/src/mundy_geom/distance/{PointPoint,PointLineSegment,LineSegmentLineSegment,
//                                                                    SphereSphere,LineSegmentSphere}.hpp
//   geom::PeriodicScaledMetric / PeriodicMetricX,Y,XY,YZ, unwrap_points_to_ref
//                               mundy/geom/src/mundy_geom/periodicity.hpp:335-845,1333-1516
//   math::rotate_quaternion     mundy/math/src/mundy_math/Quaternion.hpp:1359-1385
//   math::solve_cqpp/solve_lcp  mundy/math/src/mundy_math/convex.hpp:789-845 (KokkosBackend + MundyMathBackend)
#include <Kokkos_Core.hpp>

#include <mundy_geom/compute_aabb.hpp>
#include <mundy_geom/distance.hpp>
#include <mundy_geom/periodicity.hpp>
#include <mundy_geom/primitives.hpp>
#include <mundy_math/Quaternion.hpp>
#include <mundy_math/Vector3.hpp>
#include <mundy_math/convex.hpp>

namespace mg = mundy::geom;
namespace mm = mundy::math;
using P3 = mg::Point<double>;
using V3 = mm::Vector3<double>;
using Quat = mm::Quaternion<double>;

static inline P3 pt(const double* p) { return P3(p[0], p[1], p[2]); }
static inline void put3(double* o, const V3& v) {
  o[0] = v[0];
  o[1] = v[1];
  o[2] = v[2];
}
template <class AABB>
static inline void put_aabb(double* o, const AABB& b) {
  for (int k = 0; k < 6; ++k) o[k] = b[k];
}

extern "C" {

int ref_abi_version() { return 1; }

void ref_aabb_sphere(long n, const double* pos, const double* r, double* out) {
  for (long i = 0; i < n; ++i) put_aabb(out + 6 * i, mg::compute_aabb(mg::Sphere<double>(pt(pos + 3 * i), r[i])));
}

// quat layout (w, x, y, z), compute_aabb.hpp:106-127 (z-axis convention)
void ref_aabb_spherocylinder(long n, const double* pos, const double* quat, const double* r, const double* len,
                             double* out) {
  for (long i = 0; i < n; ++i) {
    const double* q = quat + 4 * i;
    mg::Spherocylinder<double> sc(pt(pos + 3 * i), Quat(q[0], q[1], q[2], q[3]), r[i], len[i]);
    put_aabb(out + 6 * i, mg::compute_aabb(sc));
  }
}

void ref_aabb_segment(long n, const double* p0, const double* p1, const double* r, double* out) {
  for (long i = 0; i < n; ++i) {
    mg::SpherocylinderSegment<double> seg(pt(p0 + 3 * i), pt(p1 + 3 * i), r[i]);
    put_aabb(out + 6 * i, mg::compute_aabb(seg));
  }
}

// q (x) v exactly as the reference evaluates it (impl/QuaternionImpl.hpp:191-202)
void ref_quat_rotate(long n, const double* quat, const double* v, double* out) {
  for (long i = 0; i < n; ++i) {
    const double* q = quat + 4 * i;
    const V3 r = Quat(q[0], q[1], q[2], q[3]) * V3(v[3 * i], v[3 * i + 1], v[3 * i + 2]);
    put3(out + 3 * i, r);
  }
}

void ref_rotate_quaternion(long n, double* quat, const double* omega, double dt) {
  for (long i = 0; i < n; ++i) {
    double* q = quat + 4 * i;
    Quat qq(q[0], q[1], q[2], q[3]);
    mm::rotate_quaternion(qq, V3(omega[3 * i], omega[3 * i + 1], omega[3 * i + 2]), dt);
    for (int k = 0; k < 4; ++k) q[k] = qq[k];
  }
}

// distance(LineSegment, LineSegment, cp1, cp2, s, t, sep)  LineSegmentLineSegment.hpp:189-318
void ref_segseg(long n, const double* l0, const double* l1, const double* m0, const double* m1, double* dist, double* s,
                double* t, double* cp1, double* cp2, double* sep) {
#pragma omp parallel for schedule(static)
  for (long i = 0; i < n; ++i) {
    mg::LineSegment<double> a(pt(l0 + 3 * i), pt(l1 + 3 * i)), b(pt(m0 + 3 * i), pt(m1 + 3 * i));
    P3 c1, c2;
    V3 sp;
    double ss, tt;
    dist[i] = mg::distance(a, b, c1, c2, ss, tt, sp);
    s[i] = ss;
    t[i] = tt;
    put3(cp1 + 3 * i, c1);
    put3(cp2 + 3 * i, c2);
    put3(sep + 3 * i, sp);
  }
}

// distance(Point, LineSegment, cp, t, sep)  PointLineSegment.hpp:128-172
void ref_point_seg(long n, const double* x, const double* p1, const double* p2, double* dist, double* t, double* cp,
                   double* sep) {
  for (long i = 0; i < n; ++i) {
    mg::LineSegment<double> seg(pt(p1 + 3 * i), pt(p2 + 3 * i));
    P3 c;
    V3 sp;
    double tt;
    dist[i] = mg::distance(pt(x + 3 * i), seg, c, tt, sp);
    t[i] = tt;
    put3(cp + 3 * i, c);
    put3(sep + 3 * i, sp);
  }
}

// distance(Sphere, Sphere, sep)  SphereSphere.hpp:54-76
void ref_sphere_sphere(long n, const double* c1, const double* r1, const double* c2, const double* r2, double* dist,
                       double* sep) {
  for (long i = 0; i < n; ++i) {
    V3 sp;
    dist[i] = mg::distance(mg::Sphere<double>(pt(c1 + 3 * i), r1[i]), mg::Sphere<double>(pt(c2 + 3 * i), r2[i]), sp);
    put3(sep + 3 * i, sp);
  }
}

// distance(LineSegment, Sphere, cp, t, sep)  LineSegmentSphere.hpp:86-99
void ref_seg_sphere(long n, const double* p1, const double* p2, const double* c, const double* r, double* dist,
                    double* t, double* cp, double* sep) {
  for (long i = 0; i < n; ++i) {
    mg::LineSegment<double> seg(pt(p1 + 3 * i), pt(p2 + 3 * i));
    P3 cpt;
    V3 sp;
    double tt;
    dist[i] = mg::distance(seg, mg::Sphere<double>(pt(c + 3 * i), r[i]), cpt, tt, sp);
    t[i] = tt;
    put3(cp + 3 * i, cpt);
    put3(sep + 3 * i, sp);
  }
}

// ---- periodic metrics (periodicity.hpp). mode = bit mask of the periodic axes: 7 = PeriodicScaledMetric (:743-845),
// 1 = PeriodicMetricX, 2 = PeriodicMetricY, 3 = PeriodicMetricXY, 6 = PeriodicMetricYZ (:335-741).
// op: 0 sep(a,b) | 1 wrap(a) | 2 unwrap_points_to_ref(Point a, m, ref b) | 3 unwrap_points_to_ref(LineSegment(a,c), m, ref b)
//     4 centre of unwrap_points_to_ref(Sphere(a, 1), m, b) | 5 centre of unwrap_points_to_ref(Spherocylinder(a,..), m, b)
}  // extern "C"
template <class Metric>
static void periodic_apply(const Metric& m, int op, long n, const double* a, const double* b, const double* c,
                           double* o0, double* o1) {
  for (long i = 0; i < n; ++i) {
    const P3 pa = pt(a + 3 * i);
    if (op == 0) {
      put3(o0 + 3 * i, m.sep(pa, pt(b + 3 * i)));
    } else if (op == 1) {
      put3(o0 + 3 * i, m.wrap(pa));
    } else if (op == 2) {
      put3(o0 + 3 * i, mg::unwrap_points_to_ref(pa, m, pt(b + 3 * i)));
    } else if (op == 3) {
      const auto seg = mg::unwrap_points_to_ref(mg::LineSegment<double>(pa, pt(c + 3 * i)), m, pt(b + 3 * i));
      put3(o0 + 3 * i, seg.start());
      put3(o1 + 3 * i, seg.end());
    } else if (op == 4) {
      put3(o0 + 3 * i, mg::unwrap_points_to_ref(mg::Sphere<double>(pa, 1.0), m, pt(b + 3 * i)).center());
    } else {
      const mg::Spherocylinder<double> sc(pa, Quat(1.0, 0.0, 0.0, 0.0), 0.5, 2.0);
      put3(o0 + 3 * i, mg::unwrap_points_to_ref(sc, m, pt(b + 3 * i)).center());
    }
  }
}

extern "C" {
int ref_periodic(int mode, int op, long n, const double* cell, const double* a, const double* b, const double* c,
                 double* o0, double* o1) {
  switch (mode) {
    case 7: periodic_apply(mg::PeriodicScaledMetric<double>(V3(cell[0], cell[1], cell[2])), op, n, a, b, c, o0, o1); return 0;
    case 1: periodic_apply(mg::PeriodicMetricX<double>(cell[0]), op, n, a, b, c, o0, o1); return 0;
    case 2: periodic_apply(mg::PeriodicMetricY<double>(cell[1]), op, n, a, b, c, o0, o1); return 0;
    case 3: periodic_apply(mg::PeriodicMetricXY<double>(cell[0], cell[1]), op, n, a, b, c, o0, o1); return 0;
    case 6: periodic_apply(mg::PeriodicMetricYZ<double>(cell[1], cell[2]), op, n, a, b, c, o0, o1); return 0;
    default: return -1;
  }
}

// ---- math::convex through the reference's own KokkosBackend -------------------------------------------------------
typedef void (*ref_apply_fn)(const double* x, double* y, long n, void* ctx);

}  // extern "C"

namespace {
using vec_t = Kokkos::View<double*>;
using mat_t = Kokkos::View<double**>;

struct CallbackOp {  // LinearOp path 2 of KokkosBackend::apply (convex.hpp:186-191)
  ref_apply_fn fn;
  void* ctx;
  void apply(const vec_t& x, vec_t& y) const { fn(x.data(), y.data(), static_cast<long>(x.extent(0)), ctx); }
};

template <class Prob, class Backend>
int run_pgd(const Prob& prob, Backend backend, vec_t& x, long n, unsigned max_iters, double tol, int resid_kind,
            unsigned* iters, double* residual) {
  vec_t g("g", n), xt("xt", n), gt("gt", n);
  mm::convex::PGDConfig<double> cfg{max_iters, tol};
  auto state = mm::make_pgd_state(backend, x, g, xt, gt);
  mm::convex::SolveResult<double> res;
  if (resid_kind == 0) {
    auto pgd = mm::make_pgd_solution_strategy(backend, cfg);  // BB1 + LinfNormProjectedDiffResidual
    res = mm::solve_cqpp(prob, pgd, state);
  } else {
    auto pgd = mm::make_pgd_solution_strategy(backend, mm::convex::BBStepStrategy{},
                                              mm::convex::LinfNormProjectedGradientResidual{}, cfg);
    res = mm::solve_cqpp(prob, pgd, state);
  }
  *iters = res.num_iters;
  *residual = res.residual;
  return res.converged ? 1 : 0;
}

template <class Op>
int solve_with_space(const Op& A, vec_t& q, vec_t& x, long n, int space_kind, double lo, double hi, unsigned max_iters,
                     double tol, int resid_kind, unsigned* iters, double* residual) {
  Kokkos::Serial exec;
  namespace sp = mm::convex::space;
  switch (space_kind) {
    case 0: {
      sp::Unconstrained<double> s{};
      auto prob = mm::make_kokkos_cqpp(exec, A, q, s);
      return run_pgd(prob, prob.backend(), x, n, max_iters, tol, resid_kind, iters, residual);
    }
    case 1: {
      sp::LowerBound<double> s{lo};
      auto prob = mm::make_kokkos_cqpp(exec, A, q, s);
      return run_pgd(prob, prob.backend(), x, n, max_iters, tol, resid_kind, iters, residual);
    }
    case 2: {
      sp::UpperBound<double> s{hi};
      auto prob = mm::make_kokkos_cqpp(exec, A, q, s);
      return run_pgd(prob, prob.backend(), x, n, max_iters, tol, resid_kind, iters, residual);
    }
    default: {
      sp::Bounded<double> s{lo, hi};
      auto prob = mm::make_kokkos_cqpp(exec, A, q, s);
      return run_pgd(prob, prob.backend(), x, n, max_iters, tol, resid_kind, iters, residual);
    }
  }
}
}  // namespace

extern "C" {

// space_kind: 0 unconstrained, 1 lower bound, 2 upper bound, 3 box.  resid_kind: 0 projected-diff, 1 projected-grad.
// A is dense row-major n x n. x is in/out. Returns converged (0/1).
int ref_solve_cqpp_dense(long n, const double* A, const double* q, double* x, int space_kind, double lo, double hi,
                         unsigned max_iters, double tol, int resid_kind, unsigned* iters, double* residual) {
  mat_t Av("A", n, n);
  vec_t qv("q", n), xv("x", n);
  for (long i = 0; i < n * n; ++i) Av.data()[i] = A[i];
  for (long i = 0; i < n; ++i) {
    qv(i) = q[i];
    xv(i) = x[i];
  }
  int c = solve_with_space(Av, qv, xv, n, space_kind, lo, hi, max_iters, tol, resid_kind, iters, residual);
  for (long i = 0; i < n; ++i) x[i] = xv(i);
  return c;
}

// Same, with a matrix-free operator y = A x supplied by the caller (the F11 composition: A = dt * D^T M D).
int ref_solve_cqpp_op(long n, ref_apply_fn fn, void* ctx, const double* q, double* x, int space_kind, double lo,
                      double hi, unsigned max_iters, double tol, int resid_kind, unsigned* iters, double* residual) {
  CallbackOp op{fn, ctx};
  vec_t qv("q", n), xv("x", n);
  for (long i = 0; i < n; ++i) {
    qv(i) = q[i];
    xv(i) = x[i];
  }
  int c = solve_with_space(op, qv, xv, n, space_kind, lo, hi, max_iters, tol, resid_kind, iters, residual);
  for (long i = 0; i < n; ++i) x[i] = xv(i);
  return c;
}

// solve_lcp through LCPProblem/to_cqpp (convex.hpp:840-845) on a dense matrix.
int ref_solve_lcp_dense(long n, const double* A, const double* q, double* x, unsigned max_iters, double tol,
                        unsigned* iters, double* residual) {
  Kokkos::Serial exec;
  mat_t Av("A", n, n);
  vec_t qv("q", n), xv("x", n), g("g", n), xt("xt", n), gt("gt", n);
  for (long i = 0; i < n * n; ++i) Av.data()[i] = A[i];
  for (long i = 0; i < n; ++i) {
    qv(i) = q[i];
    xv(i) = x[i];
  }
  auto lcp = mm::make_kokkos_lcp(exec, Av, qv);
  auto backend = lcp.backend();
  mm::convex::PGDConfig<double> cfg{max_iters, tol};
  auto pgd = mm::make_pgd_solution_strategy(backend, cfg);
  auto state = mm::make_pgd_state(backend, xv, g, xt, gt);
  auto res = mm::solve_lcp(lcp, pgd, state);
  for (long i = 0; i < n; ++i) x[i] = xv(i);
  *iters = res.num_iters;
  *residual = res.residual;
  return res.converged ? 1 : 0;
}

// MundyMathBackend (in-kernel fixed-size backend, convex.hpp:288-350) on the 3x3 problems of UnitTestConvex.cpp:46-143.
int ref_solve_cqpp_mm3(const double* A, const double* q, double* x, int space_kind, double lo, double hi,
                       unsigned max_iters, double tol, unsigned* iters, double* residual) {
  mm::Matrix3<double> Am(A[0], A[1], A[2], A[3], A[4], A[5], A[6], A[7], A[8]);
  mm::Vector<double, 3> qv{q[0], q[1], q[2]}, xv{x[0], x[1], x[2]}, g{}, xt{}, gt{};
  mm::convex::PGDConfig<double> cfg{max_iters, tol};
  namespace sp = mm::convex::space;
  auto run = [&](const auto& space) {
    auto prob = mm::make_mundy_math_cqpp(Am, qv, space);
    auto backend = prob.backend();
    auto pgd = mm::make_pgd_solution_strategy(backend, cfg);
    auto state = mm::make_pgd_state(backend, xv, g, xt, gt);
    auto res = mm::solve_cqpp(prob, pgd, state);
    *iters = res.num_iters;
    *residual = res.residual;
    return res.converged ? 1 : 0;
  };
  int c;
  if (space_kind == 0) {
    c = run(sp::Unconstrained<double>{});
  } else if (space_kind == 1) {
    c = run(sp::LowerBound<double>{lo});
  } else if (space_kind == 2) {
    c = run(sp::UpperBound<double>{hi});
  } else {
    c = run(sp::Bounded<double>{lo, hi});
  }
  for (int k = 0; k < 3; ++k) x[k] = xv[k];
  return c;
}

}  // extern "C"
