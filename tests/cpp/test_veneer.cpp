// C++ parity test of the veneer (include/mundy_b200/*.hpp) written the way the reference's GoogleTests call the API.
// Built by tests/test_cpp_veneer.py with g++ against libmundy_b200.so; runs on a GPU box (exit 0 = all checks passed).
// Known answers: mundy/geom/tests/unit_tests/UnitTestComputeAABB.cpp:170-261, UnitTestSegmentSegment.cpp:449-472,
// mundy/math/tests/unit_tests/UnitTestConvex.cpp:46-143, mundy/mesh/tests/unit_tests/UnitTestGenNeighborLinks.cpp:73-152,
// scrap/.../UnitTestComputeSignedSeparationDistanceAndContactNormal.cpp:300-330 (sphere-sphere -0.5, (1/3,2/3,2/3)).
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>

#include "mundy_b200/convex.hpp"
#include "mundy_b200/geom.hpp"
#include "mundy_b200/mobility.hpp"
#include "mundy_b200/neighbor_links.hpp"

using namespace mundy_b200;
using geom::AABB;
using geom::LineSegment;
using geom::Point;
using geom::Quaternion;
using geom::Sphere;
using geom::Spherocylinder;
using geom::SpherocylinderSegment;

static int g_checks = 0, g_failed = 0;
#define EXPECT_TRUE(c)                                                      \
  do {                                                                      \
    ++g_checks;                                                             \
    if (!(c)) {                                                             \
      ++g_failed;                                                           \
      std::printf("FAILED %s:%d  %s\n", __FILE__, __LINE__, #c);            \
    }                                                                       \
  } while (0)
#define EXPECT_NEAR(a, b, tol) EXPECT_TRUE(std::fabs((a) - (b)) <= (tol))

static void expect_aabb(const AABB& got, double x0, double y0, double z0, double x1, double y1, double z1) {
  const double e[6] = {x0, y0, z0, x1, y1, z1};
  double err = 0;
  for (int k = 0; k < 6; ++k) err += std::fabs(got[k] - e[k]);
  EXPECT_TRUE(err < 1e-8);  // UnitTestComputeAABB.cpp:53-56,72
}

struct ScaledIdentity {  // a user LinearOp: y = s x (exercises the generic solve_cqpp loop)
  double s;
  void apply(const DeviceVector<double>& x, DeviceVector<double>& y) const {
    check(mb200_vec_axpby(nullptr, static_cast<int64_t>(x.size()), s, x.data(), 0.0, y.data()), "axpby");
  }
};

int main() {
  if (mb200_device_count() < 1) {
    std::printf("no CUDA device: the veneer has no CPU path\n");
    return 77;
  }
  // ---- compute_aabb ----
  const double c45 = std::sqrt(0.5);
  const Quaternion x90{c45, c45, 0, 0};
  expect_aabb(geom::compute_aabb(Sphere{Point{0, 0, 0}, 1}), -1, -1, -1, 1, 1, 1);
  expect_aabb(geom::compute_aabb(Spherocylinder{Point{1, -2, 3}, Quaternion{1, 0, 0, 0}, 4, 0}), -3, -6, -1, 5, 2, 7);
  expect_aabb(geom::compute_aabb(Spherocylinder{Point{1, -2, 3}, Quaternion{1, 0, 0, 0}, 0, 4}), 1, -2, 1, 1, -2, 5);
  expect_aabb(geom::compute_aabb(Spherocylinder{Point{1, -2, 3}, Quaternion{1, 0, 0, 0}, 2, 4}), -1, -4, -1, 3, 0, 7);
  expect_aabb(geom::compute_aabb(Spherocylinder{Point{1, -2, 3}, x90, 2, 3}), -1, -5.5, 1, 3, 1.5, 5);
  expect_aabb(geom::compute_aabb(SpherocylinderSegment{Point{1, -2, 3}, Point{1, -2, 3}, 4}), -3, -6, -1, 5, 2, 7);
  expect_aabb(geom::compute_aabb(SpherocylinderSegment{Point{1, -3.5, 3}, Point{1, -0.5, 3}, 2}), -1, -5.5, 1, 3, 1.5, 5);
  // remaining primitives (UnitTestComputeAABB.cpp:137-205)
  expect_aabb(geom::compute_aabb(Point{1, -2, 3}), 1, -2, 3, 1, -2, 3);
  expect_aabb(geom::compute_aabb(LineSegment{Point{4, 5, 6}, Point{1, -2, 3}}), 1, -2, 3, 4, 5, 6);
  expect_aabb(geom::compute_aabb(geom::Ellipsoid{Point{1, -2, 3}, Quaternion{1, 0, 0, 0}, geom::Vector3{4, 5, 6}}), -3, -7,
              -3, 5, 3, 9);
  expect_aabb(geom::compute_aabb(geom::Ellipsoid{Point{1, -2, 3}, x90, geom::Vector3{4, 5, 6}}), -3, -8, -2, 5, 4, 8);
  EXPECT_TRUE(geom::intersects(AABB({0, 0, 0}, {1, 1, 1}), AABB({1, 1, 1}, {2, 2, 2})));   // touching counts
  EXPECT_TRUE(!geom::intersects(AABB({0, 0, 0}, {1, 1, 1}), AABB({1.1, 0, 0}, {2, 1, 1})));

  // ---- distance, remaining primitives ----
  {
    // UnitTestEllipsoidEllipsoid.cpp:119-235, tolerance 1e-4 (:53)
    const Quaternion id{1, 0, 0, 0}, x_to_y{c45, 0, 0, c45};
    const geom::Ellipsoid e0{Point{0, 0, 0}, id, 3.0, 1.0, 2.0};
    EXPECT_NEAR(geom::distance(geom::SharedNormalSigned{}, e0, e0), -2.0, 1e-4);
    EXPECT_NEAR(geom::distance(e0, geom::Ellipsoid{Point{0, 0, 0}, id, 6.0, 2.0, 4.0}), -3.0, 1e-4);
    for (double want : {0.2, -0.2, 0.0}) {
      const geom::Ellipsoid a{Point{-3.0 - 0.5 * want, 0, 0}, id, 3.0, 1.0, 2.0}, b{Point{3.0 + 0.5 * want, 0, 0}, id, 3.0, 1.0, 2.0};
      Point p1, p2;
      geom::Vector3 n1, n2;
      EXPECT_NEAR(geom::distance(geom::SharedNormalSigned{}, a, b, p1, p2, n1, n2), want, 1e-4);
      EXPECT_NEAR(n1[0], 1.0, 1e-4);
      EXPECT_NEAR(n2[0], -1.0, 1e-4);
      const geom::Ellipsoid up{Point{0, 4.0 + want, 0}, x_to_y, 3.0, 1.0, 2.0};
      EXPECT_NEAR(geom::distance(up, e0), want, 1e-4);
    }
    // a sphere written as an ellipsoid (:54-117)
    EXPECT_NEAR(geom::distance(Point{4, 4, 2}, geom::Ellipsoid{Point{0, 0, 0}, x90, 2.5, 2.5, 2.5}), 3.5, 1e-4);
    geom::Vector3 sep;
    EXPECT_NEAR(geom::distance(Point{4, 4, 2}, Sphere{Point{0, 0, 0}, 2.5}, sep), 3.5, 1e-14);
    // skew lines x-axis and (0,1,z): distance 1 at arch lengths 0 / 0
    Point q1, q2;
    double a1, a2;
    EXPECT_NEAR(geom::distance(geom::Line{Point{0, 0, 0}, {1, 0, 0}}, geom::Line{Point{0, 1, 0}, {0, 0, 1}}, q1, q2, a1, a2, sep),
                1.0, 1e-14);
    EXPECT_NEAR(sep[1], 1.0, 1e-14);
    EXPECT_NEAR(geom::distance(Point{3, 4, 0}, geom::Line{Point{0, 0, 0}, {1, 0, 0}}, q1, a1, sep), 4.0, 1e-14);
    EXPECT_NEAR(a1, 3.0, 1e-14);
    EXPECT_NEAR(geom::distance(geom::Line{Point{0, 0, 0}, {1, 0, 0}}, Sphere{Point{3, 4, 0}, 1.5}, q1, a1, sep), 2.5, 1e-14);
    // two unit circles in parallel planes one apart, coaxial: Euclidean distance 1
    EXPECT_NEAR(geom::distance(geom::Circle3D{Point{0, 0, 0}, id, 1.0}, geom::Circle3D{Point{0, 0, 1}, id, 1.0}), 1.0, 1e-6);
  }

  // ---- distance ----
  {
    const LineSegment a({9.64101615137754, 6, 3.18961417478521}, {9.64101615137754, 6, 8.189614174785209});
    const LineSegment b({10.39230484541326, 6, 0.6472696138825587}, {10.39230484541326, 6, 5.647269613882559});
    const double d = geom::distance(a, b), drev = geom::distance(b, a);
    EXPECT_NEAR(d, 0.7512886940357237, 1e-6);
    EXPECT_NEAR(d, drev, 1e-6);
    geom::Vector3 sep;
    const double ss = geom::distance(geom::SharedNormalSigned{}, Sphere{Point{0, 0, 0}, 1}, Sphere{Point{1, 2, 2}, 2.5}, sep);
    EXPECT_NEAR(ss, -0.5, 1e-14);
    const double sn = std::sqrt(sep[0] * sep[0] + sep[1] * sep[1] + sep[2] * sep[2]);
    EXPECT_NEAR(sn, 0.5, 1e-14);
    Point cp;
    double t;
    const double dp = geom::distance(Point{0, 1, 0}, LineSegment({-1, 0, 0}, {1, 0, 0}), cp, t, sep);
    EXPECT_NEAR(dp, 1.0, 1e-15);
    EXPECT_NEAR(t, 0.5, 1e-15);
    const double dsph = geom::distance(LineSegment({-1, 0, 0}, {1, 0, 0}), Sphere{Point{0, 2, 0}, 0.5}, cp, t, sep);
    EXPECT_NEAR(dsph, 1.5, 1e-15);
  }

  // ---- math::convex: tridiagonal 3x3 of UnitTestConvex.cpp:46-143, started from x = 99.99 ----
  {
    using namespace math;
    const std::vector<double> A = {4, -1, 0, -1, 4, -1, 0, -1, 4}, xs = {1, 2, 3};
    std::vector<double> q(3);
    for (int i = 0; i < 3; ++i) q[i] = -(A[3 * i] * xs[0] + A[3 * i + 1] * xs[1] + A[3 * i + 2] * xs[2]);
    DeviceVector<double> dA(A), dq(q);
    const convex::DenseMatrix Aop{dA.data(), 3};
    convex::B200Backend be{nullptr};
    const convex::PGDConfig cfg{1000, 1e-6};
    auto run = [&](auto prob, auto strat) {
      DeviceVector<double> x(std::vector<double>(3, 99.99)), g(3), xt(3), gt(3);
      auto state = make_pgd_state(be, x, g, xt, gt);
      const convex::SolveResult r = solve_cqpp(prob, strat, state);
      return std::make_pair(r, x.to_host());
    };
    auto [r0, x0] = run(make_b200_cqpp(nullptr, Aop, dq, convex::Unconstrained{}), make_pgd_solution_strategy(be, cfg));
    EXPECT_TRUE(r0.converged);
    for (int i = 0; i < 3; ++i) EXPECT_NEAR(x0[i], xs[i], 1e-5);
    auto [r1, x1] = run(make_b200_cqpp(nullptr, Aop, dq, convex::Box{0.0, 2.5}), make_pgd_solution_strategy(be, cfg));
    EXPECT_TRUE(r1.converged && x1[2] == 2.5 && x1[0] <= 2.5 && x1[1] <= 2.5);
    // the projected-gradient residual policy (:434-466) is written for the bound x >= 0
    auto [r2, x2] = run(make_b200_cqpp(nullptr, Aop, dq, convex::LowerBound{0.0}),
                        make_pgd_solution_strategy(be, convex::BBStepStrategy{}, convex::LinfNormProjectedGradientResidual{}, cfg));
    EXPECT_TRUE(r2.converged);
    for (int i = 0; i < 3; ++i) EXPECT_NEAR(x2[i], xs[i], 1e-5);
    // LCP through solve_lcp: w = A x + q >= 0, x >= 0, x.w = 0; the unconstrained optimum is positive => same solution
    {
      DeviceVector<double> x(std::vector<double>(3, 99.99)), g(3), xt(3), gt(3);
      auto state = make_pgd_state(be, x, g, xt, gt);
      const auto r = solve_lcp(make_b200_lcp(nullptr, Aop, dq), make_pgd_solution_strategy(be, cfg), state);
      const auto xh = x.to_host();
      EXPECT_TRUE(r.converged);
      for (int i = 0; i < 3; ++i) EXPECT_NEAR(xh[i], xs[i], 1e-5);
    }
    // a user-defined LinearOp goes through the generic loop: min 0.5 s x^2 + q x over x >= 0  ->  x = max(-q/s, 0)
    {
      const ScaledIdentity S{2.0};
      DeviceVector<double> qq(std::vector<double>{-4.0, 6.0, -1.0}), x(std::vector<double>(3, 99.99)), g, xt, gt;
      auto state = make_pgd_state(be, x, g, xt, gt);
      const auto r = solve_cqpp(make_b200_cqpp(nullptr, S, qq, convex::LowerBound{0.0}), make_pgd_solution_strategy(be, cfg), state);
      const auto xh = x.to_host();
      EXPECT_TRUE(r.converged);
      EXPECT_NEAR(xh[0], 2.0, 1e-5);
      EXPECT_NEAR(xh[1], 0.0, 1e-5);
      EXPECT_NEAR(xh[2], 0.5, 1e-5);
    }
  }

  // ---- GenNeighborLinks: two overlapping spheres -> exactly one link; builder protocol; a collision step ----
  {
    DeviceVector<double> center(std::vector<double>{0, 0, 0, 1.5, 0, 0, 10, 0, 0}), radius(std::vector<double>{1, 1, 1});
    mesh::Bodies bodies;
    bodies.n = 3, bodies.center = center.data(), bodies.radius = radius.data();
    mesh::GenNeighborLinks gen;
    gen.set_search_buffer(0.0).set_sort_search_results(true).set_enforce_source_target_symmetry(true)
        .set_source_target_rank(mesh::EntityRank::ELEM_RANK, mesh::EntityRank::ELEM_RANK)
        .set_search_method(mesh::SearchMethod::MORTON_LBVH);
    bool rank_refused = false;  // GenNeighborLinkers.hpp:413-415: element-to-element only
    try {
      gen.set_source_target_rank(mesh::EntityRank::NODE_RANK, mesh::EntityRank::ELEM_RANK);
    } catch (const std::runtime_error& e) {
      rank_refused = std::strstr(e.what(), "element to element") != nullptr;
    }
    EXPECT_TRUE(rank_refused);
    gen.acts_on(bodies).concretize();
    bool method_refused = false;
    try {
      gen.set_search_method(mesh::SearchMethod::KDTREE);
    } catch (const std::runtime_error& e) {
      method_refused = std::strstr(e.what(), "after concretization") != nullptr;
    }
    EXPECT_TRUE(method_refused);
    bool threw = false;
    try {
      gen.set_search_buffer(1.0);
    } catch (const std::runtime_error& e) {
      threw = std::strstr(e.what(), "after concretization") != nullptr;
    }
    EXPECT_TRUE(threw);
    EXPECT_TRUE(gen.generate());    // first call builds
    EXPECT_TRUE(!gen.generate());   // nothing moved: no rebuild (GenNeighborLinkers.hpp:802-819)
    EXPECT_TRUE(gen.num_links() == 1);
    const auto links = gen.links().to_host();
    EXPECT_TRUE(links.size() == 2 && links[0] == 0 && links[1] == 1);
    // one step: the overlap of 0.5 is removed to tolerance, the far sphere does not move
    const auto res = gen.collision_step(1e-3, 5e-3, 1e-8, 10000);
    EXPECT_TRUE(res.ite_count > 0 && res.max_abs_projected_sep < 1e-8);
    const auto c = center.to_host();
    EXPECT_NEAR(c[3] - c[0], 2.0, 1e-7);
    EXPECT_TRUE(c[6] == 10.0 && c[1] == 0.0 && c[2] == 0.0);
    // the STK drivers throw when the iteration cap is hit
    DeviceVector<double> center2(std::vector<double>{0, 0, 0, 1.5, 0, 0, 2.5, 0.5, 0});
    bodies.center = center2.data();
    mesh::GenNeighborLinks gen2;
    gen2.set_search_buffer(0.5).acts_on(bodies).concretize();
    threw = false;
    try {
      gen2.collision_step(1e-3, 5e-3, 1e-14, 2);
    } catch (const std::runtime_error& e) {
      threw = std::string(e.what()) == "Collision resolution did not converge!";
    }
    EXPECT_TRUE(threw);
  }

  // ---- periodic metrics (UnitTestPeriodicity.cpp:623-791 properties on single objects) and a periodic link build ----
  {
    const auto metric = geom::periodic_scaled_metric_from_unit_cell({100.0, 100.0, 100.0});
    EXPECT_TRUE(metric.num_periodic_dimensions() == 3 && metric.is_periodic(1));
    const geom::Vector3 d = metric.sep(Point{1, 50, 99}, Point{99, 50.5, 2});
    EXPECT_NEAR(d[0], -2.0, 1e-12);
    EXPECT_NEAR(d[1], 0.5, 1e-12);
    EXPECT_NEAR(d[2], 3.0, 1e-12);
    const Point w = metric.wrap(Point{-1, 250, 100});
    EXPECT_NEAR(w[0], 99.0, 1e-12);
    EXPECT_NEAR(w[1], 50.0, 1e-12);
    EXPECT_NEAR(w[2], 0.0, 1e-12);
    const Sphere us = geom::unwrap_points_to_ref(Sphere{Point{99, 1, 50}, 2.0}, metric, Point{101, 199, 250});
    EXPECT_NEAR(us.center()[0], 99.0, 1e-12);
    EXPECT_NEAR(us.center()[1], 201.0, 1e-12);
    EXPECT_NEAR(us.center()[2], 250.0, 1e-12);
    EXPECT_TRUE(us.radius() == 2.0);
    const LineSegment seg = geom::unwrap_points_to_ref(LineSegment({99, 5, 5}, {2, 5, 5}), metric, Point{98, 5, 5});
    EXPECT_NEAR(seg.end()[0] - seg.start()[0], 3.0, 1e-9);  // the spanning segment is whole again
    const auto mx = geom::PeriodicMetricX(10.0);
    const geom::Vector3 dx = mx.sep(Point{1, 1, 1}, Point{9, 9, 9});
    EXPECT_NEAR(dx[0], -2.0, 1e-12);
    EXPECT_NEAR(dx[1], 8.0, 1e-12);
    // two spheres facing each other through the boundary: no link in free space, one link in the periodic box
    DeviceVector<double> center(std::vector<double>{0.5, 5, 5, 9.6, 5, 5}), radius(std::vector<double>{1, 1});
    mesh::Bodies bodies;
    bodies.n = 2, bodies.center = center.data(), bodies.radius = radius.data();
    mesh::GenNeighborLinks free_space, periodic;
    free_space.set_search_buffer(0.0).acts_on(bodies).concretize();
    periodic.set_search_buffer(0.0).set_periodic_metric(geom::periodic_scaled_metric_from_unit_cell({10.0, 10.0, 10.0}))
        .acts_on(bodies).concretize();
    EXPECT_TRUE(free_space.generate() && free_space.num_links() == 0);
    EXPECT_TRUE(periodic.generate() && periodic.num_links() == 1);
    const auto res = periodic.collision_step(1e-3, 5e-3, 1e-8, 10000);  // overlap 1.1 through the boundary is removed
    EXPECT_TRUE(res.max_abs_projected_sep < 1e-8);
    const auto c = center.to_host();
    EXPECT_NEAR((c[0] + 10.0) - c[3], 2.0, 1e-7);
  }

  // ---- mobility operators (UnitTestPeriphery.cpp:671-775 three spheres; NgpHP1.cpp:1487-1645 solve) ----
  {
    namespace per = alens::periphery;
    const double radius = 12.34, s = 6.0, pi = 3.14159265358979323846;
    const std::vector<double> pos{0, 0, 0, -s * radius / 2, -std::sqrt(3.0) * s * radius / 2, 0,
                                  s * radius / 2, -std::sqrt(3.0) * s * radius / 2, 0};
    std::vector<double> f(9, 0.0);
    f[1] = -6.0 * pi * radius;
    DeviceVector<double> positions(pos), radii(std::vector<double>(3, radius)), forces(f);
    DeviceVector<double> v_rpy(std::vector<double>(9, 0.0)), v_rpyc(std::vector<double>(9, 0.0));
    per::apply_rpy_kernel(per::DeviceSpace{}, 1.0, positions, positions, radii, radii, forces, v_rpy);
    per::apply_rpyc_kernel(per::DeviceSpace{}, 1.0, positions, positions, radii, radii, forces, v_rpyc);
    per::apply_local_drag(per::DeviceSpace{}, 1.0, v_rpy, forces, radii);  // "add self-interaction"
    per::apply_local_drag(per::DeviceSpace{}, 1.0, v_rpyc, forces, radii);
    const auto a = v_rpy.to_host(), b = v_rpyc.to_host();
    EXPECT_NEAR(-a[1], 1.0, 1e-12);         // U1
    EXPECT_NEAR(-a[4], 0.21586, 1e-3);      // U2 (Wilson 2013, s = 6)
    EXPECT_NEAR(a[6], 0.05078, 1e-3);       // U3
    EXPECT_NEAR(-b[4], 0.21586, 1e-4);
    bool threw = false;
    try {
      DeviceVector<double> short_forces(std::vector<double>(6, 0.0));
      per::apply_rpy_kernel(per::DeviceSpace{}, 1.0, positions, positions, radii, radii, short_forces, v_rpy);
    } catch (const std::invalid_argument&) {
      threw = true;
    }
    EXPECT_TRUE(threw);
    // the operator assigns: velocity = M force, whatever velocity held
    alens::SphereRpyMobilityOp rpy_op(1.0, positions, radii, MB200_KERNEL_RPYC);
    DeviceVector<double> u(std::vector<double>(9, 123.0));
    rpy_op.update();
    rpy_op.apply(forces, u);
    const auto uh = u.to_host();
    for (int k = 0; k < 9; ++k) EXPECT_NEAR(uh[k], b[k], 1e-15);
    // two overlapping spheres pushed apart with hydrodynamic coupling: the overlap is removed, both move (Newton pair)
    DeviceVector<double> p2(std::vector<double>{0, 0, 0, 1.5, 0, 0}), r2(std::vector<double>{1.0, 1.0});
    DeviceVector<int32_t> ids(std::vector<int32_t>{0, 1});
    DeviceVector<double> sep(std::vector<double>{-0.5}), nrm(std::vector<double>{1, 0, 0}), gam(std::vector<double>{0.0});
    DeviceVector<double> cf(std::vector<double>(6, 0.0)), cu(std::vector<double>(6, 0.0));
    alens::SphereRpyMobilityOp op2(1e-3, p2, r2, MB200_KERNEL_RPYC);
    const auto res = alens::resolve_collisions(1e-2, 1e-10, 1000, op2, ids, nullptr, nullptr, cf, cu, sep, nrm, gam);
    EXPECT_TRUE(res.max_abs_projected_sep < 1e-10 && res.ite_count >= 1);
    const auto uu = cu.to_host();
    EXPECT_NEAR(uu[3] - uu[0], 0.5 / 1e-2, 1e-6);  // relative normal velocity closes the overlap in one step
    EXPECT_NEAR(uu[0] + uu[3], 0.0, 1e-9);
    alens::SphereLocalDragMobilityOp dry(1e-3, r2);
    DeviceVector<double> gam2(std::vector<double>{0.0}), sep2(std::vector<double>{-0.5});
    const auto res_dry = alens::resolve_collisions(1e-2, 1e-10, 1000, dry, ids, nullptr, nullptr, cf, cu, sep2, nrm, gam2);
    EXPECT_TRUE(res_dry.max_abs_projected_sep < 1e-10);
    EXPECT_TRUE(gam2.to_host()[0] < gam.to_host()[0]);  // the coupling drags the partner along: more force is needed
  }

  std::printf("%d checks, %d failed\n", g_checks, g_failed);
  return g_failed == 0 ? 0 : 1;
}
