// Device form of the reference's derivative-free minimiser and of the three distances built on it.
//   math::find_min_using_approximate_derivatives   mundy/math/src/mundy_math/minimize.hpp:41-50,
//                                                   impl/minimize_impl.hpp (L-BFGS memory 10, central differences,
//                                                   Wolfe line search; nine starts per pair)
//   distance(Point, Ellipsoid)        mundy/geom/src/mundy_geom/distance/PointEllipsoid.hpp:97-135
//   distance(Ellipsoid, Ellipsoid)    mundy/geom/src/mundy_geom/distance/EllipsoidEllipsoid.hpp:109-151
//   distance(Circle3D, Circle3D)      mundy/geom/src/mundy_geom/distance/Circle3DCircle3D.hpp:88-151
// One thread runs one pair: the state (two angles, ten (s, y, rho) triples) lives in registers / local memory and the
// work is fp64 arithmetic (thousands of objective evaluations per pair), not HBM traffic. The operation order is the
// reference's; sin / cos are CUDA's (<= 2 ulp from libm), so results agree with the CPU to the minimiser's own
// tolerance, not bit for bit (tests/test_gpu_primitives.py states the bound).
#pragma once
#include "common.cuh"

// Every function is __host__ __device__: tests/cpp/host_minimize_check.cu instantiates this file's HOST side and
// tests/test_primitives_cpu.py checks it bit for bit against the reference (no GPU needed). The library itself exports
// device paths only.
#define MB200_HD __host__ __device__

namespace mb200 {
namespace prim {

constexpr double kEps = 2.220446049250313e-16;     // Kokkos::Experimental::epsilon_v<double>
constexpr double kRelaxedTol = 1e-8;               // Tolerance.hpp:53-69 get_relaxed_zero_tolerance<double>()
constexpr double kPi = 3.141592653589793238462643383279502884;

struct V2 {
  double a, b;
};
MB200_HD inline V2 operator+(const V2& u, const V2& v) { return {u.a + v.a, u.b + v.b}; }
MB200_HD inline V2 operator-(const V2& u, const V2& v) { return {u.a - v.a, u.b - v.b}; }
MB200_HD inline V2 operator*(double s, const V2& v) { return {s * v.a, s * v.b}; }
MB200_HD inline V2 neg(const V2& v) { return {-v.a, -v.b}; }
MB200_HD inline double dot2(const V2& u, const V2& v) { return u.a * v.a + u.b * v.b; }

MB200_HD inline Q4 qconj(const Q4& q) { return {q.w, -q.x, -q.y, -q.z}; }

// mundy/math/src/mundy_math/impl/minimize_impl.hpp:42-45
MB200_HD inline double put_in_range(double lo, double hi, double v) { return (v < lo) ? lo : (v > hi) ? hi : v; }

// Minimiser in [0, limit] of the cubic through two values and two slopes at 0 and 1 (minimize_impl.hpp:52-80).
MB200_HD inline double cubic_minimiser(double f0, double d0, double f1, double d1, double limit = 1.0) {
  const double n = 3 * (f1 - f0) - 2 * d0 - d1;
  const double e = d0 + d1 - 2 * (f1 - f0);
  const double disc = fmax(n * n - 3 * e * d0, 0.0);
  if (disc < 0 || fabs(e) <= kEps) return 0.5;
  const double root = sqrt(disc);
  const double x1 = (root - n) / (3 * e);
  const double x2 = -(root + n) / (3 * e);
  const double y1 = f0 + d0 * x1 + n * x1 * x1 + e * x1 * x1 * x1;
  const double y2 = f0 + d0 * x2 + n * x2 * x2 + e * x2 * x2 * x2;
  return put_in_range(0, limit, (y1 < y2) ? x1 : x2);
}

// One evaluation of the search function: step length, value, central-difference slope.
struct Probe {
  double step, value, slope;
};

// Wolfe line search of minimize_impl.hpp:233-398 (strong Wolfe conditions with constants rho, sigma; a bracketing
// phase that extrapolates by factors 1.4 .. 9, then a sectioning phase that keeps the trial inside [0.1, 0.5] of the
// bracket), evaluation for evaluation in the reference's order. phi(step) is the objective along the search direction;
// its slope is the central difference (phi(x + h) - phi(x - h)) / ((x + h) - (x - h)) (:217-220).
template <class F>
MB200_HD inline double line_search(const F& phi, double phi0, double slope0, double rho, double sigma,
                                     double floor_value, int max_evals, double h) {
  auto probe = [&](double x) {
    Probe p;
    p.step = x;
    p.value = phi(x);
    p.slope = (phi(x + h) - phi(x - h)) / ((x + h) - (x - h));
    return p;
  };
  auto too_high = [&](const Probe& p) { return p.value > phi0 + rho * p.step * slope0; };  // no sufficient decrease
  if (fabs(slope0) <= fabs(phi0) * kEps) return 0;
  if (phi0 <= floor_value) return 0;
  const double step_cap = (floor_value - phi0) / (rho * slope0);
  const double flat = fabs(sigma * slope0);
  double trial = put_in_range(0, 0.65 * step_cap, (step_cap < 0) ? -1.0 : 1.0);
  Probe prev{0.0, phi0, slope0}, lo, hi;
  int evals = 0;
  for (;;) {  // bracketing: walk outwards until the minimum along the direction is enclosed
    ++evals;
    const Probe cur = probe(trial);
    if (cur.value <= floor_value) return cur.step;
    if (too_high(cur) || cur.value >= prev.value) {
      lo = prev, hi = cur;
      break;
    }
    if (fabs(cur.slope) <= flat) return cur.step;
    if (prev.step == cur.step || evals >= max_evals) return cur.step;
    if (cur.slope >= 0) {
      lo = cur, hi = prev;
      break;
    }
    const double gap = cur.step - prev.step;
    double nearest = cur.step + 1.4 * gap, farthest = cur.step + 9 * gap;
    if (step_cap > 0)
      nearest = fmin(step_cap, nearest), farthest = fmin(step_cap, farthest);
    else
      nearest = fmax(step_cap, nearest), farthest = fmax(step_cap, farthest);
    const double guess =
        (prev.step < cur.step)
            ? prev.step + (cur.step - prev.step) * cubic_minimiser(prev.value, prev.slope, cur.value, cur.slope, 1e10)
            : cur.step + (prev.step - cur.step) * cubic_minimiser(cur.value, cur.slope, prev.value, prev.slope, 1e10);
    trial = put_in_range(nearest, farthest, guess);
    prev = cur;
  }
  for (;;) {  // sectioning: shrink [lo, hi] (hi may lie below lo) around a point that satisfies both conditions
    ++evals;
    const double width = hi.step - lo.step;
    const double nearest = lo.step + 0.1 * width;
    const double farthest = hi.step - 0.5 * width;
    const Probe cur = probe(put_in_range(
        nearest, farthest, lo.step + width * cubic_minimiser(lo.value, lo.slope, hi.value, hi.slope)));
    if (cur.value <= floor_value || evals >= max_evals) return cur.step;
    if (lo.step == nearest || hi.step == farthest) return hi.step;
    const double reach = fmax(fabs(lo.step), fabs(hi.step));
    if (fabs(reach * slope0) <= fabs(phi0) * kEps) return cur.step;
    if (too_high(cur) || cur.value >= lo.value) {
      hi = cur;
    } else {
      if (fabs(cur.slope) <= flat) return cur.step;
      if ((hi.step - lo.step) * cur.slope >= 0) hi = lo;
      lo = cur;
    }
  }
}

// minimize_impl.hpp:400-541  lbfgs_search_strategy<10, 2>
struct Lbfgs {
  static constexpr int kMax = 10;
  V2 s[kMax], y[kMax];
  double rho[kMax], alpha[kMax];
  bool been_used = false;
  int size = 0;
  V2 prev_x{0, 0}, prev_der{0, 0}, dir{0, 0};

  MB200_HD V2 next_direction(const V2& x, const V2& g) {
    dir = neg(g);
    if (!been_used) {
      been_used = true;
    } else {
      const V2 ds = x - prev_x, dy = g - prev_der;
      const double temp = dot2(ds, dy);
      if (fabs(temp) > kEps) {
        const double r = 1.0 / temp;
        if (size < kMax) {
          s[size] = ds, y[size] = dy, rho[size] = r;
          ++size;
        } else {
          for (int i = 1; i < kMax; ++i) s[i - 1] = s[i], y[i - 1] = y[i], rho[i - 1] = rho[i];
          s[kMax - 1] = ds, y[kMax - 1] = dy, rho[kMax - 1] = r;
        }
      } else {
        size = 0;
      }
      if (size > 0) {
        for (int i = size - 1; i >= 0; --i) {
          alpha[i] = rho[i] * dot2(s[i], dir);
          dir = dir - alpha[i] * y[i];
        }
        double H_0 = 1.0 / rho[size - 1] / dot2(y[size - 1], y[size - 1]);
        H_0 = put_in_range(0.001, 1000.0, H_0);
        dir = H_0 * dir;
        for (int i = 0; i < size; ++i) {
          const double beta = rho[i] * dot2(y[i], dir);
          dir = dir + (alpha[i] - beta) * s[i];
        }
      }
    }
    prev_x = x;
    prev_der = g;
    return dir;
  }
};

// minimize.hpp:41-50 + minimize_impl.hpp:543-575. The callers pass their `min_objective_delta` (1e-8) in the THIRD
// position, which is min_allowable_cost; the stop strategy keeps its default objective delta of 1e-7 and the derivative
// step its default 1e-7 (PointEllipsoid.hpp:118-119, EllipsoidEllipsoid.hpp:139-140, Circle3DCircle3D.hpp:140-141).
template <class Cost>
MB200_HD inline double find_min(const Cost& cost_func, V2& x, double min_allowable_cost) {
  const double eps = 1e-7, min_delta = 1e-7;
  auto grad = [&](const V2& p) {  // central_differences, :190-211
    V2 der;
    {
      const double old = p.a;
      const double fp = cost_func(V2{old + eps, p.b});
      const double fm = cost_func(V2{old - eps, p.b});
      der.a = (fp - fm) / ((old + eps) - (old - eps));
    }
    {
      const double old = p.b;
      const double fp = cost_func(V2{p.a, old + eps});
      const double fm = cost_func(V2{p.a, old - eps});
      der.b = (fp - fm) / ((old + eps) - (old - eps));
    }
    return der;
  };
  Lbfgs strat;
  double cost = cost_func(x);
  V2 g = grad(x);
  bool used = false;  // objective_delta_stop_strategy with max_iter = 0, :144-183
  double prev_val = 0;
  while (true) {
    if (used && fabs(cost - prev_val) < min_delta) break;
    used = true;
    prev_val = cost;
    if (!(cost > min_allowable_cost)) break;
    const V2 s = strat.next_direction(x, g);
    const V2 x0 = x;
    auto f = [&](double t) { return cost_func(x0 + t * s); };
    const double alpha = line_search(f, cost, dot2(g, s), 0.01, 0.9, min_allowable_cost, 100, eps);
    x = alpha * s + x;
    g = grad(x);
    cost = cost_func(x);
  }
  return cost;
}

struct Ellipsoid {
  V3 c;
  Q4 q;
  V3 rad;
};

// mundy/geom/src/mundy_geom/primitives/Ellipsoid.hpp:418-468
MB200_HD inline V3 body_normal_to_ellipsoid(const V3& nb, const Ellipsoid& e) {
  const double r1 = e.rad.x, r2 = e.rad.y, r3 = e.rad.z;
  const double sign0 = copysign(1.0, nb.x), sign1 = copysign(1.0, nb.y), sign2 = copysign(1.0, nb.z);
  double alpha1, alpha2;
  if (sign0 * nb.x > kZeroTol) {
    const double tmp0 = 1.0 / (r1 * nb.x);
    const double tmp1 = tmp0 * r2 * nb.y;
    const double tmp2 = tmp0 * r3 * nb.z;
    alpha1 = 1.0 / (1.0 + tmp1 * tmp1);
    alpha2 = 1.0 / (1.0 + tmp2 * tmp2 * alpha1);
  } else if (sign1 * nb.y > kZeroTol) {
    const double tmp = r3 * nb.z / (r2 * nb.y);
    alpha1 = 0.0;
    alpha2 = 1.0 / (1.0 + tmp * tmp);
  } else {
    alpha1 = 0.0;
    alpha2 = 0.0;
  }
  const double sa1 = sqrt(alpha1), sa2 = sqrt(alpha2);
  const double x = 0.5 * sign0 * ((1.0 + sign0) * r1 + (1.0 - sign0) * r1) * sa1 * sa2;
  const double y = 0.5 * sign1 * ((1.0 + sign1) * r2 + (1.0 - sign1) * r2) * sqrt(1.0 - alpha1) * sa2;
  const double z = 0.5 * sign2 * ((1.0 + sign2) * r3 + (1.0 - sign2) * r3) * sqrt(1.0 - alpha2);
  return {x, y, z};
}
MB200_HD inline V3 normal_to_foot_point(const V3& n, const Ellipsoid& e) {
  const V3 nb = qrot(qconj(e.q), n);
  return qrot(e.q, body_normal_to_ellipsoid(nb, e)) + e.c;
}
MB200_HD inline V3 angles_to_normal(const V2& tp) {
  const double st = sin(tp.a), ct = cos(tp.a), sp = sin(tp.b), cp = cos(tp.b);
  return {st * cp, st * sp, ct};
}

// PointEllipsoid.hpp:97-135
MB200_HD inline double dist_point_ellipsoid(const V3& x, const Ellipsoid& e, V3& cp, V3& normal) {
  auto obj = [&](const V2& tp) {
    const V3 n = angles_to_normal(tp);
    V3 sep;
    return dist_pp(normal_to_foot_point(n, e), x, sep);
  };
  const double th[3] = {0.0, 0.5 * kPi, kPi};
  const double one_third_pi = kPi / 3.0;
  const double ph[3] = {one_third_pi, kPi, 5.0 * one_third_pi};
  double best = INFINITY;
  V2 best_sol{0, 0};
  for (int t = 0; t < 3; ++t)
    for (int p = 0; p < 3; ++p) {
      V2 sol{th[t], ph[p]};
      const double d = find_min(obj, sol, kRelaxedTol);
      if (d < best) best = d, best_sol = sol;
    }
  normal = angles_to_normal(best_sol);
  cp = normal_to_foot_point(normal, e);
  return dot(x - cp, normal);
}

// EllipsoidEllipsoid.hpp:109-151
MB200_HD inline double dist_ellipsoid_ellipsoid(const Ellipsoid& e1, const Ellipsoid& e2, V3& cp1, V3& cp2, V3& n1) {
  auto obj = [&](const V2& tp) {
    const V3 n = angles_to_normal(tp);
    const V3 m{-n.x, -n.y, -n.z};
    V3 sep;
    return dist_pp(normal_to_foot_point(n, e1), normal_to_foot_point(m, e2), sep);
  };
  const double th[3] = {0.0, 0.5 * kPi, kPi};
  const double one_third_pi = kPi / 3.0;
  const double ph[3] = {one_third_pi, kPi, 5.0 * one_third_pi};
  double best = INFINITY;
  V2 best_sol{0, 0};
  for (int t = 0; t < 3; ++t)
    for (int p = 0; p < 3; ++p) {
      V2 sol{th[t], ph[p]};
      const double d = find_min(obj, sol, kRelaxedTol);
      if (d < best) best = d, best_sol = sol;
    }
  n1 = angles_to_normal(best_sol);
  cp1 = normal_to_foot_point(n1, e1);
  cp2 = normal_to_foot_point(V3{-n1.x, -n1.y, -n1.z}, e2);
  return dot(cp2 - cp1, n1);
}

struct Circle {
  V3 c;
  Q4 q;
  double r;
};
// Circle3DCircle3D.hpp:81-86
MB200_HD inline V3 circle_point(double theta, const Circle& k) {
  const V3 pl{k.r * cos(theta), k.r * sin(theta), 0.0};
  return qrot(k.q, pl) + k.c;
}
// Circle3DCircle3D.hpp:88-151 (Euclidean)
MB200_HD inline double dist_circle_circle(const Circle& k1, const Circle& k2, V3& cp1, V3& cp2, V3& n1) {
  auto obj = [&](const V2& tt) {
    V3 sep;
    return dist_pp(circle_point(tt.a, k1), circle_point(tt.b, k2), sep);
  };
  const double one_third_pi = kPi / 3.0;
  const double th[3] = {one_third_pi, kPi, 5.0 * one_third_pi};
  double best = INFINITY;
  V2 best_sol{0, 0};
  for (int t = 0; t < 3; ++t)
    for (int p = 0; p < 3; ++p) {
      V2 sol{th[t], th[p]};
      const double d = find_min(obj, sol, kRelaxedTol);
      if (d < best) best = d, best_sol = sol;
    }
  cp1 = circle_point(best_sol.a, k1);
  cp2 = circle_point(best_sol.b, k2);
  V3 sep;
  const double nrm = dist_pp(cp1, cp2, sep);
  const double den = (nrm > kZeroTol) ? nrm : 1.0;
  n1 = {sep.x / den, sep.y / den, sep.z / den};
  return best;
}

}  // namespace prim
}  // namespace mb200
