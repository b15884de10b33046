// TEST INFRASTRUCTURE ONLY. Fourth translation unit of oracle/_ref/libmundy_ref.so: the reference's space-filling-curve
// specifications, compiled unmodified from /root/reference behind oracle/kokkos_shim/.
//   math::hilbert_3d                 mundy/math/src/mundy_math/Hilbert.hpp:48-85  (lattice visiting order)
//   zorder_knn::Less<Point, 3>       mundy/math/src/mundy_math/zmort.hpp:195-220  (z-order comparator)
//   math::zmorton_less               mundy/math/src/mundy_math/zmort.hpp:228-266
#include <Kokkos_Core.hpp>

#include <vector>

#include <mundy_math/Hilbert.hpp>
#include <mundy_math/Vector3.hpp>
#include <mundy_math/zmort.hpp>

using V3 = mundy::math::Vector3<double>;

extern "C" {

// out[3 * s^3]: lattice points in visiting order, start (0,0,0), frame (x, y, z) as in UnitTestHilbert.cpp:48-66
void ref_hilbert_3d(long s, double* out) {
  std::vector<mundy::math::Vector3d> pos(static_cast<size_t>(s * s * s));
  mundy::math::hilbert_3d(static_cast<size_t>(s), 0, pos, mundy::math::Vector3d(0.0, 0.0, 0.0),
                          mundy::math::Vector3d(1.0, 0.0, 0.0), mundy::math::Vector3d(0.0, 1.0, 0.0),
                          mundy::math::Vector3d(0.0, 0.0, 1.0));
  for (size_t i = 0; i < pos.size(); ++i)
    for (int k = 0; k < 3; ++k) out[3 * i + k] = pos[i][k];
}

// which = 0: zorder_knn::Less<V3, 3>; which = 1: mundy::math::zmorton_less.  out[i] = less(p_i, q_i)
void ref_zorder_less(long n, int which, const double* p, const double* q, int* out) {
  zorder_knn::Less<V3, 3> less;
  for (long i = 0; i < n; ++i) {
    const V3 a(p[3 * i], p[3 * i + 1], p[3 * i + 2]), b(q[3 * i], q[3 * i + 1], q[3 * i + 2]);
    out[i] = which == 0 ? (less(a, b) ? 1 : 0) : (mundy::math::zmorton_less(a, b) ? 1 : 0);
  }
}

}  // extern "C"
