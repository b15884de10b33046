// TEST INFRASTRUCTURE ONLY. Never linked into, loaded by, or called from the product (mundy_b200/, include/).
//
// Sixth translation unit of oracle/_ref/libmundy_ref.so: the reference's OWN driver-LCP kernels and BBPGD loop,
//   scrap/lcp_spheres/KokkosLCP.cpp:449-491  compute_signed_separation_distance_and_contact_normal
//                                  :493-522  compute_maximum_abs_projected_sep
//                                  :524-557  compute_diff_dots
//                                  :559-591  sum_collision_force            (D)
//                                  :593-614  compute_the_mobility_problem   (M, dry branch)
//                                  :616-647  compute_rate_of_change_of_sep  (D^T)
//                                  :649-665  update_con_gammas
//                                  :675-876  resolve_collisions             (BBPGD, HydroType::DRY)
// cut from the reference tree at build time by oracle/extract_kokkos_lcp.py (into oracle/_ref/kokkos_lcp_extract.inc,
// git-ignored) and compiled UNMODIFIED against oracle/kokkos_shim/. What is stubbed, and why it does not matter:
//   * ArborX (BVH / query / BoundingSpheres / ExcludeDuplicateConstraints) is named only in the far-neighbour branch
//     :797-804, reached when enable_hydro is true; the wrappers below always pass enable_hydro = false.
//   * apply_rpy_kernel[_neighbors] (:219-335) are called from compute_the_mobility_problem only when enable_hydro.
// Every Kokkos parallel_for / parallel_reduce runs on ONE thread here (omp_set_num_threads(1) around each call): the
// order of the Kokkos Serial backend, in which the reference's atomic_add scatter and its reductions are deterministic.
#include <Kokkos_Core.hpp>

#include <iostream>
#include <stdexcept>
#ifdef _OPENMP
#include <omp.h>
#endif

// ---- stubs for names the extracted text mentions on branches that are never taken -------------------------------------
namespace ArborX {
template <class MemorySpace>
struct BVH {
  template <class Exec, class Prims>
  BVH(const Exec&, const Prims&) {}
};
template <class... Args>
inline void query(Args&&...) {
  Kokkos::abort("ArborX::query stub reached: the far-neighbour branch is out of scope (enable_hydro = false)");
}
}  // namespace ArborX
struct ExcludeDuplicateConstraints {};
struct BoundingSpheres {
  template <class V>
  BoundingSpheres(const V&, double) {}
};
template <class... Args>
inline void apply_rpy_kernel_neighbors(Args&&...) {
  Kokkos::abort("apply_rpy_kernel_neighbors stub reached (enable_hydro = false)");
}
template <class... Args>
inline void apply_rpy_kernel(Args&&...) {
  Kokkos::abort("apply_rpy_kernel stub reached (enable_hydro = false)");
}

namespace kokkos_lcp {
#include "_ref/kokkos_lcp_extract.inc"
}  // namespace kokkos_lcp

namespace {
using namespace kokkos_lcp;
using V1 = Kokkos::View<double*, DeviceMemorySpace>;
using V2 = Kokkos::View<double**, DeviceMemorySpace>;
using I1 = Kokkos::View<int*, DeviceMemorySpace>;
using I2 = Kokkos::View<int**, DeviceMemorySpace>;

struct SerialScope {  // Kokkos Serial backend order; also silences the driver's per-iteration std::cout chatter
  int saved = 1;
  std::ios_base::iostate cout_state;
  SerialScope() : cout_state(std::cout.rdstate()) {
#ifdef _OPENMP
    saved = omp_get_max_threads();
    omp_set_num_threads(1);
#endif
    std::cout.setstate(std::ios_base::failbit);
  }
  ~SerialScope() {
#ifdef _OPENMP
    omp_set_num_threads(saved);
#endif
    std::cout.clear(cout_state);
  }
};
inline V1 wrap1(const double* p, long n) { return V1(const_cast<double*>(p), n); }
inline V2 wrap2(const double* p, long n, long k) { return V2(const_cast<double*>(p), n, k); }
inline I2 wrap_ids(const int* p, long m) { return I2(const_cast<int*>(p), m, 2); }
}  // namespace

extern "C" {

struct ref_lcp_result {
  double max_abs_projected_sep;
  int ite_count;
  double max_displacement;
};

// :449-491
void ref_lcp_signed_sep_and_normal(long n, long m, const int* ids, double radius, const double* pos, double* sep,
                                   double* normal) {
  SerialScope s;
  I1 none;
  compute_signed_separation_distance_and_contact_normal(wrap_ids(ids, m), none, none, radius, wrap2(pos, n, 3),
                                                        wrap1(sep, m), wrap2(normal, m, 3));
}
// :493-522
double ref_lcp_max_abs_projected_sep(long m, const double* gamma, const double* sep, const double* sep_dot, double dt) {
  SerialScope s;
  double r = 0.0;
  compute_maximum_abs_projected_sep(wrap1(gamma, m), wrap1(sep, m), wrap1(sep_dot, m), dt, r);
  return r;
}
// :524-557  out = {xx, xg, gg}
void ref_lcp_diff_dots(long m, const double* gamma, const double* gamma_tmp, const double* sep_dot,
                       const double* sep_dot_tmp, double dt, double* out) {
  SerialScope s;
  compute_diff_dots(wrap1(gamma, m), wrap1(gamma_tmp, m), wrap1(sep_dot, m), wrap1(sep_dot_tmp, m), dt, out[0], out[1],
                    out[2]);
}
// :559-591  F = D gamma
void ref_lcp_sum_collision_force(long n, long m, const int* ids, const double* normal, const double* gamma, double* F) {
  SerialScope s;
  I2 idv = wrap_ids(ids, m);
  I1 none;
  V2 Fv = wrap2(F, n, 3), nv = wrap2(normal, m, 3);
  V1 gv = wrap1(gamma, m);
  sum_collision_force(idv, none, none, Fv, nv, gv);
}
// :593-614  U = M F, dry. Arguments in the order of the DECLARATION (sphere_radius, viscosity); resolve_collisions
// itself calls it with the two swapped (:711), which is what ref_lcp_resolve_collisions exercises.
void ref_lcp_mobility(long n, double radius, double viscosity, const double* F, double* U) {
  SerialScope s;
  I1 none;
  V2 Uv = wrap2(U, n, 3);
  compute_the_mobility_problem(radius, viscosity, V2(), wrap2(F, n, 3), none, none, Uv, false, false);
}
// :616-647  sep_dot = D^T U
void ref_lcp_rate_of_change_of_sep(long n, long m, const int* ids, const double* normal, const double* U,
                                   double* sep_dot) {
  SerialScope s;
  I2 idv = wrap_ids(ids, m);
  I1 none;
  V1 out = wrap1(sep_dot, m);
  compute_rate_of_change_of_sep(idv, none, none, wrap2(U, n, 3), wrap2(normal, m, 3), out);
}
// :649-665
void ref_lcp_update_con_gammas(long m, double* gamma, const double* gamma_tmp, const double* sep, const double* sep_dot,
                               double dt, double alpha) {
  SerialScope s;
  update_con_gammas(wrap1(gamma, m), wrap1(gamma_tmp, m), wrap1(sep, m), wrap1(sep_dot, m), dt, alpha);
}
// :675-876, enable_hydro = false. gamma in/out (initial guess / solution); F, U out. A fixed iteration count K is run
// as (max_allowable_overlap = 0, max_col_iterations = K): `res < 0` never holds, so the loop stops at the cap.
void ref_lcp_resolve_collisions(long n, long m, const int* ids, double radius, double viscosity, double dt,
                                double max_allowable_overlap, int max_col_iterations, const double* pos,
                                const double* sep, const double* normal, double* gamma, double* F, double* U,
                                ref_lcp_result* result) {
  SerialScope s;
  I2 idv = wrap_ids(ids, m);
  I1 none_a, none_b;
  V2 posv = wrap2(pos, n, 3), Uv = wrap2(U, n, 3), Fv = wrap2(F, n, 3), nv = wrap2(normal, m, 3);
  V1 sepv = wrap1(sep, m), gv = wrap1(gamma, m);
  const CollisionResult r = resolve_collisions(viscosity, dt, max_allowable_overlap, max_col_iterations, idv, none_a,
                                               none_b, radius, posv, Uv, Fv, sepv, nv, gv, false);
  result->max_abs_projected_sep = r.max_abs_projected_sep;
  result->ite_count = r.ite_count;
  result->max_displacement = r.max_displacement;
}

}  // extern "C"
