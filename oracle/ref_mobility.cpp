// TEST INFRASTRUCTURE ONLY. Second translation unit of oracle/_ref/libmundy_ref.so: the reference's OWN mobility
// kernels, compiled unmodified from scrap/parameter_interface/alens/src/mundy_alens/periphery/Periphery.hpp behind
// oracle/kokkos_shim/ (team policies run serially in index order, as on the Kokkos Serial backend), so that tests can
// pin the restatement in oracle/mundy_oracle_mobility.inc and generate tests/golden/mobility_ref.npz.
//   apply_stokes_kernel :686-747   apply_rpy_kernel :847-942   apply_rpyc_kernel :954-1085   apply_local_drag :1316-1354
#include <iomanip>  // Periphery.hpp uses std::setprecision without including it

#include <Kokkos_Core.hpp>

#include <mundy_alens/periphery/Periphery.hpp>

namespace mp = mundy::alens::periphery;
using View1 = Kokkos::View<double*, Kokkos::LayoutLeft, Kokkos::HostSpace>;

extern "C" {

// kind: 0 stokes, 1 rpy, 2 rpyc. tvel is accumulated into (+=), as in the reference.
int ref_mobility_kernel(int kind, long nt, long ns, double viscosity, const double* tpos, const double* spos,
                        const double* srad, const double* trad, const double* sforce, double* tvel) {
  View1 tp(const_cast<double*>(tpos), 3 * nt), sp(const_cast<double*>(spos), 3 * ns);
  View1 sr(const_cast<double*>(srad), ns), tr(const_cast<double*>(trad), nt);
  View1 sf(const_cast<double*>(sforce), 3 * ns), tv(tvel, 3 * nt);
  const Kokkos::DefaultHostExecutionSpace space;
  if (kind == 0) {
    mp::apply_stokes_kernel(space, viscosity, sp, tp, sf, tv);
  } else if (kind == 1) {
    mp::apply_rpy_kernel(space, viscosity, sp, tp, sr, tr, sf, tv);
  } else if (kind == 2) {
    mp::apply_rpyc_kernel(space, viscosity, sp, tp, sr, tr, sf, tv);
  } else {
    return -1;
  }
  return 0;
}

int ref_local_drag(long n, double viscosity, const double* rad, const double* force, double* vel) {
  View1 r(const_cast<double*>(rad), n), f(const_cast<double*>(force), 3 * n), v(vel, 3 * n);
  mp::apply_local_drag(Kokkos::DefaultHostExecutionSpace(), viscosity, v, f, r);
  return 0;
}

double ref_quake_inv_sqrt(double x) { return mp::quake_inv_sqrt(x); }

}  // extern "C"
