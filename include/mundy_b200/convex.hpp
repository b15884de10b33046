// mundy_b200/convex.hpp — the math::convex seam (mundy/math/src/mundy_math/convex.hpp) on device vectors.
//
// Kept from the reference: the convex spaces (:42-117), PGDConfig / SolveResult (:519-541), the problem factories
// make_*_cqpp / make_*_lcp (:722-757), make_pgd_solution_strategy / make_pgd_state (:759-787), solve_cqpp / solve_lcp
// (:789-845), and the LinearOp contract "dense matrix OR anything with void apply(const vector_t& x, vector_t& y) const"
// (:186-191). B200Backend is the counterpart of KokkosBackend (:141-285): the same static interface, implemented by the
// device primitives mb200_vec_* of the C-ABI. Two operators have a fused device fast path: a dense row-major matrix
// (mb200_solve_cqpp_dense) and the matrix-free collision operator dt D^T M D (mb200_solve_lcp_contact); any other LinearOp
// runs the generic loop below, which performs the reference's initialize / iterate sequence (:610-681) with
// those primitives and, like the KokkosBackend, two host-visible reductions per iteration.
#ifndef MUNDY_B200_CONVEX_HPP_
#define MUNDY_B200_CONVEX_HPP_

#include <type_traits>

#include "core.hpp"

namespace mundy_b200 {
namespace math {
namespace convex {

// ---- spaces ----
struct Unconstrained {
  static constexpr int kind = MB200_SPACE_UNCONSTRAINED;
  double lo() const { return 0; }
  double hi() const { return 0; }
};
struct LowerBound {
  double bound{0};
  static constexpr int kind = MB200_SPACE_LOWER;
  double lo() const { return bound; }
  double hi() const { return 0; }
};
struct UpperBound {
  double bound{0};
  static constexpr int kind = MB200_SPACE_UPPER;
  double lo() const { return 0; }
  double hi() const { return bound; }
};
struct Box {
  double lower{0}, upper{0};
  static constexpr int kind = MB200_SPACE_BOX;
  double lo() const { return lower; }
  double hi() const { return upper; }
};

// ---- residual policies (:434-496) ----
struct LinfNormProjectedDiffResidual {
  static constexpr int kind = MB200_RESID_PROJECTED_DIFF;
};
struct LinfNormProjectedGradientResidual {
  static constexpr int kind = MB200_RESID_PROJECTED_GRAD;
};
struct BBStepStrategy {};  // :498-516 (BB1 with the 1e-14 guard), the only step policy of the reference

struct PGDConfig {
  unsigned max_iters{1000};
  double tol{1e-8};  // get_relaxed_zero_tolerance<double>() (mundy/math/src/mundy_math/Tolerance.hpp:52-69)
};
struct SolveResult {
  unsigned num_iters{0};
  double residual{0};
  bool converged{false};
};

// ---- linear operators with a fused device path ----
struct DenseMatrix {  // row-major n x n, device pointer (the KokkosBlas::gemv path, :166-184)
  const double* data;
  int64_t n;
};
struct ContactOperator {  // A = dt D^T M D, matrix-free
  mb200_contact_op op;
};

// ---- backend: static interface of KokkosBackend (:141-285) on DeviceVector<double> ----
struct B200Backend {
  using scalar_t = double;
  using vector_t = DeviceVector<double>;
  mb200_stream stream = nullptr;

  static std::size_t vector_size(const vector_t& v) { return v.size(); }
  static double* vector_data(vector_t& v) { return v.data(); }
  void deep_copy(vector_t& dst, const vector_t& src) const { axpby(1.0, src, 0.0, dst); }
  template <class LinearOp>
  void apply(const LinearOp& A, const vector_t& x, vector_t& y) const {
    if constexpr (std::is_same_v<LinearOp, DenseMatrix>) {
      check(mb200_dense_gemv(stream, A.n, A.data, x.data(), y.data()), "mb200_dense_gemv");
    } else if constexpr (std::is_same_v<LinearOp, ContactOperator>) {
      check(mb200_apply_A(stream, &A.op, x.data(), y.data()), "mb200_apply_A");
    } else {
      A.apply(x, y);
    }
  }
  void axpby(double alpha, const vector_t& x, double beta, vector_t& y) const {
    check(mb200_vec_axpby(stream, static_cast<int64_t>(y.size()), alpha, x.data(), beta, y.data()), "mb200_vec_axpby");
  }
  template <class Space>
  void wrapped_axpbyz(double alpha, const vector_t& x, double beta, const vector_t& y, vector_t& z, const Space& sp) const {
    check(mb200_vec_wrapped_axpbyz(stream, static_cast<int64_t>(z.size()), alpha, x.data(), beta, y.data(), z.data(),
                                   Space::kind, sp.lo(), sp.hi()),
          "mb200_vec_wrapped_axpbyz");
  }
  double diff_dot(const vector_t& x, const vector_t& x_old) const {
    double out = 0;
    check(mb200_vec_diff_dot(stream, static_cast<int64_t>(x.size()), x.data(), x_old.data(), &out), "mb200_vec_diff_dot");
    return out;
  }
  double diff_dot(const vector_t& x, const vector_t& x_old, const vector_t& g, const vector_t& g_old) const {
    double out = 0;
    check(mb200_vec_diff_dot2(stream, static_cast<int64_t>(x.size()), x.data(), x_old.data(), g.data(), g_old.data(), &out),
          "mb200_vec_diff_dot2");
    return out;
  }
  template <class Space>
  double residual(int residual_kind, const vector_t& x, const vector_t& g, const Space& sp) const {
    double out = 0;
    check(mb200_vec_residual(stream, static_cast<int64_t>(x.size()), x.data(), g.data(), residual_kind, Space::kind,
                             sp.lo(), sp.hi(), &out),
          "mb200_vec_residual");
    return out;
  }
};

// ---- problems (:363-428): references to caller-owned A and q, as in the reference ----
template <class LinearOp, class Space>
struct CQPPProblem {
  B200Backend backend;
  const LinearOp& A;
  const DeviceVector<double>& q;
  Space space;
};
template <class LinearOp>
struct LCPProblem {
  B200Backend backend;
  const LinearOp& A;
  const DeviceVector<double>& q;
};
template <class LinearOp>
CQPPProblem<LinearOp, LowerBound> to_cqpp(const LCPProblem<LinearOp>& p) {  // :424-428
  return {p.backend, p.A, p.q, LowerBound{0.0}};
}

template <class ResidualPolicy = LinfNormProjectedDiffResidual>
struct PGDStrategy {
  B200Backend backend;
  PGDConfig cfg;
};
struct PGDState {  // :543-590: references to the four caller-owned work vectors
  DeviceVector<double>&x, &grad, &x_tmp, &grad_tmp;
};

}  // namespace convex

template <class LinearOp, class Space>
auto make_b200_cqpp(mb200_stream stream, const LinearOp& A, const DeviceVector<double>& q, const Space& space) {
  return convex::CQPPProblem<LinearOp, Space>{convex::B200Backend{stream}, A, q, space};
}
template <class LinearOp>
auto make_b200_lcp(mb200_stream stream, const LinearOp& A, const DeviceVector<double>& q) {
  return convex::LCPProblem<LinearOp>{convex::B200Backend{stream}, A, q};
}
inline auto make_pgd_solution_strategy(const convex::B200Backend& backend, const convex::PGDConfig& cfg = {}) {
  return convex::PGDStrategy<convex::LinfNormProjectedDiffResidual>{backend, cfg};
}
template <class ResidualPolicy>
auto make_pgd_solution_strategy(const convex::B200Backend& backend, convex::BBStepStrategy, ResidualPolicy,
                                const convex::PGDConfig& cfg = {}) {
  return convex::PGDStrategy<ResidualPolicy>{backend, cfg};
}
inline convex::PGDState make_pgd_state(const convex::B200Backend&, DeviceVector<double>& x, DeviceVector<double>& grad,
                                       DeviceVector<double>& x_tmp, DeviceVector<double>& grad_tmp) {
  return convex::PGDState{x, grad, x_tmp, grad_tmp};
}

/// solve_cqpp (:789-797). state.x holds the initial guess on entry and the solution on return.
template <class LinearOp, class Space, class ResidualPolicy>
convex::SolveResult solve_cqpp(const convex::CQPPProblem<LinearOp, Space>& prob,
                               const convex::PGDStrategy<ResidualPolicy>& strat, convex::PGDState& state) {
  using namespace convex;
  const int64_t n = static_cast<int64_t>(prob.q.size());
  if (state.x.size() != prob.q.size()) throw std::invalid_argument("solve_cqpp: size of x does not match q");  // :171-172
  const mb200_pgd_config cfg{strat.cfg.max_iters, strat.cfg.tol, Space::kind, prob.space.lo(), prob.space.hi(),
                             ResidualPolicy::kind};
  mb200_solve_result r{};
  if constexpr (std::is_same_v<LinearOp, DenseMatrix>) {
    check(mb200_solve_cqpp_dense(strat.backend.stream, n, prob.A.data, prob.q.data(), state.x.data(), &cfg, &r),
          "mb200_solve_cqpp_dense");
    return SolveResult{r.num_iters, r.residual, r.converged != 0};
  } else if constexpr (std::is_same_v<LinearOp, ContactOperator> && std::is_same_v<Space, LowerBound>) {
    if (prob.space.bound == 0.0) {
      check(mb200_solve_lcp_contact(strat.backend.stream, &prob.A.op, prob.q.data(), state.x.data(), &cfg, &r),
            "mb200_solve_lcp_contact");
      return SolveResult{r.num_iters, r.residual, r.converged != 0};
    }
  }
  // generic LinearOp: PGDStrategy::initialize / iterate (:610-681) on the backend primitives
  const B200Backend& be = strat.backend;
  DeviceVector<double>&x = state.x, &g = state.grad, &xt = state.x_tmp, &gt = state.grad_tmp;
  for (DeviceVector<double>* v : {&g, &xt, &gt})
    if (v->size() != x.size()) v->resize(x.size());
  SolveResult out;
  be.deep_copy(xt, x);
  be.apply(prob.A, xt, gt);
  be.axpby(1.0, prob.q, 1.0, gt);  // g_tmp = A x_tmp + q
  out.residual = be.residual(ResidualPolicy::kind, xt, gt, prob.space);
  if (out.residual <= strat.cfg.tol) {  // <=, unlike the drivers' <  (:633)
    out.converged = true;
    return out;
  }
  double step = 1.0 / out.residual;
  while (out.num_iters < strat.cfg.max_iters) {
    be.wrapped_axpbyz(1.0, xt, -step, gt, x, prob.space);  // x = P(x_tmp - step g_tmp)
    be.apply(prob.A, x, g);
    be.axpby(1.0, prob.q, 1.0, g);
    out.residual = be.residual(ResidualPolicy::kind, x, g, prob.space);
    if (out.residual <= strat.cfg.tol) {
      out.converged = true;
      return out;
    }
    const double num = be.diff_dot(x, xt);  // BB1 (:498-516)
    double den = be.diff_dot(x, xt, g, gt);
    if (den < 1e-14 && den > -1e-14) den += 1e-14;
    step = num / den;
    be.deep_copy(xt, x);
    be.deep_copy(gt, g);
    ++out.num_iters;
  }
  return out;
}

/// solve_lcp (:840-845): the LCP as a CQPP over x >= 0.
template <class LinearOp, class ResidualPolicy>
convex::SolveResult solve_lcp(const convex::LCPProblem<LinearOp>& prob, const convex::PGDStrategy<ResidualPolicy>& strat,
                              convex::PGDState& state) {
  const auto cq = convex::to_cqpp(prob);
  return solve_cqpp(cq, strat, state);
}

}  // namespace math
}  // namespace mundy_b200
#endif  // MUNDY_B200_CONVEX_HPP_
