// E2. math::convex projected-gradient solver (mundy/math/src/mundy_math/convex.hpp) on device vectors.
// The vector primitives are the KokkosBackend static interface (:141-285); solve_cqpp (:789-797) with
// PGDStrategy::initialize/iterate (:610-681) is composed from them exactly as the reference composes it, including its
// two host-visible reductions per iteration. The matrix-free collision solve that stays on the device is lcp_solve.cu.
#include <functional>

#include "internal.cuh"

namespace mb200 {
namespace {

struct Space {  // convex.hpp:42-117
  int kind;
  double lo, hi;
  __host__ __device__ double project(double x) const {
    switch (kind) {
      case MB200_SPACE_UNCONSTRAINED: return x;
      case MB200_SPACE_LOWER: return fmax(x, lo);
      case MB200_SPACE_UPPER: return fmin(x, hi);
      default: return fmin(fmax(x, lo), hi);
    }
  }
};

// MODE: 0 both non-zero, 1 alpha ~ 0, 2 beta ~ 0, 3 both ~ 0 (the zero-tolerance short cuts of :201-248)
template <int MODE, bool WRAP>
__global__ void __launch_bounds__(256) axpbyz_kernel(int64_t n, double alpha, const double* __restrict__ x, double beta,
                                                     const double* y, double* z, Space sp) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    double v;
    if (MODE == 0) v = alpha * x[i] + beta * y[i];
    else if (MODE == 1) v = beta * y[i];
    else if (MODE == 2) v = alpha * x[i];
    else v = 0.0;
    z[i] = WRAP ? sp.project(v) : v;
  }
}

template <bool WRAP>
int launch_axpbyz(cudaStream_t st, int64_t n, double alpha, const double* x, double beta, const double* y, double* z,
                  Space sp) {
  if (n == 0) return MB200_OK;
  const bool az = fabs(alpha) < kZeroTol, bz = fabs(beta) < kZeroTol;
  const int grid = grid_for(n, 256, kSMs * 8);
  if (!az && !bz) axpbyz_kernel<0, WRAP><<<grid, 256, 0, st>>>(n, alpha, x, beta, y, z, sp);
  else if (az && !bz) axpbyz_kernel<1, WRAP><<<grid, 256, 0, st>>>(n, alpha, x, beta, y, z, sp);
  else if (!az && bz) axpbyz_kernel<2, WRAP><<<grid, 256, 0, st>>>(n, alpha, x, beta, y, z, sp);
  else axpbyz_kernel<3, WRAP><<<grid, 256, 0, st>>>(n, alpha, x, beta, y, z, sp);
  MB200_LAUNCHED();
  return MB200_OK;
}

// OP 0: sum (x - xo)^2 ; 1: sum (x - xo)(g - go) ; 2: max |x - P(x - 1e-6 g)| ; 3: max (x<1e-15 ? max(0,g) : |g|)
template <int OP>
__device__ inline double reduce_term(int64_t i, const double* x, const double* xo, const double* g, const double* go,
                                     const Space& sp) {
  if (OP == 0) {
    const double d = x[i] - xo[i];
    return d * d;
  } else if (OP == 1) {
    return (x[i] - xo[i]) * (g[i] - go[i]);
  } else if (OP == 2) {
    const double xp = sp.project(x[i] - 1e-6 * g[i]);  // :468-496
    return fabs(x[i] - xp);
  } else {
    return (x[i] < kZeroTol) ? fmax(0.0, g[i]) : fabs(g[i]);  // :434-466
  }
}

constexpr int kRedGrid = kSMs * 2;

template <int OP>
__global__ void __launch_bounds__(256) reduce_kernel(int64_t n, const double* __restrict__ x, const double* __restrict__ xo,
                                                     const double* __restrict__ g, const double* __restrict__ go,
                                                     Space sp, double* __restrict__ partials,
                                                     unsigned int* __restrict__ ticket, double* __restrict__ out) {
  __shared__ double sh[33];
  __shared__ bool is_last;
  constexpr bool IS_MAX = OP >= 2;
  const double ident = IS_MAX ? -1.7976931348623157e308 : 0.0;  // Kokkos::Max identity = lowest()
  double acc = ident;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double v = reduce_term<OP>(i, x, xo, g, go, sp);
    acc = IS_MAX ? fmax(acc, v) : acc + v;
  }
  acc = IS_MAX ? block_max(acc, sh, ident) : block_sum(acc, sh);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = acc;
    __threadfence();
    is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  acc = ident;
  const volatile double* vp = partials;
  for (int k = threadIdx.x; k < static_cast<int>(gridDim.x); k += blockDim.x) acc = IS_MAX ? fmax(acc, vp[k]) : acc + vp[k];
  acc = IS_MAX ? block_max(acc, sh, ident) : block_sum(acc, sh);
  if (threadIdx.x == 0) {
    *out = acc;
    *ticket = 0;
  }
}

struct PinnedDouble {
  double* h = nullptr;
  ~PinnedDouble() {
    if (h) cudaFreeHost(h);
  }
};

template <int OP>
int reduce_to_host(cudaStream_t st, int64_t n, const double* x, const double* xo, const double* g, const double* go,
                   Space sp, double* out_h) {
  static thread_local PinnedDouble pin;
  if (!pin.h) MB200_CUDA(cudaMallocHost(reinterpret_cast<void**>(&pin.h), sizeof(double)));
  Scratch<double> partials;
  Scratch<unsigned int> ticket;
  const int grid = grid_for(n, 256, kRedGrid);
  MB200_TRY(partials.alloc(grid + 1, st));
  MB200_TRY(ticket.alloc(1, st));
  MB200_CUDA(cudaMemsetAsync(ticket.p, 0, sizeof(unsigned int), st));
  reduce_kernel<OP><<<grid, 256, 0, st>>>(n, x, xo, g, go, sp, partials.p, ticket.p, partials.p + grid);
  MB200_LAUNCHED();
  MB200_CUDA(cudaMemcpyAsync(pin.h, partials.p + grid, sizeof(double), cudaMemcpyDeviceToHost, st));
  MB200_CUDA(cudaStreamSynchronize(st));
  *out_h = *pin.h;
  return MB200_OK;
}

// y = A x, dense row-major: one warp per row (the KokkosBlas::gemv("N", 1, A, x, 0, y) path of :166-184).
__global__ void __launch_bounds__(256) gemv_kernel(int64_t n, const double* __restrict__ A, const double* __restrict__ x,
                                                   double* __restrict__ y) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  if (row >= n) return;
  double acc = 0.0;
  for (int64_t j = lane; j < n; j += 32) acc += A[row * n + j] * x[j];
  acc = warp_sum(acc);
  if (lane == 0) y[row] = acc;
}

using ApplyFn = std::function<int(const double*, double*)>;

// solve_cqpp (:789-797) = PGDStrategy::initialize (:615-636) + iterate (:639-666)
int pgd_solve(cudaStream_t st, int64_t n, const ApplyFn& apply, const double* q, double* x, const mb200_pgd_config* cfg,
              mb200_solve_result* result_h) {
  const Space sp{cfg->space_kind, cfg->lo, cfg->hi};
  Scratch<double> g, xt, gt;
  MB200_TRY(g.alloc(n, st));
  MB200_TRY(xt.alloc(n, st));
  MB200_TRY(gt.alloc(n, st));
  auto residual = [&](const double* xx, const double* gg, double* out) {
    if (cfg->residual_kind == MB200_RESID_PROJECTED_GRAD) return reduce_to_host<3>(st, n, xx, nullptr, gg, nullptr, sp, out);
    MB200_TRY(reduce_to_host<2>(st, n, xx, nullptr, gg, nullptr, sp, out));
    *out = *out / 1e-6;
    return (int)MB200_OK;
  };
  MB200_CUDA(cudaMemcpyAsync(xt.p, x, n * sizeof(double), cudaMemcpyDeviceToDevice, st));
  MB200_TRY(apply(xt.p, gt.p));
  MB200_TRY((launch_axpbyz<false>(st, n, 1.0, q, 1.0, gt.p, gt.p, sp)));
  double res = 0.0;
  MB200_TRY(residual(xt.p, gt.p, &res));
  double step = 1.0 / res;
  uint32_t iter = 0;
  bool converged = res <= cfg->tol;
  while (!(converged || iter >= cfg->max_iters)) {
    MB200_TRY((launch_axpbyz<true>(st, n, 1.0, xt.p, -step, gt.p, x, sp)));
    MB200_TRY(apply(x, g.p));
    MB200_TRY((launch_axpbyz<false>(st, n, 1.0, q, 1.0, g.p, g.p, sp)));
    MB200_TRY(residual(x, g.p, &res));
    if (res <= cfg->tol) {
      converged = true;
      break;
    }
    double num = 0.0, den = 0.0;  // BBStepStrategy :498-516
    MB200_TRY(reduce_to_host<0>(st, n, x, xt.p, nullptr, nullptr, sp, &num));
    MB200_TRY(reduce_to_host<1>(st, n, x, xt.p, g.p, gt.p, sp, &den));
    const double eps = kZeroTol * 10;
    den += eps * (fabs(den) < eps);
    step = num / den;
    MB200_CUDA(cudaMemcpyAsync(xt.p, x, n * sizeof(double), cudaMemcpyDeviceToDevice, st));
    MB200_CUDA(cudaMemcpyAsync(gt.p, g.p, n * sizeof(double), cudaMemcpyDeviceToDevice, st));
    ++iter;
  }
  MB200_CUDA(cudaStreamSynchronize(st));
  if (result_h) {
    result_h->num_iters = iter;
    result_h->residual = res;
    result_h->converged = converged ? 1 : 0;
  }
  return MB200_OK;
}

int check_cfg(const mb200_pgd_config* cfg) {
  MB200_REQUIRE(cfg, "null config");
  MB200_REQUIRE(cfg->space_kind >= 0 && cfg->space_kind <= 3, "bad space_kind");
  MB200_REQUIRE(cfg->residual_kind == 0 || cfg->residual_kind == 1, "bad residual_kind");
  return MB200_OK;
}

}  // namespace
}  // namespace mb200

using namespace mb200;

extern "C" {

int mb200_vec_axpby(mb200_stream stream, int64_t n, double alpha, const double* x, double beta, double* y) {
  MB200_REQUIRE(n >= 0 && (n == 0 || (x && y)), "bad arguments");
  MB200_TRY(require_device());
  return launch_axpbyz<false>(as_stream(stream), n, alpha, x, beta, y, y, Space{0, 0.0, 0.0});
}

int mb200_vec_wrapped_axpbyz(mb200_stream stream, int64_t n, double alpha, const double* x, double beta,
                             const double* y, double* z, int space_kind, double lo, double hi) {
  MB200_REQUIRE(n >= 0 && (n == 0 || (x && y && z)), "bad arguments");
  MB200_REQUIRE(space_kind >= 0 && space_kind <= 3, "bad space_kind");
  MB200_TRY(require_device());
  return launch_axpbyz<true>(as_stream(stream), n, alpha, x, beta, y, z, Space{space_kind, lo, hi});
}

int mb200_vec_diff_dot(mb200_stream stream, int64_t n, const double* x, const double* x_old, double* out_h) {
  MB200_REQUIRE(n >= 0 && out_h && (n == 0 || (x && x_old)), "bad arguments");
  MB200_TRY(require_device());
  return reduce_to_host<0>(as_stream(stream), n, x, x_old, nullptr, nullptr, Space{0, 0.0, 0.0}, out_h);
}

int mb200_vec_diff_dot2(mb200_stream stream, int64_t n, const double* x, const double* x_old, const double* g,
                        const double* g_old, double* out_h) {
  MB200_REQUIRE(n >= 0 && out_h && (n == 0 || (x && x_old && g && g_old)), "bad arguments");
  MB200_TRY(require_device());
  return reduce_to_host<1>(as_stream(stream), n, x, x_old, g, g_old, Space{0, 0.0, 0.0}, out_h);
}

int mb200_vec_residual(mb200_stream stream, int64_t n, const double* x, const double* g, int residual_kind,
                       int space_kind, double lo, double hi, double* out_h) {
  MB200_REQUIRE(n >= 0 && out_h && (n == 0 || (x && g)), "bad arguments");
  MB200_REQUIRE(space_kind >= 0 && space_kind <= 3, "bad space_kind");
  MB200_TRY(require_device());
  const Space sp{space_kind, lo, hi};
  if (residual_kind == MB200_RESID_PROJECTED_GRAD)
    return reduce_to_host<3>(as_stream(stream), n, x, nullptr, g, nullptr, sp, out_h);
  MB200_TRY(reduce_to_host<2>(as_stream(stream), n, x, nullptr, g, nullptr, sp, out_h));
  *out_h = *out_h / 1e-6;
  return MB200_OK;
}

int mb200_dense_gemv(mb200_stream stream, int64_t n, const double* A, const double* x, double* y) {
  MB200_REQUIRE(n >= 0 && (n == 0 || (A && x && y)), "bad arguments");
  if (n == 0) return MB200_OK;
  MB200_TRY(require_device());
  gemv_kernel<<<grid_for(n * 32, 256), 256, 0, as_stream(stream)>>>(n, A, x, y);
  MB200_LAUNCHED();
  return MB200_OK;
}

int mb200_solve_cqpp_dense(mb200_stream stream, int64_t n, const double* A, const double* q, double* x,
                           const mb200_pgd_config* cfg, mb200_solve_result* result_h) {
  MB200_TRY(check_cfg(cfg));
  MB200_REQUIRE(n > 0 && A && q && x, "bad arguments");
  MB200_TRY(require_device());
  ApplyFn apply = [&](const double* in, double* out) { return mb200_dense_gemv(stream, n, A, in, out); };
  return pgd_solve(as_stream(stream), n, apply, q, x, cfg, result_h);
}

int mb200_solve_lcp_contact(mb200_stream stream, const mb200_contact_op* op, const double* phi0, double* gamma,
                            const mb200_pgd_config* cfg, mb200_solve_result* result_h) {
  MB200_TRY(check_cfg(cfg));
  MB200_REQUIRE(op && op->npairs > 0 && phi0 && gamma, "bad arguments");
  MB200_TRY(require_device());
  ApplyFn apply = [&](const double* in, double* out) { return mb200_apply_A(stream, op, in, out); };
  return pgd_solve(as_stream(stream), op->npairs, apply, phi0, gamma, cfg, result_h);
}

}  // extern "C"
