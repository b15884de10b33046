// E1. Driver BBPGD (scrap/lcp_spheres/KokkosLCP.cpp:675-876), device-resident.
//
// One iteration = three kernels, no host round trip (the reference blocks the host twice per iteration):
//   gamma_update : gamma = max(gamma_prev - alpha (phi0 + dt g), 0)                 contact pass, 32 B/contact
//   body_pass    : F = D gamma (ordered per-body sums over the incidence list, no atomics), U = M F
//                  -> 32-byte velocity records                                        body pass
//   rate_pass    : g = D^T U, residual max, the three BB dot products; the last CTA to finish reduces the per-CTA
//                  partials in a fixed order and advances the scalar state (alpha, residual, done)   contact pass
// Every kernel returns immediately once `done` is set, so the host enqueues iterations in batches and only reads
// the 48-byte state back between batches.
#include "internal.cuh"
#include "lcp.cuh"
#include "tile_io.cuh"

namespace mb200 {
namespace {

struct LcpState {
  double alpha;
  double res;
  double vmax;
  int32_t it;
  int32_t done;
  int32_t converged;
  uint32_t ticket;
};

struct LcpArgs {  // everything the iteration kernels need; passed by value
  int64_t n, npairs;
  const int2* pairs;
  const double* normal;
  const double* arm_i;
  const double* arm_j;
  const double* quat;
  const double* radius;
  const int32_t* kind;
  const int64_t* inc_ptr;
  const uint32_t* inc_list;
  const double* phi0;
  D4* vel;  // [n] (3-DOF) or [2n] (6-DOF: U record then W record per body)
  double mu, dt, tol;
  int32_t fixed;   // ignore tol
  int32_t it_cap;  // max iterations (or the fixed count)
  LcpState* state;
  double* partials;  // [4 * rate grid]
};

constexpr int kRateThreads = kTile;

__global__ void __launch_bounds__(256)
    gamma_update_kernel(int64_t npairs, const double* __restrict__ gamma_prev, const double* __restrict__ phi0,
                        const double* __restrict__ g, int first, double dt, const LcpState* __restrict__ state,
                        double* __restrict__ gamma) {
  if (state->done) return;
  const double alpha = state->alpha;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; c < npairs; c += stride) {
    const double gc = first ? 0.0 : g[c];  // KokkosLCP.cpp:756 uses signed_sep_dot (zeros) on the first iteration
    const double sep_new = phi0[c] + dt * gc;
    gamma[c] = fmax(gamma_prev[c] - alpha * sep_new, 0.0);
  }
}

// FINAL: also writes F,T,U,W in the caller's layout and reduces max |U| into state->vmax (single launch at the end).
template <bool SIX, bool FINAL>
__global__ void __launch_bounds__(128)
    body_pass_kernel(LcpArgs a, const double* __restrict__ gamma, double* __restrict__ F, double* __restrict__ T,
                     double* __restrict__ U, double* __restrict__ W, const double* __restrict__ U_ext,
                     const double* __restrict__ W_ext) {
  if (!FINAL && a.state->done) return;
  __shared__ double sh[33];
  const int64_t b = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  double speed = 0.0;
  if (b < a.n) {
    V3 f, t;
    body_wrench<SIX>(b, a.inc_ptr, a.inc_list, gamma, a.normal, a.arm_i, a.arm_j, f, t);
    const Mob m = mobility_of(a.radius[b], a.mu);
    const bool rod = SIX && a.quat && a.kind && a.kind[b] == MB200_SPHEROCYLINDER;
    const Q4 q = rod ? ldq(a.quat, b) : Q4{1.0, 0.0, 0.0, 0.0};
    V3 u, w;
    body_mobility(m, rod, q, f, t, u, w);
    if (!FINAL) {
      if (SIX) {
        st256(a.vel + 2 * b, D4{u.x, u.y, u.z, 0.0});
        st256(a.vel + 2 * b + 1, D4{w.x, w.y, w.z, 0.0});
      } else {
        st256(a.vel + b, D4{u.x, u.y, u.z, 0.0});
      }
    } else {
      speed = sqrt(u.x * u.x + u.y * u.y + u.z * u.z);  // KokkosLCP.cpp:861-872 (collision velocity only)
      if (F) st3(F, b, f);
      if (SIX && T) st3(T, b, t);
      if (U_ext) u = u + ld3(U_ext, b);  // U += U_col on top of the external velocity (NGPSpheresLCP.cpp:867-877)
      if (SIX && W_ext) w = w + ld3(W_ext, b);
      if (U) st3(U, b, u);
      if (SIX && W) st3(W, b, w);
    }
  }
  if (FINAL) {
    const double m = block_max(speed, sh, 0.0);
    if (threadIdx.x == 0) {
      // non-negative doubles order like their bit patterns
      atomicMax(reinterpret_cast<unsigned long long*>(&a.state->vmax), static_cast<unsigned long long>(__double_as_longlong(m)));
    }
  }
}

// MODE 0: initial pass (g_out = D^T M D gamma_0, residual only). MODE 1: iteration pass (residual + BB dots).
template <bool SIX, int MODE>
__global__ void __launch_bounds__(kRateThreads)
    rate_pass_kernel(LcpArgs a, int it, const double* __restrict__ gamma, const double* __restrict__ gamma_prev,
                     const double* __restrict__ g_prev, double* __restrict__ g_out) {
  if (a.state->done) return;
  __shared__ __align__(16) double s_n[3 * kTile];
  __shared__ __align__(16) double s_ai[SIX ? 3 * kTile : 2];
  __shared__ __align__(16) double s_aj[SIX ? 3 * kTile : 2];
  __shared__ double sh[33];
  __shared__ bool is_last;
  double res = -1.0, xx = 0.0, xg = 0.0, gg = 0.0;
  const int64_t ntiles = (a.npairs + kTile - 1) / kTile;
  const int t = threadIdx.x;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = tile * kTile;
    const int cnt = static_cast<int>(min(static_cast<int64_t>(kTile), a.npairs - base));
    load_tile<3>(s_n, a.normal, base, cnt);
    if (SIX) {
      load_tile<3>(s_ai, a.arm_i, base, cnt);
      load_tile<3>(s_aj, a.arm_j, base, cnt);
    }
    __syncthreads();
    if (t < cnt) {
      const int64_t c = base + t;
      const int2 p = __ldg(a.pairs + c);
      const V3 n{s_n[3 * t], s_n[3 * t + 1], s_n[3 * t + 2]};
      V3 ui, uj, wi{0, 0, 0}, wj{0, 0, 0}, ai{0, 0, 0}, aj{0, 0, 0};
      if (SIX) {
        const D4 r0 = ld256(a.vel + 2 * p.x), r1 = ld256(a.vel + 2 * p.x + 1);
        const D4 r2 = ld256(a.vel + 2 * p.y), r3 = ld256(a.vel + 2 * p.y + 1);
        ui = V3{r0.x, r0.y, r0.z}, wi = V3{r1.x, r1.y, r1.z};
        uj = V3{r2.x, r2.y, r2.z}, wj = V3{r3.x, r3.y, r3.z};
        ai = V3{s_ai[3 * t], s_ai[3 * t + 1], s_ai[3 * t + 2]};
        aj = V3{s_aj[3 * t], s_aj[3 * t + 1], s_aj[3 * t + 2]};
      } else {
        const D4 r0 = ld256(a.vel + p.x), r2 = ld256(a.vel + p.y);
        ui = V3{r0.x, r0.y, r0.z}, uj = V3{r2.x, r2.y, r2.z};
      }
      const double g = contact_rate<SIX>(n, ui, uj, wi, wj, ai, aj);
      g_out[c] = g;
      const double gam = gamma[c];
      const double sep_new = a.phi0[c] + a.dt * g;  // KokkosLCP.cpp:500-521
      const double v = (gam < 1e-12) ? fabs(fmin(sep_new, 0.0)) : fabs(sep_new);
      res = fmax(res, v);
      if (MODE == 1) {  // KokkosLCP.cpp:536-556
        const double xd = gam - gamma_prev[c];
        const double gd = a.dt * (g - g_prev[c]);
        xx += xd * xd;
        xg += xd * gd;
        gg += gd * gd;
      }
    }
    __syncthreads();
  }
  // per-CTA partials, then the last CTA reduces them in index order (deterministic)
  res = block_max(res, sh, -1.0);
  if (MODE == 1) {
    xx = block_sum(xx, sh);
    xg = block_sum(xg, sh);
    gg = block_sum(gg, sh);
  }
  if (t == 0) {
    double* p = a.partials + 4 * blockIdx.x;
    p[0] = res, p[1] = xx, p[2] = xg, p[3] = gg;
    __threadfence();
    const uint32_t ticket = atomicAdd(&a.state->ticket, 1u);
    is_last = (ticket == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  res = -1.0, xx = 0.0, xg = 0.0, gg = 0.0;
  const volatile double* vp = a.partials;
  for (int k = t; k < static_cast<int>(gridDim.x); k += kRateThreads) {
    res = fmax(res, vp[4 * k]);
    xx += vp[4 * k + 1], xg += vp[4 * k + 2], gg += vp[4 * k + 3];
  }
  res = block_max(res, sh, -1.0);
  if (MODE == 1) {
    xx = block_sum(xx, sh);
    xg = block_sum(xg, sh);
    gg = block_sum(gg, sh);
  }
  if (t == 0) {
    LcpState* s = a.state;
    s->ticket = 0;
    s->res = res;
    s->it = it;
    if (!a.fixed && res < a.tol) {  // strict <, KokkosLCP.cpp:744,788
      s->done = 1;
      s->converged = 1;
    } else {
      if (MODE == 0) {
        s->alpha = 1.0 / res;  // :750
      } else {
        double num, den;  // :831-848; `it` was already incremented at the loop top, so iteration 1 takes BB2
        if (it % 2 == 0) {
          num = xx, den = xg;
        } else {
          num = xg, den = gg;
        }
        if (fabs(den) < 1e-12) den += 1e-12;
        s->alpha = num / den;
      }
      if (it >= a.it_cap) s->done = 1;
    }
  }
}

// phi0 += dt * D^T (U_ext, W_ext)   scrap/parameter_interface/NGPSpheresLCP.cpp:751-756 (caller-layout velocities)
template <bool SIX>
__global__ void __launch_bounds__(256)
    ext_correction_kernel(int64_t npairs, const int2* __restrict__ pairs, const double* __restrict__ normal,
                          const double* __restrict__ arm_i, const double* __restrict__ arm_j,
                          const double* __restrict__ U_ext, const double* __restrict__ W_ext, double dt,
                          double* __restrict__ phi0) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= npairs) return;
  const int2 p = pairs[c];
  V3 wi{0, 0, 0}, wj{0, 0, 0}, ai{0, 0, 0}, aj{0, 0, 0};
  const bool six = SIX && W_ext;
  if (six) {
    wi = ld3(W_ext, p.x), wj = ld3(W_ext, p.y);
    ai = ld3(arm_i, c), aj = ld3(arm_j, c);
  }
  const V3 n = ld3(normal, c);
  const double g = six ? contact_rate<true>(n, ld3(U_ext, p.x), ld3(U_ext, p.y), wi, wj, ai, aj)
                       : contact_rate<false>(n, ld3(U_ext, p.x), ld3(U_ext, p.y), wi, wj, ai, aj);
  phi0[c] += dt * g;
}

struct PinnedState {
  LcpState* h = nullptr;
  ~PinnedState() {
    if (h) cudaFreeHost(h);
  }
};

}  // namespace

int resolve_collisions_impl(cudaStream_t st, const mb200_contact_op* op, const Incidence& inc,
                            const mb200_lcp_params* params, const double* U_ext, const double* W_ext, double* phi0,
                            double* gamma, double* F, double* T, double* U, double* W,
                            mb200_collision_result* result_h) {
  const int64_t n = op->n, C = op->npairs;
  const bool six = op->arm_i != nullptr;
  const bool fixed = params->fixed_iters >= 0;
  const int it_cap = fixed ? params->fixed_iters : params->max_iters;

  static thread_local PinnedState pinned;
  if (!pinned.h) MB200_CUDA(cudaMallocHost(reinterpret_cast<void**>(&pinned.h), sizeof(LcpState)));

  if (C == 0) {  // nothing to solve: zero force / velocity, residual identity of the max-reduction (-1)
    if (F) MB200_CUDA(cudaMemsetAsync(F, 0, 3 * n * sizeof(double), st));
    if (U) MB200_CUDA(cudaMemsetAsync(U, 0, 3 * n * sizeof(double), st));
    if (six && T) MB200_CUDA(cudaMemsetAsync(T, 0, 3 * n * sizeof(double), st));
    if (six && W) MB200_CUDA(cudaMemsetAsync(W, 0, 3 * n * sizeof(double), st));
    MB200_CUDA(cudaStreamSynchronize(st));
    if (result_h) *result_h = mb200_collision_result{-1.0, 0, 1, 0.0};
    return MB200_OK;
  }

  const int64_t ntiles = (C + kTile - 1) / kTile;
  const int rate_grid = static_cast<int>(ntiles < kSMs * 4 ? ntiles : kSMs * 4);
  const int gam_grid = grid_for(C, 256, kSMs * 8);
  const int body_grid = grid_for(n, 128);

  Scratch<double> gamma_b, g_a, g_b, partials;
  Scratch<D4> vel;
  Scratch<LcpState> state;
  MB200_TRY(gamma_b.alloc(C, st));
  MB200_TRY(g_a.alloc(C, st));
  MB200_TRY(g_b.alloc(C, st));
  MB200_TRY(partials.alloc(4 * static_cast<size_t>(rate_grid), st));
  MB200_TRY(vel.alloc(static_cast<size_t>(six ? 2 : 1) * n, st));
  MB200_TRY(state.alloc(1, st));
  MB200_CUDA(cudaMemsetAsync(state.p, 0, sizeof(LcpState), st));

  LcpArgs a;
  a.n = n, a.npairs = C;
  a.pairs = reinterpret_cast<const int2*>(op->pairs);
  a.normal = op->normal, a.arm_i = op->arm_i, a.arm_j = op->arm_j;
  a.quat = op->quat, a.radius = op->radius, a.kind = op->kind;
  a.inc_ptr = inc.ptr.p, a.inc_list = inc.list.p;
  a.phi0 = phi0, a.vel = vel.p;
  a.mu = op->viscosity, a.dt = op->dt, a.tol = params->tol;
  a.fixed = fixed ? 1 : 0, a.it_cap = it_cap;
  a.state = state.p, a.partials = partials.p;

  if (U_ext) {
    if (six)
      ext_correction_kernel<true><<<grid_for(C, 256), 256, 0, st>>>(C, a.pairs, a.normal, a.arm_i, a.arm_j, U_ext,
                                                                   W_ext, a.dt, phi0);
    else
      ext_correction_kernel<false><<<grid_for(C, 256), 256, 0, st>>>(C, a.pairs, a.normal, nullptr, nullptr, U_ext,
                                                                    nullptr, a.dt, phi0);
    MB200_LAUNCHED();
  }

  // initial residual: g_a = D^T M D gamma_0
  {
    StageScope ps(MB200_STAGE_LCP_BODY, st);
    if (six)
      body_pass_kernel<true, false><<<body_grid, 128, 0, st>>>(a, gamma, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);
    else
      body_pass_kernel<false, false><<<body_grid, 128, 0, st>>>(a, gamma, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);
    MB200_LAUNCHED();
  }
  {
    StageScope ps(MB200_STAGE_LCP_RATE, st);
    if (six)
      rate_pass_kernel<true, 0><<<rate_grid, kRateThreads, 0, st>>>(a, 0, gamma, nullptr, nullptr, g_a.p);
    else
      rate_pass_kernel<false, 0><<<rate_grid, kRateThreads, 0, st>>>(a, 0, gamma, nullptr, nullptr, g_a.p);
    MB200_LAUNCHED();
  }

  // iteration it: gamma_cur/g_cur are the B buffers when it is odd, the A buffers when even
  int it = 0;
  int batch = 4;
  LcpState hs;
  for (;;) {
    MB200_CUDA(cudaMemcpyAsync(pinned.h, state.p, sizeof(LcpState), cudaMemcpyDeviceToHost, st));
    MB200_CUDA(cudaStreamSynchronize(st));
    hs = *pinned.h;
    if (hs.done || it >= it_cap) break;
    const int stop = (it + batch < it_cap) ? it + batch : it_cap;
    while (it < stop) {
      ++it;
      const bool odd = it & 1;
      double* gam_cur = odd ? gamma_b.p : gamma;
      const double* gam_prev = odd ? gamma : gamma_b.p;
      double* g_cur = odd ? g_b.p : g_a.p;
      const double* g_prev = odd ? g_a.p : g_b.p;
      {
        StageScope ps(MB200_STAGE_LCP_GAMMA, st);
        gamma_update_kernel<<<gam_grid, 256, 0, st>>>(C, gam_prev, phi0, g_prev, it == 1 ? 1 : 0, a.dt, state.p, gam_cur);
        MB200_LAUNCHED();
      }
      {
        StageScope ps(MB200_STAGE_LCP_BODY, st);
        if (six)
          body_pass_kernel<true, false><<<body_grid, 128, 0, st>>>(a, gam_cur, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);
        else
          body_pass_kernel<false, false><<<body_grid, 128, 0, st>>>(a, gam_cur, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);
        MB200_LAUNCHED();
      }
      {
        StageScope ps(MB200_STAGE_LCP_RATE, st);
        if (six)
          rate_pass_kernel<true, 1><<<rate_grid, kRateThreads, 0, st>>>(a, it, gam_cur, gam_prev, g_prev, g_cur);
        else
          rate_pass_kernel<false, 1><<<rate_grid, kRateThreads, 0, st>>>(a, it, gam_cur, gam_prev, g_prev, g_cur);
        MB200_LAUNCHED();
      }
    }
    if (batch < 32) batch *= 2;
  }

  // final gamma lives in the B buffer when the last executed iteration is odd
  const int it_final = hs.it;
  if (it_final & 1) MB200_CUDA(cudaMemcpyAsync(gamma, gamma_b.p, C * sizeof(double), cudaMemcpyDeviceToDevice, st));
  {
    StageScope ps(MB200_STAGE_LCP_BODY, st);
    if (six)
      body_pass_kernel<true, true><<<body_grid, 128, 0, st>>>(a, gamma, F, T, U, W, nullptr, nullptr);
    else
      body_pass_kernel<false, true><<<body_grid, 128, 0, st>>>(a, gamma, F, nullptr, U, nullptr, nullptr, nullptr);
    MB200_LAUNCHED();
  }
  MB200_CUDA(cudaMemcpyAsync(pinned.h, state.p, sizeof(LcpState), cudaMemcpyDeviceToHost, st));
  MB200_CUDA(cudaStreamSynchronize(st));
  hs = *pinned.h;
  if (result_h) {
    result_h->max_abs_projected_sep = hs.res;
    result_h->ite_count = it_final;
    result_h->converged = (hs.res < params->tol) ? 1 : 0;
    result_h->max_displacement = hs.vmax * op->dt;
  }
  return MB200_OK;
}

}  // namespace mb200

using namespace mb200;

extern "C" int mb200_resolve_collisions(mb200_stream stream, const mb200_contact_op* op, const mb200_lcp_params* params,
                                        const double* U_ext, const double* W_ext, double* phi0, double* gamma,
                                        double* F, double* T, double* U, double* W,
                                        mb200_collision_result* result_h) {
  MB200_REQUIRE(op && params && op->n >= 0 && op->npairs >= 0, "bad arguments");
  MB200_REQUIRE(op->n < (int64_t(1) << 31) && op->npairs < (int64_t(1) << 30), "problem too large for int32 ids");
  MB200_REQUIRE((op->arm_i == nullptr) == (op->arm_j == nullptr), "arm_i/arm_j must both be set or NULL");
  MB200_REQUIRE(op->radius || op->n == 0, "radius is NULL");
  if (op->npairs > 0) MB200_REQUIRE(op->pairs && op->normal && phi0 && gamma, "null contact arrays");
  MB200_TRY(require_device());
  cudaStream_t st = as_stream(stream);
  Incidence inc;
  MB200_TRY(build_incidence(st, op->n, op->npairs, op->pairs, &inc));
  return resolve_collisions_impl(st, op, inc, params, U_ext, W_ext, phi0, gamma, F, T, U, W, result_h);
}
