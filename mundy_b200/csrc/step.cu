// Neighbour-link handle API and the whole collision timestep
// (loop body of scrap/parameter_interface/NGPSpheresLCP.cpp:1135-1292):
//   generate links if needed -> narrow phase -> phi0 += dt D^T U_ext -> BBPGD -> U = U_ext + U_col -> Euler.
// The step works on bodies gathered into Morton order (the order of the link list) and scatters results back to the
// caller's order at the end.
#include <cmath>
#include <vector>

#include "links.cuh"
#include "lcp.cuh"
#include "sort_scan.cuh"

namespace mb200 {
namespace {

// Morton-order copies of the per-body solver inputs; optional external wrench -> external velocity (U_ext = M F_ext).
__global__ void __launch_bounds__(256)
    gather_params_kernel(int64_t n, const uint32_t* __restrict__ perm, const double* __restrict__ radius,
                         const double* __restrict__ quat, const int32_t* __restrict__ kind, double mu,
                         const double* __restrict__ F_ext, const double* __restrict__ T_ext,
                         double* __restrict__ radius_s, double* __restrict__ quat_s, int32_t* __restrict__ kind_s,
                         double* __restrict__ Uext_s, double* __restrict__ Wext_s) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t i = perm[k];
  const double r = radius[i];
  radius_s[k] = r;
  const int kd = kind ? kind[i] : 0;
  if (kind_s) kind_s[k] = kd;
  Q4 q{1.0, 0.0, 0.0, 0.0};
  if (quat_s) {
    q = ldq(quat, i);
    double2* o = reinterpret_cast<double2*>(quat_s + 4 * k);
    o[0] = make_double2(q.w, q.x);
    o[1] = make_double2(q.y, q.z);
  }
  if (Uext_s) {
    const Mob m = mobility_of(r, mu);
    const V3 f = ld3(F_ext, i);
    const V3 t = (T_ext && Wext_s) ? ld3(T_ext, i) : V3{0.0, 0.0, 0.0};
    V3 u, w;
    body_mobility(m, quat_s && kd == MB200_SPHEROCYLINDER, q, f, t, u, w);
    st3(Uext_s, k, u);
    if (Wext_s) st3(Wext_s, k, w);
  }
}

// Back to the caller's order: total velocity, Euler update of centre / quaternion, optional outputs.
__global__ void __launch_bounds__(256)
    scatter_euler_kernel(int64_t n, const uint32_t* __restrict__ perm, double dt, const double* __restrict__ F_s,
                         const double* __restrict__ T_s, const double* __restrict__ U_s, const double* __restrict__ W_s,
                         const double* __restrict__ Uext_s, const double* __restrict__ Wext_s,
                         const int32_t* __restrict__ kind, double* __restrict__ center, double* __restrict__ quat,
                         double* __restrict__ F_col, double* __restrict__ T_col, double* __restrict__ U_out,
                         double* __restrict__ W_out) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t i = perm[k];
  V3 u = ld3(U_s, k);
  if (Uext_s) u = u + ld3(Uext_s, k);
  const V3 c = ld3(center, i);
  st3(center, i, V3{c.x + dt * u.x, c.y + dt * u.y, c.z + dt * u.z});
  if (U_out) st3(U_out, i, u);
  if (F_col) st3(F_col, i, ld3(F_s, k));
  if (W_s) {
    V3 w = ld3(W_s, k);
    if (Wext_s) w = w + ld3(Wext_s, k);
    if (W_out) st3(W_out, i, w);
    if (T_col) st3(T_col, i, ld3(T_s, k));
    if (quat && (!kind || kind[i] == MB200_SPHEROCYLINDER)) {
      const Q4 q = rotate_quaternion(ldq(quat, i), w, dt);
      double2* o = reinterpret_cast<double2*>(quat + 4 * i);
      o[0] = make_double2(q.w, q.x);
      o[1] = make_double2(q.y, q.z);
    }
  }
}

int check_bodies(const mb200_bodies* b) {
  MB200_REQUIRE(b && b->n >= 0, "bad bodies");
  MB200_REQUIRE(b->n < (int64_t(1) << 31), "too many bodies for int32 ids");
  if (b->n > 0) MB200_REQUIRE(b->center && b->radius, "null centre/radius");
  if (b->kind) MB200_REQUIRE(b->quat && b->length, "kind given without quat/length");
  if (b->quat) MB200_REQUIRE((reinterpret_cast<uintptr_t>(b->quat) & 15) == 0, "quat must be 16-byte aligned");
  return MB200_OK;
}

}  // namespace
}  // namespace mb200

using namespace mb200;

extern "C" {

int mb200_neighbor_links_create(mb200_neighbor_links** out) {
  MB200_REQUIRE(out, "null out");
  *out = new (std::nothrow) mb200_neighbor_links();
  MB200_REQUIRE(*out, "allocation failed");
  return MB200_OK;
}

int mb200_neighbor_links_destroy(mb200_neighbor_links* h) {
  delete h;
  return MB200_OK;
}

int mb200_neighbor_links_set_search_buffer(mb200_neighbor_links* h, double buffer) {
  MB200_REQUIRE(h && buffer >= 0.0, "bad handle or negative buffer");
  if (buffer != h->buffer) h->built = false;
  h->buffer = buffer;
  return MB200_OK;
}

int mb200_neighbor_links_set_search_filter(mb200_neighbor_links* h, int filter) {
  MB200_REQUIRE(h && (filter == MB200_FILTER_NONE || filter == MB200_FILTER_SHAPE_DISTANCE), "bad handle or filter");
  if (filter != h->filter) h->built = false;
  h->filter = filter;
  return MB200_OK;
}

int mb200_neighbor_links_set_periodic(mb200_neighbor_links* h, const double* cell_size_h) {
  MB200_REQUIRE(h, "null handle");
  const Periodic P = make_periodic(cell_size_h);
  for (int d = 0; d < 3; ++d) {
    MB200_REQUIRE(!cell_size_h || std::isfinite(cell_size_h[d]), "non-finite cell size");
    if (P.on[d] != h->periodic.on[d] || P.L[d] != h->periodic.L[d]) h->built = false;
  }
  h->periodic = P;
  return MB200_OK;
}

int mb200_neighbor_links_generate(mb200_neighbor_links* h, mb200_stream stream, const mb200_bodies* bodies, int force,
                                  int* rebuilt_h) {
  MB200_REQUIRE(h, "null handle");
  MB200_TRY(check_bodies(bodies));
  MB200_TRY(require_device());
  return links_generate(as_stream(stream), h, bodies, force, rebuilt_h);
}

int64_t mb200_neighbor_links_count(const mb200_neighbor_links* h) { return h ? h->npairs : 0; }

int mb200_neighbor_links_get_pairs(mb200_neighbor_links* h, mb200_stream stream, int32_t* pairs_out, int64_t capacity) {
  MB200_REQUIRE(h && h->built, "links not generated");
  if (h->npairs > capacity) {
    set_error("pairs_out too small");
    return MB200_ERR_CAPACITY;
  }
  if (h->npairs == 0) return MB200_OK;
  MB200_REQUIRE(pairs_out, "null pairs_out");
  cudaStream_t st = as_stream(stream);
  return canonical_pairs(st, h, pairs_out);
}

int mb200_search_pairs_periodic(mb200_stream stream, int64_t n, const double* center, const double* search_radius,
                                const double* cell_size_h, int32_t* pairs_out, int64_t capacity, int64_t* count_h) {
  MB200_REQUIRE(n >= 0 && n < (int64_t(1) << 31) && count_h, "bad arguments");
  *count_h = 0;
  if (n == 0) return MB200_OK;
  MB200_REQUIRE(center && search_radius, "null pointer");
  MB200_TRY(require_device());
  cudaStream_t st = as_stream(stream);
  mb200_neighbor_links tmp;
  tmp.periodic = make_periodic(cell_size_h);
  mb200_bodies b{n, const_cast<double*>(center), nullptr, search_radius, nullptr, nullptr};
  MB200_TRY(build_links(st, &tmp, &b, search_radius, nullptr));
  *count_h = tmp.npairs;
  if (tmp.npairs > capacity) {
    set_error("pairs_out too small");
    return MB200_ERR_CAPACITY;
  }
  if (tmp.npairs > 0) {
    MB200_REQUIRE(pairs_out, "null pairs_out");
    MB200_TRY(canonical_pairs(st, &tmp, pairs_out));
  }
  MB200_CUDA(cudaStreamSynchronize(st));  // tmp's buffers are freed when it goes out of scope
  return MB200_OK;
}

int mb200_search_pairs(mb200_stream stream, int64_t n, const double* center, const double* search_radius,
                       int32_t* pairs_out, int64_t capacity, int64_t* count_h) {
  return mb200_search_pairs_periodic(stream, n, center, search_radius, nullptr, pairs_out, capacity, count_h);
}

}  // extern "C"

namespace mb200 {

// The step after the link list is known: records, narrow phase, BBPGD, Euler. `dist` != NULL on the multi-GPU path.
int step_core(mb200_neighbor_links* h, cudaStream_t st, const mb200_bodies* bodies, const mb200_step_params* params,
              const double* F_ext, const double* T_ext, double* F_col, double* T_col, double* U_out, double* W_out,
              int rebuilt, long long launches0, const DistSolve* dist, mb200_step_stats* stats_h) {
  const int64_t n = bodies->n;
  const int64_t C = h->npairs;
  const bool rods = bodies->kind != nullptr;
  mb200_collision_result res{-1.0, 0, 1, 0.0};
  if (n > 0) {
    // bodies -> Morton-order records and solver parameters
    MB200_TRY(h->b0.reserve(n));
    if (rods) MB200_TRY(h->b1.reserve(n));
    MB200_TRY(h->radius_s.reserve(n));
    if (rods) {
      MB200_TRY(h->quat_s.reserve(4 * n));
      MB200_TRY(h->kind_s.reserve(n));
    }
    double* Uext_s = nullptr;
    double* Wext_s = nullptr;
    if (F_ext) {
      MB200_TRY(h->vec_a.reserve(3 * n));
      Uext_s = h->vec_a.p;
      if (rods) {
        MB200_TRY(h->vec_b.reserve(3 * n));
        Wext_s = h->vec_b.p;
      }
    }
    MB200_TRY(h->rec.reserve(static_cast<size_t>(C) + 1));
    if (rods) MB200_TRY(h->sig.reserve(static_cast<size_t>(C) + 1));
    if (h->forces_pending) {  // host entry point: the wrench arrives on the side stream (first read just below)
      MB200_CUDA(cudaStreamWaitEvent(st, h->forces_ev, 0));
      h->forces_pending = false;
    }
    {
      StageScope ps(MB200_STAGE_PACK, st);
      MB200_TRY(pack_bodies(st, bodies, reinterpret_cast<const int32_t*>(h->perm.p), h->b0.p, rods ? h->b1.p : nullptr));
      gather_params_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, h->perm.p, bodies->radius, bodies->quat, bodies->kind,
                                                            params->viscosity, F_ext, T_ext, h->radius_s.p,
                                                            rods ? h->quat_s.p : nullptr, rods ? h->kind_s.p : nullptr,
                                                            Uext_s, Wext_s);
      MB200_LAUNCHED();
    }
    {  // narrow phase on the links
      StageScope ps(MB200_STAGE_NARROW, st);
      MB200_TRY(narrow_phase_to_records(st, C, h->pairs.p, h->perm.p, h->b0.p, rods ? h->b1.p : nullptr, h->rec.p,
                                        rods ? h->sig.p : nullptr, h->periodic));
    }
    // LCP
    MB200_TRY(h->vec_c.reserve(3 * n));
    MB200_TRY(h->vec_d.reserve(3 * n));
    if (rods) {
      MB200_TRY(h->vec_e.reserve(3 * n));
      MB200_TRY(h->vec_f.reserve(3 * n));
    }
    RecordSystem sys{n, C, rods ? ARM_AXIAL : ARM_NONE, reinterpret_cast<const int2*>(h->pairs.p), h->rec.p,
                     rods ? h->sig.p : nullptr, rods ? h->b1.p : nullptr, rods ? h->quat_s.p : nullptr, h->radius_s.p,
                     rods ? h->kind_s.p : nullptr, params->viscosity, params->dt};
    mb200_lcp_params lp{params->tol, params->max_iters, params->fixed_iters};
    double* F_s = h->vec_c.p;
    double* U_s = h->vec_d.p;
    double* T_s = rods ? h->vec_e.p : nullptr;
    double* W_s = rods ? h->vec_f.p : nullptr;
    MB200_TRY(resolve_records(st, sys, h->inc, &lp, Uext_s, Wext_s, nullptr, nullptr, nullptr, F_s, T_s, U_s, W_s, &res, dist));
    if (params->throw_on_max_iters && params->fixed_iters < 0 && res.ite_count == params->max_iters && !res.converged) {
      set_error("Collision resolution did not converge!");  // scrap/lcp_spheres/NGPSpheresLCP.cpp:745-747
      if (stats_h) *stats_h = mb200_step_stats{res, C, rebuilt, static_cast<int32_t>(g_launches.load() - launches0)};
      return MB200_ERR_NOT_CONVERGED;
    }
    {
      StageScope ps(MB200_STAGE_EULER, st);
      scatter_euler_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, h->perm.p, params->dt, F_s, T_s, U_s, W_s, Uext_s,
                                                            Wext_s, bodies->kind, bodies->center, bodies->quat, F_col,
                                                            T_col, U_out, W_out);
      MB200_LAUNCHED();
    }
  }
  MB200_CUDA(cudaStreamSynchronize(st));
  if (stats_h) *stats_h = mb200_step_stats{res, C, rebuilt, static_cast<int32_t>(g_launches.load() - launches0)};
  return MB200_OK;
}


}  // namespace mb200

extern "C" int mb200_collision_step(mb200_neighbor_links* h, mb200_stream stream, const mb200_bodies* bodies,
                                    const mb200_step_params* params, const double* F_ext, const double* T_ext,
                                    double* F_col, double* T_col, double* U_out, double* W_out,
                                    mb200_step_stats* stats_h) {
  MB200_REQUIRE(h && params, "null handle/params");
  MB200_TRY(check_bodies(bodies));
  MB200_TRY(require_device());
  cudaStream_t st = as_stream(stream);
  const long long launches0 = g_launches.load();
  int rebuilt = 0;
  MB200_TRY(links_generate(st, h, bodies, params->force_rebuild, &rebuilt));
  return step_core(h, st, bodies, params, F_ext, T_ext, F_col, T_col, U_out, W_out, rebuilt, launches0, nullptr, stats_h);
}

// Host-buffer entry point: H2D of this step's inputs, the device step, D2H of the advanced state and outputs.
extern "C" int mb200_collision_step_host(mb200_neighbor_links* h, mb200_stream stream, const mb200_bodies* bh,
                                         const mb200_step_params* params, const double* F_ext_h, const double* T_ext_h,
                                         double* F_col_h, double* T_col_h, double* U_out_h, double* W_out_h,
                                         mb200_step_stats* stats_h) {
  MB200_REQUIRE(h && params && bh && bh->n >= 0, "bad arguments");
  if (bh->n > 0) MB200_REQUIRE(bh->center && bh->radius, "null centre/radius");
  if (bh->kind) MB200_REQUIRE(bh->quat && bh->length, "kind given without quat/length");
  MB200_TRY(require_device());
  cudaStream_t st = as_stream(stream);
  const int64_t n = bh->n;
  const size_t d = sizeof(double);
  const bool rods = bh->kind != nullptr;
  auto up = [&](DevBuf<double>& buf, const double* src, size_t count) -> int {
    MB200_TRY(buf.reserve(count + 4));
    MB200_CUDA(cudaMemcpyAsync(buf.p, src, count * d, cudaMemcpyHostToDevice, st));
    return MB200_OK;
  };
  MB200_TRY(up(h->hm_center, bh->center, 3 * n));
  MB200_TRY(up(h->hm_radius, bh->radius, n));
  if (rods) {
    MB200_TRY(up(h->hm_quat, bh->quat, 4 * n));
    MB200_TRY(up(h->hm_length, bh->length, n));
    MB200_TRY(h->hm_kind.reserve(n + 4));
    MB200_CUDA(cudaMemcpyAsync(h->hm_kind.p, bh->kind, n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  }
  if (n > 0 && (F_ext_h || (T_ext_h && rods))) {
    if (!h->copy_st) {
      MB200_CUDA(cudaStreamCreateWithFlags(&h->copy_st, cudaStreamNonBlocking));
      MB200_CUDA(cudaEventCreateWithFlags(&h->bodies_ev, cudaEventDisableTiming));
      MB200_CUDA(cudaEventCreateWithFlags(&h->forces_ev, cudaEventDisableTiming));
    }
    // behind the body arrays on the link (the search needs those first), concurrent with the trigger and the rebuild
    MB200_CUDA(cudaEventRecord(h->bodies_ev, st));
    MB200_CUDA(cudaStreamWaitEvent(h->copy_st, h->bodies_ev, 0));
    if (F_ext_h) {
      MB200_TRY(h->hm_fext.reserve(3 * n + 4));
      MB200_CUDA(cudaMemcpyAsync(h->hm_fext.p, F_ext_h, 3 * n * d, cudaMemcpyHostToDevice, h->copy_st));
    }
    if (T_ext_h && rods) {
      MB200_TRY(h->hm_text.reserve(3 * n + 4));
      MB200_CUDA(cudaMemcpyAsync(h->hm_text.p, T_ext_h, 3 * n * d, cudaMemcpyHostToDevice, h->copy_st));
    }
    MB200_CUDA(cudaEventRecord(h->forces_ev, h->copy_st));
    h->forces_pending = true;
  }
  if (F_col_h) MB200_TRY(h->hm_fcol.reserve(3 * n + 4));
  if (T_col_h && rods) MB200_TRY(h->hm_tcol.reserve(3 * n + 4));
  if (U_out_h) MB200_TRY(h->hm_u.reserve(3 * n + 4));
  if (W_out_h && rods) MB200_TRY(h->hm_w.reserve(3 * n + 4));
  mb200_bodies bd{n, h->hm_center.p, rods ? h->hm_quat.p : nullptr, h->hm_radius.p, rods ? h->hm_length.p : nullptr,
                  rods ? h->hm_kind.p : nullptr};
  const int status = mb200_collision_step(h, stream, &bd, params, F_ext_h ? h->hm_fext.p : nullptr,
                                          (T_ext_h && rods) ? h->hm_text.p : nullptr, F_col_h ? h->hm_fcol.p : nullptr,
                                          (T_col_h && rods) ? h->hm_tcol.p : nullptr, U_out_h ? h->hm_u.p : nullptr,
                                          (W_out_h && rods) ? h->hm_w.p : nullptr, stats_h);
  if (h->forces_pending) {  // the step returned before it consumed the wrench (n == 0 or an error): drain the side stream
    cudaStreamSynchronize(h->copy_st);
    h->forces_pending = false;
  }
  if (status != MB200_OK) return status;
  MB200_CUDA(cudaMemcpyAsync(bh->center, h->hm_center.p, 3 * n * d, cudaMemcpyDeviceToHost, st));
  if (rods) MB200_CUDA(cudaMemcpyAsync(bh->quat, h->hm_quat.p, 4 * n * d, cudaMemcpyDeviceToHost, st));
  if (F_col_h) MB200_CUDA(cudaMemcpyAsync(F_col_h, h->hm_fcol.p, 3 * n * d, cudaMemcpyDeviceToHost, st));
  if (T_col_h && rods) MB200_CUDA(cudaMemcpyAsync(T_col_h, h->hm_tcol.p, 3 * n * d, cudaMemcpyDeviceToHost, st));
  if (U_out_h) MB200_CUDA(cudaMemcpyAsync(U_out_h, h->hm_u.p, 3 * n * d, cudaMemcpyDeviceToHost, st));
  if (W_out_h && rods) MB200_CUDA(cudaMemcpyAsync(W_out_h, h->hm_w.p, 3 * n * d, cudaMemcpyDeviceToHost, st));
  MB200_CUDA(cudaStreamSynchronize(st));
  return MB200_OK;
}
