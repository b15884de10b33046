// C5/K6. Narrow phase: per pair signed separation, unit normal and lever arms.
// HBM-bound: algorithmic bytes 40 C + 32 N (spheres), 88 C + 72 N (rods). Thread per pair; the two bodies of a pair
// are gathered as 32-byte records (256-bit loads: one sector each); normals/arms leave through shared memory as
// contiguous double2 runs.
#include "internal.cuh"
#include "shapes.cuh"
#include "tile_io.cuh"

namespace mb200 {

// Caller layout -> body records. perm == NULL: identity; else record k describes caller body perm[k].
__global__ void __launch_bounds__(256)
    pack_bodies_kernel(int64_t n, const int32_t* __restrict__ perm, const double* __restrict__ center,
                       const double* __restrict__ quat, const double* __restrict__ radius,
                       const double* __restrict__ length, const int32_t* __restrict__ kind, D4* __restrict__ b0,
                       D4* __restrict__ b1) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t i = perm ? perm[k] : k;
  const V3 c = ld3(center, i);
  st256(b0 + k, D4{c.x, c.y, c.z, radius[i]});
  if (b1) {
    const int kd = kind ? kind[i] : 0;
    V3 h{0.0, 0.0, 0.0};
    if (kd == MB200_SPHEROCYLINDER) {
      const V3 axis = qrot(ldq(quat, i), V3{0.0, 0.0, 1.0});
      h = (0.5 * length[i]) * axis;
    }
    st256(b1 + k, D4{h.x, h.y, h.z, static_cast<double>(kd)});
  }
}

int pack_bodies(cudaStream_t st, const mb200_bodies* b, const int32_t* perm, D4* b0, D4* b1) {
  if (b->n == 0) return MB200_OK;
  pack_bodies_kernel<<<grid_for(b->n, 256), 256, 0, st>>>(b->n, perm, b->center, b->quat, b->radius, b->length,
                                                         b->kind, b0, b1);
  MB200_LAUNCHED();
  return MB200_OK;
}

namespace {

template <bool RODS, bool ARMS>
__global__ void __launch_bounds__(kTile)
    narrow_phase_kernel(int64_t npairs, const int2* __restrict__ pairs, const D4* __restrict__ b0,
                        const D4* __restrict__ b1, double* __restrict__ phi, double* __restrict__ normal,
                        double* __restrict__ arm_i, double* __restrict__ arm_j, Periodic P) {
  __shared__ __align__(16) double s_n[3 * kTile];
  __shared__ __align__(16) double s_ai[ARMS ? 3 * kTile : 2];
  __shared__ __align__(16) double s_aj[ARMS ? 3 * kTile : 2];
  const int64_t ntiles = (npairs + kTile - 1) / kTile;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = tile * kTile;
    const int cnt = static_cast<int>(min(static_cast<int64_t>(kTile), npairs - base));
    const int t = threadIdx.x;
    if (t < cnt) {
      const int2 p = __ldg(pairs + base + t);
      const D4 b0i = ldg256(b0 + p.x);
      D4 b0j = ldg256(b0 + p.y);
      if (P.any) b0j = unwrap_to(P, b0i, b0j);  // periodic box: j unwrapped to i's centre
      D4 b1i{0, 0, 0, 0}, b1j{0, 0, 0, 0};
      if (RODS) {
        b1i = ldg256(b1 + p.x);
        b1j = ldg256(b1 + p.y);
      }
      const PairGeom g = pair_geometry<RODS>(b0i, b1i, b0j, b1j);
      phi[base + t] = g.sep;
      s_n[3 * t] = g.n.x, s_n[3 * t + 1] = g.n.y, s_n[3 * t + 2] = g.n.z;
      if (ARMS) {
        s_ai[3 * t] = g.arm_i.x, s_ai[3 * t + 1] = g.arm_i.y, s_ai[3 * t + 2] = g.arm_i.z;
        s_aj[3 * t] = g.arm_j.x, s_aj[3 * t + 1] = g.arm_j.y, s_aj[3 * t + 2] = g.arm_j.z;
      }
    }
    __syncthreads();
    store_tile<3>(normal, s_n, base, cnt);
    if (ARMS) {
      store_tile<3>(arm_i, s_ai, base, cnt);
      store_tile<3>(arm_j, s_aj, base, cnt);
    }
    __syncthreads();
  }
}

// Same arithmetic, written as the solver's contact arrays (lcp_solve.cu): R0 = {n, phi0} and, for rods, the axial arm
// parameters {sigma_i, sigma_j}.
template <bool RODS>
__global__ void __launch_bounds__(256)
    narrow_phase_rec_kernel(int64_t npairs, const int2* __restrict__ pairs, const uint32_t* __restrict__ perm,
                            const D4* __restrict__ b0, const D4* __restrict__ b1, D4* __restrict__ rec,
                            double2* __restrict__ sig, Periodic P) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= npairs) return;
  const int2 p = __ldg(pairs + c);
  const D4 b0i = ldg256(b0 + p.x), b0j = ldg256(b0 + p.y);
  D4 b1i{0, 0, 0, 0}, b1j{0, 0, 0, 0};
  if (RODS) {
    b1i = ldg256(b1 + p.x);
    b1j = ldg256(b1 + p.y);
  }
  // The pair is evaluated with the body of lower CALLER index first (the reference's (i<j) order, as the oracle does),
  // whatever the Morton order of the two: results do not depend on the internal ordering or on the rank.
  PairGeom g;
  if (!perm || perm[p.x] < perm[p.y]) {
    g = pair_geometry<RODS>(b0i, b1i, P.any ? unwrap_to(P, b0i, b0j) : b0j, b1j);
  } else {
    const PairGeom r = pair_geometry<RODS>(b0j, b1j, P.any ? unwrap_to(P, b0j, b0i) : b0i, b1i);
    g.sep = r.sep;
    g.n = V3{-r.n.x, -r.n.y, -r.n.z};
    g.sig_i = r.sig_j;
    g.sig_j = r.sig_i;
  }
  st256(rec + c, D4{g.n.x, g.n.y, g.n.z, g.sep});
  if (RODS) sig[c] = make_double2(g.sig_i, g.sig_j);
}

}  // namespace

int narrow_phase_to_records(cudaStream_t st, int64_t npairs, const int32_t* pairs, const uint32_t* perm, const D4* b0,
                            const D4* b1, D4* rec, double2* sig, const Periodic& P) {
  if (npairs == 0) return MB200_OK;
  const int grid = grid_for(npairs, 256);
  const int2* p2 = reinterpret_cast<const int2*>(pairs);
  if (b1 && sig)
    narrow_phase_rec_kernel<true><<<grid, 256, 0, st>>>(npairs, p2, perm, b0, b1, rec, sig, P);
  else if (!b1)
    narrow_phase_rec_kernel<false><<<grid, 256, 0, st>>>(npairs, p2, perm, b0, nullptr, rec, nullptr, P);
  else {
    set_error("narrow_phase_to_records: rods need the axial arm array");
    return MB200_ERR_INVALID_ARGUMENT;
  }
  MB200_LAUNCHED();
  return MB200_OK;
}

int narrow_phase_records(cudaStream_t st, int64_t npairs, const int32_t* pairs, const D4* b0, const D4* b1,
                         double* phi, double* normal, double* arm_i, double* arm_j, const Periodic& P) {
  if (npairs == 0) return MB200_OK;
  const int grid = grid_for(npairs, kTile, kSMs * 8);
  const int2* p2 = reinterpret_cast<const int2*>(pairs);
  const bool arms = arm_i && arm_j;
  if (b1) {
    if (arms)
      narrow_phase_kernel<true, true><<<grid, kTile, 0, st>>>(npairs, p2, b0, b1, phi, normal, arm_i, arm_j, P);
    else
      narrow_phase_kernel<true, false><<<grid, kTile, 0, st>>>(npairs, p2, b0, b1, phi, normal, nullptr, nullptr, P);
  } else {
    if (arms)
      narrow_phase_kernel<false, true><<<grid, kTile, 0, st>>>(npairs, p2, b0, nullptr, phi, normal, arm_i, arm_j, P);
    else
      narrow_phase_kernel<false, false><<<grid, kTile, 0, st>>>(npairs, p2, b0, nullptr, phi, normal, nullptr, nullptr, P);
  }
  MB200_LAUNCHED();
  return MB200_OK;
}

}  // namespace mb200

using namespace mb200;

extern "C" int mb200_narrow_phase_periodic(mb200_stream stream, const mb200_bodies* bodies, int64_t npairs,
                                           const int32_t* pairs, const double* cell_size_h, double* phi,
                                           double* normal, double* arm_i, double* arm_j) {
  MB200_REQUIRE(bodies && bodies->n >= 0 && npairs >= 0, "bad sizes");
  if (npairs == 0) return MB200_OK;
  MB200_REQUIRE(bodies->center && bodies->radius && pairs && phi && normal, "null pointer");
  MB200_REQUIRE((arm_i == nullptr) == (arm_j == nullptr), "arm_i and arm_j must both be given or both be NULL");
  const bool rods = bodies->kind != nullptr;
  if (rods) MB200_REQUIRE(bodies->quat && bodies->length, "kind given without quat/length");
  MB200_TRY(require_device());
  cudaStream_t st = as_stream(stream);
  Scratch<D4> b0, b1;
  MB200_TRY(b0.alloc(bodies->n, st));
  if (rods) MB200_TRY(b1.alloc(bodies->n, st));
  MB200_TRY(pack_bodies(st, bodies, nullptr, b0.p, rods ? b1.p : nullptr));
  return narrow_phase_records(st, npairs, pairs, b0.p, rods ? b1.p : nullptr, phi, normal, arm_i, arm_j,
                              make_periodic(cell_size_h));
}

extern "C" int mb200_narrow_phase(mb200_stream stream, const mb200_bodies* bodies, int64_t npairs, const int32_t* pairs,
                                  double* phi, double* normal, double* arm_i, double* arm_j) {
  return mb200_narrow_phase_periodic(stream, bodies, npairs, pairs, nullptr, phi, normal, arm_i, arm_j);
}
