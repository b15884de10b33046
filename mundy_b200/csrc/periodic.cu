// Periodic metrics, batched: item k of every array is one call of the reference function of
// mundy/geom/src/mundy_geom/periodicity.hpp on an orthorhombic cell:
//   metric.sep(p1, p2)                      PeriodicScaledMetric :814-818 (PeriodicMetricX/Y/XY/YZ :406-411 ...)
//   metric.wrap(p)                          :821-824 with impl::safe_unit_mod1 :140-150
//   unwrap_points_to_ref(point, m, ref)     :1333-1339  (= what spheres / spherocylinders do with their centre)
//   unwrap_points_to_ref(LineSegment, m, ref) :1368-1382 (fractional form: sr + min_image(s - sr), per end point)
// HBM-bound streaming (48-72 bytes in, 24-48 out per item); inputs and outputs move through shared-memory tiles as
// contiguous double2 runs like the distance kernels.
#include "tile_io.cuh"

namespace mb200 {
namespace {

__device__ inline V3 sh3(const double* s, int t) { return V3{s[3 * t], s[3 * t + 1], s[3 * t + 2]}; }
__device__ inline void put3(double* s, int t, const V3& v) { s[3 * t] = v.x, s[3 * t + 1] = v.y, s[3 * t + 2] = v.z; }

// fractional-coordinate unwrap of one point (the LineSegment / AABB / VSegment overloads, :1368-1382)
__device__ inline V3 unwrap_frac(const Periodic& P, const V3& p, const V3& ref) {
  V3 o = p;
  const double* L = P.L;
  const double* inv = P.inv;
  // to_fractional = scale_inv (.) p; free axes have scale 1 and no rounding, i.e. sr + (s - sr) in plain coordinates
  {
    const double s = inv[0] * p.x, sr = inv[0] * ref.x;
    const double d = P.on[0] ? frac_min_image(s - sr) : (s - sr);
    o.x = L[0] * (sr + d);
  }
  {
    const double s = inv[1] * p.y, sr = inv[1] * ref.y;
    const double d = P.on[1] ? frac_min_image(s - sr) : (s - sr);
    o.y = L[1] * (sr + d);
  }
  {
    const double s = inv[2] * p.z, sr = inv[2] * ref.z;
    const double d = P.on[2] ? frac_min_image(s - sr) : (s - sr);
    o.z = L[2] * (sr + d);
  }
  return o;
}

// OP 0: sep(a, b)   OP 1: wrap(a)   OP 2: unwrap point a to ref b   OP 3: unwrap segment (a, c) to ref b -> (o0, o1)
template <int OP>
__global__ void __launch_bounds__(kTile)
    periodic_kernel(int64_t n, Periodic P, const double* __restrict__ a, const double* __restrict__ b,
                    const double* __restrict__ c, double* __restrict__ o0, double* __restrict__ o1) {
  __shared__ __align__(16) double sa[3 * kTile];
  __shared__ __align__(16) double sb[(OP != 1) ? 3 * kTile : 2];
  __shared__ __align__(16) double sc[(OP == 3) ? 3 * kTile : 2];
  const int64_t ntiles = (n + kTile - 1) / kTile;
  const int t = threadIdx.x;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = tile * kTile;
    const int cnt = static_cast<int>(min(static_cast<int64_t>(kTile), n - base));
    load_tile<3>(sa, a, base, cnt);
    if (OP != 1) load_tile<3>(sb, b, base, cnt);
    if (OP == 3) load_tile<3>(sc, c, base, cnt);
    __syncthreads();
    V3 r0{0, 0, 0}, r1{0, 0, 0};
    if (t < cnt) {
      if (OP == 0) r0 = periodic_sep(P, sh3(sa, t), sh3(sb, t));
      if (OP == 1) r0 = periodic_wrap(P, sh3(sa, t));
      if (OP == 2) r0 = periodic_unwrap(P, sh3(sa, t), sh3(sb, t));
      if (OP == 3) {
        r0 = unwrap_frac(P, sh3(sa, t), sh3(sb, t));
        r1 = unwrap_frac(P, sh3(sc, t), sh3(sb, t));
      }
    }
    __syncthreads();
    if (t < cnt) {
      put3(sa, t, r0);
      if (OP == 3) put3(sc, t, r1);
    }
    __syncthreads();
    store_tile<3>(o0, sa, base, cnt);
    if (OP == 3) store_tile<3>(o1, sc, base, cnt);
    __syncthreads();
  }
}

template <int OP>
int run(mb200_stream stream, int64_t n, const double* cell_size_h, const double* a, const double* b, const double* c,
        double* o0, double* o1) {
  MB200_REQUIRE(n >= 0 && cell_size_h, "bad size or null cell size");
  if (n == 0) return MB200_OK;
  MB200_REQUIRE(a && o0 && (OP == 1 || b) && (OP != 3 || (c && o1)), "null pointer");
  MB200_TRY(require_device());
  const Periodic P = make_periodic(cell_size_h);
  periodic_kernel<OP><<<grid_for(n, kTile, kSMs * 8), kTile, 0, as_stream(stream)>>>(n, P, a, b, c, o0, o1);
  MB200_LAUNCHED();
  return MB200_OK;
}

}  // namespace
}  // namespace mb200

using namespace mb200;

extern "C" {

int mb200_periodic_sep(mb200_stream stream, int64_t n, const double* cell_size_h, const double* p1, const double* p2,
                       double* sep) {
  return run<0>(stream, n, cell_size_h, p1, p2, nullptr, sep, nullptr);
}
int mb200_periodic_wrap(mb200_stream stream, int64_t n, const double* cell_size_h, const double* p, double* out) {
  return run<1>(stream, n, cell_size_h, p, nullptr, nullptr, out, nullptr);
}
int mb200_periodic_unwrap_points(mb200_stream stream, int64_t n, const double* cell_size_h, const double* p,
                                 const double* ref, double* out) {
  return run<2>(stream, n, cell_size_h, p, ref, nullptr, out, nullptr);
}
int mb200_periodic_unwrap_segments(mb200_stream stream, int64_t n, const double* cell_size_h, const double* start,
                                   const double* end, const double* ref, double* start_out, double* end_out) {
  return run<3>(stream, n, cell_size_h, start, ref, end, start_out, end_out);
}

}  // extern "C"
