// C. geom::distance, batched: item k of every array is one call of the reference function
// (mundy/geom/src/mundy_geom/distance/{SphereSphere,LineSegmentLineSegment,PointLineSegment,LineSegmentSphere}.hpp).
// Inputs arrive through shared-memory tiles (contiguous double2 runs), 3-vector outputs leave the same way.
#include "tile_io.cuh"

namespace mb200 {
namespace {

__device__ inline V3 sh3(const double* s, int t) { return V3{s[3 * t], s[3 * t + 1], s[3 * t + 2]}; }
__device__ inline void put3(double* s, int t, const V3& v) { s[3 * t] = v.x, s[3 * t + 1] = v.y, s[3 * t + 2] = v.z; }

// KIND 0: sphere-sphere (a=c1, b=c2, ra=r1, rb=r2)            outputs: dist, v0=sep
// KIND 1: segment-segment (a=l0, b=l1, c=m0, d=m1)             outputs: dist, s0=s, s1=t, v0=cp1, v1=cp2, v2=sep
// KIND 2: point-segment (a=x, b=p1, c=p2)                      outputs: dist, s1=t, v0=cp, v2=sep
// KIND 3: segment-sphere (a=p1, b=p2, c=centre, ra=r)          outputs: dist, s1=t, v0=cp, v2=sep
template <int KIND>
__global__ void __launch_bounds__(kTile)
    distance_kernel(int64_t n, const double* __restrict__ a, const double* __restrict__ b, const double* __restrict__ c,
                    const double* __restrict__ d, const double* __restrict__ ra, const double* __restrict__ rb,
                    double* __restrict__ dist, double* __restrict__ s0, double* __restrict__ s1,
                    double* __restrict__ v0, double* __restrict__ v1, double* __restrict__ v2) {
  __shared__ __align__(16) double sa[3 * kTile];
  __shared__ __align__(16) double sb[3 * kTile];
  __shared__ __align__(16) double sc[(KIND >= 1) ? 3 * kTile : 2];
  __shared__ __align__(16) double sd[(KIND == 1) ? 3 * kTile : 2];
  const int64_t ntiles = (n + kTile - 1) / kTile;
  const int t = threadIdx.x;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = tile * kTile;
    const int cnt = static_cast<int>(min(static_cast<int64_t>(kTile), n - base));
    load_tile<3>(sa, a, base, cnt);
    load_tile<3>(sb, b, base, cnt);
    if (KIND >= 1) load_tile<3>(sc, c, base, cnt);
    if (KIND == 1) load_tile<3>(sd, d, base, cnt);
    __syncthreads();
    V3 o0{0, 0, 0}, o1{0, 0, 0}, o2{0, 0, 0};
    if (t < cnt) {
      double dd, ss = 0.0, tt = 0.0;
      if (KIND == 0) {  // SphereSphere.hpp:65-76
        V3 sep;
        const double cc = dist_pp(sh3(sa, t), sh3(sb, t), sep);
        dd = cc - ra[base + t] - rb[base + t];
        const double f = dd / cc;
        o0 = V3{sep.x * f, sep.y * f, sep.z * f};
      } else if (KIND == 1) {
        dd = dist_segseg(sh3(sa, t), sh3(sb, t), sh3(sc, t), sh3(sd, t), o0, o1, ss, tt, o2);
      } else if (KIND == 2) {
        dd = dist_pseg(sh3(sa, t), sh3(sb, t), sh3(sc, t), o0, tt, o2);
      } else {  // LineSegmentSphere.hpp:86-99
        const double lc = dist_pseg(sh3(sc, t), sh3(sa, t), sh3(sb, t), o0, tt, o2);
        dd = lc - ra[base + t];
        const double f = dd / lc;
        o2 = V3{o2.x * f, o2.y * f, o2.z * f};
      }
      if (dist) dist[base + t] = dd;
      if (s0) s0[base + t] = ss;
      if (s1) s1[base + t] = tt;
    }
    __syncthreads();  // inputs consumed; reuse sa/sb/sc as output staging
    if (t < cnt) {
      put3(sa, t, o0);
      if (KIND == 1) put3(sb, t, o1);
      if (KIND >= 1) put3(sc, t, o2);
    }
    __syncthreads();
    if (v0) store_tile<3>(v0, sa, base, cnt);
    if (KIND == 1 && v1) store_tile<3>(v1, sb, base, cnt);
    if (KIND >= 1 && v2) store_tile<3>(v2, sc, base, cnt);
    __syncthreads();
  }
}

inline int dgrid(int64_t n) { return grid_for(n, kTile, kSMs * 4); }

}  // namespace
}  // namespace mb200

using namespace mb200;

extern "C" {

int mb200_distance_sphere_sphere(mb200_stream stream, int64_t n, const double* c1, const double* r1, const double* c2,
                                 const double* r2, double* dist, double* sep) {
  MB200_REQUIRE(n >= 0, "n < 0");
  if (n == 0) return MB200_OK;
  MB200_REQUIRE(c1 && r1 && c2 && r2, "null input");
  MB200_TRY(require_device());
  distance_kernel<0><<<dgrid(n), kTile, 0, as_stream(stream)>>>(n, c1, c2, nullptr, nullptr, r1, r2, dist, nullptr,
                                                               nullptr, sep, nullptr, nullptr);
  MB200_LAUNCHED();
  return MB200_OK;
}

int mb200_distance_segment_segment(mb200_stream stream, int64_t n, const double* l0, const double* l1,
                                   const double* m0, const double* m1, double* dist, double* s, double* t,
                                   double* cp1, double* cp2, double* sep) {
  MB200_REQUIRE(n >= 0, "n < 0");
  if (n == 0) return MB200_OK;
  MB200_REQUIRE(l0 && l1 && m0 && m1, "null input");
  MB200_TRY(require_device());
  distance_kernel<1><<<dgrid(n), kTile, 0, as_stream(stream)>>>(n, l0, l1, m0, m1, nullptr, nullptr, dist, s, t, cp1,
                                                               cp2, sep);
  MB200_LAUNCHED();
  return MB200_OK;
}

int mb200_distance_point_segment(mb200_stream stream, int64_t n, const double* x, const double* p1, const double* p2,
                                 double* dist, double* t, double* cp, double* sep) {
  MB200_REQUIRE(n >= 0, "n < 0");
  if (n == 0) return MB200_OK;
  MB200_REQUIRE(x && p1 && p2, "null input");
  MB200_TRY(require_device());
  distance_kernel<2><<<dgrid(n), kTile, 0, as_stream(stream)>>>(n, x, p1, p2, nullptr, nullptr, nullptr, dist, nullptr,
                                                               t, cp, nullptr, sep);
  MB200_LAUNCHED();
  return MB200_OK;
}

int mb200_distance_segment_sphere(mb200_stream stream, int64_t n, const double* p1, const double* p2, const double* c,
                                  const double* r, double* dist, double* t, double* cp, double* sep) {
  MB200_REQUIRE(n >= 0, "n < 0");
  if (n == 0) return MB200_OK;
  MB200_REQUIRE(p1 && p2 && c && r, "null input");
  MB200_TRY(require_device());
  distance_kernel<3><<<dgrid(n), kTile, 0, as_stream(stream)>>>(n, p1, p2, c, nullptr, r, nullptr, dist, nullptr, t, cp,
                                                               nullptr, sep);
  MB200_LAUNCHED();
  return MB200_OK;
}

}  // extern "C"
