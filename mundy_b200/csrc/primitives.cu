// the remaining primitives of mundy_geom, batched (item k of every array is one call of the
// reference function).
//   compute_aabb(Point / LineSegment / Ellipsoid)   mundy/geom/src/mundy_geom/compute_aabb.hpp:47-104
//   distance(Point, Sphere, sep)                    mundy/geom/src/mundy_geom/distance/PointSphere.hpp:68-76
//   distance(Point, Line, cp, arch)                 mundy/geom/src/mundy_geom/distance/PointLine.hpp:92-103
//   distance(Line, Line, cp1, cp2, a1, a2, sep)     mundy/geom/src/mundy_geom/distance/LineLine.hpp:131-157
//   distance(Line, Sphere, cp, arch)                mundy/geom/src/mundy_geom/distance/LineSphere.hpp:93-102
//   distance(Point, Ellipsoid) / (Ellipsoid, Ellipsoid) / (Circle3D, Circle3D): minimize.cuh
// The closed-form ones are HBM-bound (48-168 bytes per item) and move their 3-vectors through shared-memory tiles as
// contiguous double2 runs like aabb.cu / distance.cu; the minimiser-based ones are fp64-compute-bound and only share
// the staging.
#include "minimize.cuh"
#include "tile_io.cuh"

namespace mb200 {
namespace {

struct PrimArgs {
  const double* in3[4];  // 3 doubles per item
  const double* in4[2];  // quaternions (w, x, y, z)
  const double* in1[2];  // scalars
  double* out1[3];
  double* out3[3];
  double* out6;  // AABB
};

// Op: static constexpr int N3, N4, N1 (inputs), O1, O3 (outputs), bool O6;
//     static __device__ void run(const V3* a, const Q4* q, const double* s, double* o1, V3* o3, double* o6)
template <class Op>
__device__ __forceinline__ void prim_body(int64_t n, const PrimArgs& A) {
  constexpr int kIn = Op::N3 > 0 ? Op::N3 : 1;
  constexpr int kOut = Op::O6 ? 2 : (Op::O3 > 0 ? Op::O3 : 1);
  __shared__ __align__(16) double s_in[kIn][3 * kTile];
  __shared__ __align__(16) double s_out[kOut * 3 * kTile];
  const int64_t ntiles = (n + kTile - 1) / kTile;
  const int t = threadIdx.x;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = tile * kTile;
    const int cnt = static_cast<int>(min(static_cast<int64_t>(kTile), n - base));
#pragma unroll
    for (int k = 0; k < Op::N3; ++k) load_tile<3>(s_in[k], A.in3[k], base, cnt);
    __syncthreads();
    if (t < cnt) {
      V3 a[kIn];
      Q4 q[Op::N4 > 0 ? Op::N4 : 1];
      double s[Op::N1 > 0 ? Op::N1 : 1];
      double o1[Op::O1 > 0 ? Op::O1 : 1];
      V3 o3[Op::O3 > 0 ? Op::O3 : 1];
      double o6[6];
#pragma unroll
      for (int k = 0; k < Op::N3; ++k) a[k] = V3{s_in[k][3 * t], s_in[k][3 * t + 1], s_in[k][3 * t + 2]};
#pragma unroll
      for (int k = 0; k < Op::N4; ++k) q[k] = ldq(A.in4[k], base + t);
#pragma unroll
      for (int k = 0; k < Op::N1; ++k) s[k] = __ldg(A.in1[k] + base + t);
      Op::run(a, q, s, o1, o3, o6);
#pragma unroll
      for (int k = 0; k < Op::O1; ++k)
        if (A.out1[k]) A.out1[k][base + t] = o1[k];
#pragma unroll
      for (int k = 0; k < Op::O3; ++k) {
        double* so = s_out + k * 3 * kTile;
        so[3 * t] = o3[k].x, so[3 * t + 1] = o3[k].y, so[3 * t + 2] = o3[k].z;
      }
      if (Op::O6) {
#pragma unroll
        for (int k = 0; k < 6; ++k) s_out[6 * t + k] = o6[k];
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < Op::O3; ++k)
      if (A.out3[k]) store_tile<3>(A.out3[k], s_out + k * 3 * kTile, base, cnt);
    if (Op::O6) store_tile<6>(A.out6, s_out, base, cnt);
    __syncthreads();
  }
}

// Two entry points over the same body: the closed-form kernels take ptxas' own register count (29-62: five to eight
// CTAs per SM; asking for a minimum made them slower), the Ellipsoid-Ellipsoid minimiser is compiled for two CTAs per SM
// (128 registers instead of 254: 16 warps per SM hide its fp64 latency chains better, +13 % pairs/s).
template <class Op>
__global__ void __launch_bounds__(kTile) prim_kernel(int64_t n, PrimArgs A) {
  prim_body<Op>(n, A);
}
template <class Op>
__global__ void __launch_bounds__(kTile, 2) prim_kernel_2cta(int64_t n, PrimArgs A) {
  prim_body<Op>(n, A);
}

#define MB200_PRIM_SHAPE(n3, n4, n1, o1, o3, o6) \
  static constexpr int N3 = n3, N4 = n4, N1 = n1, O1 = o1, O3 = o3; \
  static constexpr bool O6 = o6, TWO_CTAS = false

struct AabbPoint {  // compute_aabb.hpp:47-54
  MB200_PRIM_SHAPE(1, 0, 0, 0, 0, true);
  static __device__ void run(const V3* a, const Q4*, const double*, double*, V3*, double* o) {
    o[0] = o[3] = a[0].x, o[1] = o[4] = a[0].y, o[2] = o[5] = a[0].z;
  }
};
struct AabbLineSegment {  // compute_aabb.hpp:57-69
  MB200_PRIM_SHAPE(2, 0, 0, 0, 0, true);
  static __device__ void run(const V3* a, const Q4*, const double*, double*, V3*, double* o) {
    o[0] = fmin(a[0].x, a[1].x), o[1] = fmin(a[0].y, a[1].y), o[2] = fmin(a[0].z, a[1].z);
    o[3] = fmax(a[0].x, a[1].x), o[4] = fmax(a[0].y, a[1].y), o[5] = fmax(a[0].z, a[1].z);
  }
};
struct AabbEllipsoid {  // compute_aabb.hpp:83-103  in3 = {centre, radii}, in4 = {orientation}
  MB200_PRIM_SHAPE(2, 1, 0, 0, 0, true);
  static __device__ void run(const V3* a, const Q4* q, const double*, double*, V3*, double* o) {
    const V3 rr = qrot(q[0], a[1]);
    const V3 lo = a[0] - rr, hi = a[0] + rr;
    o[0] = fmin(lo.x, hi.x), o[1] = fmin(lo.y, hi.y), o[2] = fmin(lo.z, hi.z);
    o[3] = fmax(lo.x, hi.x), o[4] = fmax(lo.y, hi.y), o[5] = fmax(lo.z, hi.z);
  }
};
struct PointSphere {  // PointSphere.hpp:68-76  in3 = {point, centre}, in1 = {radius}; out: dist, sep
  MB200_PRIM_SHAPE(2, 0, 1, 1, 1, false);
  static __device__ void run(const V3* a, const Q4*, const double* s, double* o1, V3* o3, double*) {
    V3 sp;
    const double cd = dist_pp(a[0], a[1], sp);
    const double sd = cd - s[0];
    const double f = sd / cd;
    o1[0] = sd;
    o3[0] = V3{sp.x * f, sp.y * f, sp.z * f};
  }
};
struct PointLine {  // PointLine.hpp:92-103  in3 = {point, line centre, line direction}; out: dist, arch, cp
  MB200_PRIM_SHAPE(3, 0, 0, 2, 1, false);
  static __device__ void run(const V3* a, const Q4*, const double*, double* o1, V3* o3, double*) {
    const V3 l2p = a[0] - a[1];
    const double arch = dot(l2p, a[2]);
    o3[0] = a[1] + arch * a[2];
    V3 sp;
    o1[0] = dist_pp(arch * a[2], l2p, sp);
    o1[1] = arch;
  }
};
struct LineSphere {  // LineSphere.hpp:93-102  in3 = {line centre, line direction, sphere centre}, in1 = {radius}
  MB200_PRIM_SHAPE(3, 0, 1, 2, 1, false);
  static __device__ void run(const V3* a, const Q4*, const double* s, double* o1, V3* o3, double*) {
    const V3 l2p = a[2] - a[0];
    const double arch = dot(l2p, a[1]);
    o3[0] = a[0] + arch * a[1];
    V3 sp;
    o1[0] = dist_pp(arch * a[1], l2p, sp) - s[0];
    o1[1] = arch;
  }
};
struct LineLine {  // LineLine.hpp:131-157  in3 = {c1, d1, c2, d2}; out1 = dist, a1, a2; out3 = cp1, cp2, sep
  MB200_PRIM_SHAPE(4, 0, 0, 3, 3, false);
  static __device__ void run(const V3* in, const Q4*, const double*, double* o1, V3* o3, double*) {
    const V3 cc = in[0] - in[2];
    const double a = dot(in[1], in[1]), b = dot(in[1], in[3]), c = dot(in[3], in[3]);
    const double d = dot(in[1], cc), e = dot(in[3], cc);
    const double D = a * c - b * b;
    o1[0] = o1[1] = o1[2] = 0.0;
    o3[0] = o3[1] = o3[2] = V3{0.0, 0.0, 0.0};
    if (!(D < kZeroTol)) {  // parallel lines: the reference returns 0 and leaves the other outputs untouched
      const double inv_D = 1.0 / D;
      o1[1] = (b * e - c * d) * inv_D;
      o1[2] = (a * e - b * d) * inv_D;
      o3[0] = in[0] + o1[1] * in[1];
      o3[1] = in[2] + o1[2] * in[3];
      o1[0] = dist_pp(o3[0], o3[1], o3[2]);
    }
  }
};
struct PointEllipsoid {  // in3 = {point, centre, radii}, in4 = {orientation}; out: dist, cp, normal
  MB200_PRIM_SHAPE(3, 1, 0, 1, 2, false);
  static __device__ void run(const V3* a, const Q4* q, const double*, double* o1, V3* o3, double*) {
    const prim::Ellipsoid e{a[1], q[0], a[2]};
    o1[0] = prim::dist_point_ellipsoid(a[0], e, o3[0], o3[1]);
  }
};
struct EllipsoidEllipsoid {  // in3 = {c1, radii1, c2, radii2}, in4 = {q1, q2}; out: dist, cp1, cp2, normal of the first
  static constexpr int N3 = 4, N4 = 2, N1 = 0, O1 = 1, O3 = 3;
  static constexpr bool O6 = false, TWO_CTAS = true;
  static __device__ void run(const V3* a, const Q4* q, const double*, double* o1, V3* o3, double*) {
    const prim::Ellipsoid e1{a[0], q[0], a[1]}, e2{a[2], q[1], a[3]};
    o1[0] = prim::dist_ellipsoid_ellipsoid(e1, e2, o3[0], o3[1], o3[2]);
  }
};
struct CircleCircle {  // in3 = {c1, c2}, in4 = {q1, q2}, in1 = {r1, r2}; out: dist, cp1, cp2, unit vector cp1 -> cp2
  MB200_PRIM_SHAPE(2, 2, 2, 1, 3, false);
  static __device__ void run(const V3* a, const Q4* q, const double* s, double* o1, V3* o3, double*) {
    const prim::Circle k1{a[0], q[0], s[0]}, k2{a[1], q[1], s[1]};
    o1[0] = prim::dist_circle_circle(k1, k2, o3[0], o3[1], o3[2]);
  }
};

template <class Op>
int launch(mb200_stream stream, int64_t n, const PrimArgs& A) {
  MB200_REQUIRE(n >= 0, "n < 0");
  if (n == 0) return MB200_OK;
  for (int k = 0; k < Op::N3; ++k) MB200_REQUIRE(A.in3[k], "null input");
  for (int k = 0; k < Op::N4; ++k) MB200_REQUIRE(A.in4[k], "null input");
  for (int k = 0; k < Op::N1; ++k) MB200_REQUIRE(A.in1[k], "null input");
  if (Op::O6) MB200_REQUIRE(A.out6, "null output");
  MB200_TRY(require_device());
  // every tile is load -> barrier -> arithmetic -> barrier -> store, so the SM overlaps tiles only across CTAs: fill it
  static const int per_sm = [] {
    int b = 0;
    if constexpr (Op::TWO_CTAS)
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, prim_kernel_2cta<Op>, kTile, 0);
    else
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, prim_kernel<Op>, kTile, 0);
    return b > 0 ? b : 1;
  }();
  if constexpr (Op::TWO_CTAS)
    prim_kernel_2cta<Op><<<grid_for(n, kTile, kSMs * per_sm), kTile, 0, as_stream(stream)>>>(n, A);
  else
    prim_kernel<Op><<<grid_for(n, kTile, kSMs * per_sm), kTile, 0, as_stream(stream)>>>(n, A);
  MB200_LAUNCHED();
  return MB200_OK;
}

}  // namespace
}  // namespace mb200

using namespace mb200;

extern "C" {

int mb200_compute_aabb_point(mb200_stream stream, int64_t n, const double* point, double* aabb) {
  return launch<AabbPoint>(stream, n, PrimArgs{{point}, {}, {}, {}, {}, aabb});
}

int mb200_compute_aabb_line_segment(mb200_stream stream, int64_t n, const double* start, const double* end,
                                    double* aabb) {
  return launch<AabbLineSegment>(stream, n, PrimArgs{{start, end}, {}, {}, {}, {}, aabb});
}

int mb200_compute_aabb_ellipsoid(mb200_stream stream, int64_t n, const double* center, const double* quat,
                                 const double* radii, double* aabb) {
  return launch<AabbEllipsoid>(stream, n, PrimArgs{{center, radii}, {quat}, {}, {}, {}, aabb});
}

int mb200_distance_point_sphere(mb200_stream stream, int64_t n, const double* x, const double* c, const double* r,
                                double* dist, double* sep) {
  return launch<PointSphere>(stream, n, PrimArgs{{x, c}, {}, {r}, {dist}, {sep}, nullptr});
}

int mb200_distance_point_line(mb200_stream stream, int64_t n, const double* x, const double* line_center,
                              const double* line_dir, double* dist, double* cp, double* arch) {
  return launch<PointLine>(stream, n, PrimArgs{{x, line_center, line_dir}, {}, {}, {dist, arch}, {cp}, nullptr});
}

int mb200_distance_line_line(mb200_stream stream, int64_t n, const double* c1, const double* d1, const double* c2,
                             const double* d2, double* dist, double* cp1, double* cp2, double* arch1, double* arch2,
                             double* sep) {
  return launch<LineLine>(stream, n, PrimArgs{{c1, d1, c2, d2}, {}, {}, {dist, arch1, arch2}, {cp1, cp2, sep}, nullptr});
}

int mb200_distance_line_sphere(mb200_stream stream, int64_t n, const double* line_center, const double* line_dir,
                               const double* c, const double* r, double* dist, double* cp, double* arch) {
  return launch<LineSphere>(stream, n, PrimArgs{{line_center, line_dir, c}, {}, {r}, {dist, arch}, {cp}, nullptr});
}

int mb200_distance_point_ellipsoid(mb200_stream stream, int64_t n, const double* x, const double* center,
                                   const double* quat, const double* radii, double* dist, double* cp,
                                   double* normal) {
  return launch<PointEllipsoid>(stream, n, PrimArgs{{x, center, radii}, {quat}, {}, {dist}, {cp, normal}, nullptr});
}

int mb200_distance_ellipsoid_ellipsoid(mb200_stream stream, int64_t n, const double* c1, const double* q1,
                                       const double* radii1, const double* c2, const double* q2,
                                       const double* radii2, double* dist, double* cp1, double* cp2, double* normal1) {
  return launch<EllipsoidEllipsoid>(stream, n,
                                    PrimArgs{{c1, radii1, c2, radii2}, {q1, q2}, {}, {dist}, {cp1, cp2, normal1}, nullptr});
}

int mb200_distance_circle3d_circle3d(mb200_stream stream, int64_t n, const double* c1, const double* q1,
                                     const double* r1, const double* c2, const double* q2, const double* r2,
                                     double* dist, double* cp1, double* cp2, double* normal1) {
  return launch<CircleCircle>(stream, n, PrimArgs{{c1, c2}, {q1, q2}, {r1, r2}, {dist}, {cp1, cp2, normal1}, nullptr});
}

}  // extern "C"
