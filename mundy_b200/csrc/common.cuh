// Shared device/host helpers for the mundy_b200 kernels (sm_100a only).
// Arithmetic helpers restate the operation ORDER of the reference's small-vector algebra so that, compiled with
// -fmad=false, per-element results are bit-identical to the reference's CPU evaluation:
//   dot   : right fold a0*b0 + (a1*b1 + a2*b2)        mundy/math/src/mundy_math/impl/VectorImpl.hpp:338-345
//   norm  : sqrt(dot(v, v))                            mundy/math/src/mundy_math/Vector.hpp:1151-1156
//   cross : mundy/math/src/mundy_math/Vector3.hpp:93-106
//   q (x) v = q * (0,v) * conj(q)/|q|^2                mundy/math/src/mundy_math/impl/QuaternionImpl.hpp:142-202
#pragma once
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>
#include <stdint.h>

#include <atomic>
#include <cstdio>
#include <string>

#include "../../include/mundy_b200.h"

namespace mb200 {

// ---- host-side error plumbing -------------------------------------------------------------------------------
void set_error(const std::string& msg);
extern std::atomic<long long> g_launches;
inline void count_launch(int k = 1) { g_launches.fetch_add(k, std::memory_order_relaxed); }

#define MB200_CUDA(expr)                                                                          \
  do {                                                                                            \
    cudaError_t _e = (expr);                                                                      \
    if (_e != cudaSuccess) {                                                                      \
      ::mb200::set_error(std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" + __FILE__ +  \
                         ":" + std::to_string(__LINE__) + ")");                                   \
      return MB200_ERR_CUDA;                                                                      \
    }                                                                                             \
  } while (0)

#define MB200_REQUIRE(cond, msg)                    \
  do {                                              \
    if (!(cond)) {                                  \
      ::mb200::set_error(std::string("invalid argument: ") + (msg)); \
      return MB200_ERR_INVALID_ARGUMENT;            \
    }                                               \
  } while (0)

#define MB200_TRY(expr)        \
  do {                         \
    int _s = (expr);           \
    if (_s != MB200_OK) return _s; \
  } while (0)

// Call after every kernel launch: counts it and catches launch-configuration errors.
#define MB200_LAUNCHED()                   \
  do {                                     \
    ::mb200::count_launch();               \
    MB200_CUDA(cudaGetLastError());        \
  } while (0)

inline cudaStream_t as_stream(mb200_stream s) { return reinterpret_cast<cudaStream_t>(s); }

// Stage timers (core.cu). A StageScope brackets the launches issued during its lifetime with an event pair on `st`
// when profiling is enabled, and costs one relaxed load otherwise. It is also an NVTX range (header-only NVTX 3: a
// no-op unless a tool such as Nsight Systems is attached), so that timelines show the stages by name.
extern std::atomic<int> g_profile_on;
extern const char* const kStageNames[];
int prof_begin(int stage, cudaStream_t st);
void prof_end(int idx, cudaStream_t st);
struct StageScope {
  int idx = -1;
  cudaStream_t st;
  StageScope(int stage, cudaStream_t s) : st(s) {
    nvtxRangePushA(kStageNames[stage]);
    if (g_profile_on.load(std::memory_order_relaxed)) idx = prof_begin(stage, s);
  }
  ~StageScope() {
    if (idx >= 0) prof_end(idx, st);
    nvtxRangePop();
  }
  StageScope(const StageScope&) = delete;
  StageScope& operator=(const StageScope&) = delete;
};

// Verifies that a device exists before the first compute call (no CPU fallback).
int require_device();

// Page-locked host scratch of the calling thread (grows, kept until the thread ends; nullptr if it cannot be had): the
// landing place of the small device -> host reads a step waits on. A copy into pageable memory is staged by the driver
// and costs tens of microseconds more per read.
void* pinned_scratch(size_t bytes);

// Stream-ordered scratch buffer (cudaMallocAsync pool; freed on the same stream when it goes out of scope).
template <typename T>
struct Scratch {
  T* p = nullptr;
  cudaStream_t st = nullptr;
  size_t count = 0;
  Scratch() = default;
  Scratch(const Scratch&) = delete;
  Scratch& operator=(const Scratch&) = delete;
  ~Scratch() { release(); }
  int alloc(size_t n, cudaStream_t s) {
    release();
    st = s;
    count = n;
    if (n == 0) return MB200_OK;
    // Sizes are rounded up to a granule of about 1/32 of the request (at least 64 KB): the per-step scratch sizes follow
    // the contact count, which moves by a few parts in 10^4 from step to step, and the stream-ordered pool can only
    // reuse a cached block that is large enough. With repeating sizes it never has to map new memory mid-run.
    size_t bytes = n * sizeof(T);
    size_t granule = size_t(64) << 10;
    while (granule * 32 < bytes) granule <<= 1;
    bytes = (bytes + granule - 1) / granule * granule;
    MB200_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&p), bytes, s));
    return MB200_OK;
  }
  void release() {
    if (p) cudaFreeAsync(p, st);
    p = nullptr;
  }
  operator T*() const { return p; }
};

// Persistent (handle-owned) device buffer that only ever grows.
template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t cap = 0;
  DevBuf() = default;
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  ~DevBuf() {
    if (p) cudaFree(p);
  }
  int reserve(size_t n) {
    if (n <= cap) return MB200_OK;
    if (p) MB200_CUDA(cudaFree(p));
    p = nullptr;
    size_t want = n + n / 4 + 64;  // headroom: cudaFree + cudaMalloc of a large buffer costs milliseconds (more with peer mappings)
    MB200_CUDA(cudaMalloc(reinterpret_cast<void**>(&p), want * sizeof(T)));
    cap = want;
    return MB200_OK;
  }
  operator T*() const { return p; }
};

constexpr int kSMs = 148;  // B200: grids of persistent kernels are sized in multiples of this

inline int grid_for(int64_t n, int block, int max_blocks = 1 << 30) {
  int64_t g = (n + block - 1) / block;
  if (g < 1) g = 1;
  if (g > max_blocks) g = max_blocks;
  return static_cast<int>(g);
}

// ---- device math ---------------------------------------------------------------------------------------------
constexpr double kZeroTol = 1e-15;  // mundy/math/src/mundy_math/Tolerance.hpp:37-50 get_zero_tolerance<double>()

struct V3 {
  double x, y, z;
};
__host__ __device__ inline V3 operator-(const V3& a, const V3& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__host__ __device__ inline V3 operator+(const V3& a, const V3& b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__host__ __device__ inline V3 operator*(double s, const V3& a) { return {s * a.x, s * a.y, s * a.z}; }
__host__ __device__ inline V3 operator*(const V3& a, double s) { return {a.x * s, a.y * s, a.z * s}; }
__host__ __device__ inline double dot(const V3& a, const V3& b) { return a.x * b.x + (a.y * b.y + a.z * b.z); }
__host__ __device__ inline double norm(const V3& a) { return sqrt(dot(a, a)); }
__host__ __device__ inline V3 cross(const V3& a, const V3& b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
__device__ inline V3 ld3(const double* __restrict__ p, int64_t i) { return {p[3 * i], p[3 * i + 1], p[3 * i + 2]}; }
__device__ inline void st3(double* __restrict__ p, int64_t i, const V3& v) {
  p[3 * i] = v.x;
  p[3 * i + 1] = v.y;
  p[3 * i + 2] = v.z;
}

struct Q4 {
  double w, x, y, z;
};
__host__ __device__ inline Q4 qmul(const Q4& a, const Q4& b) {
  Q4 r;
  r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
  r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
  r.y = a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x;
  r.z = a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w;
  return r;
}
__host__ __device__ inline V3 qrot(const Q4& q, const V3& v) {
  const Q4 vq{0.0, v.x, v.y, v.z};
  const double inv_n2 = 1.0 / (q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
  const Q4 qi{q.w * inv_n2, -q.x * inv_n2, -q.y * inv_n2, -q.z * inv_n2};
  const Q4 r = qmul(qmul(q, vq), qi);
  return {r.x, r.y, r.z};
}
__device__ inline Q4 ldq(const double* __restrict__ p, int64_t i) {
  const double2 a = *reinterpret_cast<const double2*>(p + 4 * i);
  const double2 b = *reinterpret_cast<const double2*>(p + 4 * i + 2);
  return {a.x, a.y, b.x, b.y};
}

// 32-byte records move as single 256-bit transactions (sm_100: LDG.E.ENL2.256 / STG.E.ENL2.256), i.e. exactly one
// 32-byte DRAM/L2 sector per gathered item.
struct __align__(32) D4 {
  double x, y, z, w;
};
__device__ inline D4 ldg256(const D4* p) {  // read-only path
  D4 r;
  asm("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p));
  return r;
}
__device__ inline D4 ld256(const D4* p) {
  D4 r;
  asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ inline void st256(D4* p, const D4& v) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v.x), "d"(v.y), "d"(v.z), "d"(v.w) : "memory");
}

// Orthorhombic periodic metric: geom::PeriodicScaledMetric (mundy/geom/src/mundy_geom/periodicity.hpp:743-845) when all
// three axes are periodic, PeriodicMetricX / Y / XY / YZ (:335-741) when only some are (their free axes carry scale 1
// and are not rounded, which is the identity in fp64). Passed to kernels by value.
struct Periodic {
  double L[3];    // scale_
  double inv[3];  // scale_inv_ = 1.0 / scale_
  int on[3];
  int any;
};
inline Periodic make_periodic(const double* cell_size) {  // cell_size[d] <= 0 (or NULL): free axis
  Periodic P{};
  for (int d = 0; d < 3; ++d) {
    P.on[d] = cell_size && cell_size[d] > 0.0;
    P.L[d] = P.on[d] ? cell_size[d] : 1.0;
    P.inv[d] = 1.0 / P.L[d];
    P.any |= P.on[d];
  }
  return P;
}
// frac_minimum_image<int64_t> (:799-803): x - Scalar(int64(round(x))), round = half away from zero.
__host__ __device__ inline double frac_min_image(double x) { return x - static_cast<double>(static_cast<long long>(round(x))); }
// impl::safe_unit_mod1<int64_t> (:140-150)
__host__ __device__ inline double safe_unit_mod1(double s) {
  const double k = static_cast<double>(static_cast<long long>(floor(s)));
  double t = s - k;
  if (fabs(t - 1.0) < kZeroTol) t = 0.0;
  return t;
}
// metric.sep(p1, p2) (:814-818): from_fractional(frac_minimum_image(to_fractional(p2 - p1)))
__host__ __device__ inline V3 periodic_sep(const Periodic& P, const V3& p1, const V3& p2) {
  V3 d = p2 - p1;
  if (P.on[0]) d.x = P.L[0] * frac_min_image(P.inv[0] * d.x);
  if (P.on[1]) d.y = P.L[1] * frac_min_image(P.inv[1] * d.y);
  if (P.on[2]) d.z = P.L[2] * frac_min_image(P.inv[2] * d.z);
  return d;
}
// metric.wrap(p) (:821-824): from_fractional(frac_wrap_to_unit_cell(to_fractional(p)))
__host__ __device__ inline V3 periodic_wrap(const Periodic& P, const V3& p) {
  V3 o = p;
  if (P.on[0]) o.x = P.L[0] * safe_unit_mod1(P.inv[0] * p.x);
  if (P.on[1]) o.y = P.L[1] * safe_unit_mod1(P.inv[1] * p.y);
  if (P.on[2]) o.z = P.L[2] * safe_unit_mod1(P.inv[2] * p.z);
  return o;
}
// unwrap_points_to_ref(point, metric, ref) (:1333-1339): ref + metric.sep(ref, point). Spheres and spherocylinders
// unwrap their centre this way (:1487-1516).
__host__ __device__ inline V3 periodic_unwrap(const Periodic& P, const V3& p, const V3& ref) {
  return ref + periodic_sep(P, ref, p);
}

// distance(Point, Point, sep)  mundy/geom/src/mundy_geom/distance/PointPoint.hpp:54-62
__host__ __device__ inline double dist_pp(const V3& p1, const V3& p2, V3& sep) {
  sep = p2 - p1;
  return norm(sep);
}

// distance(Point, LineSegment, cp, t, sep)  mundy/geom/src/mundy_geom/distance/PointLineSegment.hpp:128-172
__host__ __device__ inline double dist_pseg(const V3& x, const V3& p1, const V3& p2, V3& cp, double& t, V3& sep) {
  const V3 p21 = p2 - p1;
  const double num = dot(p21, x - p1);
  cp = p1;
  t = 0.0;
  if (!((num < kZeroTol) & (num > -kZeroTol))) {
    const double denom = dot(p21, p21);
    if (!(denom < kZeroTol)) {
      t = num / denom;
      if (t < 0.0) {
        cp = p1;
      } else if (t > 1.0) {
        cp = p2;
      } else {
        cp = p1 + t * p21;
      }
    }
  }
  return dist_pp(x, cp, sep);
}

// distance(LineSegment, LineSegment, cp1, cp2, s, t, sep)
// mundy/geom/src/mundy_geom/distance/LineSegmentLineSegment.hpp:189-318
__host__ __device__ inline double dist_segseg(const V3& l0, const V3& l1, const V3& m0, const V3& m1, V3& cp1, V3& cp2,
                                              double& s, double& t, V3& sep) {
  const V3 u = l1 - l0, v = m1 - m0, w = l0 - m0;
  const double a = dot(u, u), b = dot(u, v), c = dot(v, v), d = dot(u, w), e = dot(v, w);
  const double D = a * c - b * b;
  if (D < 3.1622776601683795e-08) {  // sqrt(1e-15), :215 — collinear / degenerate branch
    V3 c1, c2, c3, c4, s1, s2, s3, s4;
    double t1, t2, t3, t4;
    const double d1 = dist_pseg(l0, m0, m1, c1, t1, s1);
    const double d2 = dist_pseg(l1, m0, m1, c2, t2, s2);
    const double d3 = dist_pseg(m0, l0, l1, c3, t3, s3);
    const double d4 = dist_pseg(m1, l0, l1, c4, t4, s4);
    const double dmin = fmin(fmin(d1, d2), fmin(d3, d4));
    if (dmin == d1) {
      s = 0.0, t = t1, cp1 = l0, cp2 = c1, sep = s1;
    } else if (dmin == d2) {
      s = 1.0, t = t2, cp1 = l1, cp2 = c2, sep = s2;
    } else if (dmin == d3) {
      s = t3, t = 0.0, cp1 = c3, cp2 = m0, sep = s3;
    } else {
      s = t4, t = 1.0, cp1 = c4, cp2 = m1, sep = s4;
    }
    return dmin;
  }
  double sN = b * e - c * d, tN = a * e - b * d, sD = D, tD = D;
  if (sN < 0.0) {
    sN = 0.0, tN = e, tD = c;
  } else if (sN > sD) {
    sN = sD, tN = e + b, tD = c;
  }
  if (tN < 0.0) {
    tN = 0.0;
    if (-d < 0.0) {
      sN = 0.0;
    } else if (-d > a) {
      sN = sD;
    } else {
      sN = -d, sD = a;
    }
  } else if (tN > tD) {
    tN = tD;
    if ((-d + b) < 0.0) {
      sN = 0.0;
    } else if ((-d + b) > a) {
      sN = sD;
    } else {
      sN = (-d + b), sD = a;
    }
  }
  s = (fabs(sN) < kZeroTol) ? 0.0 : sN / sD;
  t = (fabs(tN) < kZeroTol) ? 0.0 : tN / tD;
  cp1 = l0 + s * u;
  cp2 = m0 + t * v;
  return dist_pp(cp1, cp2, sep);
}

// math::rotate_quaternion(q, omega, dt)  mundy/math/src/mundy_math/Quaternion.hpp:1359-1385 (Delong 2015 App. A)
__device__ inline Q4 rotate_quaternion(const Q4& q, const V3& om, double dt) {
  const double w = norm(om);
  if (w < kZeroTol) return q;
  const double winv = 1.0 / w;
  const double sw = sin(0.5 * w * dt);
  const double cw = cos(0.5 * w * dt);
  const double s = q.w;
  const V3 p{q.x, q.y, q.z};
  const V3 t1 = ((s * sw) * om) * winv;
  const V3 t2 = cw * p;
  const V3 t3 = (sw * winv) * cross(om, p);
  const V3 xyz = (t1 + t2) + t3;
  const double qw = s * cw - dot(om, p) * sw * winv;
  const double inv_norm = 1.0 / sqrt(qw * qw + xyz.x * xyz.x + xyz.y * xyz.y + xyz.z * xyz.z);
  return Q4{qw * inv_norm, xyz.x * inv_norm, xyz.y * inv_norm, xyz.z * inv_norm};
}

// ---- warp / block reductions (fixed order => run-to-run deterministic) -----------------------------------------
__device__ inline double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}
__device__ inline double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_down_sync(0xffffffffu, v, o));
  return v;
}
// Block-wide sum; result valid in thread 0. `sh` must hold >= 32 doubles. All threads must call.
__device__ inline double block_sum(double v, double* sh) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) sh[w] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  v = (threadIdx.x < nw) ? sh[threadIdx.x] : 0.0;
  if (w == 0) v = warp_sum(v);
  return v;
}
__device__ inline double block_max(double v, double* sh, double identity) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) sh[w] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  v = (threadIdx.x < nw) ? sh[threadIdx.x] : identity;
  if (w == 0) v = warp_max(v);
  return v;
}

}  // namespace mb200
