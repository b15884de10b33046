// Internal body records and the per-pair shape distance shared by the search filter and the narrow phase.
//
// A body is two 32-byte records (one DRAM sector each, moved with 256-bit loads):
//   B0 = {cx, cy, cz, r}          centre and radius
//   B1 = {hx, hy, hz, kind}       half centre-line vector h = (0.5 * L) * (q (x) z_hat); kind 0 sphere / 1 rod
// Endpoints are formed per pair as c - h and c + h, the same operations the reference composition performs
// (scrap/.../SpherocylinderSpherocylinderLinker.cpp:207-214 with the z-axis of mundy/geom compute_aabb.hpp:115-117),
// so per-pair results are bit-identical to the CPU oracle.
#pragma once
#include "common.cuh"

namespace mb200 {

struct PairGeom {
  double sep;  // signed surface separation
  V3 n;        // unit normal from i to j
  V3 arm_i;    // centre-line closest point minus centre (zero for spheres)
  V3 arm_j;
  // axial form of the arms: arm = sigma h, h = the body's half-axis vector, sigma = 2 clamp(s, 0, 1) - 1 with s the
  // segment parameter of the closest point (geom::distance returns the point-segment parameter unclamped,
  // PointLineSegment.hpp:156-167); 0 for spheres. Equal to the vectors above up to rounding.
  double sig_i, sig_j;
};

// Narrow-phase arithmetic for one pair. RODS=false: both bodies are spheres and b1* are not read.
template <bool RODS>
__device__ inline PairGeom pair_geometry(const D4& b0i, const D4& b1i, const D4& b0j, const D4& b1j) {
  PairGeom g;
  g.arm_i = V3{0.0, 0.0, 0.0};
  g.arm_j = V3{0.0, 0.0, 0.0};
  g.sig_i = 0.0, g.sig_j = 0.0;
  const V3 ci{b0i.x, b0i.y, b0i.z}, cj{b0j.x, b0j.y, b0j.z};
  const double ri = b0i.w, rj = b0j.w;
  const bool rod_i = RODS && (b1i.w != 0.0), rod_j = RODS && (b1j.w != 0.0);
  if (!rod_i && !rod_j) {
    // scrap/parameter_interface/NGPSpheresLCP.cpp:384-406 (driver form of geom::distance(Sphere, Sphere))
    const double sx = cj.x - ci.x, sy = cj.y - ci.y, sz = cj.z - ci.z;
    const double d = sqrt(sx * sx + sy * sy + sz * sz);
    g.sep = d - ri - rj;
    const double inv = 1.0 / d;
    g.n = V3{sx * inv, sy * inv, sz * inv};
    return g;
  }
  V3 cp1, cp2, sv;
  double d, s = 0.5, t = 0.5;
  if (rod_i && rod_j) {
    const V3 hi{b1i.x, b1i.y, b1i.z}, hj{b1j.x, b1j.y, b1j.z};
    d = dist_segseg(ci - hi, ci + hi, cj - hj, cj + hj, cp1, cp2, s, t, sv);
  } else if (!rod_i) {  // sphere i, rod j: point-segment, sep vector from the point (i) to the segment (j)
    const V3 hj{b1j.x, b1j.y, b1j.z};
    cp1 = ci;
    d = dist_pseg(ci, cj - hj, cj + hj, cp2, t, sv);
  } else {  // rod i, sphere j: sep from the point (j) to the segment (i), flipped so that n points i -> j
    const V3 hi{b1i.x, b1i.y, b1i.z};
    cp2 = cj;
    d = dist_pseg(cj, ci - hi, ci + hi, cp1, s, sv);
    sv = V3{-sv.x, -sv.y, -sv.z};
  }
  g.sep = d - (ri + rj);
  const double inv = 1.0 / d;
  g.n = sv * inv;
  if (rod_i) g.arm_i = cp1 - ci, g.sig_i = 2.0 * fmin(fmax(s, 0.0), 1.0) - 1.0;
  if (rod_j) g.arm_j = cp2 - cj, g.sig_j = 2.0 * fmin(fmax(t, 0.0), 1.0) - 1.0;
  return g;
}

// Candidate predicate of the broad phase: search spheres touch or overlap, evaluated in fp64
// exactly as written (no FMA): d2 = dx*dx + (dy*dy + dz*dz) <= (Si + Sj)^2.
__device__ inline bool is_candidate(const D4& si, const D4& sj) {
  const double dx = sj.x - si.x, dy = sj.y - si.y, dz = sj.z - si.z;
  const double d2 = dx * dx + (dy * dy + dz * dz);
  const double s = si.w + sj.w;
  return d2 <= s * s;
}

// Periodic forms: the second body is unwrapped to the first body's reference point (its centre,
// periodicity.hpp:925-933 reference_point) with unwrap_points_to_ref (:1487-1516) and the free-space functions above
// are applied to the result; the search predicate measures the centre separation with metric.sep (:814-818).
__device__ inline D4 unwrap_to(const Periodic& P, const D4& b0i, const D4& b0j) {
  const V3 c = periodic_unwrap(P, V3{b0j.x, b0j.y, b0j.z}, V3{b0i.x, b0i.y, b0i.z});
  return D4{c.x, c.y, c.z, b0j.w};
}
__device__ inline bool is_candidate_periodic(const Periodic& P, const D4& si, const D4& sj) {
  const V3 d = periodic_sep(P, V3{si.x, si.y, si.z}, V3{sj.x, sj.y, sj.z});
  const double d2 = d.x * d.x + (d.y * d.y + d.z * d.z);
  const double s = si.w + sj.w;
  return d2 <= s * s;
}

}  // namespace mb200
