// B. Broad phase: the device replacement of mesh::GenNeighborLinks' search
// (mundy/mesh/src/mundy_mesh/GenNeighborLinkers.hpp:510-543,564-664; stk::search::coarse_search is third-party).
//
//   search spheres {c, S}  ->  Morton key of the uniform-grid cell (edge >= 2 S_max)  ->  radix sort (key, body)
//   -> bodies gathered into Morton order (32-byte records)  ->  dense cell table [start, end)
//   -> a warp walks each body's <= 27 neighbour cells in ascending key order, 32 candidates per step read as one
//      contiguous run of records, exact fp64 predicate, warp ballot + popc compaction
//   -> count pass, exclusive scan, fill pass: pair list sorted by (i, j) in Morton-order indices, no atomics.
// The optional shape-distance search filter (GenNeighborLinkers.hpp:116-200 SearchFilter) is fused into the same passes.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <string>
#include <type_traits>
#include <vector>

#include "links.cuh"
#include "shapes.cuh"
#include "sort_scan.cuh"

namespace mb200 {
namespace {

constexpr int kMaxBitsPerAxis = 8;  // dense cell table of at most 2^24 entries

struct GridParams {
  double lo[3];
  double h[3];
  int nc[3];
  int bits;       // per axis
  int per[3];     // periodic axis: cells tile [0, L) exactly, h = L / nc, neighbour cells wrap around
  double inv[3];  // 1 / L on periodic axes
};

__device__ __host__ inline uint32_t spread3(uint32_t v) {  // 10 bits -> every third bit
  v &= 0x3ffu;
  v = (v | (v << 16)) & 0x030000ffu;
  v = (v | (v << 8)) & 0x0300f00fu;
  v = (v | (v << 4)) & 0x030c30c3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}
// z is the most significant dimension, as in the reference's Z-order comparator (mundy/math/.../zmort.hpp:261-265).
__device__ __host__ inline uint32_t morton3(uint32_t cx, uint32_t cy, uint32_t cz) {
  return spread3(cx) | (spread3(cy) << 1) | (spread3(cz) << 2);
}

__device__ inline int cell_coord(double x, const GridParams& gp, int d) {
  int v;
  if (gp.per[d]) {  // cell of the wrapped fractional coordinate (any periodic image gives the same cell)
    const double f = x * gp.inv[d];
    v = static_cast<int>((f - floor(f)) * gp.nc[d]);
  } else {
    v = static_cast<int>(floor((x - gp.lo[d]) / gp.h[d]));
  }
  return min(max(v, 0), gp.nc[d] - 1);
}

// Search spheres in the caller's order + per-CTA bounds partials (min xyz, max xyz, max S).
__global__ void __launch_bounds__(256)
    search_spheres_kernel(int64_t n, const double* __restrict__ center, const double* __restrict__ radius,
                          const double* __restrict__ length, const int32_t* __restrict__ kind,
                          const double* __restrict__ explicit_S, double buffer, D4* __restrict__ sr,
                          double* __restrict__ partials) {
  __shared__ double sh[33];
  const double big = 1.7976931348623157e308;
  double lo[3] = {big, big, big}, hi[3] = {-big, -big, -big}, smax = 0.0;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const V3 c = ld3(center, i);
    double S;
    if (explicit_S) {
      S = explicit_S[i];
    } else {
      // bounding radius: r (compute_bounding_radius.hpp:61-64) or 0.5*L + r (:74-82); search sphere = + buffer
      const bool rod = kind && kind[i] == MB200_SPHEROCYLINDER;
      const double rb = rod ? 0.5 * length[i] + radius[i] : radius[i];
      S = rb + buffer;
    }
    st256(sr + i, D4{c.x, c.y, c.z, S});
    lo[0] = fmin(lo[0], c.x), lo[1] = fmin(lo[1], c.y), lo[2] = fmin(lo[2], c.z);
    hi[0] = fmax(hi[0], c.x), hi[1] = fmax(hi[1], c.y), hi[2] = fmax(hi[2], c.z);
    smax = fmax(smax, S);
  }
  double out[7];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    out[k] = -block_max(-lo[k], sh, -big);
    out[3 + k] = block_max(hi[k], sh, -big);
  }
  out[6] = block_max(smax, sh, 0.0);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < 7; ++k) partials[7 * blockIdx.x + k] = out[k];
  }
}

__global__ void __launch_bounds__(256) cell_keys_kernel(int64_t n, const D4* __restrict__ sr, GridParams gp,
                                                        uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const D4 s = ldg256(sr + i);
  keys[i] = morton3(cell_coord(s.x, gp, 0), cell_coord(s.y, gp, 1), cell_coord(s.z, gp, 2));
  vals[i] = static_cast<uint32_t>(i);
}

__global__ void __launch_bounds__(256) gather_records_kernel(int64_t n, const uint32_t* __restrict__ perm,
                                                             const D4* __restrict__ src, D4* __restrict__ dst) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k < n) st256(dst + k, ldg256(src + perm[k]));
}

// cell_range[key] = {start, end} over the sorted keys (table pre-zeroed: empty cells have start == end == 0).
// A. AABB stage on the step's path: geom::compute_aabb(Spherocylinder) / (Sphere) (mundy/geom/src/mundy_geom/
// compute_aabb.hpp:72-80,106-127: corners min/max(c -/+ d) -/+ r with d = (L/2)(q (x) z_hat) = the record's half-axis
// vector), inflated by the search buffer and rounded OUTWARD to fp32: a conservative pre-reject for the shape-distance
// filter. Two bodies whose surface separation is <= 2 * buffer have overlapping buffer-inflated capsules, hence
// overlapping boxes; the outward fp32 rounding (>= 6e-8 relative) dwarfs the fp64 rounding of the separation itself.
// The exact fp64 predicate and the exact shape distance stay the deciders. Record: {lo.xyz, hi.xyz, -, -} as 8 floats.
__global__ void __launch_bounds__(256) box_records_kernel(int64_t n, const D4* __restrict__ b0, const D4* __restrict__ b1,
                                                          double buffer, float4* __restrict__ box) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const D4 c = ldg256(b0 + k), h = ldg256(b1 + k);
  const double e = c.w + buffer;
  const double ex = fabs(h.x) + e, ey = fabs(h.y) + e, ez = fabs(h.z) + e;
  box[2 * k] = make_float4(__double2float_rd(c.x - ex), __double2float_rd(c.y - ey), __double2float_rd(c.z - ez),
                           __double2float_ru(c.x + ex));
  box[2 * k + 1] = make_float4(__double2float_ru(c.y + ey), __double2float_ru(c.z + ez), 0.f, 0.f);
}
struct Box6f {
  float lx, ly, lz, hx, hy, hz;
};
__device__ inline Box6f load_box(const float4* box, int64_t k) {
  const D4 r = ldg256(reinterpret_cast<const D4*>(box) + k);  // one 256-bit load
  Box6f b;
  b.lx = __int_as_float(__double2loint(r.x)), b.ly = __int_as_float(__double2hiint(r.x));
  b.lz = __int_as_float(__double2loint(r.y)), b.hx = __int_as_float(__double2hiint(r.y));
  b.hy = __int_as_float(__double2loint(r.z)), b.hz = __int_as_float(__double2hiint(r.z));
  return b;
}
// closed intervals: touching counts (primitives/AABB.hpp:419-431)
__device__ inline bool boxes_overlap(const Box6f& a, const Box6f& b) {
  return !(a.hx < b.lx || b.hx < a.lx || a.hy < b.ly || b.hy < a.ly || a.hz < b.lz || b.hz < a.lz);
}

__global__ void __launch_bounds__(256) cell_ranges_kernel(int64_t n, const uint32_t* __restrict__ keys,
                                                          int2* __restrict__ cell_range) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const uint32_t key = keys[k];
  if (k == 0 || keys[k - 1] != key) cell_range[key].x = static_cast<int>(k);
  if (k == n - 1 || keys[k + 1] != key) cell_range[key].y = static_cast<int>(k + 1);
}

// One warp per kFindGroup consecutive bodies (Morton order), one body at a time; the per-cell table (neighbour cells,
// ranks, prefixes) is shared by the bodies of a cell.
//   FILL = false: counts[i] = number of kept pairs (i, j>i); the first `slot_cap` of them are also cached in
//                 slots[i * slot_cap ...] so that the fill pass does not have to search again.
//   FILL = true : writes the pairs at pairs[rowptr[i] ...] in ascending j: copied from the slots when they all fit,
//                 found again otherwise.
// The candidates of the <= 27 neighbour cells with key >= own key are walked as ONE flattened index range in
// ascending key order (= ascending j): 64 candidates per trip, both 256-bit record loads of a lane issued before the
// first test, exact fp64 predicate, warp ballot + popc compaction.
// PER: periodic box. Neighbour cells wrap around on periodic axes (an axis with fewer than three cells visits each of
// its cells once), the predicate uses the minimum-image separation and the shape filter unwraps the second body.
constexpr int kFindGroup = 8;  // consecutive bodies per warp in find_pairs_kernel
#ifndef MB200_FIND_MIN_CTAS
#define MB200_FIND_MIN_CTAS 4  // 64 registers: four CTAs (32 warps) per SM; measured 4.70 / 3.98 / 3.67 ms per rebuild at 2 / 3 / 4
#endif
// BOX (shape filter, free space): the candidate walk tests the fp32 boxes above instead of the search spheres; the exact
// sphere predicate moves to the queued survivors, in front of the shape distance.
template <bool FILL, bool FILTER, bool PER, bool BOX>
__global__ void __launch_bounds__(256, MB200_FIND_MIN_CTAS)
    find_pairs_kernel(int64_t n, const D4* __restrict__ sr, const float4* __restrict__ box, const D4* __restrict__ b0,
                      const D4* __restrict__ b1,
                      const uint32_t* __restrict__ perm, const int2* __restrict__ cell_range, GridParams gp,
                      double cutoff, uint32_t* __restrict__ counts, const int64_t* __restrict__ rowptr,
                      int2* __restrict__ pairs, int32_t* __restrict__ slots, int slot_cap,
                      const uint8_t* __restrict__ own, Periodic P) {
  const int lane = threadIdx.x & 31;
  const int wib = threadIdx.x >> 5;
  // A warp takes kFindGroup consecutive bodies. Consecutive bodies mostly share their cell, and with it the neighbour
  // cells, their ranks and prefixes: that set-up (27 cell keys and ranges, 54 shuffles) is redone only when the cell
  // changes. Inside a cell only the own cell's range [i + 1, end) moves: every later prefix shrinks by i - base_i.
  const int64_t i_begin = ((static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5) * kFindGroup;
  if (i_begin >= n) return;
  const int64_t i_end = min(n, i_begin + kFindGroup);
  const uint32_t lt_mask = (1u << lane) - 1u;
  __shared__ int s_pre[8][28], s_beg[8][28];
  __shared__ int s_queue[FILTER ? 8 : 1][FILTER ? 96 : 1];
  const int* wpre = s_pre[wib];
  const int* wbeg = s_beg[wib];
  int* q = s_queue[FILTER ? wib : 0];
  uint32_t cur_key = 0xffffffffu;
  int64_t base_i = 0;
  int nvalid = 0, total_base = 0, rpre_base = 0x7fffffff;
  for (int64_t i = i_begin; i < i_end; ++i) {
  int64_t out = FILL ? rowptr[i] : 0;
  if (FILL) {
    const int cnt = static_cast<int>(rowptr[i + 1] - out);
    if (cnt <= slot_cap) {  // everything was cached by the count pass
      for (int k = lane; k < cnt; k += 32)
        pairs[out + k] = make_int2(static_cast<int>(i), slots[i * slot_cap + k]);
      continue;
    }
  }
  const D4 si = ldg256(sr + i);
  Box6f bi{};
  if (BOX) bi = load_box(box, i);
  // multi-GPU: links between two ghost bodies belong to other ranks
  const bool own_i = own ? own[i] != 0 : true;
  // (the shape filter's records of body i are re-read where they are used: one broadcast L1 hit per batch instead of 17
  // registers held across the candidate loop)
  const int cx = cell_coord(si.x, gp, 0);
  const int cy = cell_coord(si.y, gp, 1);
  const int cz = cell_coord(si.z, gp, 2);
  const uint32_t my_key = morton3(cx, cy, cz);
  if (my_key != cur_key) {
  // lane l < 27 owns neighbour cell l
  uint32_t nkey = 0xffffffffu;
  int beg = 0, len = 0;
  if (lane < 27) {
    int x = cx + (lane % 3) - 1, y = cy + ((lane / 3) % 3) - 1, z = cz + (lane / 9) - 1;
    bool inside = true;
    if (PER) {
      const int off[3] = {(lane % 3) - 1, ((lane / 3) % 3) - 1, (lane / 9) - 1};
      int* xyz[3] = {&x, &y, &z};
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        if (gp.per[d]) {
          // offsets that would revisit a cell of a short axis are dropped: -1 needs two cells, +1 needs three
          inside = inside && (off[d] == 0 || (off[d] < 0 ? gp.nc[d] >= 2 : gp.nc[d] >= 3));
          *xyz[d] = (*xyz[d] + gp.nc[d]) % gp.nc[d];
        }
      }
    }
    if (inside && x >= 0 && y >= 0 && z >= 0 && x < gp.nc[0] && y < gp.nc[1] && z < gp.nc[2]) {
      const uint32_t k = morton3(x, y, z);
      if (k >= my_key) {  // bodies with a smaller key have a smaller index: never j > i
        const int2 r = __ldg(cell_range + k);
        beg = (k == my_key) ? static_cast<int>(i) + 1 : r.x;
        len = r.y - beg;
        // the own cell (smallest key: rank 0) stays in the table even when i is its last body, so that the bodies that
        // reuse this table find it at rank 0
        if (len > 0 || k == my_key) nkey = k;
        if (len < 0) len = 0;
      }
    }
  }
  const bool valid = nkey != 0xffffffffu;
  // ascending-key rank among the valid cells and the number of candidates in front of this cell
  int rank = 0, pre = 0;
#pragma unroll
  for (int m = 0; m < 27; ++m) {
    const uint32_t km = __shfl_sync(0xffffffffu, nkey, m);
    const int lm = __shfl_sync(0xffffffffu, len, m);
    if (km < nkey) {
      rank += 1;
      pre += lm;
    }
  }
  nvalid = __popc(__ballot_sync(0xffffffffu, valid));
  __syncwarp();  // the previous table is no longer read
  if (valid) {
    s_pre[wib][rank] = pre;
    s_beg[wib][rank] = beg;
  }
  int tc = valid ? pre + len : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) tc = max(tc, __shfl_xor_sync(0xffffffffu, tc, o));
  __syncwarp();
  total_base = tc;
  rpre_base = (lane < nvalid) ? wpre[lane] : 0x7fffffff;
  cur_key = my_key;
  base_i = i;
  }
  // this body against the table of base_i (same cell): the own cell starts delta later and is delta shorter
  const int delta = static_cast<int>(i - base_i);
  const int total_cand = total_base - delta;
  // flattened index tb0 + lane -> body index. Lane r keeps the prefix of the cell of rank r; the cells that start inside
  // a batch of 32 consecutive indices are marked in one warp-wide OR, so a lane finds its cell with two population counts
  // instead of a binary search through shared memory. (An emptied own cell shares prefix 0 with rank 1 and is skipped.)
  const int rpre = (lane == 0) ? 0 : ((lane < nvalid) ? rpre_base - delta : 0x7fffffff);
  auto candidate = [&](int tb0) {
    const uint32_t below = __ballot_sync(0xffffffffu, rpre <= tb0);  // rank 0 starts at 0: never empty
    const int rel = rpre - tb0;
    const uint32_t starts = __reduce_or_sync(0xffffffffu, (rel > 0 && rel < 32) ? (1u << rel) : 0u);
    const int lo = __popc(below) - 1 + __popc(starts & ((2u << lane) - 1u));
    const int pre_lo = (lo == 0) ? 0 : wpre[lo] - delta;
    const int beg_lo = (lo == 0) ? static_cast<int>(i) + 1 : wbeg[lo];
    return beg_lo + (tb0 + lane - pre_lo);
  };

  uint32_t total = 0;
  // With the shape filter the survivors of the sphere test (about 1 in 10) are queued per warp and the expensive
  // shape distance runs on dense batches of 32, in queue order (= ascending j), instead of on a few divergent lanes.
  int qn = 0;
  auto emit = [&](bool keep, int j) {
    const uint32_t m = __ballot_sync(0xffffffffu, keep);
    if (keep) {
      const int pos = static_cast<int>(total) + __popc(m & lt_mask);
      if (FILL) pairs[out + pos] = make_int2(static_cast<int>(i), j);
      else if (pos < slot_cap) slots[i * slot_cap + pos] = j;
    }
    total += __popc(m);
  };
  auto evaluate = [&](int take) {
    const int j = (lane < take) ? q[lane] : -1;
    bool keep = false;
    if (j >= 0) {
      // evaluate in the caller's (i<j) order so that the decision is bit-identical to the reference order
      const D4 b0i = ldg256(b0 + i), b1i = ldg256(b1 + i);
      const uint32_t pi = __ldg(perm + i);
      const D4 b0j = ldg256(b0 + j), b1j = ldg256(b1 + j);
      // operands are put in caller order first, so that the (long) distance code runs once per batch and not once per
      // order under divergence
      const bool fwd = pi < perm[j];
      const D4 a0 = fwd ? b0i : b0j, a1 = fwd ? b1i : b1j, c0 = fwd ? b0j : b0i, c1 = fwd ? b1j : b1i;
      // BOX: the survivor of the box test has not seen the exact candidate predicate yet
      if (!BOX || is_candidate(si, ldg256(sr + j))) {
        const PairGeom g = PER ? pair_geometry<true>(a0, a1, unwrap_to(P, a0, c0), c1) : pair_geometry<true>(a0, a1, c0, c1);
        keep = g.sep <= cutoff;
      }
    }
    emit(keep, j);
    // the qn - take < 64 remaining entries move to the front
    const int r0 = (lane + take < qn) ? q[lane + take] : 0;
    const int r1 = (lane + 32 + take < qn) ? q[lane + 32 + take] : 0;
    __syncwarp();
    q[lane] = r0;
    q[lane + 32] = r1;
    __syncwarp();
    qn -= take;
  };
  for (int t0 = 0; t0 < total_cand; t0 += 64) {
    const int ta = t0 + lane, tb = t0 + 32 + lane;
    const int ca = candidate(t0), cb = candidate(t0 + 32);  // warp-wide: called by all lanes
    const int ja = (ta < total_cand) ? ca : -1;
    const int jb = (tb < total_cand) ? cb : -1;
    bool ka = false, kb = false;
    if (BOX) {
      if (ja >= 0) ka = boxes_overlap(bi, load_box(box, ja));
      if (jb >= 0) kb = boxes_overlap(bi, load_box(box, jb));
    } else {
      D4 sa{0, 0, 0, 0}, sb{0, 0, 0, 0};
      if (ja >= 0) sa = ldg256(sr + ja);
      if (jb >= 0) sb = ldg256(sr + jb);
      ka = (ja >= 0) && (PER ? is_candidate_periodic(P, si, sa) : is_candidate(si, sa));
      kb = (jb >= 0) && (PER ? is_candidate_periodic(P, si, sb) : is_candidate(si, sb));
    }
    ka = ka && (own_i || own[ja]);
    kb = kb && (own_i || own[jb]);
    if (FILTER) {
      const uint32_t ma = __ballot_sync(0xffffffffu, ka), mb = __ballot_sync(0xffffffffu, kb);
      if (ka) q[qn + __popc(ma & lt_mask)] = ja;
      if (kb) q[qn + __popc(ma) + __popc(mb & lt_mask)] = jb;
      qn += __popc(ma) + __popc(mb);
      __syncwarp();
      const bool last = t0 + 64 >= total_cand;  // the tail of the queue is flushed by the same (single) call site
      while (qn >= 32 || (last && qn > 0)) evaluate(min(qn, 32));
    } else {
      emit(ka, ja);
      emit(kb, jb);
    }
  }
  if (!FILL && lane == 0) counts[i] = total;
  }
}


// ---- tiled form of the same search ----------------------------------------------------------------------------------
// The walk above redoes the flattened-index -> body look-up and the record loads of a candidate for every body that
// meets it (about 4 warp instructions per candidate and body). Bodies of one cell share their candidates, so here the
// loops are interchanged: a warp takes 32 consecutive bodies, splits them into runs of equal cell key, and for every
// run walks the run's candidates ONCE in tiles of 32 (lane = candidate j: one look-up, one record load per tile); inside
// a tile it passes over the run's bodies, whose records are broadcast from shared memory, so that a body costs one
// overlap test, one ballot and the queue append per tile (about 0.5 warp instructions per candidate and body).
// For a fixed body the hits still arrive in ascending j (tiles ascend, lanes ascend inside a tile), so pair lists
// come out in the same (i, j) order as from the walk. The own cell's candidates start behind the FIRST body of the run;
// a later body of the run skips those that do not follow it (j > i).
// FILTER: hits are queued as (body, j) and the exact tests run on dense batches of 32 from one call site; the position
// of a kept pair in its body's list = the body's running count (shared memory) + its rank among the kept lanes of the
// same body in the batch (__match_any_sync).
constexpr int kTileBodies = 32;  // consecutive bodies per warp
#ifndef MB200_TILED_MIN_CTAS
#define MB200_TILED_MIN_CTAS 5  // 48 registers, 40 warps per SM: rebuild 1.59 / 1.50 / 1.47 ms at 3 / 4 / 5 (10^6 rods)
#endif

// Per-warp working set in shared memory (the inner loop keeps only the candidate's record and a few counters in
// registers; everything else lives here).
struct TileWarp {
  D4 rec[kTileBodies];       // per body of the warp: its box (BOX) or its search sphere
  int pre[28], beg[28];      // the run's neighbour cells by ascending key: candidates in front of the cell, first body
  int cnt[kTileBodies];      // FILTER: kept pairs so far, per body of the warp
  int qj[64];                // FILTER: queued (body, j) awaiting the exact tests: a ring of 64
  uint8_t qr[64];
};
// What the exact tests need, once per CTA (the out-of-line call takes one pointer instead of a dozen arguments).
struct TileFilterArgs {
  const D4 *sr, *b0, *b1;
  const uint32_t* perm;
  const int64_t* rowptr;
  int2* pairs;
  int32_t* slots;
  double cutoff;
  int slot_cap;
  Periodic P;
};

// Exact decisions for `take` queued pairs from position qh of the ring (shape filter): the candidate predicate where
// the pre-reject was only the boxes, then the shape distance in the caller's (i<j) order; kept pairs go to the body's
// next positions (its running count + the rank among the kept lanes of the same body). Out of line: the distance code
// is ~3000 instructions and there are two call sites; its registers are its own.
template <bool FILL, bool PER, bool BOX>
__device__ __noinline__ void filter_batch(TileWarp* w, const TileFilterArgs* fa, int i0, int qh, int take) {
  const int lane = threadIdx.x & 31;
  int j = -1, r = 32 + lane;  // idle lanes: distinct body tags
  if (lane < take) {
    const int slot = (qh + lane) & 63;
    j = w->qj[slot];
    r = w->qr[slot];
  }
  const int i = i0 + (r & 31);
  bool keep = false;
  if (j >= 0 && (!BOX || is_candidate(ldg256(fa->sr + i), ldg256(fa->sr + j)))) {
    const D4 b0i = ldg256(fa->b0 + i), b1i = ldg256(fa->b1 + i);
    const D4 b0j = ldg256(fa->b0 + j), b1j = ldg256(fa->b1 + j);
    const bool fwd = __ldg(fa->perm + i) < __ldg(fa->perm + j);
    const D4 a0 = fwd ? b0i : b0j, a1 = fwd ? b1i : b1j, c0 = fwd ? b0j : b0i, c1 = fwd ? b1j : b1i;
    D4 c0u = c0;
    if constexpr (PER) c0u = unwrap_to(fa->P, a0, c0);
    keep = pair_geometry<true>(a0, a1, c0u, c1).sep <= fa->cutoff;
  }
  const uint32_t grp = __match_any_sync(0xffffffffu, r);
  const uint32_t km = __ballot_sync(0xffffffffu, keep) & grp;
  const int pos = w->cnt[r & 31] + __popc(km & ((1u << lane) - 1u));
  if (keep) {
    if (FILL) fa->pairs[fa->rowptr[i] + pos] = make_int2(i, j);
    else if (pos < fa->slot_cap) fa->slots[static_cast<int64_t>(i) * fa->slot_cap + pos] = j;
  }
  __syncwarp();
  if (keep && (km >> lane) == 1u) w->cnt[r] = pos + 1;  // the last kept lane of the body
  __syncwarp();
}

template <bool FILL, bool FILTER, bool PER, bool BOX>
__global__ void __launch_bounds__(256, MB200_TILED_MIN_CTAS)
    find_pairs_tiled_kernel(int64_t n, const D4* __restrict__ sr, const float4* __restrict__ box,
                            const D4* __restrict__ b0, const D4* __restrict__ b1, const uint32_t* __restrict__ perm,
                            const int2* __restrict__ cell_range, GridParams gp, double cutoff,
                            uint32_t* __restrict__ counts, const int64_t* __restrict__ rowptr, int2* __restrict__ pairs,
                            int32_t* __restrict__ slots, int slot_cap, const uint8_t* __restrict__ own, Periodic P) {
  __shared__ __align__(32) TileWarp s_warp[8];
  __shared__ TileFilterArgs s_fa;
  if (FILTER) {
    if (threadIdx.x == 0) {
      s_fa.sr = sr, s_fa.b0 = b0, s_fa.b1 = b1, s_fa.perm = perm, s_fa.rowptr = rowptr, s_fa.pairs = pairs;
      s_fa.slots = slots, s_fa.cutoff = cutoff, s_fa.slot_cap = slot_cap, s_fa.P = P;
    }
    __syncthreads();
  }
  const int lane = threadIdx.x & 31;
  TileWarp* w = &s_warp[threadIdx.x >> 5];
  const int64_t i_begin = ((static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5) * kTileBodies;
  if (i_begin >= n) return;
  const int i0 = static_cast<int>(i_begin);  // body indices fit an int (pair lists hold ints)
  const int nb = static_cast<int>(min(static_cast<int64_t>(kTileBodies), n - i_begin));  // bodies of this warp
  const uint32_t lt_mask = (1u << lane) - 1u;

  // the warp's bodies: search sphere -> cell; records into shared memory
  const bool have = lane < nb;
  uint32_t mkey = 0xffffffffu, mcell = 0;  // Morton key and packed cell coordinates of this lane's body
  {
    D4 s_mine{0, 0, 0, 0};
    if (have) {
      s_mine = ldg256(sr + i0 + lane);
      const int cx = cell_coord(s_mine.x, gp, 0), cy = cell_coord(s_mine.y, gp, 1), cz = cell_coord(s_mine.z, gp, 2);
      mkey = morton3(cx, cy, cz);
      mcell = static_cast<uint32_t>(cx) | (static_cast<uint32_t>(cy) << 10) | (static_cast<uint32_t>(cz) << 20);
      w->rec[lane] = BOX ? ldg256(reinterpret_cast<const D4*>(box) + i0 + lane) : s_mine;
    }
  }
  w->cnt[lane] = 0;
  // multi-GPU: bit r set = body r of the warp is owned (links between two ghosts belong to other ranks)
  const uint32_t own_bits = own ? __ballot_sync(0xffffffffu, have && own[i0 + lane] != 0) : 0xffffffffu;
  // FILL: bodies whose pairs the count pass cached are copied; the others are found again
  uint32_t todo = __ballot_sync(0xffffffffu, have);
  int64_t my_out = 0;
  if (FILL) {
    int cnt = 0;
    if (have) {
      my_out = rowptr[i0 + lane];
      cnt = static_cast<int>(rowptr[i0 + lane + 1] - my_out);
    }
    todo = __ballot_sync(0xffffffffu, have && cnt > slot_cap);
    for (int r = 0; r < nb; ++r) {
      const int c = __shfl_sync(0xffffffffu, cnt, r);
      if (c > slot_cap) continue;
      const int64_t o = __shfl_sync(0xffffffffu, my_out, r);
      const int64_t i = i_begin + r;
      for (int k = lane; k < c; k += 32) pairs[o + k] = make_int2(static_cast<int>(i), slots[i * slot_cap + k]);
    }
    if (todo == 0) return;
  }
  uint32_t heads;
  {
    const uint32_t prev_key = __shfl_up_sync(0xffffffffu, mkey, 1);
    heads = __ballot_sync(0xffffffffu, have && (lane == 0 || prev_key != mkey));
  }
  __syncwarp();

  int qn = 0, qh = 0;  // FILTER: queue length and head
  int mycnt = 0;       // !FILTER: kept pairs of the body this lane stands for

  while (heads) {
    const int ra = __ffs(heads) - 1;
    heads &= heads - 1;
    const int rb = heads ? __ffs(heads) - 1 : nb;  // run = bodies [ra, rb) of the warp: one cell
    // bodies of the run that are searched (FILL: only those the count pass could not cache)
    const uint32_t run_bits = ((rb >= 32) ? 0xffffffffu : ((1u << rb) - 1u)) & ~((1u << ra) - 1u) & todo;
    if (run_bits == 0) continue;
    const int base_i = i0 + ra;
    const uint32_t my_key = __shfl_sync(0xffffffffu, mkey, ra);
    int total_cand, nvalid;
    {
      // table of the neighbour cells with key >= own key: lane l < 27 owns cell l (as in the walk above)
      const uint32_t cc = __shfl_sync(0xffffffffu, mcell, ra);
      const int cx = cc & 1023u, cy = (cc >> 10) & 1023u, cz = cc >> 20;
      uint32_t nkey = 0xffffffffu;
      int beg = 0, len = 0;
      if (lane < 27) {
        int x = cx + (lane % 3) - 1, y = cy + ((lane / 3) % 3) - 1, z = cz + (lane / 9) - 1;
        bool inside = true;
        if (PER) {
          const int off[3] = {(lane % 3) - 1, ((lane / 3) % 3) - 1, (lane / 9) - 1};
          int* xyz[3] = {&x, &y, &z};
#pragma unroll
          for (int d = 0; d < 3; ++d) {
            if (gp.per[d]) {
              inside = inside && (off[d] == 0 || (off[d] < 0 ? gp.nc[d] >= 2 : gp.nc[d] >= 3));
              *xyz[d] = (*xyz[d] + gp.nc[d]) % gp.nc[d];
            }
          }
        }
        if (inside && x >= 0 && y >= 0 && z >= 0 && x < gp.nc[0] && y < gp.nc[1] && z < gp.nc[2]) {
          const uint32_t k = morton3(x, y, z);
          if (k >= my_key) {
            const int2 rg = __ldg(cell_range + k);
            beg = (k == my_key) ? base_i + 1 : rg.x;
            len = rg.y - beg;
            if (len > 0 || k == my_key) nkey = k;
            if (len < 0) len = 0;
          }
        }
      }
      const bool valid = nkey != 0xffffffffu;
      int rank = 0, pre = 0;
#pragma unroll
      for (int m = 0; m < 27; ++m) {
        const uint32_t km = __shfl_sync(0xffffffffu, nkey, m);
        const int lm = __shfl_sync(0xffffffffu, len, m);
        if (km < nkey) {
          rank += 1;
          pre += lm;
        }
      }
      nvalid = __popc(__ballot_sync(0xffffffffu, valid));
      __syncwarp();  // the previous run's table is no longer read
      if (valid) {
        w->pre[rank] = pre;
        w->beg[rank] = beg;
      }
      total_cand = valid ? pre + len : 0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) total_cand = max(total_cand, __shfl_xor_sync(0xffffffffu, total_cand, o));
      __syncwarp();
    }
    const int rpre = (lane == 0) ? 0 : ((lane < nvalid) ? w->pre[lane] : 0x7fffffff);

    for (int t0 = 0; t0 < total_cand; t0 += 32) {
      uint32_t pending = run_bits;  // bodies of the run still to meet this tile
      do {
        // The tile: flattened index t0 + lane -> body j, its record, and the bodies of the warp it may pair with (those
        // in front of it: only the own cell's candidates can fail i < j; for a ghost candidate, the owned ones).
        // Computed again after a batch of exact tests, so that none of it has to survive the call.
        int j = -1;
        {
          const uint32_t below = __ballot_sync(0xffffffffu, rpre <= t0);
          const int rel = rpre - t0;
          const uint32_t starts = __reduce_or_sync(0xffffffffu, (rel > 0 && rel < 32) ? (1u << rel) : 0u);
          const int lo = __popc(below) - 1 + __popc(starts & ((2u << lane) - 1u));
          const int pre_lo = (lo == 0) ? 0 : w->pre[lo];
          const int beg_lo = (lo == 0) ? base_i + 1 : w->beg[lo];
          if (t0 + lane < total_cand) j = beg_lo + (t0 + lane - pre_lo);
        }
        Box6f bj{};
        D4 sj{0, 0, 0, 0};
        uint32_t allow = 0;
        if (j >= 0) {
          if (BOX) bj = load_box(box, j);
          else sj = ldg256(sr + j);
          const int ahead = j - i0;  // bodies [0, ahead) of the warp precede j
          allow = (ahead >= 32) ? 0xffffffffu : ((1u << ahead) - 1u);
          if (own && !own[j]) allow &= own_bits;
        }
        allow &= pending;
        // one body of the warp against the tile: overlap test, ballot, hits appended in lane order (ascending j)
        auto meet = [&](int r, bool hit) -> bool {  // true: the queue holds a full batch
          if (BOX) {
            const D4 rec = w->rec[r];  // broadcast
            Box6f bi;
            bi.lx = __int_as_float(__double2loint(rec.x)), bi.ly = __int_as_float(__double2hiint(rec.x));
            bi.lz = __int_as_float(__double2loint(rec.y)), bi.hx = __int_as_float(__double2hiint(rec.y));
            bi.hy = __int_as_float(__double2loint(rec.z)), bi.hz = __int_as_float(__double2hiint(rec.z));
            hit = hit && boxes_overlap(bi, bj);
          } else {
            const D4 si = w->rec[r];
            hit = hit && (PER ? is_candidate_periodic(P, si, sj) : is_candidate(si, sj));
          }
          const uint32_t m = __ballot_sync(0xffffffffu, hit);
          if (m == 0) return false;
          if (FILTER) {
            if (hit) {
              const int slot = (qh + qn + __popc(m & lt_mask)) & 63;
              w->qj[slot] = j;
              w->qr[slot] = static_cast<uint8_t>(r);
            }
            qn += __popc(m);
            return qn >= 32;
          } else {
            const int base = __shfl_sync(0xffffffffu, mycnt, r);
            const int64_t out_r = FILL ? __shfl_sync(0xffffffffu, my_out, r) : 0;
            if (hit) {
              const int pos = base + __popc(m & lt_mask);
              if (FILL) pairs[out_r + pos] = make_int2(i0 + r, j);
              else if (pos < slot_cap) slots[static_cast<int64_t>(i0 + r) * slot_cap + pos] = j;
            }
            if (lane == r) mycnt += __popc(m);
            return false;
          }
        };
        // Past the own cell every candidate may pair with every pending body: when those are consecutive bodies
        // (always, except in a fill pass that redoes some bodies of a run) the loop is a plain count with no
        // per-lane mask to test. Lanes past the end of the list carry a record that meets nothing.
        const int r_lo = __ffs(pending) - 1, r_hi = 32 - __clz(pending);
        const bool consecutive = pending == ((r_hi >= 32 ? 0xffffffffu : ((1u << r_hi) - 1u)) & ~((1u << r_lo) - 1u));
        const bool plain = consecutive && __all_sync(0xffffffffu, j < 0 || allow == pending);
        const uint32_t before = pending;
        pending = 0;
        if (plain) {
          if (j < 0) {
            if (BOX) bj.lx = __int_as_float(0x7f800000);  // +inf: below no upper bound
            else sj.w = __longlong_as_double(0x7ff8000000000000ll);  // NaN: d2 <= s * s is false
          }
          for (int r = r_lo; r < r_hi; ++r) {
            if (meet(r, true)) {  // a full batch: leave the tile, run the exact tests, come back for the other bodies
              pending = (r + 1 >= 32) ? 0u : (before & ~((2u << r) - 1u));
              break;
            }
          }
        } else {
          uint32_t rbits = __reduce_or_sync(0xffffffffu, allow);  // bodies some candidate of the tile may pair with
          while (rbits) {
            const int r = __ffs(rbits) - 1;
            rbits &= rbits - 1;
            if (meet(r, (allow >> r) & 1u)) {
              pending = rbits;
              break;
            }
          }
        }
        if (FILTER && qn >= 32) {
          __syncwarp();
          filter_batch<FILL, PER, BOX>(w, &s_fa, i0, qh, 32);
          qh = (qh + 32) & 63;
          qn -= 32;
        }
      } while (pending);
    }
  }
  if (FILTER) {
    __syncwarp();
    if (qn > 0) filter_batch<FILL, PER, BOX>(w, &s_fa, i0, qh, qn);
    if (!FILL && have) counts[i0 + lane] = static_cast<uint32_t>(w->cnt[lane]);
  } else {
    if (!FILL && have) counts[i0 + lane] = static_cast<uint32_t>(mycnt);
  }
}

}  // namespace

namespace {

// any |c - c_at_build| > 0.5 * buffer (GenNeighborLinkers.hpp:603-615); with the shape filter the two centre-line
// endpoints are tracked instead of the centre (a rotation moves a rod's surface without moving its bounding sphere).
__global__ void __launch_bounds__(256)
    moved_kernel(int64_t n, const double* __restrict__ center, const double* __restrict__ ref_center,
                 const double* __restrict__ quat, const double* __restrict__ length, const int32_t* __restrict__ kind,
                 const D4* __restrict__ ref_b1, double thresh, int32_t* __restrict__ flag, Periodic P) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const V3 c = ld3(center, i), c0 = ld3(ref_center, i);
  bool moved;
  if (ref_b1 && kind && kind[i] == MB200_SPHEROCYLINDER) {
    const D4 r = ldg256(ref_b1 + i);
    const V3 h0{r.x, r.y, r.z};
    const V3 h = (0.5 * length[i]) * qrot(ldq(quat, i), V3{0.0, 0.0, 1.0});
    // periodic box: displacement = metric.sep(point at build time, point now) (a wrapped centre has not moved)
    const V3 da = P.any ? periodic_sep(P, c0 - h0, c - h) : (c - h) - (c0 - h0);
    const V3 db = P.any ? periodic_sep(P, c0 + h0, c + h) : (c + h) - (c0 + h0);
    moved = sqrt(da.x * da.x + da.y * da.y + da.z * da.z) > thresh || sqrt(db.x * db.x + db.y * db.y + db.z * db.z) > thresh;
  } else {
    const V3 dd = P.any ? periodic_sep(P, c0, c) : V3{c.x - c0.x, c.y - c0.y, c.z - c0.z};
    moved = sqrt(dd.x * dd.x + dd.y * dd.y + dd.z * dd.z) > thresh;
  }
  if (moved) *flag = 1;
}

__global__ void __launch_bounds__(256) canonical_keys_kernel(int64_t npairs, const int2* __restrict__ pairs,
                                                             const uint32_t* __restrict__ perm, int shift,
                                                             uint64_t* __restrict__ keys) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= npairs) return;
  const int2 p = pairs[c];
  const uint32_t a = perm[p.x], b = perm[p.y];
  const uint64_t lo = a < b ? a : b, hi = a < b ? b : a;
  keys[c] = (lo << shift) | hi;
}
__global__ void __launch_bounds__(256) canonical_unpack_kernel(int64_t npairs, const uint64_t* __restrict__ keys,
                                                               int shift, int2* __restrict__ out) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= npairs) return;
  const uint64_t k = keys[c];
  out[c] = make_int2(static_cast<int>(k >> shift), static_cast<int>(k & ((uint64_t(1) << shift) - 1)));
}

int bits_for(int64_t n) {
  int bits = 1;
  while ((int64_t(1) << bits) < n) ++bits;
  return bits;
}

}  // namespace

__global__ void __launch_bounds__(256) gather_u8_kernel(int64_t n, const uint32_t* __restrict__ perm,
                                                        const uint8_t* __restrict__ src, uint8_t* __restrict__ dst) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k < n) dst[k] = src[perm[k]];
}

int build_links(cudaStream_t st, mb200_neighbor_links* h, const mb200_bodies* b, const double* explicit_S,
                const uint8_t* owned) {
  const int64_t n = b->n;
  h->n = n;
  h->npairs = 0;
  h->rods = b->kind != nullptr;
  h->built = true;
  const bool filter = !explicit_S && h->filter == MB200_FILTER_SHAPE_DISTANCE && h->rods;
  MB200_TRY(h->perm.reserve(n > 0 ? n : 1));
  MB200_TRY(h->ref_center.reserve(3 * n + 1));
  if (n == 0) return build_incidence(st, 0, 0, nullptr, &h->inc);
  MB200_CUDA(cudaMemcpyAsync(h->ref_center.p, b->center, 3 * n * sizeof(double), cudaMemcpyDeviceToDevice, st));

  // 1. search spheres + bounds
  Scratch<D4> sr, sr_s, b0c, b1c, b0s, b1s;
  Scratch<double> partials;
  const int bgrid = grid_for(n, 256, kSMs * 2);
  MB200_TRY(sr.alloc(n, st));
  MB200_TRY(sr_s.alloc(n, st));
  MB200_TRY(partials.alloc(7 * static_cast<size_t>(bgrid), st));
  search_spheres_kernel<<<bgrid, 256, 0, st>>>(n, b->center, b->radius, b->length, b->kind, explicit_S, h->buffer,
                                               sr.p, partials.p);
  MB200_LAUNCHED();
  const size_t hp_count = 7 * static_cast<size_t>(bgrid);
  double* hp = static_cast<double*>(pinned_scratch(hp_count * sizeof(double)));
  if (!hp) return MB200_ERR_CUDA;
  MB200_CUDA(cudaMemcpyAsync(hp, partials.p, hp_count * sizeof(double), cudaMemcpyDeviceToHost, st));
  MB200_CUDA(cudaStreamSynchronize(st));
  GridParams gp;
  double lo[3] = {hp[0], hp[1], hp[2]}, hi[3] = {hp[3], hp[4], hp[5]}, smax = hp[6];
  for (int k = 1; k < bgrid; ++k) {
    for (int d = 0; d < 3; ++d) {
      lo[d] = std::min(lo[d], hp[7 * k + d]);
      hi[d] = std::max(hi[d], hp[7 * k + 3 + d]);
    }
    smax = std::max(smax, hp[7 * k + 6]);
  }
  for (int d = 0; d < 3; ++d) MB200_REQUIRE(std::isfinite(lo[d]) && std::isfinite(hi[d]), "non-finite body centre");
  MB200_REQUIRE(std::isfinite(smax), "non-finite search radius");
  // cell edge >= largest possible centre distance of a candidate pair (2 S_max), with a relative margin so that the
  // floor() binning can never separate a candidate pair by more than one cell
  double edge = 2.0 * smax * (1.0 + 1e-9);
  if (!(edge > 0.0)) edge = 1.0;
  const Periodic P = h->periodic;
  for (;;) {
    bool ok = true;
    for (int d = 0; d < 3; ++d) {
      if (P.on[d]) continue;
      const double cells = std::floor((hi[d] - lo[d]) / edge) + 1.0;
      ok = ok && cells <= double(1 << kMaxBitsPerAxis);
      gp.nc[d] = ok ? static_cast<int>(cells) : 0;
    }
    if (ok) break;
    edge *= 2.0;  // larger cells stay correct, only slower
  }
  for (int d = 0; d < 3; ++d) {
    gp.per[d] = P.on[d];
    gp.inv[d] = P.inv[d];
    if (P.on[d]) {  // whole cells tile the period: nc = floor(L / edge) cells of edge L / nc >= edge
      const double cells = std::floor(P.L[d] / edge);
      gp.nc[d] = static_cast<int>(std::min(std::max(cells, 1.0), double(1 << kMaxBitsPerAxis)));
      gp.h[d] = P.L[d] / gp.nc[d];
      gp.lo[d] = 0.0;
    } else {
      gp.h[d] = edge;
      gp.lo[d] = lo[d];
    }
  }
  gp.bits = bits_for(std::max(gp.nc[0], std::max(gp.nc[1], gp.nc[2])));
  const int key_bits = 3 * gp.bits;
  const int64_t table = int64_t(1) << key_bits;

  // 2. Morton keys, sort, gather into Morton order
  Scratch<uint32_t> keys, keys_alt, vals_alt;
  MB200_TRY(keys.alloc(n, st));
  MB200_TRY(keys_alt.alloc(n, st));
  MB200_TRY(vals_alt.alloc(n, st));
  cell_keys_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, sr.p, gp, keys.p, h->perm.p);
  MB200_LAUNCHED();
  MB200_TRY(radix_sort_u32(st, n, keys.p, keys_alt.p, h->perm.p, vals_alt.p, key_bits));
  gather_records_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, h->perm.p, sr.p, sr_s.p);
  MB200_LAUNCHED();
  Scratch<float4> boxes;  // fp32 AABB pre-reject records (shape filter, free space)
  static const bool box_off = getenv("MB200_NO_BOX_PREREJECT") != nullptr;
  const bool use_box = filter && !h->periodic.any && !box_off;
  Scratch<uint8_t> own_s;  // ownership flags in Morton order (multi-GPU only)
  if (owned) {
    MB200_TRY(own_s.alloc(n, st));
    gather_u8_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, h->perm.p, owned, own_s.p);
    MB200_LAUNCHED();
  }
  if (filter) {
    MB200_TRY(b0s.alloc(n, st));
    MB200_TRY(b1s.alloc(n, st));
    MB200_TRY(pack_bodies(st, b, reinterpret_cast<const int32_t*>(h->perm.p), b0s.p, b1s.p));
    MB200_TRY(h->ref_b1.reserve(n));
    MB200_TRY(b0c.alloc(n, st));
    MB200_TRY(pack_bodies(st, b, nullptr, b0c.p, h->ref_b1.p));
    if (use_box) {
      MB200_TRY(boxes.alloc(2 * static_cast<size_t>(n), st));
      box_records_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, b0s.p, b1s.p, h->buffer, boxes.p);
      MB200_LAUNCHED();
    }
  }

  // 3. cell table
  Scratch<int2> cell_range;
  MB200_TRY(cell_range.alloc(table, st));
  MB200_CUDA(cudaMemsetAsync(cell_range.p, 0, table * sizeof(int2), st));
  cell_ranges_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, keys.p, cell_range.p);
  MB200_LAUNCHED();

  // 4. count -> scan -> fill
  Scratch<uint32_t> counts;
  Scratch<int64_t> rowptr;
  Scratch<int32_t> slots;
  const int slot_cap = filter ? 16 : 48;  // kept pairs cached per body by the count pass (more => recomputed by the fill pass)
  MB200_TRY(slots.alloc(static_cast<size_t>(n) * slot_cap, st));
  MB200_TRY(counts.alloc(n, st));
  MB200_TRY(rowptr.alloc(n + 1, st));
  const double cutoff = 2.0 * h->buffer;
  // MB200_FIND=walk selects the per-body walk (the tiled form's predecessor, kept for A/B runs; same pair lists)
  static const bool walk = getenv("MB200_FIND") && std::string(getenv("MB200_FIND")) == "walk";
  const int pgrid = walk ? grid_for(((n + kFindGroup - 1) / kFindGroup) * 32, 256)
                         : grid_for(((n + kTileBodies - 1) / kTileBodies) * 32, 256);
  auto launch_find = [&](auto fill_tag, const int64_t* rp, int2* out) {
    constexpr bool FILL = decltype(fill_tag)::value;
    uint32_t* cnt = FILL ? nullptr : counts.p;
#define MB200_FIND_ARGS(FILTER)                                                                                      \
  n, sr_s.p, boxes.p, FILTER ? b0s.p : nullptr, FILTER ? b1s.p : nullptr, FILTER ? h->perm.p : nullptr,              \
      cell_range.p, gp, cutoff, cnt, rp, out, slots.p, slot_cap, own_s.p, P
#define MB200_FIND(FILTER, PER, BOX)                                                                                 \
  do {                                                                                                               \
    if (walk)                                                                                                        \
      find_pairs_kernel<FILL, FILTER, PER, BOX><<<pgrid, 256, 0, st>>>(MB200_FIND_ARGS(FILTER));                     \
    else                                                                                                             \
      find_pairs_tiled_kernel<FILL, FILTER, PER, BOX><<<pgrid, 256, 0, st>>>(MB200_FIND_ARGS(FILTER));               \
  } while (0)
    if (filter) {
      if (P.any) MB200_FIND(true, true, false);
      else if (use_box) MB200_FIND(true, false, true);
      else MB200_FIND(true, false, false);
    } else {
      if (P.any) MB200_FIND(false, true, false); else MB200_FIND(false, false, false);
    }
#undef MB200_FIND
#undef MB200_FIND_ARGS
  };
  launch_find(std::false_type{}, nullptr, nullptr);
  MB200_LAUNCHED();
  MB200_TRY(exclusive_scan_u32_to_i64(st, n, counts.p, rowptr.p));
  int64_t* total_h = static_cast<int64_t*>(pinned_scratch(sizeof(int64_t)));
  if (!total_h) return MB200_ERR_CUDA;
  MB200_CUDA(cudaMemcpyAsync(total_h, rowptr.p + n, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  MB200_CUDA(cudaStreamSynchronize(st));
  const int64_t total = *total_h;
  if (total >= (int64_t(1) << 30)) {
    set_error("pair list exceeds 2^30 entries");
    return MB200_ERR_CAPACITY;
  }
  h->npairs = total;
  MB200_TRY(h->pairs.reserve(2 * total + 2));
  if (total > 0) {
    int2* out = reinterpret_cast<int2*>(h->pairs.p);
    launch_find(std::true_type{}, rowptr.p, out);
    MB200_LAUNCHED();
  }
  // 5. body-major incidence lists for the contact operator (the list is sorted by (i, j) and its rows are known)
  return build_incidence_sorted(st, n, total, h->pairs.p, rowptr.p, &h->inc);
}

// Device flag (h->flag) = 1 when some body moved more than half the buffer since the last build. No host sync.
int links_moved_flag(cudaStream_t st, mb200_neighbor_links* h, const mb200_bodies* b, int64_t ref_offset) {
  MB200_TRY(h->flag.reserve(1));
  MB200_CUDA(cudaMemsetAsync(h->flag.p, 0, sizeof(int32_t), st));
  if (b->n == 0) return MB200_OK;
  MB200_REQUIRE(ref_offset >= 0 && ref_offset + b->n <= h->n, "bodies outside the set the links were built from");
  const bool shape = h->filter == MB200_FILTER_SHAPE_DISTANCE && h->rods;
  moved_kernel<<<grid_for(b->n, 256), 256, 0, st>>>(b->n, b->center, h->ref_center.p + 3 * ref_offset, b->quat, b->length,
                                                   b->kind, shape ? h->ref_b1.p + ref_offset : nullptr, 0.5 * h->buffer,
                                                   h->flag.p, h->periodic);
  MB200_LAUNCHED();
  return MB200_OK;
}

int links_generate(cudaStream_t st, mb200_neighbor_links* h, const mb200_bodies* b, int force, int* rebuilt_h) {
  bool rebuild = force || !h->built || h->n != b->n;
  if (!rebuild && b->n > 0) {
    MB200_TRY(h->flag.reserve(1));
    StageScope ps(MB200_STAGE_TRIGGER, st);
    MB200_CUDA(cudaMemsetAsync(h->flag.p, 0, sizeof(int32_t), st));
    const bool shape = h->filter == MB200_FILTER_SHAPE_DISTANCE && h->rods;
    moved_kernel<<<grid_for(b->n, 256), 256, 0, st>>>(b->n, b->center, h->ref_center.p, b->quat, b->length, b->kind,
                                                     shape ? h->ref_b1.p : nullptr, 0.5 * h->buffer, h->flag.p,
                                                   h->periodic);
    MB200_LAUNCHED();
    int32_t* moved = static_cast<int32_t*>(pinned_scratch(sizeof(int32_t)));
    if (!moved) return MB200_ERR_CUDA;
    MB200_CUDA(cudaMemcpyAsync(moved, h->flag.p, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    MB200_CUDA(cudaStreamSynchronize(st));
    rebuild = *moved != 0;
  }
  if (rebuild) {
    StageScope ps(MB200_STAGE_BROADPHASE, st);
    MB200_TRY(build_links(st, h, b, nullptr, nullptr));
  }
  if (rebuilt_h) *rebuilt_h = rebuild ? 1 : 0;
  return MB200_OK;
}

}  // namespace mb200

namespace mb200 {

int canonical_pairs(cudaStream_t st, mb200_neighbor_links* h, int32_t* out) {
  const int64_t C = h->npairs;
  if (C == 0) return MB200_OK;
  const int shift = bits_for(h->n);
  Scratch<uint64_t> keys, keys_alt;
  MB200_TRY(keys.alloc(C, st));
  MB200_TRY(keys_alt.alloc(C, st));
  canonical_keys_kernel<<<grid_for(C, 256), 256, 0, st>>>(C, reinterpret_cast<const int2*>(h->pairs.p), h->perm.p,
                                                         shift, keys.p);
  MB200_LAUNCHED();
  MB200_TRY(radix_sort_u64(st, C, keys.p, keys_alt.p, nullptr, nullptr, 2 * shift));
  canonical_unpack_kernel<<<grid_for(C, 256), 256, 0, st>>>(C, keys.p, shift, reinterpret_cast<int2*>(out));
  MB200_LAUNCHED();
  return MB200_OK;
}

}  // namespace mb200
