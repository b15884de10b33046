// Multi-GPU collision step (one process per GPU). See include/mundy_b200.h "Multi-GPU" for the decomposition.
//
// Every rank passes ONLY the bodies it owns; nothing is replicated and nothing scales with the world size:
// Per rebuild (decided by an all-reduced moved flag over the OWNED bodies, so that every rank rebuilds in the same step):
//   owned bounding box + largest search radius -> all-gathered (7 doubles per rank) -> for every peer, the owned bodies
//   whose search sphere reaches that peer's box (periodic boxes: nearest image) are packed and sent to it (NCCL send/recv
//   between ranks that have something for each other) -> local set = the received ghosts and the owned bodies in GLOBAL
//   id order (ranks' blocks in rank order, a rank's bodies in its own order: the same pair orientation on every rank) ->
//   the single-GPU link build with ghost-ghost links dropped -> per contact "counted here" flags -> send / receive lists
//   of the velocity halo, derived on each side from its own links (a cut link is found by both owners, so the two lists
//   agree without a request exchange).
// Per step without a rebuild: the owners send the current centres / quaternions of the bodies they ghosted out.
// Then the single-GPU step core with the solver's halo / reduction hooks; the owned block is copied back to the caller.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "comm.cuh"
#include "links.cuh"
#include "sort_scan.cuh"

namespace mb200 {
namespace {

// doubles per exchanged ghost body: centre 3, radius, quaternion 4, length, kind, external force 3, external torque 3. The
// same record at a rebuild (everything is unpacked) and per step (centre, quaternion and the external wrench: the
// solver's phi0 correction needs the external velocity of BOTH bodies of a contact).
constexpr int kBodyRec = 16;

__device__ inline double search_radius_of(int64_t g, const double* radius, const double* length, const int32_t* kind,
                                          double buffer) {
  const bool rod = kind && kind[g] == MB200_SPHEROCYLINDER;
  return (rod ? 0.5 * length[g] + radius[g] : radius[g]) + buffer;
}

__global__ void __launch_bounds__(256) wrap_points_kernel(int64_t n, Periodic P, const double* __restrict__ in,
                                                          double* __restrict__ out) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k < n) st3(out, k, periodic_wrap(P, ld3(in, k)));
}

// per-CTA partials {min xyz, max xyz, max S} over the owned bodies
__global__ void __launch_bounds__(256)
    owned_bounds_kernel(int64_t n_own, const double* __restrict__ center, const double* __restrict__ radius,
                        const double* __restrict__ length, const int32_t* __restrict__ kind, double buffer,
                        double* __restrict__ partials) {
  __shared__ double sh[33];
  const double big = 1.7976931348623157e308;
  double lo[3] = {big, big, big}, hi[3] = {-big, -big, -big}, smax = 0.0;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; k < n_own; k += stride) {
    const V3 c = ld3(center, k);
    lo[0] = fmin(lo[0], c.x), lo[1] = fmin(lo[1], c.y), lo[2] = fmin(lo[2], c.z);
    hi[0] = fmax(hi[0], c.x), hi[1] = fmax(hi[1], c.y), hi[2] = fmax(hi[2], c.z);
    smax = fmax(smax, search_radius_of(k, radius, length, kind, buffer));
  }
  double out[7];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    out[d] = -block_max(-lo[d], sh, -big);
    out[3 + d] = block_max(hi[d], sh, -big);
  }
  out[6] = block_max(smax, sh, 0.0);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < 7; ++k) partials[7 * blockIdx.x + k] = out[k];
  }
}

// Every rank's owned bounding box and largest search radius (all-gathered); an empty rank has lo > hi.
struct RankBoxes {
  double lo[kMaxRanks][3], hi[kMaxRanks][3], smax[kMaxRanks];
};
struct PeerList {
  int npeers;
  int peer[kMaxRanks];
};

// An occupancy grid of every rank's owned centres (G^3 cells over the bounding box of all bodies, one bit per cell,
// all-gathered: at most 256 KB per rank). A rank's bounding box is a poor picture of a space-filling-curve block (a few
// bodies of the next octant stretch it across a whole neighbour); the grid follows the block's shape. The cell edge is
// about one search radius (G <= 128): a ghost layer selected through the grid is at most one cell thicker than needed.
constexpr int kOccMaxG = 128;
struct OccGrid {
  double lo[3], inv_cell[3];
  int G, words;  // cells per axis; 64-bit words per rank (G^3 / 64 rounded up)
};
__device__ inline int occ_cell(const OccGrid& og, double x, int d) {
  const int c = static_cast<int>(floor((x - og.lo[d]) * og.inv_cell[d]));
  return min(max(c, 0), og.G - 1);
}
__global__ void __launch_bounds__(256) occupancy_kernel(int64_t n, OccGrid og, const double* __restrict__ center,
                                                        unsigned long long* __restrict__ mask) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const V3 c = ld3(center, k);
  const int cell = (occ_cell(og, c.z, 2) * og.G + occ_cell(og, c.y, 1)) * og.G + occ_cell(og, c.x, 0);
  const unsigned long long bit = 1ull << (cell & 63);
  if (!(mask[cell >> 6] & bit)) atomicOr(mask + (cell >> 6), bit);
}
// Does rank r own a body in a cell that the cube [c - reach, c + reach] touches? (periodic axes: the images too)
__device__ inline bool occ_reaches(const OccGrid& og, const unsigned long long* __restrict__ mask, const V3& c,
                                   double reach, const Periodic& P) {
  int lo[3], hi[3];
  const double cc[3] = {c.x, c.y, c.z};
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    if (P.on[d]) {  // unclamped cell range, wrapped below; a reach of half the period or more covers the axis
      lo[d] = static_cast<int>(floor((cc[d] - reach - og.lo[d]) * og.inv_cell[d]));
      hi[d] = static_cast<int>(floor((cc[d] + reach - og.lo[d]) * og.inv_cell[d]));
      if (hi[d] - lo[d] >= og.G - 1) lo[d] = 0, hi[d] = og.G - 1;
    } else {
      lo[d] = occ_cell(og, cc[d] - reach, d);
      hi[d] = occ_cell(og, cc[d] + reach, d);
    }
  }
  for (int z = lo[2]; z <= hi[2]; ++z)
    for (int y = lo[1]; y <= hi[1]; ++y)
      for (int x = lo[0]; x <= hi[0]; ++x) {
        const int G = og.G;
        const int xw = ((x % G) + G) % G, yw = ((y % G) + G) % G, zw = ((z % G) + G) % G;
        const int cell = (zw * G + yw) * G + xw;
        if ((__ldg(mask + (cell >> 6)) >> (cell & 63)) & 1ull) return true;
      }
  return false;
}

// Distance from coordinate x to the interval [lo, hi] along one axis; periodic axis of length L: nearest image.
__device__ inline double axis_gap(double x, double lo, double hi, bool periodic, double L) {
  double g = fmax(fmax(lo - x, x - hi), 0.0);
  if (periodic) {
    g = fmin(g, fmax(fmax(lo - (x + L), (x + L) - hi), 0.0));
    g = fmin(g, fmax(fmax(lo - (x - L), (x - L) - hi), 0.0));
  }
  return g;
}

// flags[q * n_own + k] = 1 when owned body k can be a neighbour of a body owned by the q-th peer: its centre lies within
// S_k + S_max(peer) of that peer's bounding box AND of an occupied cell of that peer's occupancy grid (conservative; the
// exact predicate runs in the peer's link build).
// Periodic boxes: centres are first wrapped into the unit cell (the boxes and grids are those of wrapped centres).
__global__ void __launch_bounds__(256)
    select_for_peers_kernel(int64_t n_own, PeerList pl, RankBoxes rb, OccGrid og,
                            const unsigned long long* __restrict__ masks, const double* __restrict__ center,
                            const double* __restrict__ radius, const double* __restrict__ length,
                            const int32_t* __restrict__ kind, double buffer, Periodic P, uint32_t* __restrict__ flags) {
  const int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t >= n_own * pl.npeers) return;
  const int q = static_cast<int>(t / n_own);
  const int64_t k = t % n_own;
  const int r = pl.peer[q];
  V3 c = ld3(center, k);
  if (P.any) c = periodic_wrap(P, c);
  const double reach = (search_radius_of(k, radius, length, kind, buffer) + rb.smax[r]) * (1.0 + 1e-9);
  const double gx = axis_gap(c.x, rb.lo[r][0], rb.hi[r][0], P.on[0], P.L[0]);
  const double gy = axis_gap(c.y, rb.lo[r][1], rb.hi[r][1], P.on[1], P.L[1]);
  const double gz = axis_gap(c.z, rb.lo[r][2], rb.hi[r][2], P.on[2], P.L[2]);
  const bool empty = rb.lo[r][0] > rb.hi[r][0];
  bool near = !empty && gx <= reach && gy <= reach && gz <= reach;
  if (near) near = occ_reaches(og, masks + static_cast<int64_t>(r) * og.words, c, reach, P);
  flags[t] = near ? 1u : 0u;
}

__device__ inline void write_body_record(double* r, int64_t k, const double* __restrict__ center,
                                         const double* __restrict__ quat, const double* __restrict__ radius,
                                         const double* __restrict__ length, const int32_t* __restrict__ kind,
                                         const double* __restrict__ F_ext, const double* __restrict__ T_ext) {
  const V3 c = ld3(center, k);
  r[0] = c.x, r[1] = c.y, r[2] = c.z, r[3] = radius[k];
  Q4 q{1.0, 0.0, 0.0, 0.0};
  if (quat) q = ldq(quat, k);
  r[4] = q.w, r[5] = q.x, r[6] = q.y, r[7] = q.z;
  r[8] = length ? length[k] : 0.0;
  r[9] = kind ? static_cast<double>(kind[k]) : 0.0;
  const V3 f = F_ext ? ld3(F_ext, k) : V3{0.0, 0.0, 0.0}, tq = T_ext ? ld3(T_ext, k) : V3{0.0, 0.0, 0.0};
  r[10] = f.x, r[11] = f.y, r[12] = f.z, r[13] = tq.x, r[14] = tq.y, r[15] = tq.z;
}

// Selected bodies -> the list of their owned indices (peer-major, ascending owned index inside a peer).
__global__ void __launch_bounds__(256)
    ghost_list_kernel(int64_t count, int64_t n_own, const uint32_t* __restrict__ flags, const int64_t* __restrict__ pos,
                      uint32_t* __restrict__ idx) {
  const int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t >= count || !flags[t]) return;
  idx[pos[t]] = static_cast<uint32_t>(t % n_own);
}

// Send records of the listed bodies: at a rebuild and, with the current state and wrench, at every step.
__global__ void __launch_bounds__(256)
    pack_ghosts_kernel(int64_t count, const uint32_t* __restrict__ idx, const double* __restrict__ center,
                       const double* __restrict__ quat, const double* __restrict__ radius,
                       const double* __restrict__ length, const int32_t* __restrict__ kind,
                       const double* __restrict__ F_ext, const double* __restrict__ T_ext, double* __restrict__ rec) {
  const int64_t o = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (o >= count) return;
  write_body_record(rec + kBodyRec * o, idx[o], center, quat, radius, length, kind, F_ext, T_ext);
}

__device__ inline int64_t ghost_local_index(const GhostMap& gm, int64_t g) {
  int q = 0;
  while (q + 1 < gm.npeers && g >= gm.recv_off[q + 1]) ++q;
  return gm.local_beg[q] + (g - gm.recv_off[q]);
}

template <bool FULL>  // FULL: everything (rebuild); else the state and the external wrench (every step)
__global__ void __launch_bounds__(256)
    unpack_ghosts_kernel(int64_t count, GhostMap gm, const double* __restrict__ rec, double* __restrict__ lc,
                         double* __restrict__ lq, double* __restrict__ lr, double* __restrict__ ll,
                         int32_t* __restrict__ lk, double* __restrict__ lf, double* __restrict__ lt,
                         uint8_t* __restrict__ owned) {
  const int64_t g = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (g >= count) return;
  const int64_t k = ghost_local_index(gm, g);
  const double* r = rec + kBodyRec * g;
  st3(lc, k, V3{r[0], r[1], r[2]});
  if (lq) {
    double2* o = reinterpret_cast<double2*>(lq + 4 * k);
    o[0] = make_double2(r[4], r[5]);
    o[1] = make_double2(r[6], r[7]);
  }
  if (lf) st3(lf, k, V3{r[10], r[11], r[12]});
  if (lt) st3(lt, k, V3{r[13], r[14], r[15]});
  if (FULL) {
    lr[k] = r[3];
    if (ll) ll[k] = r[8];
    if (lk) lk[k] = static_cast<int32_t>(r[9]);
    owned[k] = 0;
  }
}

// owned bodies -> their block of the local set. STATIC: also the fields that do not change between rebuilds.
template <bool STATIC>
__global__ void __launch_bounds__(256)
    copy_owned_kernel(int64_t n_own, int64_t o_beg, const double* __restrict__ gc, const double* __restrict__ gq,
                      const double* __restrict__ gr, const double* __restrict__ gl, const int32_t* __restrict__ gk,
                      const double* __restrict__ gf, const double* __restrict__ gt, double* __restrict__ lc,
                      double* __restrict__ lq, double* __restrict__ lr, double* __restrict__ ll, int32_t* __restrict__ lk,
                      double* __restrict__ lf, double* __restrict__ lt, uint8_t* __restrict__ owned) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= n_own) return;
  const int64_t l = o_beg + k;
  st3(lc, l, ld3(gc, k));
  if (gq) {
    const Q4 q = ldq(gq, k);
    double2* o = reinterpret_cast<double2*>(lq + 4 * l);
    o[0] = make_double2(q.w, q.x);
    o[1] = make_double2(q.y, q.z);
  }
  if (gf) st3(lf, l, ld3(gf, k));
  if (gt) st3(lt, l, ld3(gt, k));
  if (STATIC) {
    lr[l] = gr[k];
    if (gl) ll[l] = gl[k];
    if (gk) lk[l] = gk[k];
    owned[l] = 1;
  }
}

__global__ void __launch_bounds__(256) owned_morton_kernel(int64_t n, const uint32_t* __restrict__ perm,
                                                           const uint8_t* __restrict__ owned, uint8_t* __restrict__ out) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k < n) out[k] = owned[perm[k]];
}

// Per-body lists of ghost copies from the send list: count, (scan), fill. Entry e of the send list goes to peer q(e).
struct SendOffsets {
  int npeers;
  int peer[16];
  int64_t off[17];
};
template <bool FILL>
__global__ void __launch_bounds__(256)
    push_lists_kernel(int64_t count, SendOffsets so, const uint32_t* __restrict__ send_idx,
                      const uint32_t* __restrict__ remote_row, uint32_t* __restrict__ counts,
                      const int64_t* __restrict__ ptr, uint32_t* __restrict__ dst) {
  const int64_t e = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (e >= count) return;
  const uint32_t b = send_idx[e];
  const uint32_t slot = atomicAdd(counts + b, 1u);
  if (FILL) {
    int q = 0;
    while (q + 1 < so.npeers && e >= so.off[q + 1]) ++q;
    dst[ptr[b] + slot] = (static_cast<uint32_t>(so.peer[q]) << 28) | remote_row[e];
  }
}
// The bodies with a push list, ascending (Morton) index: the body pass visits them first so that the velocity halo
// leaves while the interior bodies are still being processed.
__global__ void __launch_bounds__(256) boundary_flags_kernel(int64_t n, const int64_t* __restrict__ push_ptr,
                                                             uint32_t* __restrict__ flag) {
  const int64_t b = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (b < n) flag[b] = push_ptr[b + 1] > push_ptr[b] ? 1u : 0u;
}
// ... and their class in owned_s: 0 ghost, 1 owned interior, 2 owned with ghost copies (one byte tells the body pass all)
__global__ void __launch_bounds__(256) boundary_list_kernel(int64_t n, const uint32_t* __restrict__ flag,
                                                            const int64_t* __restrict__ pos, uint32_t* __restrict__ list,
                                                            uint8_t* __restrict__ owned_s) {
  const int64_t b = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (b < n && flag[b]) {
    list[pos[b]] = static_cast<uint32_t>(b);
    owned_s[b] = 2;
  }
}

__global__ void __launch_bounds__(256) invert_perm_kernel(int64_t n, const uint32_t* __restrict__ perm,
                                                          uint32_t* __restrict__ inv) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k < n) inv[perm[k]] = static_cast<uint32_t>(k);
}

// Local index -> owning rank: the local set is the ranks' blocks in rank order.
struct Blocks {
  int world;
  int64_t beg[kMaxRanks + 1];
};
__device__ inline int owner_of(const Blocks& bl, int64_t l) {
  int r = 0;
  while (r + 1 < bl.world && l >= bl.beg[r + 1]) ++r;
  return r;
}

// Per link: counted here? For a cut link: mark the owned end for sending to the other owner, the ghost end for receiving.
// Local order is global id order, so "the lower global id" is the lower local index.
__global__ void __launch_bounds__(256)
    classify_links_kernel(int64_t npairs, const int2* __restrict__ pairs, const uint32_t* __restrict__ perm, Blocks bl,
                          int me, int64_t o_beg, uint8_t* __restrict__ counted, uint32_t* __restrict__ sendmask,
                          uint32_t* __restrict__ recvflag) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= npairs) return;
  const int2 p = pairs[c];
  const uint32_t li = perm[p.x], lj = perm[p.y];
  const int oi = owner_of(bl, li), oj = owner_of(bl, lj);
  counted[c] = ((li < lj ? oi : oj) == me) ? 1 : 0;
  if (oi != oj) {
    if (oi == me) {
      atomicOr(sendmask + (li - o_beg), 1u << oj);
      recvflag[lj] = 1u;
    } else {
      atomicOr(sendmask + (lj - o_beg), 1u << oi);
      recvflag[li] = 1u;
    }
  }
}

// flags2[q * n_own + k] = owned body k goes to the q-th peer
__global__ void __launch_bounds__(256) send_flags_kernel(int64_t n_own, int npeers, const int* __restrict__ peers,
                                                         const uint32_t* __restrict__ sendmask, uint32_t* __restrict__ flags2) {
  const int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t >= n_own * npeers) return;
  const int q = static_cast<int>(t / n_own);
  const int64_t k = t % n_own;
  flags2[t] = (sendmask[k] >> peers[q]) & 1u;
}
// idx[pos[t]] = Morton index of the flagged local body (local index = base + t % period)
__global__ void __launch_bounds__(256)
    compact_halo_kernel(int64_t count, int64_t period, int64_t base, const uint32_t* __restrict__ flags,
                        const int64_t* __restrict__ pos, const uint32_t* __restrict__ inv_perm, uint32_t* __restrict__ idx) {
  const int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t >= count || !flags[t]) return;
  idx[pos[t]] = inv_perm[base + t % period];
}

int d2h(cudaStream_t st, void* dst, const void* src, size_t bytes) {
  MB200_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st));
  return MB200_OK;
}

// Grouped send / receive of per-peer slices (`width` doubles per item) between the ranks that have something to exchange.
int exchange(const NcclApi* nccl, mb200_comm* comm, cudaStream_t st, int npeers, const int* peer, const int64_t* send_off,
             const int64_t* recv_off, const double* sendbuf, double* recvbuf, int width) {
  MB200_NCCL(nccl->GroupStart());
  for (int q = 0; q < npeers; ++q) {
    const int64_t cs = send_off[q + 1] - send_off[q], cr = recv_off[q + 1] - recv_off[q];
    if (cs > 0)
      MB200_NCCL(nccl->Send(sendbuf + send_off[q] * width, static_cast<size_t>(cs) * width, ncclDouble, peer[q], comm->comm, st));
    if (cr > 0)
      MB200_NCCL(nccl->Recv(recvbuf + recv_off[q] * width, static_cast<size_t>(cr) * width, ncclDouble, peer[q], comm->comm, st));
  }
  MB200_NCCL(nccl->GroupEnd());
  return MB200_OK;
}

// MB200_DIST_TIMING=1: host wall time of the phases of a distributed rebuild (synchronising; diagnostics only)
struct PhaseTimer {
  bool on;
  cudaStream_t st;
  int rank;
  std::chrono::steady_clock::time_point t0;
  PhaseTimer(cudaStream_t s, int r) : on(getenv("MB200_DIST_TIMING") != nullptr), st(s), rank(r) {
    if (on) cudaStreamSynchronize(st), t0 = std::chrono::steady_clock::now();
  }
  void mark(const char* what) {
    if (!on) return;
    cudaStreamSynchronize(st);
    const auto t1 = std::chrono::steady_clock::now();
    std::fprintf(stderr, "[mb200 rank %d] rebuild %-22s %8.3f ms\n", rank, what,
                 std::chrono::duration<double, std::milli>(t1 - t0).count());
    t0 = t1;
  }
};

// (Re)build the local body set, its links and the halo lists.
int dist_build(cudaStream_t st, mb200_neighbor_links* h, mb200_comm* comm, const mb200_bodies* g, bool with_f, bool with_t) {
  auto& D = h->dist;
  const NcclApi* nccl = nccl_api();
  const int W = comm->world, me = comm->rank;
  const int64_t n_own = g->n;
  const bool rods = g->kind != nullptr;
  D.n_own = n_own;
  D.npeers = W - 1;
  MB200_REQUIRE(W <= kMaxRanks, "at most 16 ranks");
  PhaseTimer pt(st, me);
  for (int q = 0, r = 0; r < W; ++r)
    if (r != me) D.peer[q++] = r;

  // 1. owned bounds (of the wrapped centres when the box is periodic: the same wrapped cloud the peers test against)
  const int bgrid = grid_for(n_own, 256, kSMs * 2);
  Scratch<double> partials, wrapped;
  MB200_TRY(partials.alloc(7 * static_cast<size_t>(bgrid), st));
  const double* bc = g->center;
  if (h->periodic.any && n_own > 0) {
    MB200_TRY(wrapped.alloc(3 * n_own, st));
    wrap_points_kernel<<<grid_for(n_own, 256), 256, 0, st>>>(n_own, h->periodic, g->center, wrapped.p);
    MB200_LAUNCHED();
    bc = wrapped.p;
  }
  owned_bounds_kernel<<<bgrid, 256, 0, st>>>(n_own, bc, g->radius, g->length, g->kind, h->buffer, partials.p);
  MB200_LAUNCHED();
  std::vector<double> hp(7 * static_cast<size_t>(bgrid));
  MB200_TRY(d2h(st, hp.data(), partials.p, hp.size() * sizeof(double)));
  MB200_CUDA(cudaStreamSynchronize(st));
  double mine[8] = {hp[0], hp[1], hp[2], hp[3], hp[4], hp[5], hp[6], static_cast<double>(n_own)};
  for (int k = 1; k < bgrid; ++k) {
    for (int d = 0; d < 3; ++d) {
      mine[d] = std::min(mine[d], hp[7 * k + d]);
      mine[3 + d] = std::max(mine[3 + d], hp[7 * k + 3 + d]);
    }
    mine[6] = std::max(mine[6], hp[7 * k + 6]);
  }
  std::vector<double> all(8 * static_cast<size_t>(W));
  MB200_TRY(comm_allgather_host(comm, st, mine, all.data(), sizeof(mine)));
  RankBoxes rb;
  std::vector<int64_t> own_of(W);
  int64_t n_global = 0;
  for (int r = 0; r < W; ++r) {
    for (int d = 0; d < 3; ++d) rb.lo[r][d] = all[8 * r + d], rb.hi[r][d] = all[8 * r + 3 + d];
    rb.smax[r] = all[8 * r + 6];
    own_of[r] = static_cast<int64_t>(all[8 * r + 7]);
    n_global += own_of[r];
  }
  D.n_global = n_global;
  // occupancy grids over the bounding box of all bodies (identical on every rank), all-gathered
  OccGrid og;
  double glo[3], ghi[3], extent = 0.0, smax_all = 0.0;
  for (int r = 0; r < W; ++r) smax_all = std::max(smax_all, rb.smax[r]);
  for (int d = 0; d < 3; ++d) {
    double lo = 1.7976931348623157e308, hi = -1.7976931348623157e308;
    for (int r = 0; r < W; ++r)
      if (rb.lo[r][0] <= rb.hi[r][0]) lo = std::min(lo, rb.lo[r][d]), hi = std::max(hi, rb.hi[r][d]);
    if (h->periodic.on[d]) lo = 0.0, hi = h->periodic.L[d];  // wrapped centres fill the unit cell
    if (!(hi > lo)) lo = 0.0, hi = 1.0;
    glo[d] = lo, ghi[d] = hi;
    extent = std::max(extent, hi - lo);
  }
  og.G = 8;
  if (smax_all > 0.0 && std::isfinite(extent / smax_all))
    og.G = static_cast<int>(std::min<double>(kOccMaxG, std::max(8.0, std::ceil(extent / smax_all))));
  og.words = (og.G * og.G * og.G + 63) / 64;
  for (int d = 0; d < 3; ++d) {
    og.lo[d] = glo[d];
    og.inv_cell[d] = og.G / ((ghi[d] - glo[d]) * (1.0 + 1e-12));
  }
  Scratch<unsigned long long> occ;
  MB200_TRY(occ.alloc(static_cast<size_t>(W + 1) * og.words, st));
  MB200_CUDA(cudaMemsetAsync(occ.p, 0, og.words * sizeof(unsigned long long), st));
  if (n_own > 0) {
    occupancy_kernel<<<grid_for(n_own, 256), 256, 0, st>>>(n_own, og, bc, occ.p);
    MB200_LAUNCHED();
  }
  MB200_NCCL(nccl->AllGather(occ.p, occ.p + og.words, og.words, ncclUint64, comm->comm, st));

  pt.mark("1 boxes + grids");
  // 2. which owned bodies does each peer need? one scan over (peer, owned body)
  PeerList pl;
  pl.npeers = D.npeers;
  for (int q = 0; q < kMaxRanks; ++q) pl.peer[q] = q < D.npeers ? D.peer[q] : 0;
  const int64_t m2 = n_own * D.npeers;
  Scratch<uint32_t> gflags;
  Scratch<int64_t> gpos;
  MB200_TRY(gflags.alloc(m2 > 0 ? m2 : 1, st));
  MB200_TRY(gpos.alloc(m2 + 1, st));
  if (m2 > 0) {
    select_for_peers_kernel<<<grid_for(m2, 256), 256, 0, st>>>(n_own, pl, rb, og, occ.p + og.words, g->center, g->radius,
                                                              g->length, g->kind, h->buffer, h->periodic, gflags.p);
    MB200_LAUNCHED();
  }
  MB200_TRY(exclusive_scan_u32_to_i64(st, m2, gflags.p, gpos.p));
  std::vector<int64_t> gcut(D.npeers + 1);
  for (int q = 0; q <= D.npeers; ++q) MB200_TRY(d2h(st, &gcut[q], gpos.p + q * n_own, sizeof(int64_t)));
  MB200_CUDA(cudaStreamSynchronize(st));
  for (int q = 0; q <= D.npeers; ++q) D.bsend_off[q] = gcut[q];
  const int64_t nbs = D.bsend_off[D.npeers];
  // every rank's counts to every rank -> what this rank receives
  std::vector<int64_t> to(W, 0), matrix(static_cast<size_t>(W) * W);
  for (int q = 0; q < D.npeers; ++q) to[D.peer[q]] = D.bsend_off[q + 1] - D.bsend_off[q];
  MB200_TRY(comm_allgather_host(comm, st, to.data(), matrix.data(), sizeof(int64_t) * W));
  D.brecv_off[0] = 0;
  for (int q = 0; q < D.npeers; ++q) D.brecv_off[q + 1] = D.brecv_off[q] + matrix[static_cast<size_t>(D.peer[q]) * W + me];
  const int64_t ng = D.brecv_off[D.npeers];
  const int64_t nl = n_own + ng;
  MB200_REQUIRE(nl < (int64_t(1) << 31), "local body set too large for int32 ids");
  // the local set: the ranks' blocks in rank order (global id order)
  std::vector<int64_t> blk(W + 1, 0);
  for (int r = 0; r < W; ++r) blk[r + 1] = blk[r] + (r == me ? n_own : matrix[static_cast<size_t>(r) * W + me]);
  D.n_local = nl, D.o_beg = blk[me];
  // largest local set of any rank: the size of everybody's velocity arena (identical on every rank)
  int64_t nl_max = 0;
  for (int r = 0; r < W; ++r) {
    int64_t s = own_of[r];
    for (int q = 0; q < W; ++q)
      if (q != r) s += matrix[static_cast<size_t>(q) * W + r];
    nl_max = std::max(nl_max, s);
  }

  pt.mark("2 select + counts");
  // 3. ghosts: pack, exchange, assemble the local arrays
  MB200_TRY(D.bsend_idx.reserve(nbs + 1));
  MB200_TRY(D.bsend_rec.reserve(kBodyRec * static_cast<size_t>(nbs) + 1));
  MB200_TRY(D.brecv_rec.reserve(kBodyRec * static_cast<size_t>(ng) + 1));
  if (m2 > 0) {
    ghost_list_kernel<<<grid_for(m2, 256), 256, 0, st>>>(m2, n_own, gflags.p, gpos.p, D.bsend_idx.p);
    MB200_LAUNCHED();
  }
  if (nbs > 0) {
    pack_ghosts_kernel<<<grid_for(nbs, 256), 256, 0, st>>>(nbs, D.bsend_idx.p, g->center, g->quat, g->radius, g->length,
                                                          g->kind, nullptr, nullptr, D.bsend_rec.p);
    MB200_LAUNCHED();
  }
  MB200_TRY(exchange(nccl, comm, st, D.npeers, D.peer, D.bsend_off, D.brecv_off, D.bsend_rec.p, D.brecv_rec.p, kBodyRec));
  MB200_TRY(D.owned.reserve(nl));
  MB200_TRY(D.center.reserve(3 * nl));
  MB200_TRY(D.radius.reserve(nl));
  if (rods) {
    MB200_TRY(D.quat.reserve(4 * nl));
    MB200_TRY(D.length.reserve(nl));
    MB200_TRY(D.kind.reserve(nl));
  }
  if (with_f) MB200_TRY(D.fext.reserve(3 * nl));
  if (with_t) MB200_TRY(D.text.reserve(3 * nl));
  D.gmap.npeers = D.npeers;
  for (int q = 0; q <= D.npeers; ++q) D.gmap.recv_off[q] = D.brecv_off[q];
  for (int q = 0; q < D.npeers; ++q) D.gmap.local_beg[q] = blk[D.peer[q]];
  if (n_own > 0) {
    copy_owned_kernel<true><<<grid_for(n_own, 256), 256, 0, st>>>(n_own, D.o_beg, g->center, g->quat, g->radius, g->length,
                                                                  g->kind, nullptr, nullptr, D.center.p, D.quat.p,
                                                                  D.radius.p, D.length.p, D.kind.p, nullptr, nullptr,
                                                                  D.owned.p);
    MB200_LAUNCHED();
  }
  if (ng > 0) {
    unpack_ghosts_kernel<true><<<grid_for(ng, 256), 256, 0, st>>>(ng, D.gmap, D.brecv_rec.p, D.center.p, D.quat.p, D.radius.p,
                                                                  D.length.p, D.kind.p, nullptr, nullptr, D.owned.p);
    MB200_LAUNCHED();
  }
  mb200_bodies lb{nl, D.center.p, rods ? D.quat.p : nullptr, D.radius.p, rods ? D.length.p : nullptr,
                  rods ? D.kind.p : nullptr};

  pt.mark("3 ghost exchange");
  // 4. links with at least one owned end
  MB200_TRY(build_links(st, h, &lb, nullptr, D.owned.p));
  const int64_t C = h->npairs;
  pt.mark("4 local link build");

  // 5. counted flags and halo marks
  MB200_TRY(D.counted.reserve(C + 1));
  MB200_TRY(D.inv_perm.reserve(nl + 1));
  if (nl > 0) {
    invert_perm_kernel<<<grid_for(nl, 256), 256, 0, st>>>(nl, h->perm.p, D.inv_perm.p);
    MB200_LAUNCHED();
  }
  Scratch<uint32_t> sendmask, recvflag;
  MB200_TRY(sendmask.alloc(n_own > 0 ? n_own : 1, st));
  MB200_TRY(recvflag.alloc(nl > 0 ? nl : 1, st));
  MB200_CUDA(cudaMemsetAsync(sendmask.p, 0, (n_own > 0 ? n_own : 1) * sizeof(uint32_t), st));
  MB200_CUDA(cudaMemsetAsync(recvflag.p, 0, (nl > 0 ? nl : 1) * sizeof(uint32_t), st));
  Blocks bl;
  bl.world = W;
  for (int r = 0; r <= kMaxRanks; ++r) bl.beg[r] = blk[r <= W ? r : W];
  if (C > 0) {
    classify_links_kernel<<<grid_for(C, 256), 256, 0, st>>>(C, reinterpret_cast<const int2*>(h->pairs.p), h->perm.p, bl,
                                                           me, D.o_beg, D.counted.p, sendmask.p, recvflag.p);
    MB200_LAUNCHED();
  }

  // 6. receive lists: ghosts flagged above, in local (= global id) order; each peer's ghosts are one contiguous block
  Scratch<int64_t> rpos;
  MB200_TRY(rpos.alloc(nl + 1, st));
  MB200_TRY(exclusive_scan_u32_to_i64(st, nl, recvflag.p, rpos.p));
  std::vector<int64_t> rcut(W + 1);
  for (int r = 0; r <= W; ++r) MB200_TRY(d2h(st, &rcut[r], rpos.p + blk[r], sizeof(int64_t)));
  // 7. send lists: one scan over (peer, owned body)
  Scratch<uint32_t> flags2;
  Scratch<int64_t> spos;
  Scratch<int> peers_d;
  MB200_TRY(flags2.alloc(m2 > 0 ? m2 : 1, st));
  MB200_TRY(spos.alloc(m2 + 1, st));
  MB200_TRY(peers_d.alloc(16, st));
  MB200_CUDA(cudaMemcpyAsync(peers_d.p, D.peer, sizeof(int) * 16, cudaMemcpyHostToDevice, st));
  if (m2 > 0) {
    send_flags_kernel<<<grid_for(m2, 256), 256, 0, st>>>(n_own, D.npeers, peers_d.p, sendmask.p, flags2.p);
    MB200_LAUNCHED();
  }
  MB200_TRY(exclusive_scan_u32_to_i64(st, m2, flags2.p, spos.p));
  std::vector<int64_t> scut(D.npeers + 1);
  for (int q = 0; q <= D.npeers; ++q) MB200_TRY(d2h(st, &scut[q], spos.p + q * n_own, sizeof(int64_t)));
  MB200_CUDA(cudaStreamSynchronize(st));
  for (int q = 0; q <= D.npeers; ++q) D.send_off[q] = scut[q];
  D.recv_off[0] = 0;
  for (int q = 0; q < D.npeers; ++q) {
    const int r = D.peer[q];
    D.recv_off[q + 1] = D.recv_off[q] + (rcut[r + 1] - rcut[r]);
  }
  const int64_t ns = D.send_off[D.npeers], nr = D.recv_off[D.npeers];
  MB200_REQUIRE(nr == rcut[W] - (rcut[me + 1] - rcut[me]), "inconsistent receive lists");
  MB200_TRY(D.send_idx.reserve(ns + 1));
  MB200_TRY(D.recv_idx.reserve(nr + 1));
  MB200_TRY(D.sendbuf.reserve(2 * ns + 1));
  MB200_TRY(D.recvbuf.reserve(2 * nr + 1));
  MB200_TRY(D.red.reserve(4 + 4 * W));
  if (m2 > 0) {
    compact_halo_kernel<<<grid_for(m2, 256), 256, 0, st>>>(m2, n_own, D.o_beg, flags2.p, spos.p, D.inv_perm.p, D.send_idx.p);
    MB200_LAUNCHED();
  }
  // owned bodies are never flagged for receiving, so the compacted receive order is peer-block order
  if (nl > 0) {
    compact_halo_kernel<<<grid_for(nl, 256), 256, 0, st>>>(nl, nl, 0, recvflag.p, rpos.p, D.inv_perm.p, D.recv_idx.p);
    MB200_LAUNCHED();
  }
  pt.mark("5-7 halo lists");
  // 8. peer-memory exchange: map the ranks' velocity arenas (sized for the largest local set of any rank, with headroom,
  //    so that every rank re-maps in the same step), then tell every peer in which rows of MY arena its records belong:
  //    my receive list, in the order of its send list
  {
    const int64_t need = 2 * nl_max;  // re-mapped (collectively, with headroom) only when some rank outgrows the arenas
    if (!comm->peer.tried || need > comm->peer.rows) MB200_TRY(peer_setup(comm, st, need + need / 4 + 4096));
  }
  D.n_boundary = 0;
  if (comm->peer.ok) {
    MB200_TRY(D.owned_s.reserve(nl + 1));
    if (nl > 0) {
      owned_morton_kernel<<<grid_for(nl, 256), 256, 0, st>>>(nl, h->perm.p, D.owned.p, D.owned_s.p);
      MB200_LAUNCHED();
    }
    MB200_TRY(D.remote_row.reserve(ns + 1));
    MB200_NCCL(nccl->GroupStart());
    for (int q = 0; q < D.npeers; ++q) {
      const int64_t cs = D.send_off[q + 1] - D.send_off[q], cr = D.recv_off[q + 1] - D.recv_off[q];
      if (cr > 0) MB200_NCCL(nccl->Send(D.recv_idx.p + D.recv_off[q], cr, ncclUint32, D.peer[q], comm->comm, st));
      if (cs > 0) MB200_NCCL(nccl->Recv(D.remote_row.p + D.send_off[q], cs, ncclUint32, D.peer[q], comm->comm, st));
    }
    MB200_NCCL(nccl->GroupEnd());
    // body-major form for the fused push of the body pass
    MB200_REQUIRE(2 * nl_max < (int64_t(1) << 28), "too many bodies for the 28-bit ghost row index");
    Scratch<uint32_t> pcount;
    MB200_TRY(pcount.alloc(nl + 1, st));
    MB200_TRY(D.push_ptr.reserve(nl + 1));
    MB200_TRY(D.push_dst.reserve(ns + 1));
    SendOffsets so;
    so.npeers = D.npeers;
    for (int q = 0; q < 16; ++q) so.peer[q] = D.peer[q];
    for (int q = 0; q < 17; ++q) so.off[q] = D.send_off[q];
    MB200_CUDA(cudaMemsetAsync(pcount.p, 0, (nl + 1) * sizeof(uint32_t), st));
    if (ns > 0) {
      push_lists_kernel<false><<<grid_for(ns, 256), 256, 0, st>>>(ns, so, D.send_idx.p, D.remote_row.p, pcount.p, nullptr, nullptr);
      MB200_LAUNCHED();
    }
    MB200_TRY(exclusive_scan_u32_to_i64(st, nl, pcount.p, D.push_ptr.p));
    MB200_CUDA(cudaMemsetAsync(pcount.p, 0, (nl + 1) * sizeof(uint32_t), st));
    if (ns > 0) {
      push_lists_kernel<true><<<grid_for(ns, 256), 256, 0, st>>>(ns, so, D.send_idx.p, D.remote_row.p, pcount.p, D.push_ptr.p,
                                                                 D.push_dst.p);
      MB200_LAUNCHED();
    }
    // the bodies that have ghost copies somewhere, ascending: the body pass handles them first
    Scratch<int64_t> bpos;
    MB200_TRY(bpos.alloc(nl + 1, st));
    MB200_TRY(D.boundary.reserve(nl + 1));
    if (nl > 0) {
      boundary_flags_kernel<<<grid_for(nl, 256), 256, 0, st>>>(nl, D.push_ptr.p, pcount.p);
      MB200_LAUNCHED();
    }
    MB200_TRY(exclusive_scan_u32_to_i64(st, nl, pcount.p, bpos.p));
    if (nl > 0) {
      boundary_list_kernel<<<grid_for(nl, 256), 256, 0, st>>>(nl, pcount.p, bpos.p, D.boundary.p, D.owned_s.p);
      MB200_LAUNCHED();
    }
    int64_t nb = 0;
    MB200_TRY(d2h(st, &nb, bpos.p + nl, sizeof(int64_t)));
    MB200_CUDA(cudaStreamSynchronize(st));
    D.n_boundary = nb;
  }
  pt.mark("8 peer rows + push lists");
  D.built = true;
  return MB200_OK;
}

}  // namespace
}  // namespace mb200

using namespace mb200;

extern "C" int mb200_collision_step_dist(mb200_neighbor_links* h, mb200_comm* comm, mb200_stream stream,
                                         const mb200_bodies* g, const mb200_step_params* params, const double* F_ext,
                                         const double* T_ext, double* F_col, double* U_out, mb200_step_stats* stats_h) {
  MB200_REQUIRE(h && comm && g && params, "null argument");
  MB200_REQUIRE(g->n >= 0 && g->n < (int64_t(1) << 31), "bad owned body count");
  if (g->n > 0) MB200_REQUIRE(g->center && g->radius, "null centre/radius");
  if (g->kind) MB200_REQUIRE(g->quat && g->length, "kind given without quat/length");
  MB200_TRY(require_device());
  const NcclApi* nccl = nccl_api();
  if (!nccl) return MB200_ERR_COMM;
  cudaStream_t st = as_stream(stream);
  const long long launches0 = g_launches.load();
  auto& D = h->dist;
  const int64_t n_own = g->n;
  const bool rods = g->kind != nullptr;

  // rebuild decision: the same on every rank (all-reduced flag). Only OWNED bodies are tested: every body is owned by
  // somebody, so a ghost that moved too far raises the flag on its owner's side.
  bool rebuild = params->force_rebuild || !D.built || !h->built || D.n_own != n_own;
  {
    StageScope ps(MB200_STAGE_TRIGGER, st);
    MB200_TRY(D.flag_all.reserve(2));
    if (!rebuild) {
      MB200_TRY(links_moved_flag(st, h, g, D.o_beg));
      MB200_CUDA(cudaMemcpyAsync(D.flag_all.p, h->flag.p, sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
    } else {
      const int32_t one = 1;
      MB200_CUDA(cudaMemcpyAsync(D.flag_all.p, &one, sizeof(int32_t), cudaMemcpyHostToDevice, st));
    }
    MB200_NCCL(nccl->AllReduce(D.flag_all.p, D.flag_all.p + 1, 1, ncclInt32, ncclMax, comm->comm, st));
    int32_t* any = static_cast<int32_t*>(pinned_scratch(sizeof(int32_t)));
    if (!any) return MB200_ERR_CUDA;
    MB200_CUDA(cudaMemcpyAsync(any, D.flag_all.p + 1, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    MB200_CUDA(cudaStreamSynchronize(st));
    rebuild = *any != 0;
  }
  if (rebuild) {
    StageScope ps(MB200_STAGE_BROADPHASE, st);
    MB200_TRY(dist_build(st, h, comm, g, F_ext != nullptr, T_ext != nullptr));
  }
  const int64_t nl = D.n_local;
  if (F_ext) MB200_TRY(D.fext.reserve(3 * nl + 1));
  if (T_ext) MB200_TRY(D.text.reserve(3 * nl + 1));
  if (F_col) MB200_TRY(D.fcol.reserve(3 * nl + 1));
  if (U_out) MB200_TRY(D.uout.reserve(3 * nl + 1));
  {
    // this step's owned state and forcing into the local arrays; without a rebuild the ghosts' current state comes from
    // their owners (the same lists as at the rebuild)
    StageScope ps(MB200_STAGE_STATE_COMM, st);
    if (n_own > 0) {
      copy_owned_kernel<false><<<grid_for(n_own, 256), 256, 0, st>>>(n_own, D.o_beg, g->center, g->quat, nullptr, nullptr,
                                                                     nullptr, F_ext, T_ext, D.center.p, D.quat.p, nullptr,
                                                                     nullptr, nullptr, D.fext.p, D.text.p, nullptr);
      MB200_LAUNCHED();
    }
    // the ghosts' current state (not needed right after a rebuild, which has just fetched it) and external wrench (the
    // phi0 correction reads the external velocity of both bodies of a contact): same lists as at the rebuild
    if (!rebuild || F_ext || T_ext) {
      const int64_t nbs = D.bsend_off[D.npeers], ng = D.brecv_off[D.npeers];
      if (nbs > 0) {
        pack_ghosts_kernel<<<grid_for(nbs, 256), 256, 0, st>>>(nbs, D.bsend_idx.p, g->center, g->quat, g->radius, g->length,
                                                              g->kind, F_ext, T_ext, D.bsend_rec.p);
        MB200_LAUNCHED();
      }
      MB200_TRY(exchange(nccl, comm, st, D.npeers, D.peer, D.bsend_off, D.brecv_off, D.bsend_rec.p, D.brecv_rec.p, kBodyRec));
      if (ng > 0) {
        unpack_ghosts_kernel<false><<<grid_for(ng, 256), 256, 0, st>>>(
            ng, D.gmap, D.brecv_rec.p, D.center.p, rods ? D.quat.p : nullptr, nullptr, nullptr, nullptr,
            F_ext ? D.fext.p : nullptr, T_ext ? D.text.p : nullptr, nullptr);
        MB200_LAUNCHED();
      }
    }
  }
  mb200_bodies lb{nl, D.center.p, rods ? D.quat.p : nullptr, D.radius.p, rods ? D.length.p : nullptr,
                  rods ? D.kind.p : nullptr};
  DistSolve ds;
  ds.comm = comm, ds.counted = D.counted.p, ds.npeers = D.npeers;
  for (int q = 0; q < 16; ++q) ds.peer[q] = D.peer[q];
  for (int q = 0; q < 17; ++q) ds.send_off[q] = D.send_off[q], ds.recv_off[q] = D.recv_off[q];
  ds.send_idx = D.send_idx.p, ds.recv_idx = D.recv_idx.p, ds.sendbuf = D.sendbuf.p, ds.recvbuf = D.recvbuf.p;
  ds.red = D.red.p;
  const bool peer_mode = comm->peer.ok;
  ds.peer_dev = peer_mode ? comm->peer.dev : nullptr;
  ds.vel_arena = peer_mode ? comm->peer.vel : nullptr;
  ds.push_ptr = peer_mode ? D.push_ptr.p : nullptr;
  ds.push_dst = peer_mode ? D.push_dst.p : nullptr;
  ds.owned_s = peer_mode ? D.owned_s.p : nullptr;
  ds.boundary = peer_mode ? D.boundary.p : nullptr;
  ds.n_boundary = peer_mode ? D.n_boundary : 0;
  ds.halo_seq = &comm->peer.halo_seq, ds.red_seq = &comm->peer.red_seq;
  mb200_step_params p = *params;
  p.throw_on_max_iters = 0;  // reported through stats (every rank must finish the collective sequence)
  MB200_TRY(step_core(h, st, &lb, &p, F_ext ? D.fext.p : nullptr, T_ext ? D.text.p : nullptr, F_col ? D.fcol.p : nullptr,
                      nullptr, U_out ? D.uout.p : nullptr, nullptr, rebuild ? 1 : 0, launches0, &ds, stats_h));
  {
    // the advanced owned block back to the caller
    StageScope ps(MB200_STAGE_STATE_COMM, st);
    const size_t d = sizeof(double);
    if (n_own > 0) {
      MB200_CUDA(cudaMemcpyAsync(g->center, D.center.p + 3 * D.o_beg, 3 * n_own * d, cudaMemcpyDeviceToDevice, st));
      if (rods) MB200_CUDA(cudaMemcpyAsync(g->quat, D.quat.p + 4 * D.o_beg, 4 * n_own * d, cudaMemcpyDeviceToDevice, st));
      if (F_col) MB200_CUDA(cudaMemcpyAsync(F_col, D.fcol.p + 3 * D.o_beg, 3 * n_own * d, cudaMemcpyDeviceToDevice, st));
      if (U_out) MB200_CUDA(cudaMemcpyAsync(U_out, D.uout.p + 3 * D.o_beg, 3 * n_own * d, cudaMemcpyDeviceToDevice, st));
    }
  }
  MB200_CUDA(cudaStreamSynchronize(st));
  if (stats_h) {
    stats_h->kernel_launches = static_cast<int32_t>(g_launches.load() - launches0);
    if (params->throw_on_max_iters && params->fixed_iters < 0 && stats_h->lcp.ite_count == params->max_iters &&
        !stats_h->lcp.converged) {
      set_error("Collision resolution did not converge!");
      return MB200_ERR_NOT_CONVERGED;
    }
  }
  return MB200_OK;
}

extern "C" int mb200_dist_local_counts(const mb200_neighbor_links* h, int64_t* n_local_h, int64_t* n_ghost_rows_h) {
  MB200_REQUIRE(h && n_local_h && n_ghost_rows_h, "null argument");
  *n_local_h = h->dist.n_local;
  *n_ghost_rows_h = h->dist.built ? h->dist.recv_off[h->dist.npeers] : 0;
  return MB200_OK;
}

extern "C" int mb200_comm_exchange_mode(const mb200_comm* comm) { return (comm && comm->peer.ok) ? 1 : 0; }
