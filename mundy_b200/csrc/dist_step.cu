// Multi-GPU collision step (one process per GPU, NCCL). See include/mundy_b200.h "Multi-GPU" for the decomposition.
//
// Per rebuild (decided by an all-reduced moved flag, so that every rank rebuilds in the same step):
//   owned bounds -> select owned + foreign bodies within search reach (stable compaction: local order = global index
//   order, hence the same pair orientation on every rank) -> the single-GPU link build with ghost-ghost links dropped
//   -> per contact "counted here" flags -> send / receive lists of the velocity halo, derived on each side from its own
//   links (a cut link is found by both owners, so the two lists agree without a request exchange).
// Per step: gather the local bodies from the replica, the single-GPU step core with the solver's halo / all-gather hooks,
//   write the owned block back, all-gather the advanced state into every replica.
#include <vector>

#include "comm.cuh"
#include "links.cuh"
#include "sort_scan.cuh"

namespace mb200 {
namespace {

__device__ inline double search_radius_of(int64_t g, const double* radius, const double* length, const int32_t* kind,
                                          double buffer) {
  const bool rod = kind && kind[g] == MB200_SPHEROCYLINDER;
  return (rod ? 0.5 * length[g] + radius[g] : radius[g]) + buffer;
}

// per-CTA partials {min xyz, max xyz, max S} over the owned block
__global__ void __launch_bounds__(256)
    owned_bounds_kernel(int64_t o0, int64_t n_own, const double* __restrict__ center, const double* __restrict__ radius,
                        const double* __restrict__ length, const int32_t* __restrict__ kind, double buffer,
                        double* __restrict__ partials) {
  __shared__ double sh[33];
  const double big = 1.7976931348623157e308;
  double lo[3] = {big, big, big}, hi[3] = {-big, -big, -big}, smax = 0.0;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; k < n_own; k += stride) {
    const V3 c = ld3(center, o0 + k);
    lo[0] = fmin(lo[0], c.x), lo[1] = fmin(lo[1], c.y), lo[2] = fmin(lo[2], c.z);
    hi[0] = fmax(hi[0], c.x), hi[1] = fmax(hi[1], c.y), hi[2] = fmax(hi[2], c.z);
    smax = fmax(smax, search_radius_of(o0 + k, radius, length, kind, buffer));
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

struct Box6 {
  double lo[3], hi[3];
};

// flag[g] = 1 for owned bodies and for foreign bodies whose search sphere can touch an owned one (conservative:
// centre inside the owned bounding box inflated by S_g + S_max_owned)
__global__ void __launch_bounds__(256)
    select_local_kernel(int64_t n, int64_t o0, int64_t n_own, const double* __restrict__ center,
                        const double* __restrict__ radius, const double* __restrict__ length,
                        const int32_t* __restrict__ kind, double buffer, Box6 box, double smax_own,
                        uint32_t* __restrict__ flag) {
  const int64_t g = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (g >= n) return;
  uint32_t f = 1;
  if (g < o0 || g >= o0 + n_own) {
    const V3 c = ld3(center, g);
    const double reach = (search_radius_of(g, radius, length, kind, buffer) + smax_own) * (1.0 + 1e-9);
    f = (c.x >= box.lo[0] - reach && c.x <= box.hi[0] + reach && c.y >= box.lo[1] - reach && c.y <= box.hi[1] + reach &&
         c.z >= box.lo[2] - reach && c.z <= box.hi[2] + reach)
            ? 1u
            : 0u;
  }
  flag[g] = f;
}

__global__ void __launch_bounds__(256)
    compact_local_kernel(int64_t n, int64_t o0, int64_t n_own, const uint32_t* __restrict__ flag,
                         const int64_t* __restrict__ pos, uint32_t* __restrict__ gidx, uint8_t* __restrict__ owned) {
  const int64_t g = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (g >= n || !flag[g]) return;
  const int64_t k = pos[g];
  gidx[k] = static_cast<uint32_t>(g);
  owned[k] = (g >= o0 && g < o0 + n_own) ? 1 : 0;
}

// local arrays <- replica. STATIC: also the fields that do not change between rebuilds.
template <bool STATIC>
__global__ void __launch_bounds__(256)
    gather_local_kernel(int64_t n_local, const uint32_t* __restrict__ gidx, const double* __restrict__ gc,
                        const double* __restrict__ gq, const double* __restrict__ gr, const double* __restrict__ gl,
                        const int32_t* __restrict__ gk, const double* __restrict__ gf, const double* __restrict__ gt,
                        double* __restrict__ lc, double* __restrict__ lq, double* __restrict__ lr, double* __restrict__ ll,
                        int32_t* __restrict__ lk, double* __restrict__ lf, double* __restrict__ lt) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= n_local) return;
  const int64_t g = gidx[k];
  st3(lc, k, ld3(gc, g));
  if (gq) {
    const Q4 q = ldq(gq, g);
    double2* o = reinterpret_cast<double2*>(lq + 4 * k);
    o[0] = make_double2(q.w, q.x);
    o[1] = make_double2(q.y, q.z);
  }
  if (gf) st3(lf, k, ld3(gf, g));
  if (gt) st3(lt, k, ld3(gt, g));
  if (STATIC) {
    lr[k] = gr[g];
    if (gl) ll[k] = gl[g];
    if (gk) lk[k] = gk[g];
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

__global__ void __launch_bounds__(256) invert_perm_kernel(int64_t n, const uint32_t* __restrict__ perm,
                                                          uint32_t* __restrict__ inv) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k < n) inv[perm[k]] = static_cast<uint32_t>(k);
}

// Per link: counted here? For a cut link: mark the owned end for sending to the other owner, the ghost end for receiving.
__global__ void __launch_bounds__(256)
    classify_links_kernel(int64_t npairs, const int2* __restrict__ pairs, const uint32_t* __restrict__ perm,
                          const uint32_t* __restrict__ gidx, int64_t n_own, int me, int64_t o_beg,
                          uint8_t* __restrict__ counted, uint32_t* __restrict__ sendmask, uint32_t* __restrict__ recvflag) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= npairs) return;
  const int2 p = pairs[c];
  const uint32_t li = perm[p.x], lj = perm[p.y];
  const int64_t gi = gidx[li], gj = gidx[lj];
  const int oi = static_cast<int>(gi / n_own), oj = static_cast<int>(gj / n_own);
  counted[c] = ((gi < gj ? oi : oj) == me) ? 1 : 0;
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

// (Re)build the local body set, its links and the halo lists.
int dist_build(cudaStream_t st, mb200_neighbor_links* h, mb200_comm* comm, const mb200_bodies* g, bool with_f, bool with_t) {
  auto& D = h->dist;
  const int W = comm->world, me = comm->rank;
  const int64_t N = g->n, n_own = N / W, o0 = me * n_own;
  const bool rods = g->kind != nullptr;
  D.n_global = N, D.n_own = n_own;

  // 1. owned bounds
  const int bgrid = grid_for(n_own, 256, kSMs * 2);
  Scratch<double> partials;
  MB200_TRY(partials.alloc(7 * static_cast<size_t>(bgrid), st));
  owned_bounds_kernel<<<bgrid, 256, 0, st>>>(o0, n_own, g->center, g->radius, g->length, g->kind, h->buffer, partials.p);
  MB200_LAUNCHED();
  std::vector<double> hp(7 * static_cast<size_t>(bgrid));
  MB200_TRY(d2h(st, hp.data(), partials.p, hp.size() * sizeof(double)));
  MB200_CUDA(cudaStreamSynchronize(st));
  Box6 box;
  double smax = 0.0;
  for (int d = 0; d < 3; ++d) box.lo[d] = hp[d], box.hi[d] = hp[3 + d];
  for (int k = 0; k < bgrid; ++k) {
    for (int d = 0; d < 3; ++d) {
      box.lo[d] = std::min(box.lo[d], hp[7 * k + d]);
      box.hi[d] = std::max(box.hi[d], hp[7 * k + 3 + d]);
    }
    smax = std::max(smax, hp[7 * k + 6]);
  }

  // 2. select + stable compaction
  Scratch<uint32_t> flag;
  Scratch<int64_t> pos;
  MB200_TRY(flag.alloc(N, st));
  MB200_TRY(pos.alloc(N + 1, st));
  select_local_kernel<<<grid_for(N, 256), 256, 0, st>>>(N, o0, n_own, g->center, g->radius, g->length, g->kind, h->buffer,
                                                       box, smax, flag.p);
  MB200_LAUNCHED();
  MB200_TRY(exclusive_scan_u32_to_i64(st, N, flag.p, pos.p));
  std::vector<int64_t> blk(W + 1);  // local index at which each rank's block of bodies starts
  for (int r = 0; r <= W; ++r) MB200_TRY(d2h(st, &blk[r], pos.p + r * n_own, sizeof(int64_t)));
  MB200_CUDA(cudaStreamSynchronize(st));
  const int64_t nl = blk[W];
  D.n_local = nl, D.o_beg = blk[me];
  MB200_TRY(D.gidx.reserve(nl));
  MB200_TRY(D.owned.reserve(nl));
  compact_local_kernel<<<grid_for(N, 256), 256, 0, st>>>(N, o0, n_own, flag.p, pos.p, D.gidx.p, D.owned.p);
  MB200_LAUNCHED();

  // 3. local bodies
  MB200_TRY(D.center.reserve(3 * nl));
  MB200_TRY(D.radius.reserve(nl));
  if (rods) {
    MB200_TRY(D.quat.reserve(4 * nl));
    MB200_TRY(D.length.reserve(nl));
    MB200_TRY(D.kind.reserve(nl));
  }
  if (with_f) MB200_TRY(D.fext.reserve(3 * nl));
  if (with_t) MB200_TRY(D.text.reserve(3 * nl));
  gather_local_kernel<true><<<grid_for(nl, 256), 256, 0, st>>>(nl, D.gidx.p, g->center, g->quat, g->radius, g->length,
                                                              g->kind, nullptr, nullptr, D.center.p, D.quat.p, D.radius.p,
                                                              D.length.p, D.kind.p, nullptr, nullptr);
  MB200_LAUNCHED();
  mb200_bodies lb{nl, D.center.p, rods ? D.quat.p : nullptr, D.radius.p, rods ? D.length.p : nullptr,
                  rods ? D.kind.p : nullptr};

  // 4. links with at least one owned end
  MB200_TRY(build_links(st, h, &lb, nullptr, D.owned.p));
  const int64_t C = h->npairs;

  // 5. counted flags and halo marks
  MB200_TRY(D.counted.reserve(C + 1));
  MB200_TRY(D.inv_perm.reserve(nl));
  invert_perm_kernel<<<grid_for(nl, 256), 256, 0, st>>>(nl, h->perm.p, D.inv_perm.p);
  MB200_LAUNCHED();
  Scratch<uint32_t> sendmask, recvflag;
  MB200_TRY(sendmask.alloc(n_own, st));
  MB200_TRY(recvflag.alloc(nl, st));
  MB200_CUDA(cudaMemsetAsync(sendmask.p, 0, n_own * sizeof(uint32_t), st));
  MB200_CUDA(cudaMemsetAsync(recvflag.p, 0, nl * sizeof(uint32_t), st));
  if (C > 0) {
    classify_links_kernel<<<grid_for(C, 256), 256, 0, st>>>(C, reinterpret_cast<const int2*>(h->pairs.p), h->perm.p,
                                                           D.gidx.p, n_own, me, D.o_beg, D.counted.p, sendmask.p, recvflag.p);
    MB200_LAUNCHED();
  }

  // 6. receive lists: ghosts flagged above, in local (= global index) order; each peer's ghosts are one contiguous block
  D.npeers = W - 1;
  MB200_REQUIRE(D.npeers <= 16, "at most 17 ranks");
  for (int q = 0, r = 0; r < W; ++r)
    if (r != me) D.peer[q++] = r;
  Scratch<int64_t> rpos;
  MB200_TRY(rpos.alloc(nl + 1, st));
  MB200_TRY(exclusive_scan_u32_to_i64(st, nl, recvflag.p, rpos.p));
  std::vector<int64_t> rcut(W + 1);
  for (int r = 0; r <= W; ++r) MB200_TRY(d2h(st, &rcut[r], rpos.p + blk[r], sizeof(int64_t)));
  // 7. send lists: one scan over (peer, owned body)
  Scratch<uint32_t> flags2;
  Scratch<int64_t> spos;
  Scratch<int> peers_d;
  const int64_t m2 = n_own * D.npeers;
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
  compact_halo_kernel<<<grid_for(nl, 256), 256, 0, st>>>(nl, nl, 0, recvflag.p, rpos.p, D.inv_perm.p, D.recv_idx.p);
  MB200_LAUNCHED();
  // 8. peer-memory exchange: map the ranks' velocity arenas once (any local set fits in 2 records per global body), then
  //    tell every peer in which rows of MY arena its records belong: my receive list, in the order of its send list
  MB200_TRY(peer_setup(comm, st, 2 * N));
  if (comm->peer.ok) {
    const NcclApi* nccl = nccl_api();
    MB200_TRY(D.owned_s.reserve(nl));
    owned_morton_kernel<<<grid_for(nl, 256), 256, 0, st>>>(nl, h->perm.p, D.owned.p, D.owned_s.p);
    MB200_LAUNCHED();
    MB200_TRY(D.remote_row.reserve(ns + 1));
    MB200_NCCL(nccl->GroupStart());
    for (int q = 0; q < D.npeers; ++q) {
      const int64_t cs = D.send_off[q + 1] - D.send_off[q], cr = D.recv_off[q + 1] - D.recv_off[q];
      if (cr > 0) MB200_NCCL(nccl->Send(D.recv_idx.p + D.recv_off[q], cr, ncclUint32, D.peer[q], comm->comm, st));
      if (cs > 0) MB200_NCCL(nccl->Recv(D.remote_row.p + D.send_off[q], cs, ncclUint32, D.peer[q], comm->comm, st));
    }
    MB200_NCCL(nccl->GroupEnd());
    // body-major form for the fused push of the body pass
    MB200_REQUIRE(2 * N < (int64_t(1) << 28), "too many bodies for the 28-bit ghost row index");
    Scratch<uint32_t> pcount;
    MB200_TRY(pcount.alloc(nl, st));
    MB200_TRY(D.push_ptr.reserve(nl + 1));
    MB200_TRY(D.push_dst.reserve(ns + 1));
    SendOffsets so;
    so.npeers = D.npeers;
    for (int q = 0; q < 16; ++q) so.peer[q] = D.peer[q];
    for (int q = 0; q < 17; ++q) so.off[q] = D.send_off[q];
    MB200_CUDA(cudaMemsetAsync(pcount.p, 0, nl * sizeof(uint32_t), st));
    if (ns > 0) {
      push_lists_kernel<false><<<grid_for(ns, 256), 256, 0, st>>>(ns, so, D.send_idx.p, D.remote_row.p, pcount.p, nullptr, nullptr);
      MB200_LAUNCHED();
    }
    MB200_TRY(exclusive_scan_u32_to_i64(st, nl, pcount.p, D.push_ptr.p));
    MB200_CUDA(cudaMemsetAsync(pcount.p, 0, nl * sizeof(uint32_t), st));
    if (ns > 0) {
      push_lists_kernel<true><<<grid_for(ns, 256), 256, 0, st>>>(ns, so, D.send_idx.p, D.remote_row.p, pcount.p, D.push_ptr.p,
                                                                 D.push_dst.p);
      MB200_LAUNCHED();
    }
  }
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
  MB200_REQUIRE(g->n > 0 && g->n % comm->world == 0, "the number of bodies must be a positive multiple of the world size");
  MB200_REQUIRE(g->n < (int64_t(1) << 31) && g->center && g->radius, "bad global bodies");
  if (g->kind) MB200_REQUIRE(g->quat && g->length, "kind given without quat/length");
  MB200_REQUIRE(!h->periodic.any, "a periodic box is not supported on the multi-GPU path yet (ghost selection is by "
                                  "free-space bounding boxes)");
  MB200_TRY(require_device());
  const NcclApi* nccl = nccl_api();
  if (!nccl) return MB200_ERR_COMM;
  cudaStream_t st = as_stream(stream);
  const long long launches0 = g_launches.load();
  auto& D = h->dist;
  const int W = comm->world, me = comm->rank;
  const int64_t N = g->n, n_own = N / W, o0 = me * n_own;
  const bool rods = g->kind != nullptr;

  // rebuild decision: the same on every rank (all-reduced flag)
  bool rebuild = params->force_rebuild || !D.built || !h->built || D.n_global != N;
  {
    StageScope ps(MB200_STAGE_TRIGGER, st);
    MB200_TRY(D.flag_all.reserve(2));
    if (!rebuild) {
      gather_local_kernel<false><<<grid_for(D.n_local, 256), 256, 0, st>>>(
          D.n_local, D.gidx.p, g->center, g->quat, nullptr, nullptr, nullptr, nullptr, nullptr, D.center.p, D.quat.p,
          nullptr, nullptr, nullptr, nullptr, nullptr);
      MB200_LAUNCHED();
      mb200_bodies lb{D.n_local, D.center.p, rods ? D.quat.p : nullptr, D.radius.p, rods ? D.length.p : nullptr,
                      rods ? D.kind.p : nullptr};
      MB200_TRY(links_moved_flag(st, h, &lb));
      MB200_CUDA(cudaMemcpyAsync(D.flag_all.p, h->flag.p, sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
    } else {
      const int32_t one = 1;
      MB200_CUDA(cudaMemcpyAsync(D.flag_all.p, &one, sizeof(int32_t), cudaMemcpyHostToDevice, st));
    }
    MB200_NCCL(nccl->AllReduce(D.flag_all.p, D.flag_all.p + 1, 1, ncclInt32, ncclMax, comm->comm, st));
    int32_t any = 0;
    MB200_CUDA(cudaMemcpyAsync(&any, D.flag_all.p + 1, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    MB200_CUDA(cudaStreamSynchronize(st));
    rebuild = any != 0;
  }
  if (rebuild) {
    StageScope ps(MB200_STAGE_BROADPHASE, st);
    MB200_TRY(dist_build(st, h, comm, g, F_ext != nullptr, T_ext != nullptr));
  }
  const int64_t nl = D.n_local;
  if (F_ext) MB200_TRY(D.fext.reserve(3 * nl));
  if (T_ext) MB200_TRY(D.text.reserve(3 * nl));
  if (F_col) MB200_TRY(D.fcol.reserve(3 * nl));
  if (U_out) MB200_TRY(D.uout.reserve(3 * nl));
  {
    StageScope ps(MB200_STAGE_PACK, st);
    gather_local_kernel<false><<<grid_for(nl, 256), 256, 0, st>>>(nl, D.gidx.p, g->center, g->quat, nullptr, nullptr, nullptr,
                                                                 F_ext, T_ext, D.center.p, D.quat.p, nullptr, nullptr,
                                                                 nullptr, D.fext.p, D.text.p);
    MB200_LAUNCHED();
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
  ds.halo_seq = &comm->peer.halo_seq, ds.red_seq = &comm->peer.red_seq;
  mb200_step_params p = *params;
  p.throw_on_max_iters = 0;  // reported through stats (every rank must finish the collective sequence)
  MB200_TRY(step_core(h, st, &lb, &p, F_ext ? D.fext.p : nullptr, T_ext ? D.text.p : nullptr, F_col ? D.fcol.p : nullptr,
                      nullptr, U_out ? D.uout.p : nullptr, nullptr, rebuild ? 1 : 0, launches0, &ds, stats_h));
  {
    // owned block back into the replica, then every rank's block to every replica (in-place all-gather)
    StageScope ps(MB200_STAGE_STATE_COMM, st);
    const size_t d = sizeof(double);
    MB200_CUDA(cudaMemcpyAsync(g->center + 3 * o0, D.center.p + 3 * D.o_beg, 3 * n_own * d, cudaMemcpyDeviceToDevice, st));
    if (rods) MB200_CUDA(cudaMemcpyAsync(g->quat + 4 * o0, D.quat.p + 4 * D.o_beg, 4 * n_own * d, cudaMemcpyDeviceToDevice, st));
    if (F_col) MB200_CUDA(cudaMemcpyAsync(F_col + 3 * o0, D.fcol.p + 3 * D.o_beg, 3 * n_own * d, cudaMemcpyDeviceToDevice, st));
    if (U_out) MB200_CUDA(cudaMemcpyAsync(U_out + 3 * o0, D.uout.p + 3 * D.o_beg, 3 * n_own * d, cudaMemcpyDeviceToDevice, st));
    if (W > 1) {
      MB200_NCCL(nccl->AllGather(g->center + 3 * o0, g->center, 3 * n_own, ncclDouble, comm->comm, st));
      if (rods) MB200_NCCL(nccl->AllGather(g->quat + 4 * o0, g->quat, 4 * n_own, ncclDouble, comm->comm, st));
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
