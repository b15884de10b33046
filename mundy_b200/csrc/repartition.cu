// Multi-GPU: re-partitioning of the ranks' owned blocks on the device. Bodies drift; blocks that were cut
// from a space-filling curve once stop being compact and their ghost layers grow. Collective over the communicator:
//
//   global bounding box of the owned centres (6 doubles per rank, all-gathered)
//   -> position of every owned centre along the reference's Hilbert curve (math::hilbert_3d, mundy/math/src/mundy_math/
//      Hilbert.hpp:48-85) on a 2^10 lattice: 30-bit keys
//   -> histogram of the keys' top 18 bits, summed over the ranks (one all-reduce of 2 MB): every rank derives the same
//      W - 1 splitters, the bucket boundaries nearest to equal counts
//   -> local radix sort by key, one record per body packed in key order, NCCL send / receive of the per-destination
//      slices -> stable sort of what arrived (by key; ties by source rank, then source order) -> the caller's arrays.
//
// The reference decomposes with stk::balance (RCB, third party) and migrates entities with STK's change_entity_owner;
// neither is on the path. Bodies that share a bucket (16^3 lattice cells) never split between ranks, so the blocks are
// equal up to the population of one bucket (a few tens of bodies at 10^6 per rank).
#include <cmath>
#include <mutex>
#include <vector>

#include "comm.cuh"
#include "links.cuh"
#include "sort_scan.cuh"

namespace mb200 {
namespace {

constexpr int kHilbertBits = 10;                      // lattice cells per axis: 2^10
constexpr int kKeyBits = 3 * kHilbertBits;            // 30
constexpr int kBucketBits = 18;                       // histogram resolution of the splitters
constexpr int kBucketShift = kKeyBits - kBucketBits;  // 12: a bucket = 16^3 lattice cells, contiguous along the curve
constexpr int64_t kBuckets = int64_t(1) << kBucketBits;
// doubles per travelling body: centre 3, quaternion 4, radius, length, kind, key, origin (source rank << 32 | index)
constexpr int kRecBase = 12;

// The recursion of math::hilbert_3d as an automaton: state = the frame (dr1, dr2, dr3), one of the 12 signed axis
// permutations reachable from (x, y, z); input = the octant of the cell inside the current cube (one bit per axis);
// output = which of the eight recursive calls covers that octant (3 key bits), and the frame that call is made with.
__constant__ uint8_t c_child[12 * 8];
__constant__ uint8_t c_next[12 * 8];

struct Frame {
  int d[3][3];
  bool operator==(const Frame& o) const {
    for (int a = 0; a < 3; ++a)
      for (int b = 0; b < 3; ++b)
        if (d[a][b] != o.d[a][b]) return false;
    return true;
  }
};

// Built by running the recursion's own arithmetic on unit sub-cubes (Hilbert.hpp:48-85): the k-th recursive call is made
// at current_position + (c1 dr1 + c2 dr2 + c3 dr3) snew with a frame picked from (+-dr1, +-dr2, +-dr3); the curve enters
// a cube at its minimum corner shifted by snew along every axis a director points down (:63-66).
int build_hilbert_tables(uint8_t* child, uint8_t* next) {
  struct Call {
    int c[3];     // coefficients of (dr1, dr2, dr3) locating the sub-cube
    int sg[3];    // its frame: sign and index of the parent's director, per new director
    int pick[3];
  };
  static const Call calls[8] = {
      {{0, 0, 0}, {+1, +1, +1}, {1, 2, 0}}, {{1, 0, 0}, {+1, +1, +1}, {2, 0, 1}}, {{1, 1, 0}, {+1, +1, +1}, {2, 0, 1}},
      {{0, 1, 0}, {-1, -1, +1}, {0, 1, 2}}, {{0, 1, 1}, {-1, -1, +1}, {0, 1, 2}}, {{1, 1, 1}, {-1, +1, -1}, {2, 0, 1}},
      {{1, 0, 1}, {-1, +1, -1}, {2, 0, 1}}, {{0, 0, 1}, {+1, -1, -1}, {1, 2, 0}}};
  std::vector<Frame> states;
  std::vector<Frame> todo;
  Frame eye{};
  for (int a = 0; a < 3; ++a) eye.d[a][a] = 1;
  todo.push_back(eye);
  std::vector<std::vector<Frame>> kids;
  std::vector<std::vector<int>> kid_k;
  while (!todo.empty()) {
    const Frame f = todo.back();
    todo.pop_back();
    bool seen = false;
    for (const Frame& s : states) seen = seen || s == f;
    if (seen) continue;
    states.push_back(f);
    int start[3];
    for (int ax = 0; ax < 3; ++ax) {
      start[ax] = 0;
      for (int a = 0; a < 3; ++a) start[ax] -= f.d[a][ax] < 0 ? f.d[a][ax] : 0;
    }
    std::vector<Frame> fk(8);
    std::vector<int> kk(8, -1);
    for (int k = 0; k < 8; ++k) {
      int off[3];
      for (int ax = 0; ax < 3; ++ax)
        off[ax] = start[ax] + calls[k].c[0] * f.d[0][ax] + calls[k].c[1] * f.d[1][ax] + calls[k].c[2] * f.d[2][ax];
      for (int ax = 0; ax < 3; ++ax)
        if (off[ax] < 0 || off[ax] > 1) return -1;
      const int octant = off[0] | (off[1] << 1) | (off[2] << 2);
      if (kk[octant] != -1) return -1;
      Frame c{};
      for (int a = 0; a < 3; ++a)
        for (int ax = 0; ax < 3; ++ax) c.d[a][ax] = calls[k].sg[a] * f.d[calls[k].pick[a]][ax];
      kk[octant] = k;
      fk[octant] = c;
      todo.push_back(c);
    }
    kids.push_back(fk);
    kid_k.push_back(kk);
  }
  if (states.size() != 12) return -1;
  for (size_t s = 0; s < 12; ++s)
    for (int o = 0; o < 8; ++o) {
      child[8 * s + o] = static_cast<uint8_t>(kid_k[s][o]);
      int idx = -1;
      for (size_t t = 0; t < 12; ++t)
        if (states[t] == kids[s][o]) idx = static_cast<int>(t);
      if (idx < 0) return -1;
      next[8 * s + o] = static_cast<uint8_t>(idx);
    }
  return 0;
}

int upload_hilbert_tables() {  // once per device
  static std::mutex mu;
  static bool done[64] = {};
  int dev = 0;
  MB200_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(mu);
  if (done[dev & 63]) return MB200_OK;
  uint8_t child[96], next[96];
  if (build_hilbert_tables(child, next) != 0) {
    set_error("internal: the Hilbert recursion did not close on 12 frames");
    return MB200_ERR_INVALID_ARGUMENT;
  }
  MB200_CUDA(cudaMemcpyToSymbol(c_child, child, sizeof(child)));
  MB200_CUDA(cudaMemcpyToSymbol(c_next, next, sizeof(next)));
  done[dev & 63] = true;
  return MB200_OK;
}

__device__ inline uint32_t hilbert_key(int cx, int cy, int cz) {
  uint32_t key = 0;
  int state = 0;
#pragma unroll
  for (int level = kHilbertBits - 1; level >= 0; --level) {
    const int octant = ((cx >> level) & 1) | (((cy >> level) & 1) << 1) | (((cz >> level) & 1) << 2);
    key = (key << 3) | c_child[8 * state + octant];
    state = c_next[8 * state + octant];
  }
  return key;
}

// per-CTA partials {min xyz, max xyz} of the owned centres
__global__ void __launch_bounds__(256)
    centre_bounds_kernel(int64_t n, const double* __restrict__ center, double* __restrict__ partials) {
  __shared__ double sh[33];
  const double big = 1.7976931348623157e308;
  double lo[3] = {big, big, big}, hi[3] = {-big, -big, -big};
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; k < n; k += stride) {
    const V3 c = ld3(center, k);
    lo[0] = fmin(lo[0], c.x), lo[1] = fmin(lo[1], c.y), lo[2] = fmin(lo[2], c.z);
    hi[0] = fmax(hi[0], c.x), hi[1] = fmax(hi[1], c.y), hi[2] = fmax(hi[2], c.z);
  }
  double out[6];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    out[d] = -block_max(-lo[d], sh, -big);
    out[3 + d] = block_max(hi[d], sh, -big);
  }
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < 6; ++k) partials[6 * blockIdx.x + k] = out[k];
  }
}

struct Lattice {
  double lo[3], span;
};
// cell = min(trunc((c - lo) / span * 2^bits), 2^bits - 1) per axis (mundy_b200/dist.py hilbert_order, same operation order)
__global__ void __launch_bounds__(256)
    hilbert_keys_kernel(int64_t n, const double* __restrict__ center, Lattice L, uint32_t* __restrict__ keys,
                        uint32_t* __restrict__ idx, unsigned long long* __restrict__ hist) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const V3 c = ld3(center, k);
  const double scale = static_cast<double>(1 << kHilbertBits);
  const int top = (1 << kHilbertBits) - 1;
  const int cx = min(max(static_cast<int>((c.x - L.lo[0]) / L.span * scale), 0), top);
  const int cy = min(max(static_cast<int>((c.y - L.lo[1]) / L.span * scale), 0), top);
  const int cz = min(max(static_cast<int>((c.z - L.lo[2]) / L.span * scale), 0), top);
  const uint32_t key = hilbert_key(cx, cy, cz);
  keys[k] = key;
  idx[k] = static_cast<uint32_t>(k);
  if (hist) atomicAdd(hist + (key >> kBucketShift), 1ull);
}

// first[r] = first position of the sorted keys whose bucket is >= split[r]   (r = 0 .. W; split[0] = 0, split[W] = inf)
__global__ void first_of_rank_kernel(int64_t n, const uint32_t* __restrict__ keys, int W, const uint32_t* __restrict__ split,
                                     int64_t* __restrict__ first) {
  const int r = threadIdx.x;
  if (r > W) return;
  if (r == W) {
    first[r] = n;
    return;
  }
  const uint32_t s = split[r];
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if ((keys[mid] >> kBucketShift) < s) lo = mid + 1;
    else hi = mid;
  }
  first[r] = lo;
}

__global__ void __launch_bounds__(256)
    pack_bodies_sorted_kernel(int64_t n, const uint32_t* __restrict__ idx, const uint32_t* __restrict__ keys,
                              const double* __restrict__ center, const double* __restrict__ quat,
                              const double* __restrict__ radius, const double* __restrict__ length,
                              const int32_t* __restrict__ kind, const double* __restrict__ extra, int extra_dim, int rank,
                              double* __restrict__ rec) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t i = idx[k];
  double* r = rec + k * (kRecBase + extra_dim);
  r[0] = center[3 * i], r[1] = center[3 * i + 1], r[2] = center[3 * i + 2];
  if (quat) r[3] = quat[4 * i], r[4] = quat[4 * i + 1], r[5] = quat[4 * i + 2], r[6] = quat[4 * i + 3];
  else r[3] = 1.0, r[4] = 0.0, r[5] = 0.0, r[6] = 0.0;
  r[7] = radius[i];
  r[8] = length ? length[i] : 0.0;
  r[9] = kind ? static_cast<double>(kind[i]) : 0.0;
  r[10] = static_cast<double>(keys[k]);
  r[11] = static_cast<double>((static_cast<long long>(rank) << 32) | i);
  for (int e = 0; e < extra_dim; ++e) r[kRecBase + e] = extra[i * extra_dim + e];
}

__global__ void __launch_bounds__(256)
    arrived_keys_kernel(int64_t n, const double* __restrict__ rec, int width, uint32_t* __restrict__ keys,
                        uint32_t* __restrict__ idx) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= n) return;
  keys[k] = static_cast<uint32_t>(rec[k * width + 10]);
  idx[k] = static_cast<uint32_t>(k);
}

__global__ void __launch_bounds__(256)
    unpack_bodies_kernel(int64_t n, const uint32_t* __restrict__ idx, const double* __restrict__ rec, int extra_dim,
                         double* __restrict__ center, double* __restrict__ quat, double* __restrict__ radius,
                         double* __restrict__ length, int32_t* __restrict__ kind, double* __restrict__ extra,
                         int64_t* __restrict__ origin, uint32_t* __restrict__ key_out) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const double* r = rec + static_cast<int64_t>(idx[k]) * (kRecBase + extra_dim);
  center[3 * k] = r[0], center[3 * k + 1] = r[1], center[3 * k + 2] = r[2];
  if (quat) quat[4 * k] = r[3], quat[4 * k + 1] = r[4], quat[4 * k + 2] = r[5], quat[4 * k + 3] = r[6];
  radius[k] = r[7];
  if (length) length[k] = r[8];
  if (kind) kind[k] = static_cast<int32_t>(r[9]);
  if (key_out) key_out[k] = static_cast<uint32_t>(r[10]);
  if (origin) origin[k] = static_cast<int64_t>(r[11]);
  for (int e = 0; e < extra_dim; ++e) extra[k * extra_dim + e] = r[kRecBase + e];
}

}  // namespace
}  // namespace mb200

using namespace mb200;

extern "C" int mb200_hilbert_keys(mb200_stream stream, int64_t n, const double* center, const double* lo_h, double span,
                                  uint32_t* keys_out) {
  MB200_REQUIRE(n >= 0 && lo_h && span > 0.0 && std::isfinite(span), "bad arguments");
  if (n == 0) return MB200_OK;
  MB200_REQUIRE(center && keys_out, "null pointer");
  MB200_TRY(require_device());
  MB200_TRY(upload_hilbert_tables());
  cudaStream_t st = as_stream(stream);
  Scratch<uint32_t> idx;
  MB200_TRY(idx.alloc(n, st));
  Lattice L{{lo_h[0], lo_h[1], lo_h[2]}, span};
  hilbert_keys_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, center, L, keys_out, idx.p, nullptr);
  MB200_LAUNCHED();
  return MB200_OK;
}

extern "C" int mb200_dist_repartition(mb200_comm* comm, mb200_stream stream, mb200_neighbor_links* links,
                                      const mb200_bodies* owned_in, const double* extra_in, int extra_dim,
                                      int64_t capacity, mb200_bodies* owned_out, double* extra_out,
                                      int64_t* origin_out, uint32_t* keys_out) {
  MB200_REQUIRE(comm && owned_in && owned_out, "null argument");
  MB200_REQUIRE(owned_in->n >= 0 && owned_in->n < (int64_t(1) << 31) && capacity >= 0 && extra_dim >= 0 && extra_dim <= 64,
                "bad sizes");
  if (owned_in->n > 0) MB200_REQUIRE(owned_in->center && owned_in->radius, "null centre/radius");
  if (owned_in->kind) MB200_REQUIRE(owned_in->quat && owned_in->length, "kind given without quat/length");
  MB200_REQUIRE(extra_dim == 0 || (extra_out && (extra_in || owned_in->n == 0)), "extra_dim > 0 without extra arrays");
  MB200_REQUIRE(capacity == 0 || (owned_out->center && owned_out->radius), "null output centre/radius");
  MB200_TRY(require_device());
  MB200_TRY(upload_hilbert_tables());
  const NcclApi* nccl = nccl_api();
  if (!nccl) return MB200_ERR_COMM;
  cudaStream_t st = as_stream(stream);
  const int W = comm->world, me = comm->rank;
  const int64_t n = owned_in->n;
  const int width = kRecBase + extra_dim;
  if (links) links->dist.built = false;  // the next distributed step rebuilds, whatever the counts turn out to be

  // 1. global bounding box of the centres
  double mine[6] = {1.7976931348623157e308, 1.7976931348623157e308, 1.7976931348623157e308,
                    -1.7976931348623157e308, -1.7976931348623157e308, -1.7976931348623157e308};
  if (n > 0) {
    const int bgrid = grid_for(n, 256, kSMs * 2);
    Scratch<double> partials;
    MB200_TRY(partials.alloc(6 * static_cast<size_t>(bgrid), st));
    centre_bounds_kernel<<<bgrid, 256, 0, st>>>(n, owned_in->center, partials.p);
    MB200_LAUNCHED();
    std::vector<double> hp(6 * static_cast<size_t>(bgrid));
    MB200_CUDA(cudaMemcpyAsync(hp.data(), partials.p, hp.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
    MB200_CUDA(cudaStreamSynchronize(st));
    for (int k = 0; k < bgrid; ++k)
      for (int d = 0; d < 3; ++d) {
        mine[d] = std::fmin(mine[d], hp[6 * k + d]);
        mine[3 + d] = std::fmax(mine[3 + d], hp[6 * k + 3 + d]);
      }
  }
  std::vector<double> all(6 * static_cast<size_t>(W));
  MB200_TRY(comm_allgather_host(comm, st, mine, all.data(), sizeof(mine)));
  Lattice L{};
  double hi[3];
  for (int d = 0; d < 3; ++d) {
    L.lo[d] = all[d], hi[d] = all[3 + d];
    for (int r = 1; r < W; ++r) L.lo[d] = std::fmin(L.lo[d], all[6 * r + d]), hi[d] = std::fmax(hi[d], all[6 * r + 3 + d]);
  }
  bool any = L.lo[0] <= hi[0];
  for (int d = 0; d < 3 && any; ++d) MB200_REQUIRE(std::isfinite(L.lo[d]) && std::isfinite(hi[d]), "non-finite body centre");
  L.span = any ? std::fmax(hi[0] - L.lo[0], std::fmax(hi[1] - L.lo[1], hi[2] - L.lo[2])) : 1.0;
  if (!(L.span > 0.0)) L.span = 1.0;
  if (!any) L.lo[0] = L.lo[1] = L.lo[2] = 0.0;

  // 2. keys and the global bucket histogram
  Scratch<uint32_t> keys, keys_alt, idx, idx_alt;
  Scratch<unsigned long long> hist;
  const size_t nn = static_cast<size_t>(n > 0 ? n : 1);
  MB200_TRY(keys.alloc(nn, st));
  MB200_TRY(keys_alt.alloc(nn, st));
  MB200_TRY(idx.alloc(nn, st));
  MB200_TRY(idx_alt.alloc(nn, st));
  MB200_TRY(hist.alloc(kBuckets, st));
  MB200_CUDA(cudaMemsetAsync(hist.p, 0, kBuckets * sizeof(unsigned long long), st));
  if (n > 0) {
    hilbert_keys_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, owned_in->center, L, keys.p, idx.p, hist.p);
    MB200_LAUNCHED();
  }
  MB200_NCCL(nccl->AllReduce(hist.p, hist.p, kBuckets, ncclUint64, ncclSum, comm->comm, st));
  std::vector<unsigned long long> hh(kBuckets);
  MB200_CUDA(cudaMemcpyAsync(hh.data(), hist.p, kBuckets * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
  MB200_CUDA(cudaStreamSynchronize(st));

  // 3. splitters: rank r takes the buckets [split[r], split[r + 1]); split[r] = the bucket boundary whose count of
  //    bodies below it is nearest to r N / W from above (the same arithmetic on every rank)
  unsigned long long total = 0;
  for (int64_t b = 0; b < kBuckets; ++b) total += hh[b];
  std::vector<uint32_t> split(W + 1, 0);
  {
    unsigned long long below = 0;
    int r = 1;
    for (int64_t b = 0; b <= kBuckets && r < W; ++b) {
      while (r < W && below >= (total * static_cast<unsigned long long>(r)) / static_cast<unsigned long long>(W)) {
        split[r] = static_cast<uint32_t>(b);
        ++r;
      }
      if (b < kBuckets) below += hh[b];
    }
    for (; r < W; ++r) split[r] = static_cast<uint32_t>(kBuckets);
    split[W] = static_cast<uint32_t>(kBuckets);
  }

  // 4. local sort by key; send slices
  if (n > 0) MB200_TRY(radix_sort_u32(st, n, keys.p, keys_alt.p, idx.p, idx_alt.p, kKeyBits));
  Scratch<uint32_t> split_d;
  Scratch<int64_t> first_d;
  MB200_TRY(split_d.alloc(W + 1, st));
  MB200_TRY(first_d.alloc(W + 1, st));
  MB200_CUDA(cudaMemcpyAsync(split_d.p, split.data(), (W + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
  first_of_rank_kernel<<<1, 32, 0, st>>>(n, keys.p, W, split_d.p, first_d.p);
  MB200_LAUNCHED();
  std::vector<int64_t> first(W + 1);
  MB200_CUDA(cudaMemcpyAsync(first.data(), first_d.p, (W + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  MB200_CUDA(cudaStreamSynchronize(st));
  std::vector<int64_t> send_cnt(W), cnt_all(static_cast<size_t>(W) * W);
  for (int r = 0; r < W; ++r) send_cnt[r] = first[r + 1] - first[r];
  MB200_TRY(comm_allgather_host(comm, st, send_cnt.data(), cnt_all.data(), W * sizeof(int64_t)));
  std::vector<int64_t> recv_off(W + 1, 0);
  for (int r = 0; r < W; ++r) recv_off[r + 1] = recv_off[r] + cnt_all[static_cast<size_t>(r) * W + me];
  const int64_t n_out = recv_off[W];

  // 5. pack in key order, exchange (own slice: a device copy)
  Scratch<double> sendbuf, recvbuf;
  MB200_TRY(sendbuf.alloc(nn * width, st));
  MB200_TRY(recvbuf.alloc(static_cast<size_t>(n_out > 0 ? n_out : 1) * width, st));
  if (n > 0) {
    pack_bodies_sorted_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, idx.p, keys.p, owned_in->center, owned_in->quat,
                                                               owned_in->radius, owned_in->length, owned_in->kind,
                                                               extra_in, extra_dim, me, sendbuf.p);
    MB200_LAUNCHED();
  }
  MB200_NCCL(nccl->GroupStart());
  for (int r = 0; r < W; ++r) {
    if (r == me) continue;
    const int64_t cs = send_cnt[r], cr = recv_off[r + 1] - recv_off[r];
    if (cs > 0)
      MB200_NCCL(nccl->Send(sendbuf.p + first[r] * width, static_cast<size_t>(cs) * width, ncclDouble, r, comm->comm, st));
    if (cr > 0)
      MB200_NCCL(nccl->Recv(recvbuf.p + recv_off[r] * width, static_cast<size_t>(cr) * width, ncclDouble, r, comm->comm, st));
  }
  MB200_NCCL(nccl->GroupEnd());
  if (send_cnt[me] > 0)
    MB200_CUDA(cudaMemcpyAsync(recvbuf.p + recv_off[me] * width, sendbuf.p + first[me] * width,
                               static_cast<size_t>(send_cnt[me]) * width * sizeof(double), cudaMemcpyDeviceToDevice, st));
  owned_out->n = n_out;
  if (n_out > capacity) {
    MB200_CUDA(cudaStreamSynchronize(st));
    set_error("mb200_dist_repartition: this rank receives " + std::to_string(n_out) + " bodies, capacity " +
              std::to_string(capacity));
    return MB200_ERR_CAPACITY;
  }

  // 6. what arrived, in key order (stable: ties by source rank, then by the source's order)
  if (n_out > 0) {
    Scratch<uint32_t> k2, k2_alt, i2, i2_alt;
    MB200_TRY(k2.alloc(n_out, st));
    MB200_TRY(k2_alt.alloc(n_out, st));
    MB200_TRY(i2.alloc(n_out, st));
    MB200_TRY(i2_alt.alloc(n_out, st));
    arrived_keys_kernel<<<grid_for(n_out, 256), 256, 0, st>>>(n_out, recvbuf.p, width, k2.p, i2.p);
    MB200_LAUNCHED();
    MB200_TRY(radix_sort_u32(st, n_out, k2.p, k2_alt.p, i2.p, i2_alt.p, kKeyBits));
    unpack_bodies_kernel<<<grid_for(n_out, 256), 256, 0, st>>>(n_out, i2.p, recvbuf.p, extra_dim, owned_out->center,
                                                              owned_out->quat, const_cast<double*>(owned_out->radius),
                                                              const_cast<double*>(owned_out->length),
                                                              const_cast<int32_t*>(owned_out->kind), extra_out, origin_out,
                                                              keys_out);
    MB200_LAUNCHED();
  }
  MB200_CUDA(cudaStreamSynchronize(st));
  return MB200_OK;
}
