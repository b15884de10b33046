// Stable LSD radix sort, 8-bit digits, hand-written (histogram -> scan -> ranked scatter per pass).
// Used for the Morton ordering of bodies (32-bit cell keys + body index), for the body-major incidence lists of the
// contact operator (key = body, value = 2*contact+side) and for canonical (i,j) pair ordering (64-bit keys).
//
// Tile = 256 threads x 8 rounds. Ranking inside a tile is stable in tile order (round-major, then warp, then lane):
// per (round, warp) a MATCH.ANY groups equal digits, the group leader records the group size in shared memory, one
// thread per digit prefix-sums the 64 (round, warp) counters, and each key lands at
//   global_base[digit][tile] + prefix[round][warp][digit] + rank_in_group.
#include "sort_scan.cuh"

namespace mb200 {
namespace {

constexpr int kRsThreads = 256;
constexpr int kRsRounds = 8;
constexpr int kRsTile = kRsThreads * kRsRounds;
constexpr int kRsWarps = kRsThreads / 32;

template <typename K>
__global__ void __launch_bounds__(kRsThreads) rs_hist_kernel(int64_t n, const K* __restrict__ keys, int shift,
                                                             uint32_t* __restrict__ block_hist, int64_t ntiles) {
  __shared__ uint32_t h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  const int64_t base = static_cast<int64_t>(blockIdx.x) * kRsTile;
#pragma unroll
  for (int r = 0; r < kRsRounds; ++r) {
    const int64_t idx = base + r * kRsThreads + threadIdx.x;
    if (idx < n) atomicAdd(&h[static_cast<uint32_t>(keys[idx] >> shift) & 0xffu], 1u);
  }
  __syncthreads();
  block_hist[static_cast<int64_t>(threadIdx.x) * ntiles + blockIdx.x] = h[threadIdx.x];
}

template <typename K, bool HAS_VALS>
__global__ void __launch_bounds__(kRsThreads)
    rs_scatter_kernel(int64_t n, const K* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                      K* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int shift,
                      const int64_t* __restrict__ scanned_hist, int64_t ntiles) {
  __shared__ uint16_t cnt[kRsRounds * kRsWarps][256];  // 32 KB
  __shared__ int64_t gbase[256];
  {
    uint4* z = reinterpret_cast<uint4*>(&cnt[0][0]);
    constexpr int nz = sizeof(cnt) / sizeof(uint4);
    for (int k = threadIdx.x; k < nz; k += kRsThreads) z[k] = make_uint4(0, 0, 0, 0);
  }
  gbase[threadIdx.x] = scanned_hist[static_cast<int64_t>(threadIdx.x) * ntiles + blockIdx.x];
  __syncthreads();

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const int64_t base = static_cast<int64_t>(blockIdx.x) * kRsTile;
  K key[kRsRounds];
  uint32_t rank[kRsRounds];  // rank inside the (round, warp) group of equal digits
#pragma unroll
  for (int r = 0; r < kRsRounds; ++r) {
    const int64_t idx = base + r * kRsThreads + threadIdx.x;
    const bool active = idx < n;
    key[r] = active ? keys_in[idx] : K(0);
    const uint32_t d = active ? (static_cast<uint32_t>(key[r] >> shift) & 0xffu) : 0x100u;
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    rank[r] = __popc(peers & lt_mask);
    if (active && rank[r] == 0) cnt[r * kRsWarps + warp][d] = static_cast<uint16_t>(__popc(peers));
  }
  __syncthreads();
  {  // exclusive prefix over the 64 (round, warp) counters of digit threadIdx.x
    uint32_t run = 0;
#pragma unroll 8
    for (int e = 0; e < kRsRounds * kRsWarps; ++e) {
      const uint32_t c = cnt[e][threadIdx.x];
      cnt[e][threadIdx.x] = static_cast<uint16_t>(run);
      run += c;
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < kRsRounds; ++r) {
    const int64_t idx = base + r * kRsThreads + threadIdx.x;
    if (idx < n) {
      const uint32_t d = static_cast<uint32_t>(key[r] >> shift) & 0xffu;
      const int64_t pos = gbase[d] + cnt[r * kRsWarps + warp][d] + rank[r];
      keys_out[pos] = key[r];
      if (HAS_VALS) vals_out[pos] = vals_in[idx];
    }
  }
}

template <typename K>
int radix_sort_impl(cudaStream_t st, int64_t n, K* keys, K* keys_alt, uint32_t* vals, uint32_t* vals_alt, int bits) {
  if (n <= 1 || bits <= 0) return MB200_OK;
  const int64_t ntiles = (n + kRsTile - 1) / kRsTile;
  Scratch<uint32_t> hist;
  Scratch<int64_t> scanned;
  MB200_TRY(hist.alloc(256 * ntiles, st));
  MB200_TRY(scanned.alloc(256 * ntiles + 1, st));
  const int passes = (bits + 7) / 8;
  K* kin = keys;
  K* kout = keys_alt;
  uint32_t* vin = vals;
  uint32_t* vout = vals_alt;
  for (int p = 0; p < passes; ++p) {
    const int shift = 8 * p;
    rs_hist_kernel<K><<<static_cast<int>(ntiles), kRsThreads, 0, st>>>(n, kin, shift, hist.p, ntiles);
    MB200_LAUNCHED();
    MB200_TRY(exclusive_scan_u32_to_i64(st, 256 * ntiles, hist.p, scanned.p));
    if (vals) {
      rs_scatter_kernel<K, true><<<static_cast<int>(ntiles), kRsThreads, 0, st>>>(n, kin, vin, kout, vout, shift,
                                                                                 scanned.p, ntiles);
    } else {
      rs_scatter_kernel<K, false><<<static_cast<int>(ntiles), kRsThreads, 0, st>>>(n, kin, nullptr, kout, nullptr,
                                                                                  shift, scanned.p, ntiles);
    }
    MB200_LAUNCHED();
    K* tk = kin;
    kin = kout;
    kout = tk;
    uint32_t* tv = vin;
    vin = vout;
    vout = tv;
  }
  if (kin != keys) {  // odd number of passes: result is in the alt buffers
    MB200_CUDA(cudaMemcpyAsync(keys, kin, n * sizeof(K), cudaMemcpyDeviceToDevice, st));
    if (vals) MB200_CUDA(cudaMemcpyAsync(vals, vin, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
  }
  return MB200_OK;
}

}  // namespace

int radix_sort_u32(cudaStream_t st, int64_t n, uint32_t* keys, uint32_t* keys_alt, uint32_t* vals, uint32_t* vals_alt,
                   int bits) {
  return radix_sort_impl<uint32_t>(st, n, keys, keys_alt, vals, vals_alt, bits > 32 ? 32 : bits);
}
int radix_sort_u64(cudaStream_t st, int64_t n, uint64_t* keys, uint64_t* keys_alt, uint32_t* vals, uint32_t* vals_alt,
                   int bits) {
  return radix_sort_impl<uint64_t>(st, n, keys, keys_alt, vals, vals_alt, bits > 64 ? 64 : bits);
}

}  // namespace mb200
