// Hand-written scan and LSD radix sort (no CUB / thrust on the product path).
#include "sort_scan.cuh"

namespace mb200 {
namespace {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;

// Block-wide exclusive scan of one value per thread; returns the exclusive prefix, *total = block sum.
template <typename T>
__device__ inline T block_exclusive_scan(T v, T* sh /* >= 32 */, T* total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  T inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    T up = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += up;
  }
  __syncthreads();
  if (lane == 31) sh[w] = inc;
  __syncthreads();
  if (w == 0) {
    const int nw = blockDim.x >> 5;
    T s = (lane < nw) ? sh[lane] : T(0);
    T sinc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      T up = __shfl_up_sync(0xffffffffu, sinc, o);
      if (lane >= o) sinc += up;
    }
    sh[lane] = sinc - s;  // exclusive warp offsets
    if (lane == 31) sh[32] = sinc;
  }
  __syncthreads();
  const T off = sh[w];
  *total = sh[32];
  return off + inc - v;
}

template <typename TIn, typename TOut>
__global__ void __launch_bounds__(kScanThreads) scan_tile_sums(int64_t n, const TIn* __restrict__ in,
                                                               TOut* __restrict__ sums) {
  __shared__ TOut sh[33];
  const int64_t base = static_cast<int64_t>(blockIdx.x) * kScanTile;
  TOut acc = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    const int64_t idx = base + k * kScanThreads + threadIdx.x;
    if (idx < n) acc += static_cast<TOut>(in[idx]);
  }
  TOut total;
  block_exclusive_scan<TOut>(acc, sh, &total);
  if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

// Single CTA: exclusive scan of sums[0..m) in place, sums[m] = grand total.
template <typename TOut>
__global__ void __launch_bounds__(kScanThreads) scan_block_sums(int64_t m, TOut* __restrict__ sums) {
  __shared__ TOut sh[33];
  TOut carry = 0;
  for (int64_t base = 0; base < m; base += kScanThreads) {
    const int64_t idx = base + threadIdx.x;
    const TOut v = idx < m ? sums[idx] : TOut(0);
    TOut total;
    const TOut ex = block_exclusive_scan<TOut>(v, sh, &total);
    if (idx < m) sums[idx] = carry + ex;
    carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) sums[m] = carry;
}

template <typename TIn, typename TOut>
__global__ void __launch_bounds__(kScanThreads) scan_apply(int64_t n, const TIn* __restrict__ in,
                                                           const TOut* __restrict__ sums, TOut* __restrict__ out,
                                                           int64_t nblocks) {
  __shared__ TOut sh[33];
  const int64_t base = static_cast<int64_t>(blockIdx.x) * kScanTile;
  // blocked arrangement inside the tile: thread t owns items [t*kScanItems, (t+1)*kScanItems)
  TIn v[kScanItems];
  TOut acc = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    const int64_t idx = base + threadIdx.x * kScanItems + k;
    v[k] = idx < n ? in[idx] : TIn(0);
    acc += static_cast<TOut>(v[k]);
  }
  TOut total;
  TOut ex = block_exclusive_scan<TOut>(acc, sh, &total) + sums[blockIdx.x];
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    const int64_t idx = base + threadIdx.x * kScanItems + k;
    if (idx < n) out[idx] = ex;
    ex += static_cast<TOut>(v[k]);
  }
  if (blockIdx.x == nblocks - 1 && threadIdx.x == 0) out[n] = sums[nblocks];
}

template <typename TIn, typename TOut>
int exclusive_scan_impl(cudaStream_t st, int64_t n, const TIn* in, TOut* out) {
  if (n == 0) {
    MB200_CUDA(cudaMemsetAsync(out, 0, sizeof(TOut), st));
    return MB200_OK;
  }
  const int64_t nblocks = (n + kScanTile - 1) / kScanTile;
  Scratch<TOut> sums;
  MB200_TRY(sums.alloc(nblocks + 1, st));
  scan_tile_sums<TIn, TOut><<<static_cast<int>(nblocks), kScanThreads, 0, st>>>(n, in, sums.p);
  MB200_LAUNCHED();
  scan_block_sums<TOut><<<1, kScanThreads, 0, st>>>(nblocks, sums.p);
  MB200_LAUNCHED();
  scan_apply<TIn, TOut><<<static_cast<int>(nblocks), kScanThreads, 0, st>>>(n, in, sums.p, out, nblocks);
  MB200_LAUNCHED();
  return MB200_OK;
}

}  // namespace

int exclusive_scan_u32_to_i64(cudaStream_t st, int64_t n, const uint32_t* in, int64_t* out) {
  return exclusive_scan_impl<uint32_t, int64_t>(st, n, in, out);
}

}  // namespace mb200
