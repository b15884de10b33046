// Coalesced, 128-bit-vectorised movement of xyz-interleaved (K doubles per item) tiles between global and shared
// memory. The reference's layouts are AoS (3/4/6 doubles per item); reading them "one item per thread" makes every
// warp load touch 24-48 sectors. Instead a CTA moves the tile as a flat, contiguous run of doubles (double2 per
// thread, fully coalesced), and threads then pick their item out of shared memory.
#pragma once
#include "common.cuh"

namespace mb200 {

constexpr int kTile = 256;  // items per CTA tile == threads per CTA for the per-item kernels

// Copies K*cnt doubles src[K*base ...] -> sh[0 ...]. K*base*8 is 16-byte aligned whenever base is a multiple of kTile.
template <int K>
__device__ inline void load_tile(double* __restrict__ sh, const double* __restrict__ src, int64_t base, int cnt) {
  const double* g = src + K * base;
  const int total = K * cnt;
  const bool aligned = (reinterpret_cast<uintptr_t>(g) & 15) == 0;
  if (aligned) {
    const int nv = total >> 1;
    const double2* g2 = reinterpret_cast<const double2*>(g);
    double2* s2 = reinterpret_cast<double2*>(sh);
    for (int k = threadIdx.x; k < nv; k += blockDim.x) s2[k] = __ldg(g2 + k);
    if ((total & 1) && threadIdx.x == 0) sh[total - 1] = g[total - 1];
  } else {
    for (int k = threadIdx.x; k < total; k += blockDim.x) sh[k] = g[k];
  }
}

template <int K>
__device__ inline void store_tile(double* __restrict__ dst, const double* __restrict__ sh, int64_t base, int cnt) {
  double* g = dst + K * base;
  const int total = K * cnt;
  const bool aligned = (reinterpret_cast<uintptr_t>(g) & 15) == 0;
  if (aligned) {
    const int nv = total >> 1;
    double2* g2 = reinterpret_cast<double2*>(g);
    const double2* s2 = reinterpret_cast<const double2*>(sh);
    for (int k = threadIdx.x; k < nv; k += blockDim.x) g2[k] = s2[k];
    if ((total & 1) && threadIdx.x == 0) g[total - 1] = sh[total - 1];
  } else {
    for (int k = threadIdx.x; k < total; k += blockDim.x) g[k] = sh[k];
  }
}

}  // namespace mb200
