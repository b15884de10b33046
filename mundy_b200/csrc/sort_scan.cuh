// Hand-written device primitives used by the broad phase and the incidence build:
//   * exclusive_scan_i64 / exclusive_scan_i32 : 3-phase block scan
//   * radix_sort_pairs / radix_sort_keys      : stable LSD radix sort, 8-bit digits, over the low `bits` bits
#pragma once
#include "common.cuh"

namespace mb200 {

// out[k] = sum_{m<k} in[m] for k in [0, n]; out has n + 1 entries (out[n] = total). in may alias out (n entries used).
int exclusive_scan_u32_to_i64(cudaStream_t st, int64_t n, const uint32_t* in, int64_t* out);

// Stable sort of (keys, vals) by the low `bits` bits of the key. Results end up in keys/vals (the alt buffers are
// scratch of the same size). vals/vals_alt may be NULL (keys only).
int radix_sort_u32(cudaStream_t st, int64_t n, uint32_t* keys, uint32_t* keys_alt, uint32_t* vals, uint32_t* vals_alt,
                   int bits);
int radix_sort_u64(cudaStream_t st, int64_t n, uint64_t* keys, uint64_t* keys_alt, uint32_t* vals, uint32_t* vals_alt,
                   int bits);

}  // namespace mb200
