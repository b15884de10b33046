// D. Matrix-free contact operator entry points (caller layout) and the body-major incidence build.
#include "internal.cuh"
#include "lcp.cuh"
#include "sort_scan.cuh"

namespace mb200 {
namespace {

// A body index outside [0, n) (caller-supplied pair lists reach this through mb200_resolve_collisions / mb200_apply_*)
// raises *bad and is counted on body 0, so that nothing is written out of bounds; the entry point then fails.
__global__ void __launch_bounds__(256) incidence_keys_kernel(int64_t m, int64_t n, const int32_t* __restrict__ pairs_flat,
                                                             uint32_t* __restrict__ keys, uint32_t* __restrict__ vals,
                                                             uint32_t* __restrict__ counts, int32_t* __restrict__ bad) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= m) return;
  int64_t b = pairs_flat[k];
  if (b < 0 || b >= n) {
    *bad = 1;
    b = 0;
  }
  keys[k] = static_cast<uint32_t>(b);
  vals[k] = static_cast<uint32_t>(k);  // 2 * contact + side
  atomicAdd(counts + b, 1u);
}

// COO -> CRS with empty slots (entity < 0): they get the sentinel key n and sort behind every entity.
__global__ void __launch_bounds__(256) coo_keys_kernel(int64_t m, int64_t n, const int32_t* __restrict__ linked,
                                                       uint32_t* __restrict__ keys, uint32_t* __restrict__ vals,
                                                       uint32_t* __restrict__ counts, int32_t* __restrict__ bad) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= m) return;
  const int32_t b = linked[k];
  if (b >= n) *bad = 1;
  const bool empty = b < 0 || b >= n;
  keys[k] = empty ? static_cast<uint32_t>(n) : static_cast<uint32_t>(b);
  vals[k] = static_cast<uint32_t>(k);  // 2 * link + slot
  if (!empty) atomicAdd(counts + b, 1u);
}
__global__ void __launch_bounds__(256) invert_list_kernel(int64_t m, const uint32_t* __restrict__ list,
                                                          uint32_t* __restrict__ epos) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k < m) epos[list[k]] = static_cast<uint32_t>(k);
}
__global__ void __launch_bounds__(256) crs_split_kernel(int64_t m, const uint32_t* __restrict__ vals,
                                                        int32_t* __restrict__ links, int32_t* __restrict__ slots) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= m) return;
  const uint32_t v = vals[k];
  links[k] = static_cast<int32_t>(v >> 1);
  if (slots) slots[k] = static_cast<int32_t>(v & 1u);
}

template <bool SIX>
__global__ void __launch_bounds__(128)
    apply_D_kernel(int64_t n, const int64_t* __restrict__ ptr, const uint32_t* __restrict__ list,
                   const double* __restrict__ gamma, const double* __restrict__ normal,
                   const double* __restrict__ arm_i, const double* __restrict__ arm_j, double* __restrict__ F,
                   double* __restrict__ T) {
  const int64_t b = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (b >= n) return;
  V3 f, t;
  body_wrench<SIX>(b, ptr, list, gamma, normal, arm_i, arm_j, f, t);
  st3(F, b, f);
  if (SIX) st3(T, b, t);
}

__global__ void __launch_bounds__(256)
    apply_M_kernel(int64_t n, const double* __restrict__ quat, const double* __restrict__ radius,
                   const int32_t* __restrict__ kind, double mu, const double* __restrict__ F,
                   const double* __restrict__ T, double* __restrict__ U, double* __restrict__ W) {
  const int64_t b = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (b >= n) return;
  const Mob m = mobility_of(radius[b], mu);
  const bool rod = quat && kind && kind[b] == MB200_SPHEROCYLINDER;
  const Q4 q = rod ? ldq(quat, b) : Q4{1.0, 0.0, 0.0, 0.0};
  const V3 f = ld3(F, b);
  const V3 t = T ? ld3(T, b) : V3{0.0, 0.0, 0.0};
  V3 u, w;
  body_mobility(m, rod, q, f, t, u, w);
  st3(U, b, u);
  if (W) st3(W, b, w);
}

template <bool SIX>
__global__ void __launch_bounds__(256)
    apply_DT_kernel(int64_t npairs, const int2* __restrict__ pairs, const double* __restrict__ normal,
                    const double* __restrict__ arm_i, const double* __restrict__ arm_j, const double* __restrict__ U,
                    const double* __restrict__ W, double scale, double* __restrict__ gdot) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= npairs) return;
  const int2 p = pairs[c];
  const V3 n = ld3(normal, c);
  V3 wi{0, 0, 0}, wj{0, 0, 0}, ai{0, 0, 0}, aj{0, 0, 0};
  if (SIX) {
    wi = ld3(W, p.x), wj = ld3(W, p.y);
    ai = ld3(arm_i, c), aj = ld3(arm_j, c);
  }
  const double g = contact_rate<SIX>(n, ld3(U, p.x), ld3(U, p.y), wi, wj, ai, aj);
  gdot[c] = (scale == 1.0) ? g : g * scale;
}

int check_op(const mb200_contact_op* op) {
  MB200_REQUIRE(op && op->n >= 0 && op->npairs >= 0, "bad contact op sizes");
  if (op->npairs > 0) MB200_REQUIRE(op->pairs && op->normal, "contact op: null pairs/normal");
  MB200_REQUIRE((op->arm_i == nullptr) == (op->arm_j == nullptr), "contact op: arm_i/arm_j must both be set or NULL");
  MB200_REQUIRE(op->n < (int64_t(1) << 31) && op->npairs < (int64_t(1) << 30), "contact op too large for int32 ids");
  return MB200_OK;
}

}  // namespace

int build_incidence(cudaStream_t st, int64_t n, int64_t npairs, const int32_t* pairs, Incidence* inc) {
  const int64_t m = 2 * npairs;
  MB200_TRY(inc->ptr.alloc(n + 1, st));
  MB200_TRY(inc->list.alloc(m > 0 ? m : 1, st));
  Scratch<uint32_t> counts, keys, keys_alt, vals_alt;
  MB200_TRY(counts.alloc(n > 0 ? n : 1, st));
  MB200_TRY(inc->bad.alloc(1, st));
  MB200_CUDA(cudaMemsetAsync(counts.p, 0, (n > 0 ? n : 1) * sizeof(uint32_t), st));
  MB200_CUDA(cudaMemsetAsync(inc->bad.p, 0, sizeof(int32_t), st));
  if (m > 0) {
    MB200_REQUIRE(n > 0, "pairs given for an empty body set");
    MB200_TRY(keys.alloc(m, st));
    MB200_TRY(keys_alt.alloc(m, st));
    MB200_TRY(vals_alt.alloc(m, st));
    incidence_keys_kernel<<<grid_for(m, 256), 256, 0, st>>>(m, n, pairs, keys.p, inc->list.p, counts.p, inc->bad.p);
    MB200_LAUNCHED();
    int bits = 1;
    while ((int64_t(1) << bits) < n) ++bits;
    MB200_TRY(radix_sort_u32(st, m, keys.p, keys_alt.p, inc->list.p, vals_alt.p, bits));
    MB200_TRY(inc->epos.alloc(m, st));
    invert_list_kernel<<<grid_for(m, 256), 256, 0, st>>>(m, inc->list.p, inc->epos.p);
    MB200_LAUNCHED();
  }
  MB200_TRY(exclusive_scan_u32_to_i64(st, n, counts.p, inc->ptr.p));
  return MB200_OK;
}

// The same lists for a pair list that is sorted by (i, j) with i < j and whose row pointers are known (the link build's
// own output): no sort. In ascending (contact, side) order a body's entries are first the contacts in which it is the
// SECOND body (they belong to rows i < b, hence to smaller contact indices), then its own row, which is contiguous. The
// row part is placed directly; the second-body part is claimed slot by slot with an atomic counter and then put in
// ascending order per body (a few entries each), so the result does not depend on the order of the atomics.
namespace {

__global__ void __launch_bounds__(256) second_counts_kernel(int64_t npairs, const int2* __restrict__ pairs,
                                                            uint32_t* __restrict__ jcount) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c < npairs) atomicAdd(jcount + pairs[c].y, 1u);
}
__global__ void __launch_bounds__(256) degree_kernel(int64_t n, const int64_t* __restrict__ rowptr,
                                                     const uint32_t* __restrict__ jcount, uint32_t* __restrict__ deg) {
  const int64_t b = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (b < n) deg[b] = jcount[b] + static_cast<uint32_t>(rowptr[b + 1] - rowptr[b]);
}
__global__ void __launch_bounds__(256)
    incidence_fill_kernel(int64_t npairs, const int2* __restrict__ pairs, const int64_t* __restrict__ rowptr,
                          const int64_t* __restrict__ ptr, const uint32_t* __restrict__ jcount,
                          uint32_t* __restrict__ jfill, uint32_t* __restrict__ list, uint32_t* __restrict__ epos) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= npairs) return;
  const int2 p = pairs[c];
  const int64_t pos_i = ptr[p.x] + jcount[p.x] + (c - rowptr[p.x]);
  list[pos_i] = static_cast<uint32_t>(2 * c);
  epos[2 * c] = static_cast<uint32_t>(pos_i);
  list[ptr[p.y] + atomicAdd(jfill + p.y, 1u)] = static_cast<uint32_t>(2 * c + 1);
}
// ascending order of a body's second-body entries (insertion sort; heap sort for long lists), then their positions
__global__ void __launch_bounds__(128)
    incidence_order_kernel(int64_t n, const int64_t* __restrict__ ptr, const uint32_t* __restrict__ jcount,
                           uint32_t* __restrict__ list, uint32_t* __restrict__ epos) {
  const int64_t b = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (b >= n) return;
  const int k = static_cast<int>(jcount[b]);
  uint32_t* a = list + ptr[b];
  if (k <= 24) {
    for (int x = 1; x < k; ++x) {
      const uint32_t v = a[x];
      int y = x - 1;
      while (y >= 0 && a[y] > v) {
        a[y + 1] = a[y];
        --y;
      }
      a[y + 1] = v;
    }
  } else {
    auto sift = [&](int root, int end) {
      for (;;) {
        int child = 2 * root + 1;
        if (child > end) break;
        if (child + 1 <= end && a[child] < a[child + 1]) ++child;
        if (a[root] >= a[child]) break;
        const uint32_t t = a[root];
        a[root] = a[child], a[child] = t;
        root = child;
      }
    };
    for (int start = (k - 2) / 2; start >= 0; --start) sift(start, k - 1);
    for (int end = k - 1; end > 0; --end) {
      const uint32_t t = a[0];
      a[0] = a[end], a[end] = t;
      sift(0, end - 1);
    }
  }
  const int64_t base = ptr[b];
  for (int x = 0; x < k; ++x) epos[a[x]] = static_cast<uint32_t>(base + x);
}

}  // namespace

int build_incidence_sorted(cudaStream_t st, int64_t n, int64_t npairs, const int32_t* pairs, const int64_t* rowptr,
                           Incidence* inc) {
  const int64_t m = 2 * npairs;
  MB200_TRY(inc->ptr.alloc(n + 1, st));
  MB200_TRY(inc->list.alloc(m > 0 ? m : 1, st));
  MB200_TRY(inc->epos.alloc(m > 0 ? m : 1, st));
  MB200_TRY(inc->bad.alloc(1, st));
  MB200_CUDA(cudaMemsetAsync(inc->bad.p, 0, sizeof(int32_t), st));
  Scratch<uint32_t> jc, deg;  // jc: [0, n) second-body counts, [n, 2n) fill counters
  const size_t nn = static_cast<size_t>(n > 0 ? n : 1);
  MB200_TRY(jc.alloc(2 * nn, st));
  MB200_TRY(deg.alloc(nn, st));
  MB200_CUDA(cudaMemsetAsync(jc.p, 0, 2 * nn * sizeof(uint32_t), st));
  const int2* p2 = reinterpret_cast<const int2*>(pairs);
  if (npairs > 0) {
    second_counts_kernel<<<grid_for(npairs, 256), 256, 0, st>>>(npairs, p2, jc.p);
    MB200_LAUNCHED();
  }
  if (n > 0) {
    degree_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, rowptr, jc.p, deg.p);
    MB200_LAUNCHED();
  } else {
    MB200_CUDA(cudaMemsetAsync(deg.p, 0, sizeof(uint32_t), st));
  }
  MB200_TRY(exclusive_scan_u32_to_i64(st, n, deg.p, inc->ptr.p));
  if (npairs > 0) {
    incidence_fill_kernel<<<grid_for(npairs, 256), 256, 0, st>>>(npairs, p2, rowptr, inc->ptr.p, jc.p, jc.p + nn,
                                                                inc->list.p, inc->epos.p);
    MB200_LAUNCHED();
    incidence_order_kernel<<<grid_for(n, 128), 128, 0, st>>>(n, inc->ptr.p, jc.p, inc->list.p, inc->epos.p);
    MB200_LAUNCHED();
  }
  return MB200_OK;
}

// For the entry points that take the caller's pair list: synchronises and fails if an index was out of range.
int check_incidence(cudaStream_t st, const Incidence& inc) {
  int32_t bad = 0;
  MB200_CUDA(cudaMemcpyAsync(&bad, inc.bad.p, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  MB200_CUDA(cudaStreamSynchronize(st));
  MB200_REQUIRE(bad == 0, "a pair names a body index outside [0, n)");
  return MB200_OK;
}

}  // namespace mb200

using namespace mb200;

extern "C" {

// the COO -> CRS synchroniser of the link data
// (mundy/mesh/src/mundy_mesh/impl/NgpCOOToCRSSynchronizer.hpp:253-266: count with atomic_add per linked entity :405-468,
// partial sum, scatter fill), for pair links, as one stable radix sort: deterministic, ascending (link, slot) per entity.
int mb200_links_coo_to_crs(mb200_stream stream, int64_t n_entities, int64_t n_links, const int32_t* linked,
                           int64_t* offsets, int32_t* crs_links, int32_t* crs_slots) {
  MB200_REQUIRE(n_entities >= 0 && n_links >= 0 && n_entities < (int64_t(1) << 31) - 1 && n_links < (int64_t(1) << 30),
                "bad sizes");
  MB200_REQUIRE(offsets, "null offsets");
  MB200_TRY(require_device());
  cudaStream_t st = as_stream(stream);
  const int64_t n = n_entities, m = 2 * n_links;
  Scratch<uint32_t> counts, keys, keys_alt, vals, vals_alt;
  Scratch<int32_t> bad;
  MB200_TRY(counts.alloc(n > 0 ? n : 1, st));
  MB200_TRY(bad.alloc(1, st));
  MB200_CUDA(cudaMemsetAsync(counts.p, 0, (n > 0 ? n : 1) * sizeof(uint32_t), st));
  MB200_CUDA(cudaMemsetAsync(bad.p, 0, sizeof(int32_t), st));
  if (m > 0) {
    MB200_REQUIRE(linked && crs_links, "null link arrays");
    MB200_TRY(keys.alloc(m, st));
    MB200_TRY(keys_alt.alloc(m, st));
    MB200_TRY(vals.alloc(m, st));
    MB200_TRY(vals_alt.alloc(m, st));
    coo_keys_kernel<<<grid_for(m, 256), 256, 0, st>>>(m, n, linked, keys.p, vals.p, counts.p, bad.p);
    MB200_LAUNCHED();
    int bits = 1;
    while ((int64_t(1) << bits) < n + 1) ++bits;
    MB200_TRY(radix_sort_u32(st, m, keys.p, keys_alt.p, vals.p, vals_alt.p, bits));
    crs_split_kernel<<<grid_for(m, 256), 256, 0, st>>>(m, vals.p, crs_links, crs_slots);
    MB200_LAUNCHED();
  }
  MB200_TRY(exclusive_scan_u32_to_i64(st, n, counts.p, offsets));
  int32_t bad_h = 0;
  MB200_CUDA(cudaMemcpyAsync(&bad_h, bad.p, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  MB200_CUDA(cudaStreamSynchronize(st));
  MB200_REQUIRE(bad_h == 0, "a link references an entity index >= n_entities");
  return MB200_OK;
}

int mb200_apply_D(mb200_stream stream, const mb200_contact_op* op, const double* gamma, double* F, double* T) {
  MB200_TRY(check_op(op));
  if (op->n == 0) return MB200_OK;
  const bool six = op->arm_i != nullptr;
  MB200_REQUIRE(F && (!six || T) && (gamma || op->npairs == 0), "null pointer");
  MB200_TRY(require_device());
  cudaStream_t st = as_stream(stream);
  Incidence inc;
  MB200_TRY(build_incidence(st, op->n, op->npairs, op->pairs, &inc));
  MB200_TRY(check_incidence(st, inc));
  if (six)
    apply_D_kernel<true><<<grid_for(op->n, 128), 128, 0, st>>>(op->n, inc.ptr.p, inc.list.p, gamma, op->normal,
                                                              op->arm_i, op->arm_j, F, T);
  else
    apply_D_kernel<false><<<grid_for(op->n, 128), 128, 0, st>>>(op->n, inc.ptr.p, inc.list.p, gamma, op->normal,
                                                               nullptr, nullptr, F, nullptr);
  MB200_LAUNCHED();
  return MB200_OK;
}

int mb200_apply_M(mb200_stream stream, const mb200_contact_op* op, const double* F, const double* T, double* U,
                  double* W) {
  MB200_REQUIRE(op && op->n >= 0, "bad contact op");
  if (op->n == 0) return MB200_OK;
  MB200_REQUIRE(op->radius && F && U && ((T == nullptr) == (W == nullptr)), "null pointer");
  MB200_TRY(require_device());
  apply_M_kernel<<<grid_for(op->n, 256), 256, 0, as_stream(stream)>>>(op->n, op->quat, op->radius, op->kind,
                                                                     op->viscosity, F, T, U, W);
  MB200_LAUNCHED();
  return MB200_OK;
}

int mb200_apply_DT(mb200_stream stream, const mb200_contact_op* op, const double* U, const double* W, double* gdot) {
  MB200_TRY(check_op(op));
  if (op->npairs == 0) return MB200_OK;
  const bool six = op->arm_i != nullptr;
  MB200_REQUIRE(U && gdot && (!six || W), "null pointer");
  MB200_TRY(require_device());
  const int2* p2 = reinterpret_cast<const int2*>(op->pairs);
  if (six)
    apply_DT_kernel<true><<<grid_for(op->npairs, 256), 256, 0, as_stream(stream)>>>(
        op->npairs, p2, op->normal, op->arm_i, op->arm_j, U, W, 1.0, gdot);
  else
    apply_DT_kernel<false><<<grid_for(op->npairs, 256), 256, 0, as_stream(stream)>>>(
        op->npairs, p2, op->normal, nullptr, nullptr, U, nullptr, 1.0, gdot);
  MB200_LAUNCHED();
  return MB200_OK;
}

int mb200_apply_A(mb200_stream stream, const mb200_contact_op* op, const double* x, double* y) {
  MB200_TRY(check_op(op));
  if (op->npairs == 0) return MB200_OK;
  MB200_REQUIRE(x && y && op->radius, "null pointer");
  MB200_TRY(require_device());
  cudaStream_t st = as_stream(stream);
  const bool six = op->arm_i != nullptr;
  Scratch<double> F, T, U, W;
  MB200_TRY(F.alloc(3 * op->n, st));
  MB200_TRY(U.alloc(3 * op->n, st));
  if (six) {
    MB200_TRY(T.alloc(3 * op->n, st));
    MB200_TRY(W.alloc(3 * op->n, st));
  }
  MB200_TRY(mb200_apply_D(stream, op, x, F.p, six ? T.p : nullptr));
  MB200_TRY(mb200_apply_M(stream, op, F.p, six ? T.p : nullptr, U.p, six ? W.p : nullptr));
  const int2* p2 = reinterpret_cast<const int2*>(op->pairs);
  if (six)
    apply_DT_kernel<true><<<grid_for(op->npairs, 256), 256, 0, st>>>(op->npairs, p2, op->normal, op->arm_i, op->arm_j,
                                                                    U.p, W.p, op->dt, y);
  else
    apply_DT_kernel<false><<<grid_for(op->npairs, 256), 256, 0, st>>>(op->npairs, p2, op->normal, nullptr, nullptr,
                                                                     U.p, nullptr, op->dt, y);
  MB200_LAUNCHED();
  return MB200_OK;
}

}  // extern "C"
