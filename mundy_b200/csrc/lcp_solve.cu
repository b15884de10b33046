// E1. Driver BBPGD (scrap/lcp_spheres/KokkosLCP.cpp:675-876), device-resident.
//
// Data layout in HBM (all internal; Morton order when called from the step), contact-major, one array per field group:
//   R0[c]    = {nx, ny, nz, phi0}            32 bytes, one DRAM sector, moved with one 256-bit load
//   sig[c]   = {sigma_i, sigma_j}            ARM_AXIAL (the step's spherocylinders): the contact point of a rod lies on
//                                            its centre line, so arm = sigma h with h = (L/2) (q (x) z_hat) the body's
//                                            half-axis vector: 16 bytes per contact instead of 48
//   R1[c], R2[c] = {arm_i, -}, {arm_j, -}    ARM_GENERAL (caller-supplied arms, mb200_resolve_collisions)
//   pairs[c] = (i, j) int32                  the link list itself
//   state[c] = {gamma, g}                    16 bytes, updated IN PLACE by the rate pass (the thread that reads a
//                                            contact's previous state is the one that overwrites it; no roll-forward
//                                            copies, K13)
//   active   = one bit per contact: gamma > 0 or phi0 + dt g < 0, i.e. the contact CAN carry an impulse in the next
//              iteration whatever the (non-negative) step length; written by the rate pass with one ballot per warp
//   bmask    = the same bits per BODY: 128 bits each, bit k = the k-th entry of the body's incidence list, so that a body
//              reads the activity of its contacts with one 16-byte load that depends on nothing; the rate pass flips the
//              two entry bits of a contact (atomicXor) only when its activity changes. Entries beyond the 128th of a body
//              (dense packings without a search filter) are always evaluated.
//   body: mob[b] = {axis, +r (rod) | -r (sphere)}, half[b] = {h, kind}; velocity records {U, 0} (+ {V, 0}; V = omega x h
//         for ARM_AXIAL, omega for ARM_GENERAL): 64 MB at 10^6 rods, gathered by the rate pass out of L2
//
// One iteration = two kernels, no host round trip (the reference blocks the host twice per iteration):
//   body_pass : a thread per body visits the set bits of its `bmask` words in ascending (contact, side) order:
//               contacts whose bit is clear have gamma = 0 in this iteration and are neither looked up nor loaded
//               (about four in five at cfg3); for the others gamma = max(gamma_prev - alpha (phi0 + dt g_prev), 0)
//               (KokkosLCP.cpp:655-664) is re-derived from the state and the wrench summed in registers: F = D gamma
//               without atomics, bit-identical to a serial loop. U = M F; velocity records written (bodies whose
//               records are zero and stay zero are not rewritten).
//   rate_pass : per contact: the same gamma, g = D^T U, new state, active bit, residual max and the three BB dot
//               products; the last CTA to finish reduces the per-CTA partials in index order and advances alpha / done.
// Every kernel returns at once when `done` is set, so the host enqueues iterations in batches and reads the 48-byte
// solver state back between batches only.
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "comm.cuh"
#include "internal.cuh"
#include "lcp.cuh"

namespace mb200 {
namespace {

struct LcpState {
  double alpha;
  double res;
  double vmax;
  int32_t it;
  int32_t done;
  int32_t converged;
  uint32_t ticket;
  uint32_t ticket2;  // body pass's CTA counter (peer mode)
  int32_t error;     // multi-GPU: a peer did not answer in time
};

struct LcpArgs {  // everything the iteration kernels need; passed by value
  int64_t n, npairs;
  const int2* pairs;
  const D4* rec;       // R0 [, R1, R2]
  const double2* sig;  // ARM_AXIAL
  const D4* half;      // ARM_AXIAL
  const D4* mob;       // [n] mobility records
  const int64_t* inc_ptr;
  const uint32_t* inc_list;
  const uint32_t* inc_epos;  // [2 npairs] entry (2 c + side) -> its position in inc_list
  double2* state;    // [npairs]
  uint32_t* active;  // [ceil(npairs / 32)]
  unsigned long long* bmask;  // [2 n] the active bits per body, in incidence order (first 128 entries of each body)
  uint8_t* vzero;    // [n] 1 = the body's velocity records hold zeros
  D4* vel;           // [n] (3-DOF) or [2n] (6-DOF)
  double mu, dt, tol;
  int32_t fixed;   // ignore tol
  int32_t it_cap;  // max iterations (or the fixed count)
  LcpState* state_blk;
  double* partials;  // [4 * rate grid]
  // multi-GPU (null on one GPU): contacts this rank adds to the global sums; where the last CTA leaves the local totals
  const uint8_t* counted;
  double* red_local;
  // peer-memory exchange (null: NCCL): table of the ranks' mapped addresses; sequence numbers of the velocity halo this
  // rate pass consumes and of the reduction it produces
  const PeerDev* peer;
  unsigned long long halo_seq, red_seq;
  // peer mode: [n] 1 = owned body. The body pass leaves the ghost rows of the velocity arena alone: their owners store
  // into them from their own GPUs, possibly before this rank's body pass has run.
  const uint8_t* owned;
  // peer mode: per-body list of the ghost rows (rank << 28 | row in that rank's arena) that mirror an owned body
  const int64_t* push_ptr;
  const uint32_t* push_dst;
  // peer mode: the bodies that have ghost copies (ascending), handled by the first CTAs of the body pass so that the
  // velocity halo is on its way while the interior bodies are still being processed
  const uint32_t* boundary;
  int64_t n_boundary;
};

// ---- system-scope signalling between GPUs -----------------------------------------------------------------------
__device__ inline void st_relaxed_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ inline unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ inline double ld_relaxed_sys(const double* p) {
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ inline unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
constexpr unsigned long long kPeerTimeoutNs = 20ull * 1000ull * 1000ull * 1000ull;
// Spin until *flag >= want (a peer's release store). false after kPeerTimeoutNs: the caller raises state->error.
__device__ inline bool wait_flag(const unsigned long long* flag, unsigned long long want) {
  if (ld_acquire_sys(flag) >= want) return true;
  const unsigned long long t0 = global_ns();
  for (;;) {
    if (ld_acquire_sys(flag) >= want) return true;
    if (global_ns() - t0 > kPeerTimeoutNs) return false;
    __nanosleep(64);
  }
}

__device__ inline double2 ldg128(const double2* p) {
  double2 r;
  asm("ld.global.nc.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
  return r;
}
__device__ inline double2 ld128(const double2* p) {  // coherent: the rate pass overwrites the array it reads
  double2 r;
  asm volatile("ld.global.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p) : "memory");
  return r;
}

// gamma of one contact for the current iteration (KokkosLCP.cpp:655-664). MODE 0: the state already holds it.
template <int MODE>
__device__ inline double gamma_of(const double2& sp, double phi0, double alpha, double dt, int first) {
  if (MODE == 0) return sp.x;
  const double gc = first ? 0.0 : sp.y;  // :756 passes signed_sep_dot (zeros) on the first iteration
  const double sep_new = phi0 + dt * gc;
  return fmax(sp.x - alpha * sep_new, 0.0);
}

// ---- prologue / epilogue kernels ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    mobility_records_kernel(int64_t n, const double* __restrict__ radius, const double* __restrict__ quat,
                            const int32_t* __restrict__ kind, D4* __restrict__ mob) {
  const int64_t b = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (b >= n) return;
  const double r = radius[b];
  const bool rod = quat && kind && kind[b] == MB200_SPHEROCYLINDER;
  V3 a{0.0, 0.0, 1.0};
  if (rod) a = qrot(ldq(quat, b), V3{0.0, 0.0, 1.0});
  st256(mob + b, D4{a.x, a.y, a.z, rod ? r : -r});
}

// {h, kind} per body from the caller's quaternions and lengths (mb200_resolve_collisions_axial)
__global__ void __launch_bounds__(256)
    half_records_kernel(int64_t n, const double* __restrict__ quat, const double* __restrict__ length,
                        const int32_t* __restrict__ kind, D4* __restrict__ half) {
  const int64_t b = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (b >= n) return;
  const int kd = kind[b];
  V3 h{0.0, 0.0, 0.0};
  if (kd == MB200_SPHEROCYLINDER) h = (0.5 * length[b]) * qrot(ldq(quat, b), V3{0.0, 0.0, 1.0});
  st256(half + b, D4{h.x, h.y, h.z, static_cast<double>(kd)});
}

// caller layout -> records
template <int ARM>
__global__ void __launch_bounds__(256)
    pack_contacts_kernel(int64_t npairs, const double* __restrict__ normal, const double* __restrict__ arm_i,
                         const double* __restrict__ arm_j, const double* __restrict__ phi0, D4* __restrict__ rec) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= npairs) return;
  const V3 n = ld3(normal, c);
  st256(rec + c, D4{n.x, n.y, n.z, phi0[c]});
  if (ARM == ARM_GENERAL) {
    const V3 ai = ld3(arm_i, c), aj = ld3(arm_j, c);
    st256(rec + npairs + c, D4{ai.x, ai.y, ai.z, 0.0});
    st256(rec + 2 * npairs + c, D4{aj.x, aj.y, aj.z, 0.0});
  }
}

__global__ void __launch_bounds__(256)
    unpack_contacts_kernel(int64_t npairs, const D4* __restrict__ rec, const double2* __restrict__ state,
                           double* __restrict__ phi0, double* __restrict__ gamma) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= npairs) return;
  if (phi0) phi0[c] = ldg256(rec + c).w;
  if (gamma) gamma[c] = state[c].x;
}

constexpr int kMaskBits = 128;  // incidence entries per body covered by bmask
// The two entry bits of contact c = (i, j) change with its activity.
__device__ inline void flip_entries(const uint32_t* __restrict__ epos, const int64_t* __restrict__ ptr,
                                    unsigned long long* bmask, int64_t c, const int2& p) {
  const int64_t bi = static_cast<int64_t>(__ldg(epos + 2 * c)) - __ldg(ptr + p.x);
  const int64_t bj = static_cast<int64_t>(__ldg(epos + 2 * c + 1)) - __ldg(ptr + p.y);
  if (bi < kMaskBits) atomicXor(bmask + 2 * static_cast<int64_t>(p.x) + (bi >> 6), 1ull << (bi & 63));
  if (bj < kMaskBits) atomicXor(bmask + 2 * static_cast<int64_t>(p.y) + (bj >> 6), 1ull << (bj & 63));
}

// Initial state {gamma0, 0}; active bit = gamma0 > 0 (what the initial body pass needs). Whole warps: one ballot each.
// bmask was zeroed before.
__global__ void __launch_bounds__(256) init_state_kernel(int64_t npairs, const double* __restrict__ gamma0,
                                                         double2* __restrict__ state, uint32_t* __restrict__ active,
                                                         const int2* __restrict__ pairs, const uint32_t* __restrict__ epos,
                                                         const int64_t* __restrict__ ptr, unsigned long long* bmask) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const bool valid = c < npairs;
  const double g0 = (valid && gamma0) ? gamma0[c] : 0.0;
  if (valid) state[c] = make_double2(g0, 0.0);
  const bool flag = valid && g0 > 0.0;
  const uint32_t bits = __ballot_sync(0xffffffffu, flag);
  if ((threadIdx.x & 31) == 0 && c - (c & 31) < npairs) active[c >> 5] = bits;
  if (flag) flip_entries(epos, ptr, bmask, c, pairs[c]);
}

// gdot of one contact from the gathered velocity records. ARM_AXIAL: vi = Ui + sigma_i Vi, V = omega x h (the body pass
// stores it); ARM_GENERAL: vi = Ui + omega_i x arm_i (scrap/UnitTestMundy.cpp:783-824).
template <int ARM>
__device__ inline double rate_of(const V3& n, const V3& ui, const V3& uj, const V3& wi, const V3& wj, const V3& ai,
                                 const V3& aj, double si, double sj) {
  if (ARM == ARM_NONE) return contact_rate<false>(n, ui, uj, wi, wj, ai, aj);
  if (ARM == ARM_GENERAL) return contact_rate<true>(n, ui, uj, wi, wj, ai, aj);
  const V3 vi{ui.x + si * wi.x, ui.y + si * wi.y, ui.z + si * wi.z};
  const V3 vj{uj.x + sj * wj.x, uj.y + sj * wj.y, uj.z + sj * wj.z};
  return -(n.x * vi.x + n.y * vi.y + n.z * vi.z) - ((-n.x) * vj.x + (-n.y) * vj.y + (-n.z) * vj.z);
}

// phi0 += dt * D^T (U_ext, W_ext)   scrap/parameter_interface/NGPSpheresLCP.cpp:751-756
template <int ARM>
__global__ void __launch_bounds__(256)
    ext_correction_kernel(int64_t npairs, const int2* __restrict__ pairs, D4* rec, const double2* __restrict__ sig,
                          const D4* __restrict__ half, const double* __restrict__ U_ext,
                          const double* __restrict__ W_ext, double dt) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= npairs) return;
  D4* r0p = rec + c;
  D4 r0 = ld256(r0p);
  const V3 n{r0.x, r0.y, r0.z};
  const int2 p = pairs[c];
  V3 wi{0, 0, 0}, wj{0, 0, 0}, ai{0, 0, 0}, aj{0, 0, 0};
  double si = 0.0, sj = 0.0;
  double g;
  if (ARM != ARM_NONE && W_ext) {
    wi = ld3(W_ext, p.x), wj = ld3(W_ext, p.y);
    if (ARM == ARM_GENERAL) {
      const D4 r1 = ld256(rec + npairs + c), r2 = ld256(rec + 2 * npairs + c);
      ai = V3{r1.x, r1.y, r1.z}, aj = V3{r2.x, r2.y, r2.z};
    } else {
      const double2 sg = sig[c];
      si = sg.x, sj = sg.y;
      const D4 hi = ldg256(half + p.x), hj = ldg256(half + p.y);
      wi = cross(wi, V3{hi.x, hi.y, hi.z}), wj = cross(wj, V3{hj.x, hj.y, hj.z});
    }
    g = rate_of<ARM>(n, ld3(U_ext, p.x), ld3(U_ext, p.y), wi, wj, ai, aj, si, sj);
  } else {
    g = rate_of<ARM_NONE>(n, ld3(U_ext, p.x), ld3(U_ext, p.y), wi, wj, ai, aj, si, sj);
  }
  r0.w += dt * g;
  st256(r0p, r0);
}

// ---- iteration kernels ---------------------------------------------------------------------------------------------
constexpr int kBodyThreads = 128;
#ifndef MB200_BODY_CHUNK
#define MB200_BODY_CHUNK 2
#endif
// Incidence entries whose loads are in flight together, per thread. Measured at cfg3 (body pass, us): 1: 66.5, 2: 65.6,
// 3: 77, 4: 76 (registers 72 / 72 / 96 / 116: occupancy matters more than the second pair of requests, a body has 1.6
// active contacts on average); requesting the body's own records ahead of the contact records: 69.4 (96 registers).
constexpr int kChunk = MB200_BODY_CHUNK;

// MODE 0: gamma is read from the state (initial pass and final pass). MODE 1: gamma re-derived from the previous state.
// FINAL: writes F,T,U,W in the solver's body order and reduces max |U| into state->vmax (single launch at the end).
// A thread per body. Its incidence entries are taken kChunk at a time: entry words, then the active bits, then the
// records of the active ones are each loaded as a group of independent requests before their first use, and the
// contributions are added in entry order.
// (Forcing more resident CTAs through __launch_bounds__ loses: 9 CTAs / 56 registers 79 us, 11 / 40 registers 123 us
// against 66 us at ptxas' own 72 registers: the spills go through the same load/store path as the gathers.)
// HALO: the peer-memory multi-GPU form (ownership classes, boundary bodies first, remote stores, early halo flag); a
// template parameter so that the single-GPU kernel does not carry its registers (72 against 92).
// Resident CTAs per SM, axial arms (the kernel waits on gathers: warps in flight matter more than a few spilled words).
// Single-GPU form: 7 (72 registers, 44 bytes of spills): 69.2 -> 66.5 us at cfg3 (8 CTAs / 64 registers: 66.9 us; left to
// the compiler it takes 80 registers, 6 CTAs). Peer-mode form (left alone: 84 registers, 5 CTAs): also 7, 72 registers with
// 68 bytes of spills: 69.9 -> 68.8 us in a one-rank run of the peer-mode kernels (6 CTAs / 80 registers: 69.9 us).
#ifndef MB200_BODY_CTAS_HALO
#define MB200_BODY_CTAS_HALO 7
#endif
template <int ARM, int MODE, bool FINAL, bool HALO>
__global__ void __launch_bounds__(kBodyThreads, (ARM == ARM_AXIAL) ? (HALO ? MB200_BODY_CTAS_HALO : 7) : 0)
    body_pass_kernel(LcpArgs a, int first, double* __restrict__ F, double* __restrict__ T, double* __restrict__ U,
                     double* __restrict__ W, const double* __restrict__ U_ext, const double* __restrict__ W_ext) {
  if (!FINAL && a.state_blk->done) return;
  constexpr int NC = (ARM == ARM_NONE) ? 3 : 6;
  constexpr bool SIX = ARM != ARM_NONE;
  __shared__ double sh[33];
  const double alpha = (MODE == 1) ? a.state_blk->alpha : 0.0;
  // A clear bit means gamma = 0 and phi0 + dt g >= 0: the projected step max(0 - alpha (..), 0) is zero for every
  // alpha >= 0. A negative (or NaN) step length cannot rely on it: then every entry is evaluated.
  const bool use_mask = (MODE == 0) || (alpha >= 0.0);
  const int t = threadIdx.x;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * kBodyThreads;
  const double* sigp = reinterpret_cast<const double*>(a.sig);
  double speed = 0.0;
  // peer mode: work items [0, nb_pad) are the boundary bodies (padded to whole CTAs), the rest all bodies
  constexpr bool halo = HALO && !FINAL;
  const int64_t nb_ctas = halo ? (a.n_boundary + kBodyThreads - 1) / kBodyThreads : 0;
  const int64_t nb_pad = nb_ctas * kBodyThreads;

  // Bodies are visited from the last to the first: the rate pass walks the contacts upwards, so the state, normal and
  // arm records of the highest contacts (the last ~100 MB it touched) are still in the 126 MB L2 when this kernel starts
  // with the highest bodies, and the velocity records this kernel writes last (the lowest bodies) are there for the
  // rate pass that follows.
  for (int64_t wi = static_cast<int64_t>(blockIdx.x) * kBodyThreads + t; wi < nb_pad + a.n; wi += stride) {
    int64_t b;
    if (wi < nb_pad) {
      if (wi >= a.n_boundary) continue;
      b = __ldg(a.boundary + wi);
    } else {
      b = a.n - 1 - (wi - nb_pad);
    }
    // everything the body needs first is requested together (the skip tests below must not serialise the loads)
    const int64_t beg = __ldg(a.inc_ptr + b), end = __ldg(a.inc_ptr + b + 1);
    unsigned long long m0, m1;  // the body's active bits: one load, independent of the pointers
    asm("ld.global.nc.v2.u64 {%0,%1}, [%2];" : "=l"(m0), "=l"(m1) : "l"(a.bmask + 2 * b));
    uint8_t own = 1;  // peer mode: 0 ghost, 1 owned, 2 owned with ghost copies on other ranks
    if (halo) own = __ldg(a.owned + b);
    // peer mode: ghost rows belong to their owners, who store into them from their own GPUs; a body with ghost copies
    // is done by a boundary CTA. Both are skipped by emptying the entry range (a select, not a branch on `own`: a
    // branch here would hold the pointer and mask loads back until the flag byte has arrived, one more memory latency
    // per body) and leave at the epilogue.
    const bool skip = halo && (!own || (wi >= nb_pad && own == 2));
    double acc[NC];
#pragma unroll
    for (int k = 0; k < NC; ++k) acc[k] = 0.0;
    bool any = false;
    // kChunk entries at a time: entry words, then the records of all of them, are loaded as groups of independent
    // requests before their first use; contributions are added in entry order
    auto add_entries = [&](const int64_t (&pos)[kChunk]) {
      uint32_t e[kChunk];
      double2 sp[kChunk];
      D4 r0[kChunk], ra[kChunk];
      double sg[kChunk];
#pragma unroll
      for (int u = 0; u < kChunk; ++u) e[u] = (pos[u] >= 0) ? __ldg(a.inc_list + pos[u]) : 0xffffffffu;
#pragma unroll
      for (int u = 0; u < kChunk; ++u) {
        if (e[u] != 0xffffffffu) {
          const int64_t c = e[u] >> 1;
          const int side = e[u] & 1u;
          sp[u] = ldg128(a.state + c);
          r0[u] = ldg256(a.rec + c);
          if (ARM == ARM_AXIAL) sg[u] = __ldg(sigp + 2 * c + side);
          if (ARM == ARM_GENERAL) ra[u] = ldg256(a.rec + (side ? 2 : 1) * a.npairs + c);
        }
      }
#pragma unroll
      for (int u = 0; u < kChunk; ++u) {
        if (e[u] != 0xffffffffu) {
          any = true;
          const bool is_j = e[u] & 1u;
          const double g = gamma_of<MODE>(sp[u], r0[u].w, alpha, a.dt, first);
          // n_I = n, n_J = -n: F += -gamma n_side (KokkosLCP.cpp:566-590)
          const double mx = is_j ? -r0[u].x : r0[u].x, my = is_j ? -r0[u].y : r0[u].y, mz = is_j ? -r0[u].z : r0[u].z;
          const double wx = -g * mx, wy = -g * my, wz = -g * mz;
          acc[0] += wx, acc[1] += wy, acc[2] += wz;
          if (ARM == ARM_AXIAL) {  // T = sum (sigma h) x w = h x (sum sigma w): the cross product waits for the epilogue
            acc[3] += sg[u] * wx, acc[4] += sg[u] * wy, acc[5] += sg[u] * wz;
          }
          if (ARM == ARM_GENERAL) {  // T += -gamma (r_side x n_side)  (scrap/UnitTestMundy.cpp:652-674)
            acc[3] += -g * (ra[u].y * mz - ra[u].z * my);
            acc[4] += -g * (ra[u].z * mx - ra[u].x * mz);
            acc[5] += -g * (ra[u].x * my - ra[u].y * mx);
          }
        }
      }
    };
    // Positions of the entries to evaluate, ascending: the set bits of the body's mask words, then the entries past the
    // mask's reach; every entry when the mask cannot be used.
    int64_t kover = beg + kMaskBits;
    if (!use_mask) m0 = 0, m1 = 0, kover = beg;
    if (halo) m0 = skip ? 0ull : m0, m1 = skip ? 0ull : m1, kover = skip ? end : kover;
    auto next_pos = [&]() -> int64_t {
      if (m0) {
        const int64_t p = beg + (__ffsll(static_cast<long long>(m0)) - 1);
        m0 &= m0 - 1;
        return p;
      }
      if (m1) {
        const int64_t p = beg + 64 + (__ffsll(static_cast<long long>(m1)) - 1);
        m1 &= m1 - 1;
        return p;
      }
      return (kover < end) ? kover++ : -1;
    };
    for (;;) {
      int64_t pos[kChunk];
      pos[0] = next_pos();
      if (pos[0] < 0) break;
#pragma unroll
      for (int u = 1; u < kChunk; ++u) pos[u] = (pos[u - 1] >= 0) ? next_pos() : -1;
      add_entries(pos);
    }

    // epilogue: U = M F; velocity records (and, multi-GPU, their ghost copies on the peers) or the solver's outputs
    if (halo && skip) continue;
    if (!FINAL) {
      if (!any) {
        if (a.vzero[b]) continue;  // zero and known to be: nothing to rewrite
        a.vzero[b] = 1;
      } else if (a.vzero[b]) {
        a.vzero[b] = 0;
      }
    }
    V3 u{0.0, 0.0, 0.0}, w{0.0, 0.0, 0.0}, v2{0.0, 0.0, 0.0};
    V3 f{acc[0], acc[1], acc[2]}, tq{0.0, 0.0, 0.0};
    if (any || FINAL) {
      const D4 mr = ldg256(a.mob + b);
      const bool rod = SIX && mr.w > 0.0;
      const Mob m = mobility_of(fabs(mr.w), a.mu);
      V3 h{0.0, 0.0, 0.0};
      if (ARM == ARM_AXIAL) {
        const D4 hr = ldg256(a.half + b);
        h = V3{hr.x, hr.y, hr.z};
        tq = cross(h, V3{acc[3], acc[4], acc[5]});
      }
      if (ARM == ARM_GENERAL) tq = V3{acc[3], acc[4], acc[5]};
      body_mobility_axis(m, rod, V3{mr.x, mr.y, mr.z}, f, tq, u, w);
      v2 = (ARM == ARM_AXIAL) ? cross(w, h) : w;
    }
    if (!FINAL) {
      if (SIX) {
        st256(a.vel + 2 * b, D4{u.x, u.y, u.z, 0.0});
        st256(a.vel + 2 * b + 1, D4{v2.x, v2.y, v2.z, 0.0});
      } else {
        st256(a.vel + b, D4{u.x, u.y, u.z, 0.0});
      }
      if (halo && own == 2) {  // fused velocity halo: the records also go straight into the peers' ghost rows (NVLink)
        // (the list bounds are fetched here, by the few bodies that have ghost copies, and the loop is kept rolled:
        // held across the entry loop, or unrolled, this rarely taken path cost every body of the pass 4 us per launch)
        const int64_t pk = __ldg(a.push_ptr + b), pke = __ldg(a.push_ptr + b + 1);
#pragma unroll 1
        for (int64_t k = pk; k < pke; ++k) {
          const uint32_t d = __ldg(a.push_dst + k);
          D4* dst = a.peer->vel[d >> 28];
          const int64_t row = d & 0x0fffffffu;
          if (SIX) {
            st256(dst + 2 * row, D4{u.x, u.y, u.z, 0.0});
            st256(dst + 2 * row + 1, D4{v2.x, v2.y, v2.z, 0.0});
          } else {
            st256(dst + row, D4{u.x, u.y, u.z, 0.0});
          }
        }
      }
    } else {
      speed = fmax(speed, sqrt(u.x * u.x + u.y * u.y + u.z * u.z));  // KokkosLCP.cpp:861-872 (collision velocity)
      if (F) st3(F, b, f);
      if (SIX && T) st3(T, b, tq);
      if (U_ext) u = u + ld3(U_ext, b);  // U += U_col on top of the external velocity (NGPSpheresLCP.cpp:867-877)
      if (SIX && W_ext) w = w + ld3(W_ext, b);
      if (U) st3(U, b, u);
      if (SIX && W) st3(W, b, w);
    }
  }
  if constexpr (halo) {
  if (blockIdx.x < nb_ctas) {
    // the boundary CTAs' remote stores are ordered before the flag the last of them publishes to all ranks (release,
    // system scope); the consumers are the peers' rate_pass_kernels
    // (bar.sync orders the CTA's stores before thread 0's system-scope fence: the grid-sync pattern, one fence per CTA)
    __syncthreads();
    if (t == 0) {
      __threadfence_system();
      const uint32_t ticket = atomicAdd(&a.state_blk->ticket2, 1u);
      sh[32] = (ticket == nb_ctas - 1) ? 1.0 : 0.0;
      if (ticket == nb_ctas - 1) a.state_blk->ticket2 = 0;
    }
    __syncthreads();
    if (sh[32] != 0.0 && t < a.peer->world && t != a.peer->me) {  // one lane per rank: the flags leave in parallel
      __threadfence_system();
      st_relaxed_sys(&a.peer->ctrl[t]->halo_flag[a.peer->me], a.halo_seq);
    }
  } else if (nb_ctas == 0 && blockIdx.x == 0) {
    // nothing of this rank is ghosted anywhere: the peers still wait for its flag
    if (t < a.peer->world && t != a.peer->me) st_relaxed_sys(&a.peer->ctrl[t]->halo_flag[a.peer->me], a.halo_seq);
  }
  }
  if (FINAL) {
    const double m = block_max(speed, sh, 0.0);
    if (threadIdx.x == 0) {  // non-negative doubles order like their bit patterns
      atomicMax(reinterpret_cast<unsigned long long*>(&a.state_blk->vmax),
                static_cast<unsigned long long>(__double_as_longlong(m)));
    }
  }
}

constexpr int kRateThreads = 256;
constexpr int kRateStages = 3;  // tiles of kRateThreads contacts in flight per CTA

// ---- bulk asynchronous copies global -> shared (the TMA engine's 1-D form; SASS: UBLKCP) with mbarrier completion ----
__device__ inline uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ inline void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ inline void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ inline void mbar_arrive(uint64_t* bar) {  // release at CTA scope (the default)
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ inline void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(addr), "r"(parity)
        : "memory");
  } while (!ok);
}
// size: a multiple of 16 bytes; both addresses 16-byte aligned
__device__ inline void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// One pipeline stage of the rate pass: the contiguous per-contact arrays of kRateThreads consecutive contacts.
template <int ARM>
struct RateStage {
  static constexpr int kState = 0;
  static constexpr int kRec0 = kState + kRateThreads * 16;
  static constexpr int kArm = kRec0 + kRateThreads * 32;  // AXIAL: sigma pairs; GENERAL: R1 then R2
  static constexpr int kPairs = kArm + (ARM == ARM_AXIAL ? kRateThreads * 16 : (ARM == ARM_GENERAL ? kRateThreads * 64 : 0));
  static constexpr int kBytes = kPairs + kRateThreads * 8;
};
template <int ARM>
constexpr int rate_smem_bytes() {
  return kRateStages * RateStage<ARM>::kBytes;
}

// End of an operator application: residual test, next BB step length, iteration cap (KokkosLCP.cpp:735-750, 785-848).
__device__ inline void advance_state(LcpState* s, double res, double xx, double xg, double gg, int it, int mode,
                                     int fixed, double tol, int it_cap) {
  s->res = res;
  s->it = it;
  if (!fixed && res < tol) {  // strict <, KokkosLCP.cpp:744,788
    s->done = 1;
    s->converged = 1;
    return;
  }
  if (mode == 0) {
    s->alpha = 1.0 / res;  // :750
  } else {
    double num, den;  // :831-848; `it` was already incremented at the loop top, so iteration 1 takes BB2
    if (it % 2 == 0) {
      num = xx, den = xg;
    } else {
      num = xg, den = gg;
    }
    if (fabs(den) < 1e-12) den += 1e-12;
    s->alpha = num / den;
  }
  if (it >= it_cap) s->done = 1;
}

// Multi-GPU: the ranks' totals {res, xx, xg, gg} gathered in rank order -> the same state on every rank.
__global__ void advance_kernel(LcpState* s, const double* __restrict__ red_all, int world, int it, int mode, int fixed,
                               double tol, int it_cap) {
  if (threadIdx.x != 0 || s->done) return;
  double res = -1.0, xx = 0.0, xg = 0.0, gg = 0.0;
  for (int r = 0; r < world; ++r) {
    res = fmax(res, red_all[4 * r]);
    xx += red_all[4 * r + 1], xg += red_all[4 * r + 2], gg += red_all[4 * r + 3];
  }
  advance_state(s, res, xx, xg, gg, it, mode, fixed, tol, it_cap);
}

// Halo of the body velocity records: owners' records -> send buffer, received records -> ghost rows.
__global__ void __launch_bounds__(256) halo_pack_kernel(int64_t count, int recs, const uint32_t* __restrict__ idx,
                                                        const D4* __restrict__ vel, D4* __restrict__ buf,
                                                        const LcpState* __restrict__ state) {
  if (state->done) return;
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= count * recs) return;
  const int64_t b = idx[k / recs];
  buf[k] = vel[b * recs + (k % recs)];
}
__global__ void __launch_bounds__(256) halo_unpack_kernel(int64_t count, int recs, const uint32_t* __restrict__ idx,
                                                          const D4* __restrict__ buf, D4* __restrict__ vel,
                                                          const LcpState* __restrict__ state) {
  if (state->done) return;
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= count * recs) return;
  const int64_t b = idx[k / recs];
  vel[b * recs + (k % recs)] = buf[k];
}

// Velocity record of a body for the rate pass. Peer mode: ghost rows are written by other GPUs over NVLink while this
// kernel is resident, so they are read with coherent loads (ld.global.nc must not see data that changes during the
// kernel). The loads are plain asm, free to be scheduled like any other load (a volatile asm with a memory clobber kept
// the four gathers of a contact in program order behind each other and behind the state store: +5 us per launch); what
// keeps them behind the halo wait is their ADDRESS: the kernel derives the base pointer it hands down here from a value
// produced after the wait (rate_pass_kernel).
template <bool PEER>
__device__ inline D4 ld_vel(const D4* p) {
  if (!PEER) return ldg256(p);
  D4 r;
  asm("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p));
  return r;
}

// MODE 0: initial pass (state holds gamma_0; g_0 = D^T M D gamma_0, residual only). MODE 1: iteration pass.
// Multi-GPU with peer memory (PEER): every CTA first waits until all other ranks have published velocity halo
// a.halo_seq; the last CTA stores this rank's four totals into every rank's control block, publishes a.red_seq, waits for
// the other ranks' totals and reduces them in rank order, so that every rank advances to the same alpha / done without
// a collective call or a host round trip.
// One contact of the rate pass: gamma of this iteration, g = D^T U from the gathered velocity records, the new state,
// the contact's terms of the residual and of the three BB sums. Returns the contact's activity for the next iteration.
template <int ARM, int MODE, bool PEER>
__device__ inline bool rate_contact(const LcpArgs& a, const D4* vel, bool counted, int64_t c,
                                    const double2& sp, const D4& r0, const D4& r1, const D4& r2, const double2& sg,
                                    const int2& p, double alpha, int first, double& res, double& xx, double& xg,
                                    double& gg) {
  const V3 n{r0.x, r0.y, r0.z};
  V3 ui, uj, wi{0, 0, 0}, wj{0, 0, 0};
  const V3 ai{r1.x, r1.y, r1.z}, aj{r2.x, r2.y, r2.z};
  if (ARM != ARM_NONE) {
    const D4 v0 = ld_vel<PEER>(vel + 2 * p.x), v1 = ld_vel<PEER>(vel + 2 * p.x + 1);
    const D4 v2 = ld_vel<PEER>(vel + 2 * p.y), v3 = ld_vel<PEER>(vel + 2 * p.y + 1);
    ui = V3{v0.x, v0.y, v0.z}, wi = V3{v1.x, v1.y, v1.z};
    uj = V3{v2.x, v2.y, v2.z}, wj = V3{v3.x, v3.y, v3.z};
  } else {
    const D4 v0 = ld_vel<PEER>(vel + p.x), v2 = ld_vel<PEER>(vel + p.y);
    ui = V3{v0.x, v0.y, v0.z}, uj = V3{v2.x, v2.y, v2.z};
  }
  const double gam = gamma_of<MODE>(sp, r0.w, alpha, a.dt, first);
  const double g = rate_of<ARM>(n, ui, uj, wi, wj, ai, aj, sg.x, sg.y);
  a.state[c] = make_double2(gam, g);
  const double sep_new = r0.w + a.dt * g;  // KokkosLCP.cpp:500-521
  if (counted) {  // a cut contact is summed by the rank that owns its first body
    const double v = (gam < 1e-12) ? fabs(fmin(sep_new, 0.0)) : fabs(sep_new);
    res = fmax(res, v);
    if (MODE == 1) {  // KokkosLCP.cpp:536-556
      const double xd = gam - sp.x;
      const double gd = a.dt * (g - sp.y);
      xx += xd * xd;
      xg += xd * gd;
      gg += gd * gd;
    }
  }
  // can the contact carry an impulse in the next iteration? (the one after the initial pass takes g = 0, :756)
  return gam > 0.0 || (MODE == 0 ? r0.w : sep_new) < 0.0;
}

// The four contiguous per-contact streams (state, {n, phi0}, arm data, pair ids) do not go through the load/store unit:
// one thread hands whole tiles of kRateThreads contacts to the bulk-copy engine (cp.async.bulk -> shared memory, completion
// counted on an mbarrier), kRateStages tiles ahead of the tile being worked on, so that the memory-level parallelism of the
// streams no longer depends on registers or resident warps; the LSU carries only the velocity gathers (out of L2) and the
// state stores. Tiles go round-robin over the persistent grid (tile = blockIdx.x + k gridDim.x), the same contact order
// per thread as a grid-stride loop, which is the order the deterministic sums are defined in (mb200_lcp_reduction_shape).
// The last, partial tile takes plain loads.
// BULK = false is the same kernel with every tile on plain (coalesced 128/256-bit) loads and no shared-memory staging.
// Measured at cfg3: the pass is bound by the load/store unit's wavefronts of the VELOCITY GATHERS
// (about 84 of the 106 wavefronts a warp issues per 32 contacts), which the bulk copies do not relieve, and eight warps
// released together by one barrier issue their gathers in bursts: 78 us against 66 us. The plain form is the default;
// MB200_RATE_PIPELINE=bulk selects the staged one.
// (64 registers, four CTAs per SM. Five CTAs / 48 registers: 78 us, six / 40 registers: 88 us against 63 us: spills.)
template <int ARM, int MODE, bool PEER, bool BULK>
__global__ void __launch_bounds__(kRateThreads) rate_pass_kernel(LcpArgs a, int it, int first) {
  if (a.state_blk->done) return;
  using St = RateStage<ARM>;
  extern __shared__ __align__(128) unsigned char stage_mem[];
  __shared__ __align__(8) uint64_t full_bar[kRateStages];   // bytes of a tile have landed (bulk copies complete_tx)
  __shared__ __align__(8) uint64_t empty_bar[kRateStages];  // every thread has taken its contact out of the stage
  __shared__ double sh[33];
  __shared__ bool is_last;
  const double alpha = (MODE == 1) ? a.state_blk->alpha : 0.0;
  double res = -1.0, xx = 0.0, xg = 0.0, gg = 0.0;
  const int t = threadIdx.x;
  const int64_t G = gridDim.x;
  const int64_t ntiles = (a.npairs + kRateThreads - 1) / kRateThreads;
  const int64_t nfull = BULK ? a.npairs / kRateThreads : 0;
  auto issue = [&](int64_t tile, int s) {  // thread 0: tile -> stage s
    unsigned char* base = stage_mem + s * St::kBytes;
    const int64_t c0 = tile * kRateThreads;
    mbar_expect_tx(&full_bar[s], St::kBytes);
    bulk_g2s(base + St::kState, a.state + c0, kRateThreads * 16, &full_bar[s]);
    bulk_g2s(base + St::kRec0, a.rec + c0, kRateThreads * 32, &full_bar[s]);
    if (ARM == ARM_AXIAL) bulk_g2s(base + St::kArm, a.sig + c0, kRateThreads * 16, &full_bar[s]);
    if (ARM == ARM_GENERAL) {
      bulk_g2s(base + St::kArm, a.rec + a.npairs + c0, kRateThreads * 32, &full_bar[s]);
      bulk_g2s(base + St::kArm + kRateThreads * 32, a.rec + 2 * a.npairs + c0, kRateThreads * 32, &full_bar[s]);
    }
    bulk_g2s(base + St::kPairs, a.pairs + c0, kRateThreads * 8, &full_bar[s]);
  };
  if (BULK && t == 0) {
#pragma unroll
    for (int s = 0; s < kRateStages; ++s) mbar_init(&full_bar[s], 1), mbar_init(&empty_bar[s], kRateThreads);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (BULK) __syncthreads();
  if (BULK && t == 0) {  // the streams do not depend on the peers' halo: they start before the wait below
#pragma unroll
    for (int s = 0; s < kRateStages; ++s) {
      const int64_t tile = blockIdx.x + s * G;
      if (tile < nfull) issue(tile, s);
    }
  }
  if (PEER) {
    if (t < a.peer->world && t != a.peer->me && !wait_flag(&a.peer->ctrl[a.peer->me]->halo_flag[t], a.halo_seq)) {
      a.state_blk->error = 1;
      a.state_blk->done = 1;
    }
    __syncthreads();
  }
  // peer mode: the base of the velocity records as a value the compiler cannot know before this point (a zero it has to
  // take from a volatile asm placed after the wait): every gather's address depends on it, none can move above the wait
  const D4* vel = a.vel;
  if (PEER) {
    unsigned long long zero;
    asm volatile("mov.u64 %0, 0;" : "=l"(zero)::"memory");
    vel = a.vel + zero;
  }
  for (int64_t k = 0;; ++k) {
    const int64_t tile = blockIdx.x + k * G;
    if (tile >= ntiles) break;
    const int s = static_cast<int>(k % kRateStages);
    const int64_t cw = tile * kRateThreads + (t & ~31);  // whole warps walk 32 consecutive contacts: one word of bits
    const int64_t c = cw + (t & 31);
    const bool valid = c < a.npairs;
    double2 sp = make_double2(0.0, 0.0), sg = make_double2(0.0, 0.0);
    D4 r0{0.0, 0.0, 0.0, 0.0}, r1{0.0, 0.0, 0.0, 0.0}, r2{0.0, 0.0, 0.0, 0.0};
    int2 p = make_int2(0, 0);
    if (BULK && tile < nfull) {
      mbar_wait(&full_bar[s], static_cast<uint32_t>((k / kRateStages) & 1));
      const unsigned char* base = stage_mem + s * St::kBytes;
      sp = reinterpret_cast<const double2*>(base + St::kState)[t];
      r0 = reinterpret_cast<const D4*>(base + St::kRec0)[t];
      if (ARM == ARM_AXIAL) sg = reinterpret_cast<const double2*>(base + St::kArm)[t];
      if (ARM == ARM_GENERAL) {
        r1 = reinterpret_cast<const D4*>(base + St::kArm)[t];
        r2 = reinterpret_cast<const D4*>(base + St::kArm + kRateThreads * 32)[t];
      }
      p = reinterpret_cast<const int2*>(base + St::kPairs)[t];
      // Hand the stage back as soon as the values have ARRIVED in registers. The stage is refilled by the bulk-copy
      // engine (the async proxy), which does not queue behind this CTA's shared-memory loads: those share the load/store
      // unit's queue with the velocity gathers and can sit in it for longer than a refill takes, and neither a barrier
      // nor a release-arrive waits for their data (observed: whole warps of tile k computing with the records of tile
      // k + kRateStages). So every loaded word is folded into one value that a (never taken) branch tests: the test
      // cannot issue before the loads have returned, and the arrive is ordered after it.
      {
        long long fold = __double_as_longlong(sp.x) ^ __double_as_longlong(sp.y) ^ __double_as_longlong(r0.x) ^
                         __double_as_longlong(r0.y) ^ __double_as_longlong(r0.z) ^ __double_as_longlong(r0.w) ^
                         (static_cast<long long>(p.x) << 32 | static_cast<unsigned>(p.y));
        if (ARM == ARM_AXIAL) fold ^= __double_as_longlong(sg.x) ^ __double_as_longlong(sg.y);
        if (ARM == ARM_GENERAL)
          fold ^= __double_as_longlong(r1.x) ^ __double_as_longlong(r1.y) ^ __double_as_longlong(r1.z) ^
                  __double_as_longlong(r2.x) ^ __double_as_longlong(r2.y) ^ __double_as_longlong(r2.z);
        if (fold == 0x7ff8badc0ffee123ll) asm volatile("nanosleep.u32 1;" ::: "memory");
      }
      mbar_arrive(&empty_bar[s]);
      if (t == 0) {
        const int64_t next = tile + kRateStages * G;
        if (next < nfull) {
          mbar_wait(&empty_bar[s], static_cast<uint32_t>((k / kRateStages) & 1));
          issue(next, s);
        }
      }
    } else if (valid) {
      sp = ld128(a.state + c);
      r0 = ldg256(a.rec + c);
      p = __ldg(a.pairs + c);
      if (ARM == ARM_AXIAL) sg = ldg128(a.sig + c);
      if (ARM == ARM_GENERAL) r1 = ldg256(a.rec + a.npairs + c), r2 = ldg256(a.rec + 2 * a.npairs + c);
    }
    // multi-GPU: is the contact counted in this rank's sums? Requested here, with the streams, not where it is used
    bool counted = true;
    if (a.counted && valid) counted = __ldg(a.counted + c) != 0;
    const uint32_t old_bits = a.active[cw >> 5];  // one address per warp: a broadcast
    const bool flag =
        valid && rate_contact<ARM, MODE, PEER>(a, vel, counted, c, sp, r0, r1, r2, sg, p, alpha, first, res, xx, xg, gg);
    const uint32_t bits = __ballot_sync(0xffffffffu, flag);
    if ((t & 31) == 0) a.active[cw >> 5] = bits;
    if (valid && flag != (((old_bits >> (t & 31)) & 1u) != 0)) flip_entries(a.inc_epos, a.inc_ptr, a.bmask, c, p);
  }
  // per-CTA partials, then the last CTA reduces them in index order (deterministic)
  res = block_max(res, sh, -1.0);
  if (MODE == 1) {
    xx = block_sum(xx, sh);
    xg = block_sum(xg, sh);
    gg = block_sum(gg, sh);
  }
  if (t == 0) {
    double* p = a.partials + 4 * blockIdx.x;
    p[0] = res, p[1] = xx, p[2] = xg, p[3] = gg;
    __threadfence();
    const uint32_t ticket = atomicAdd(&a.state_blk->ticket, 1u);
    is_last = (ticket == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  res = -1.0, xx = 0.0, xg = 0.0, gg = 0.0;
  const volatile double* vp = a.partials;
  for (int k = t; k < static_cast<int>(gridDim.x); k += kRateThreads) {
    res = fmax(res, vp[4 * k]);
    xx += vp[4 * k + 1], xg += vp[4 * k + 2], gg += vp[4 * k + 3];
  }
  res = block_max(res, sh, -1.0);
  if (MODE == 1) {
    xx = block_sum(xx, sh);
    xg = block_sum(xg, sh);
    gg = block_sum(gg, sh);
  }
  if (PEER) {
    // one lane per rank: this rank's totals into rank r's control block, fence, flag; then wait for rank r's totals here
    __shared__ double s_tot[4];
    __shared__ int s_bad;
    if (t == 0) s_tot[0] = res, s_tot[1] = xx, s_tot[2] = xg, s_tot[3] = gg, s_bad = 0;
    __syncthreads();
    const PeerDev* pd = a.peer;
    const int me = pd->me, W = pd->world, par = static_cast<int>(a.red_seq & 1ull);
    if (t < W) {
      double* dst = pd->ctrl[t]->red[par][me];
      dst[0] = s_tot[0], dst[1] = s_tot[1], dst[2] = s_tot[2], dst[3] = s_tot[3];
      __threadfence_system();
      st_relaxed_sys(&pd->ctrl[t]->red_flag[par][me], a.red_seq);
      if (!wait_flag(&pd->ctrl[me]->red_flag[par][t], a.red_seq)) s_bad = 1;
    }
    __syncthreads();
    if (t == 0) {
      LcpState* s = a.state_blk;
      s->ticket = 0;
      if (s_bad) {
        s->error = 1;
        s->done = 1;
      } else {
        res = -1.0, xx = 0.0, xg = 0.0, gg = 0.0;
        for (int r = 0; r < W; ++r) {  // rank order: the same sums on every rank
          const double* src = pd->ctrl[me]->red[par][r];
          res = fmax(res, ld_relaxed_sys(src));
          xx += ld_relaxed_sys(src + 1), xg += ld_relaxed_sys(src + 2), gg += ld_relaxed_sys(src + 3);
        }
        advance_state(s, res, xx, xg, gg, it, MODE, a.fixed, a.tol, a.it_cap);
      }
    }
    return;
  }
  if (t == 0) {
    LcpState* s = a.state_blk;
    s->ticket = 0;
    if (a.red_local) {  // multi-GPU over NCCL: the totals go through an all-gather, advance_kernel finishes
      a.red_local[0] = res, a.red_local[1] = xx, a.red_local[2] = xg, a.red_local[3] = gg;
    } else {
      advance_state(s, res, xx, xg, gg, it, MODE, a.fixed, a.tol, a.it_cap);
    }
  }
}

struct PinnedState {  // 48 bytes of page-locked host memory per host thread, usable from every device
  LcpState* h = nullptr;
  ~PinnedState() {
    if (h) cudaFreeHost(h);
  }
};

template <typename K>
int resident_grid(K kernel, int threads) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0) != cudaSuccess || per_sm < 1) per_sm = 4;
  return per_sm * kSMs;
}

template <int ARM>
void launch_body(const LcpArgs& a, int grid, cudaStream_t st, int mode, bool fin_pass, int first, double* F, double* T,
                 double* U, double* W) {
  const bool halo = a.push_ptr != nullptr;
  if (fin_pass)
    body_pass_kernel<ARM, 0, true, false><<<grid, kBodyThreads, 0, st>>>(a, first, F, T, U, W, nullptr, nullptr);
  else if (mode == 0 && halo)
    body_pass_kernel<ARM, 0, false, true><<<grid, kBodyThreads, 0, st>>>(a, first, F, T, U, W, nullptr, nullptr);
  else if (mode == 0)
    body_pass_kernel<ARM, 0, false, false><<<grid, kBodyThreads, 0, st>>>(a, first, F, T, U, W, nullptr, nullptr);
  else if (halo)
    body_pass_kernel<ARM, 1, false, true><<<grid, kBodyThreads, 0, st>>>(a, first, F, T, U, W, nullptr, nullptr);
  else
    body_pass_kernel<ARM, 1, false, false><<<grid, kBodyThreads, 0, st>>>(a, first, F, T, U, W, nullptr, nullptr);
}
bool rate_pipeline_bulk() {
  static const bool bulk = [] {
    const char* e = getenv("MB200_RATE_PIPELINE");
    return e && std::strcmp(e, "bulk") == 0;
  }();
  return bulk;
}

template <int ARM, bool BULK>
void launch_rate_as(const LcpArgs& a, int grid, cudaStream_t st, int mode, int it, int first) {
  constexpr int smem = BULK ? rate_smem_bytes<ARM>() : 0;
  if (a.peer) {
    if (mode == 0)
      rate_pass_kernel<ARM, 0, true, BULK><<<grid, kRateThreads, smem, st>>>(a, it, first);
    else
      rate_pass_kernel<ARM, 1, true, BULK><<<grid, kRateThreads, smem, st>>>(a, it, first);
  } else {
    if (mode == 0)
      rate_pass_kernel<ARM, 0, false, BULK><<<grid, kRateThreads, smem, st>>>(a, it, first);
    else
      rate_pass_kernel<ARM, 1, false, BULK><<<grid, kRateThreads, smem, st>>>(a, it, first);
  }
}
template <int ARM>
void launch_rate(const LcpArgs& a, int grid, cudaStream_t st, int mode, int it, int first) {
  if (rate_pipeline_bulk())
    launch_rate_as<ARM, true>(a, grid, st, mode, it, first);
  else
    launch_rate_as<ARM, false>(a, grid, st, mode, it, first);
}
// Persistent grid of the rate pass: resident CTAs per SM (with the pipeline's shared memory, if any) x 148. Also opts
// the staged instantiations into their dynamic shared memory size (above the 48 KB default). Per device (the attribute
// is per context).
template <int ARM, bool BULK>
int rate_grid_cap_as() {
  constexpr int smem = BULK ? rate_smem_bytes<ARM>() : 0;
  static std::mutex mu;
  static int cap[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  std::lock_guard<std::mutex> lk(mu);
  if (cap[dev] == 0) {
    if (BULK) {
      cudaFuncSetAttribute(rate_pass_kernel<ARM, 0, false, BULK>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
      cudaFuncSetAttribute(rate_pass_kernel<ARM, 1, false, BULK>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
      cudaFuncSetAttribute(rate_pass_kernel<ARM, 0, true, BULK>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
      cudaFuncSetAttribute(rate_pass_kernel<ARM, 1, true, BULK>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    }
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rate_pass_kernel<ARM, 1, true, BULK>, kRateThreads, smem) !=
            cudaSuccess ||
        per_sm < 1)
      per_sm = 1;
    cap[dev] = per_sm * kSMs;
  }
  return cap[dev];
}
template <int ARM>
int rate_grid_cap() {
  return rate_pipeline_bulk() ? rate_grid_cap_as<ARM, true>() : rate_grid_cap_as<ARM, false>();
}

}  // namespace

// Grid of the rate pass: fixed by the occupancy of the kernel (a persistent grid), hence the same for every launch of a
// solve; it is the shape of the deterministic reductions (mb200_lcp_reduction_shape).
int rate_grid_for(int64_t npairs, int arm) {
  const int cap = arm == ARM_NONE ? rate_grid_cap<ARM_NONE>()
                                  : (arm == ARM_AXIAL ? rate_grid_cap<ARM_AXIAL>() : rate_grid_cap<ARM_GENERAL>());
  return grid_for(npairs, kRateThreads, cap);
}

int resolve_records(cudaStream_t st, const RecordSystem& sys, const Incidence& inc, const mb200_lcp_params* params,
                    const double* U_ext, const double* W_ext, const double* gamma0, double* gamma_out,
                    double* phi0_out, double* F, double* T, double* U, double* W, mb200_collision_result* result_h,
                    const DistSolve* dist) {
  const int64_t n = sys.n, C = sys.npairs;
  const int arm = sys.arm;
  const bool six = arm != ARM_NONE;
  const bool fixed = params->fixed_iters >= 0;
  const int it_cap = fixed ? params->fixed_iters : params->max_iters;

  static thread_local PinnedState pinned;
  if (!pinned.h)
    MB200_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&pinned.h), sizeof(LcpState), cudaHostAllocPortable));

  if (C == 0 && !dist) {  // nothing to solve: zero force / velocity, residual identity of the max-reduction (-1)
    if (F) MB200_CUDA(cudaMemsetAsync(F, 0, 3 * n * sizeof(double), st));
    if (six && T) MB200_CUDA(cudaMemsetAsync(T, 0, 3 * n * sizeof(double), st));
    if (U) MB200_CUDA(cudaMemsetAsync(U, 0, 3 * n * sizeof(double), st));
    if (six && W) MB200_CUDA(cudaMemsetAsync(W, 0, 3 * n * sizeof(double), st));
    MB200_CUDA(cudaStreamSynchronize(st));
    if (result_h) *result_h = mb200_collision_result{-1.0, 0, 1, 0.0};
    return MB200_OK;
  }

  const int rate_grid = rate_grid_for(C, arm);
  const int64_t n_boundary = (dist && dist->peer_dev) ? dist->n_boundary : 0;
  const int body_grid = grid_for(n + (n_boundary + kBodyThreads - 1) / kBodyThreads * kBodyThreads, kBodyThreads);
  const int cgrid = grid_for(C, 256);
  const size_t words = static_cast<size_t>((C + 31) / 32) + 1;

  Scratch<double> partials;
  Scratch<double2> s0;
  Scratch<D4> vel, mob;
  Scratch<LcpState> state;
  Scratch<uint32_t> active;
  Scratch<unsigned long long> bmask;
  Scratch<uint8_t> vzero;
  MB200_TRY(s0.alloc(C > 0 ? C : 1, st));
  MB200_TRY(active.alloc(words, st));
  MB200_TRY(bmask.alloc(2 * static_cast<size_t>(n > 0 ? n : 1), st));
  MB200_CUDA(cudaMemsetAsync(bmask.p, 0, 2 * static_cast<size_t>(n > 0 ? n : 1) * sizeof(unsigned long long), st));
  MB200_TRY(vzero.alloc(n > 0 ? n : 1, st));
  MB200_TRY(partials.alloc(4 * static_cast<size_t>(rate_grid), st));
  const bool peer_mode = dist && dist->peer_dev;
  if (!peer_mode) MB200_TRY(vel.alloc(static_cast<size_t>(six ? 2 : 1) * n, st));
  MB200_TRY(mob.alloc(n, st));
  MB200_TRY(state.alloc(1, st));
  MB200_CUDA(cudaMemsetAsync(state.p, 0, sizeof(LcpState), st));
  MB200_CUDA(cudaMemsetAsync(vzero.p, 0, n > 0 ? n : 1, st));

  LcpArgs a;
  a.n = n, a.npairs = C;
  a.pairs = sys.pairs, a.rec = sys.rec, a.sig = sys.sig, a.half = sys.half, a.mob = mob.p;
  a.inc_ptr = inc.ptr.p, a.inc_list = inc.list.p, a.inc_epos = inc.epos.p;
  a.state = s0.p, a.active = active.p, a.bmask = bmask.p, a.vzero = vzero.p;
  a.vel = peer_mode ? dist->vel_arena : vel.p;
  a.peer = peer_mode ? dist->peer_dev : nullptr;
  a.halo_seq = 0, a.red_seq = 0;
  a.owned = peer_mode ? dist->owned_s : nullptr;
  a.push_ptr = peer_mode ? dist->push_ptr : nullptr;
  a.push_dst = peer_mode ? dist->push_dst : nullptr;
  a.boundary = peer_mode ? dist->boundary : nullptr;
  a.n_boundary = n_boundary;
  a.mu = sys.mu, a.dt = sys.dt, a.tol = params->tol;
  a.fixed = fixed ? 1 : 0, a.it_cap = it_cap;
  a.state_blk = state.p, a.partials = partials.p;
  a.counted = dist ? dist->counted : nullptr;
  a.red_local = dist ? dist->red : nullptr;
  const NcclApi* nccl = dist ? nccl_api() : nullptr;
  if (dist && !nccl) return MB200_ERR_COMM;
  const int recs = six ? 2 : 1;
  const unsigned long long halo_seq0 = peer_mode ? *dist->halo_seq : 0, red_seq0 = peer_mode ? *dist->red_seq : 0;
  // multi-GPU: ghost rows of the velocity records come from their owners after every body pass
  auto halo = [&]() -> int {
    if (!dist || peer_mode) return MB200_OK;  // peer mode: the body pass has pushed the records itself
    StageScope ps(MB200_STAGE_LCP_COMM, st);
    const int64_t ns = dist->send_off[dist->npeers], nr = dist->recv_off[dist->npeers];
    if (ns > 0) {
      halo_pack_kernel<<<grid_for(ns * recs, 256), 256, 0, st>>>(ns, recs, dist->send_idx, vel.p, dist->sendbuf, state.p);
      MB200_LAUNCHED();
    }
    MB200_NCCL(nccl->GroupStart());
    for (int k = 0; k < dist->npeers; ++k) {
      const int64_t cs = dist->send_off[k + 1] - dist->send_off[k], cr = dist->recv_off[k + 1] - dist->recv_off[k];
      if (cs > 0)
        MB200_NCCL(nccl->Send(dist->sendbuf + dist->send_off[k] * recs, static_cast<size_t>(cs) * recs * 4, ncclDouble,
                              dist->peer[k], dist->comm->comm, st));
      if (cr > 0)
        MB200_NCCL(nccl->Recv(dist->recvbuf + dist->recv_off[k] * recs, static_cast<size_t>(cr) * recs * 4, ncclDouble,
                              dist->peer[k], dist->comm->comm, st));
    }
    MB200_NCCL(nccl->GroupEnd());
    if (nr > 0) {
      halo_unpack_kernel<<<grid_for(nr * recs, 256), 256, 0, st>>>(nr, recs, dist->recv_idx, dist->recvbuf, vel.p, state.p);
      MB200_LAUNCHED();
    }
    return MB200_OK;
  };
  // multi-GPU: all-gather of the ranks' totals (rank order => identical sums everywhere), then the state update
  auto reduce = [&](int it_now, int mode) -> int {
    if (!dist || peer_mode) return MB200_OK;  // peer mode: the rate pass's last CTA has done it
    StageScope ps(MB200_STAGE_LCP_COMM, st);
    MB200_NCCL(nccl->AllGather(dist->red, dist->red + 4, 4, ncclDouble, dist->comm->comm, st));
    advance_kernel<<<1, 32, 0, st>>>(state.p, dist->red + 4, dist->comm->world, it_now, mode, a.fixed, a.tol, a.it_cap);
    MB200_LAUNCHED();
    return MB200_OK;
  };
  // MODE 0 (state holds gamma) / 1 (gamma re-derived), FINAL (outputs instead of velocity records)
  auto body = [&](int mode, bool fin_pass, int first_it, double* Fo, double* To, double* Uo, double* Wo) -> int {
    StageScope ps(MB200_STAGE_LCP_BODY, st);
    if (arm == ARM_NONE)
      launch_body<ARM_NONE>(a, body_grid, st, mode, fin_pass, first_it, Fo, To, Uo, Wo);
    else if (arm == ARM_AXIAL)
      launch_body<ARM_AXIAL>(a, body_grid, st, mode, fin_pass, first_it, Fo, To, Uo, Wo);
    else
      launch_body<ARM_GENERAL>(a, body_grid, st, mode, fin_pass, first_it, Fo, To, Uo, Wo);
    MB200_LAUNCHED();
    return MB200_OK;
  };
  auto rate = [&](int mode, int it_now, int first_it) -> int {
    StageScope ps(MB200_STAGE_LCP_RATE, st);
    if (arm == ARM_NONE)
      launch_rate<ARM_NONE>(a, rate_grid, st, mode, it_now, first_it);
    else if (arm == ARM_AXIAL)
      launch_rate<ARM_AXIAL>(a, rate_grid, st, mode, it_now, first_it);
    else
      launch_rate<ARM_GENERAL>(a, rate_grid, st, mode, it_now, first_it);
    MB200_LAUNCHED();
    return MB200_OK;
  };

  {
    StageScope ps(MB200_STAGE_LCP_SETUP, st);
    mobility_records_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, sys.radius, sys.quat, sys.kind, mob.p);
    MB200_LAUNCHED();
    init_state_kernel<<<grid_for(C > 0 ? C : 1, 256), 256, 0, st>>>(C, gamma0, s0.p, active.p, sys.pairs, inc.epos.p,
                                                                    inc.ptr.p, bmask.p);
    MB200_LAUNCHED();
    if (U_ext && C > 0) {
      if (arm == ARM_NONE)
        ext_correction_kernel<ARM_NONE><<<cgrid, 256, 0, st>>>(C, sys.pairs, sys.rec, nullptr, nullptr, U_ext, nullptr, a.dt);
      else if (arm == ARM_AXIAL)
        ext_correction_kernel<ARM_AXIAL><<<cgrid, 256, 0, st>>>(C, sys.pairs, sys.rec, sys.sig, sys.half, U_ext, W_ext, a.dt);
      else
        ext_correction_kernel<ARM_GENERAL><<<cgrid, 256, 0, st>>>(C, sys.pairs, sys.rec, nullptr, nullptr, U_ext, W_ext, a.dt);
      MB200_LAUNCHED();
    }
  }

  // initial residual: state = {gamma_0, g_0 = D^T M D gamma_0}
  if (peer_mode) a.halo_seq = ++(*dist->halo_seq);
  MB200_TRY(body(0, false, 0, nullptr, nullptr, nullptr, nullptr));
  MB200_TRY(halo());
  if (peer_mode) a.red_seq = ++(*dist->red_seq);
  MB200_TRY(rate(0, 0, 0));
  MB200_TRY(reduce(0, 0));

  int it = 0;
  int batch = 4;
  LcpState hs;
  for (;;) {
    MB200_CUDA(cudaMemcpyAsync(pinned.h, state.p, sizeof(LcpState), cudaMemcpyDeviceToHost, st));
    MB200_CUDA(cudaStreamSynchronize(st));
    hs = *pinned.h;
    if (hs.done || it >= it_cap) break;
    const int stop = (it + batch < it_cap) ? it + batch : it_cap;
    while (it < stop) {
      ++it;
      const int first = it == 1 ? 1 : 0;
      if (peer_mode) a.halo_seq = ++(*dist->halo_seq);
      MB200_TRY(body(1, false, first, nullptr, nullptr, nullptr, nullptr));
      MB200_TRY(halo());
      if (peer_mode) a.red_seq = ++(*dist->red_seq);
      MB200_TRY(rate(1, it, first));
      MB200_TRY(reduce(it, 1));
    }
    if (batch < 32) batch *= 2;
  }

  // the state of the last executed iteration is in place
  const int it_final = hs.it;
  if (peer_mode) {
    // Launches enqueued after `done` was set returned at once and published nothing: the sequence numbers continue
    // from the last exchange that really happened (identically on every rank: all ranks see the same `done`), so that
    // consecutive executed reductions always alternate their parity buffers.
    *dist->halo_seq = halo_seq0 + 1 + static_cast<unsigned long long>(it_final);
    *dist->red_seq = red_seq0 + 1 + static_cast<unsigned long long>(it_final);
  }
  MB200_TRY(body(0, true, 0, F, six ? T : nullptr, U, six ? W : nullptr));
  if ((gamma_out || phi0_out) && C > 0) {
    unpack_contacts_kernel<<<cgrid, 256, 0, st>>>(C, sys.rec, s0.p, phi0_out, gamma_out);
    MB200_LAUNCHED();
  }
  MB200_CUDA(cudaMemcpyAsync(pinned.h, state.p, sizeof(LcpState), cudaMemcpyDeviceToHost, st));
  MB200_CUDA(cudaStreamSynchronize(st));
  hs = *pinned.h;
  if (hs.error) {
    set_error("multi-GPU exchange timed out: a peer rank did not publish its velocity halo / reduction totals");
    return MB200_ERR_COMM;
  }
  if (result_h) {
    result_h->max_abs_projected_sep = hs.res;
    result_h->ite_count = it_final;
    result_h->converged = (hs.res < params->tol) ? 1 : 0;
    result_h->max_displacement = hs.vmax * sys.dt;
  }
  return MB200_OK;
}

// Caller-layout front end: pack the contact arrays into records, solve, write phi0 / gamma back.
int resolve_collisions_impl(cudaStream_t st, const mb200_contact_op* op, const Incidence& inc,
                            const mb200_lcp_params* params, const double* U_ext, const double* W_ext, double* phi0,
                            double* gamma, double* F, double* T, double* U, double* W,
                            mb200_collision_result* result_h) {
  const int64_t C = op->npairs;
  const bool six = op->arm_i != nullptr;
  Scratch<D4> rec;
  MB200_TRY(rec.alloc(static_cast<size_t>(six ? 3 : 1) * (C > 0 ? C : 1), st));
  const int2* pairs = reinterpret_cast<const int2*>(op->pairs);
  if (C > 0) {
    if (six)
      pack_contacts_kernel<ARM_GENERAL><<<grid_for(C, 256), 256, 0, st>>>(C, op->normal, op->arm_i, op->arm_j, phi0, rec.p);
    else
      pack_contacts_kernel<ARM_NONE><<<grid_for(C, 256), 256, 0, st>>>(C, op->normal, nullptr, nullptr, phi0, rec.p);
    MB200_LAUNCHED();
  }
  RecordSystem sys{op->n, C, six ? ARM_GENERAL : ARM_NONE, pairs, rec.p, nullptr, nullptr, op->quat, op->radius, op->kind,
                   op->viscosity, op->dt};
  return resolve_records(st, sys, inc, params, U_ext, W_ext, gamma, gamma, U_ext ? phi0 : nullptr, F, T, U, W, result_h,
                         nullptr);
}

}  // namespace mb200

using namespace mb200;

extern "C" int mb200_resolve_collisions_axial(mb200_stream stream, const mb200_contact_op* op, const double* sigma,
                                              const double* length, const mb200_lcp_params* params,
                                              const double* U_ext, const double* W_ext, double* phi0, double* gamma,
                                              double* F, double* T, double* U, double* W,
                                              mb200_collision_result* result_h) {
  MB200_REQUIRE(op && params && op->n >= 0 && op->npairs >= 0, "bad arguments");
  MB200_REQUIRE(op->n < (int64_t(1) << 31) && op->npairs < (int64_t(1) << 30), "problem too large for int32 ids");
  MB200_REQUIRE(!op->arm_i && !op->arm_j, "axial form: arm_i / arm_j must be NULL");
  MB200_REQUIRE(op->n == 0 || (op->radius && op->quat && op->kind && length), "axial form needs radius, quat, kind, length");
  MB200_REQUIRE((reinterpret_cast<uintptr_t>(sigma) & 15) == 0, "sigma must be 16-byte aligned");
  if (op->npairs > 0) MB200_REQUIRE(op->pairs && op->normal && phi0 && gamma && sigma, "null contact arrays");
  MB200_TRY(require_device());
  cudaStream_t st = as_stream(stream);
  const int64_t n = op->n, C = op->npairs;
  Incidence inc;
  MB200_TRY(build_incidence(st, n, C, op->pairs, &inc));
  MB200_TRY(check_incidence(st, inc));
  Scratch<D4> rec, half;
  MB200_TRY(rec.alloc(C > 0 ? C : 1, st));
  MB200_TRY(half.alloc(n > 0 ? n : 1, st));
  if (n > 0) {
    half_records_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, op->quat, length, op->kind, half.p);
    MB200_LAUNCHED();
  }
  if (C > 0) {
    pack_contacts_kernel<ARM_NONE><<<grid_for(C, 256), 256, 0, st>>>(C, op->normal, nullptr, nullptr, phi0, rec.p);
    MB200_LAUNCHED();
  }
  RecordSystem sys{n, C, ARM_AXIAL, reinterpret_cast<const int2*>(op->pairs), rec.p,
                   reinterpret_cast<const double2*>(sigma), half.p, op->quat, op->radius, op->kind, op->viscosity, op->dt};
  return resolve_records(st, sys, inc, params, U_ext, W_ext, gamma, gamma, U_ext ? phi0 : nullptr, F, T, U, W, result_h,
                         nullptr);
}

extern "C" int mb200_lcp_reduction_shape(int64_t npairs, int arm_mode, int32_t* grid_h, int32_t* threads_h) {
  MB200_REQUIRE(npairs >= 0 && arm_mode >= ARM_NONE && arm_mode <= ARM_GENERAL && grid_h && threads_h, "bad arguments");
  MB200_TRY(require_device());
  *grid_h = rate_grid_for(npairs, arm_mode);
  *threads_h = kRateThreads;
  return MB200_OK;
}

extern "C" int mb200_resolve_collisions(mb200_stream stream, const mb200_contact_op* op, const mb200_lcp_params* params,
                                        const double* U_ext, const double* W_ext, double* phi0, double* gamma,
                                        double* F, double* T, double* U, double* W,
                                        mb200_collision_result* result_h) {
  MB200_REQUIRE(op && params && op->n >= 0 && op->npairs >= 0, "bad arguments");
  MB200_REQUIRE(op->n < (int64_t(1) << 31) && op->npairs < (int64_t(1) << 30), "problem too large for int32 ids");
  MB200_REQUIRE((op->arm_i == nullptr) == (op->arm_j == nullptr), "arm_i/arm_j must both be set or NULL");
  MB200_REQUIRE(op->radius || op->n == 0, "radius is NULL");
  if (op->npairs > 0) MB200_REQUIRE(op->pairs && op->normal && phi0 && gamma, "null contact arrays");
  MB200_TRY(require_device());
  cudaStream_t st = as_stream(stream);
  Incidence inc;
  MB200_TRY(build_incidence(st, op->n, op->npairs, op->pairs, &inc));
  MB200_TRY(check_incidence(st, inc));
  return resolve_collisions_impl(st, op, inc, params, U_ext, W_ext, phi0, gamma, F, T, U, W, result_h);
}
