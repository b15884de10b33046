// E1. Driver BBPGD (scrap/lcp_spheres/KokkosLCP.cpp:675-876), device-resident.
//
// Data layout in HBM (all internal; Morton order when called from the step):
//   contact records, 32 bytes each (one DRAM sector, moved with one 256-bit load):
//       R0 = {nx, ny, nz, phi0}                       3-DOF and 6-DOF
//       R1 = {arm_i.x, arm_i.y, arm_i.z, (i, j)}      6-DOF only; the int32 pair rides in the 4th slot
//       R2 = {arm_j.x, arm_j.y, arm_j.z, 0}           6-DOF only
//     stored as three arrays, one per record kind: rec[k C + c] (6-DOF) or rec[c] (3-DOF; the pair comes from the int2
//     list), so that a warp reading consecutive contacts touches 8 lines per load instead of 24
//   contact state {gamma, g} as double2, ping-pong between two buffers (no roll-forward copies, K13)
//   body mobility record {axis.x, axis.y, axis.z, +r (rod) | -r (sphere)}, axis = q (x) z_hat, built once per solve
//   body velocity records {U, 0} (+ {W, 0} for 6-DOF): 64 MB at 10^6 rods, gathered by the contact pass out of L2
//
// One iteration = two kernels, no host round trip (the reference blocks the host twice per iteration):
//   body_pass : for every body, gamma of its incident contacts is re-derived from the previous state
//               (gamma = max(gamma_prev - alpha (phi0 + dt g_prev), 0), KokkosLCP.cpp:655-664), the contact wrenches
//               are formed by all threads of the CTA in parallel (independent gathers), staged in shared memory and
//               then summed per body in ascending (contact, side) order: F = D gamma without atomics, bit-identical to
//               a serial loop; U = M F; velocity records written.
//   rate_pass : per contact: the same gamma, g = D^T U, new state, residual max and the three BB dot products;
//               the last CTA to finish reduces the per-CTA partials in index order and advances alpha / done.
// Every kernel returns at once when `done` is set, so the host enqueues iterations in batches and reads the 48-byte
// solver state back between batches only.
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
  uint32_t ticket2;  // halo_push_kernel's CTA counter
  int32_t error;     // multi-GPU: a peer did not answer in time
};

struct LcpArgs {  // everything the iteration kernels need; passed by value
  int64_t n, npairs;
  const int2* pairs;  // 3-DOF only
  const D4* rec;      // contact records
  const D4* mob;      // [n] mobility records
  const int64_t* inc_ptr;
  const uint32_t* inc_list;
  // entry-major mirror of the static contact data (built once per solve, after the phi0 correction): for incidence
  // entry k = (contact c, side s): ent0[k] = {n_s, phi0_c}, n_s = +/- n; ent1[k] = {r_s x n_s, -} (6-DOF only)
  const D4* ent0;
  const D4* ent1;
  D4* vel;  // [n] (3-DOF) or [2n] (6-DOF)
  double mu, dt, tol;
  int32_t fixed;   // ignore tol
  int32_t it_cap;  // max iterations (or the fixed count)
  LcpState* state;
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

__device__ inline int2 pair_of(const D4& r1) {
  const long long b = __double_as_longlong(r1.w);
  return make_int2(static_cast<int>(b & 0xffffffffll), static_cast<int>(b >> 32));
}
__device__ inline double pair_bits(int i, int j) {
  return __longlong_as_double((static_cast<long long>(j) << 32) | static_cast<long long>(static_cast<uint32_t>(i)));
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

// caller layout -> records and initial state {gamma0, 0}
template <bool SIX>
__global__ void __launch_bounds__(256)
    pack_contacts_kernel(int64_t npairs, const int2* __restrict__ pairs, const double* __restrict__ normal,
                         const double* __restrict__ arm_i, const double* __restrict__ arm_j,
                         const double* __restrict__ phi0, D4* __restrict__ rec) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= npairs) return;
  const V3 n = ld3(normal, c);
  if (SIX) {
    const int2 p = pairs[c];
    const V3 ai = ld3(arm_i, c), aj = ld3(arm_j, c);
    st256(rec + c, D4{n.x, n.y, n.z, phi0[c]});
    st256(rec + npairs + c, D4{ai.x, ai.y, ai.z, pair_bits(p.x, p.y)});
    st256(rec + 2 * npairs + c, D4{aj.x, aj.y, aj.z, 0.0});
  } else {
    st256(rec + c, D4{n.x, n.y, n.z, phi0[c]});
  }
}

template <bool SIX>
__global__ void __launch_bounds__(256)
    unpack_contacts_kernel(int64_t npairs, const D4* __restrict__ rec, const double2* __restrict__ state,
                           double* __restrict__ phi0, double* __restrict__ gamma) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= npairs) return;
  if (phi0) phi0[c] = ldg256(rec + c).w;
  if (gamma) gamma[c] = state[c].x;
}

__global__ void __launch_bounds__(256) init_state_kernel(int64_t npairs, const double* __restrict__ gamma0,
                                                         double2* __restrict__ state) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c < npairs) state[c] = make_double2(gamma0 ? gamma0[c] : 0.0, 0.0);
}

// phi0 += dt * D^T (U_ext, W_ext)   scrap/parameter_interface/NGPSpheresLCP.cpp:751-756
template <bool SIX>
__global__ void __launch_bounds__(256)
    ext_correction_kernel(int64_t npairs, const int2* __restrict__ pairs, D4* rec, const double* __restrict__ U_ext,
                          const double* __restrict__ W_ext, double dt) {
  const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= npairs) return;
  D4* r0p = rec + c;
  D4 r0 = ld256(r0p);
  const V3 n{r0.x, r0.y, r0.z};
  V3 wi{0, 0, 0}, wj{0, 0, 0}, ai{0, 0, 0}, aj{0, 0, 0};
  int2 p;
  const bool six = SIX && W_ext;
  if (SIX) {
    const D4 r1 = ld256(rec + npairs + c), r2 = ld256(rec + 2 * npairs + c);
    p = pair_of(r1);
    ai = V3{r1.x, r1.y, r1.z}, aj = V3{r2.x, r2.y, r2.z};
    if (six) wi = ld3(W_ext, p.x), wj = ld3(W_ext, p.y);
  } else {
    p = pairs[c];
  }
  const double g = six ? contact_rate<true>(n, ld3(U_ext, p.x), ld3(U_ext, p.y), wi, wj, ai, aj)
                       : contact_rate<false>(n, ld3(U_ext, p.x), ld3(U_ext, p.y), wi, wj, ai, aj);
  r0.w += dt * g;
  st256(r0p, r0);
}

// ---- iteration kernels ---------------------------------------------------------------------------------------------
constexpr int kBodyThreads = 128;
constexpr int kBodiesPerCta = 64;
constexpr int kCap = 512;  // incidence entries staged per round (avg ~6.4 per rod at cfg3 => one round per CTA)
constexpr int kUnroll = 4;  // entries in flight per thread

__device__ inline void cp_async4(void* smem_dst, const void* gsrc) {
  const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ inline void cp_async8(void* smem_dst, const void* gsrc) {
  const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ inline void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// Static per-entry data in incidence order, so that the body pass streams it (two coalesced 256-bit loads per entry)
// and only the 16-byte state of the contact is gathered.
template <bool SIX>
__global__ void __launch_bounds__(256) entry_mirror_kernel(int64_t m, int64_t npairs, const uint32_t* __restrict__ list,
                                                           const D4* __restrict__ rec, D4* __restrict__ ent0,
                                                           D4* __restrict__ ent1) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= m) return;
  const uint32_t e = __ldg(list + k);
  const int64_t c = e >> 1;
  const bool is_j = e & 1u;
  const D4 r0 = ldg256(rec + c);
  const double mx = is_j ? -r0.x : r0.x, my = is_j ? -r0.y : r0.y, mz = is_j ? -r0.z : r0.z;
  st256(ent0 + k, D4{mx, my, mz, r0.w});
  if (SIX) {
    const D4 ra = ldg256(rec + (is_j ? 2 : 1) * npairs + c);
    st256(ent1 + k, D4{ra.y * mz - ra.z * my, ra.z * mx - ra.x * mz, ra.x * my - ra.y * mx, 0.0});
  }
}

// End of a body's pass: U = M F from the summed wrench; velocity records (and, multi-GPU, their ghost copies on the
// peers) or, in the FINAL pass, the solver's outputs.
template <bool SIX, bool FINAL>
__device__ inline void body_epilogue(const LcpArgs& a, int64_t b, const double* acc, const D4& mr, bool own, int64_t pk,
                                     int64_t pke, double& speed, double* __restrict__ F, double* __restrict__ T,
                                     double* __restrict__ U, double* __restrict__ W, const double* __restrict__ U_ext,
                                     const double* __restrict__ W_ext) {
  {
      const bool rod = SIX && mr.w > 0.0;
      const Mob m = mobility_of(fabs(mr.w), a.mu);
      const V3 f{acc[0], acc[1], acc[2]};
      const V3 tq = SIX ? V3{acc[3], acc[4], acc[5]} : V3{0.0, 0.0, 0.0};
      V3 u, w;
      body_mobility_axis(m, rod, V3{mr.x, mr.y, mr.z}, f, tq, u, w);
      if (!FINAL) {
        if (!own) {
          // ghost row: its owner stores into it from its own GPU (below, on that rank)
        } else {
          if (SIX) {
            st256(a.vel + 2 * b, D4{u.x, u.y, u.z, 0.0});
            st256(a.vel + 2 * b + 1, D4{w.x, w.y, w.z, 0.0});
          } else {
            st256(a.vel + b, D4{u.x, u.y, u.z, 0.0});
          }
          if (a.push_ptr) {  // fused velocity halo: the records also go straight into the peers' ghost rows (NVLink)
            for (int64_t k = pk; k < pke; ++k) {
              const uint32_t d = __ldg(a.push_dst + k);
              D4* dst = a.peer->vel[d >> 28];
              const int64_t row = d & 0x0fffffffu;
              if (SIX) {
                st256(dst + 2 * row, D4{u.x, u.y, u.z, 0.0});
                st256(dst + 2 * row + 1, D4{w.x, w.y, w.z, 0.0});
              } else {
                st256(dst + row, D4{u.x, u.y, u.z, 0.0});
              }
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
}

// MODE 0: gamma is read from the state (initial pass and final pass). MODE 1: gamma re-derived from the previous state.
// FINAL: also writes F,T,U,W in the solver's body order and reduces max |U| into state->vmax (single launch at the end).
// Persistent CTAs walk blocks of kBodiesPerCta bodies. The chain of dependent loads of a block (incidence pointers ->
// incidence list -> contact records) is software-pipelined across blocks with cp.async: while block k is processed,
// the list of block k + G and the pointers of block k + 2G stream into shared memory, so that only the record gathers
// remain on the critical path.
template <bool SIX, int MODE, bool FINAL>
__global__ void __launch_bounds__(kBodyThreads)
    body_pass_kernel(LcpArgs a, int first, const double2* __restrict__ st_prev, double* __restrict__ F,
                     double* __restrict__ T, double* __restrict__ U, double* __restrict__ W,
                     const double* __restrict__ U_ext, const double* __restrict__ W_ext) {
  if (!FINAL && a.state->done) return;
  constexpr int NC = SIX ? 6 : 3;
  __shared__ double s_w[NC][kCap];
  __shared__ double sh[33];
  __shared__ __align__(8) int64_t s_ptr[3][kBodiesPerCta + 1];
  __shared__ uint32_t s_list[2][kCap];
  const double alpha = (MODE == 1) ? a.state->alpha : 0.0;
  const int t = threadIdx.x;
  const int64_t nblk = (a.n + kBodiesPerCta - 1) / kBodiesPerCta;
  const int64_t G = gridDim.x;
  int64_t blk = blockIdx.x;
  double speed = 0.0;

  auto bodies_in = [&](int64_t k) {
    return static_cast<int>(min(static_cast<int64_t>(kBodiesPerCta), a.n - k * kBodiesPerCta));
  };
  auto fetch_ptrs = [&](int64_t k, int buf) {  // async: inc_ptr[k * B .. k * B + nb]
    if (k < nblk && t <= bodies_in(k)) cp_async8(&s_ptr[buf][t], a.inc_ptr + k * kBodiesPerCta + t);
  };
  auto fetch_list = [&](int64_t k, int pbuf, int lbuf) {  // async: first kCap incidence entries of block k
    if (k >= nblk) return;
    const int64_t eb = s_ptr[pbuf][0];
    const int cnt = static_cast<int>(min(static_cast<int64_t>(kCap), s_ptr[pbuf][bodies_in(k)] - eb));
    for (int j = t; j < cnt; j += kBodyThreads) cp_async4(&s_list[lbuf][j], a.inc_list + eb + j);
  };

  if (blk < nblk) {  // prologue: pointers of the first two blocks, list of the first
    fetch_ptrs(blk, 0);
    fetch_ptrs(blk + G, 1);
    cp_async_wait_all();
    __syncthreads();
    fetch_list(blk, 0, 0);
    cp_async_wait_all();
    __syncthreads();
  }

  for (int stage = 0; blk < nblk; blk += G, ++stage) {
    const int pcur = stage % 3, pnext = (stage + 1) % 3, pnn = (stage + 2) % 3, lcur = stage & 1;
    fetch_ptrs(blk + 2 * G, pnn);
    fetch_list(blk + G, pnext, lcur ^ 1);

    const int64_t b0 = blk * kBodiesPerCta;
    const int nb = bodies_in(blk);
    D4 mr{0.0, 0.0, 1.0, -1.0};
    int64_t pk = 0, pke = 0;
    bool own = true;
    if (t < nb) {
      mr = ldg256(a.mob + b0 + t);  // needed only at the end: issued now to overlap with the gathers
      if (!FINAL && a.owned) own = __ldg(a.owned + b0 + t) != 0;
      if (!FINAL && a.push_ptr) pk = __ldg(a.push_ptr + b0 + t), pke = __ldg(a.push_ptr + b0 + t + 1);
    }
    const int64_t e_begin = s_ptr[pcur][0], e_end = s_ptr[pcur][nb];
    int64_t my_beg = 0, my_end = 0;
    if (t < nb) my_beg = s_ptr[pcur][t], my_end = s_ptr[pcur][t + 1];
    double acc[NC];
#pragma unroll
    for (int k = 0; k < NC; ++k) acc[k] = 0.0;

    for (int64_t rb = e_begin; rb < e_end; rb += kCap) {
      const int cnt = static_cast<int>(min(static_cast<int64_t>(kCap), e_end - rb));
      const bool staged = rb == e_begin;  // later rounds (more than kCap entries in the block) read the list directly
      // phase 1: every thread forms the wrench of up to kUnroll (contact, side) entries; the loads of all of them are
      // issued before the first use (independent gathers: memory-level parallelism instead of occupancy)
      for (int k0 = 0; k0 < cnt; k0 += kBodyThreads * kUnroll) {
        uint32_t e[kUnroll];
        double2 sp[kUnroll];
        D4 r0[kUnroll], ra[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          const int k = k0 + u * kBodyThreads + t;
          e[u] = (k < cnt) ? (staged ? s_list[lcur][k] : __ldg(a.inc_list + rb + k)) : 0xffffffffu;
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          if (e[u] != 0xffffffffu) {
            const int64_t k = rb + k0 + u * kBodyThreads + t;
            sp[u] = ldg128(st_prev + (e[u] >> 1));  // the only gather: 16 bytes of state per entry
            r0[u] = ldg256(a.ent0 + k);             // streamed
            if (SIX) ra[u] = ldg256(a.ent1 + k);
          }
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          if (e[u] != 0xffffffffu) {
            const int k = k0 + u * kBodyThreads + t;
            const double g = gamma_of<MODE>(sp[u], r0[u].w, alpha, a.dt, first);
            // n_I = n, n_J = -n: F += -gamma n_side, T += -gamma (r_side x n_side)  (scrap/UnitTestMundy.cpp:652-674)
            s_w[0][k] = -g * r0[u].x;
            s_w[1][k] = -g * r0[u].y;
            s_w[2][k] = -g * r0[u].z;
            if (SIX) {
              s_w[3][k] = -g * ra[u].x;
              s_w[4][k] = -g * ra[u].y;
              s_w[5][k] = -g * ra[u].z;
            }
          }
        }
      }
      __syncthreads();
      // phase 2: one thread per body adds its entries in ascending (contact, side) order
      if (t < nb) {
        const int lo = static_cast<int>(max(my_beg, rb) - rb);
        const int hi = static_cast<int>(min(my_end, rb + cnt) - rb);
        for (int k = lo; k < hi; ++k) {
#pragma unroll
          for (int j = 0; j < NC; ++j) acc[j] += s_w[j][k];
        }
      }
      __syncthreads();
    }

    if (t < nb)
      body_epilogue<SIX, FINAL>(a, b0 + t, acc, mr, own, pk, pke, speed, F, T, U, W, U_ext, W_ext);
    cp_async_wait_all();  // next block's list and the block after's pointers have landed
    __syncthreads();
  }
  if (!FINAL && a.push_ptr) {
    // every CTA's remote stores are ordered before the flag the last CTA publishes to all ranks (release, system scope);
    // the consumers are the peers' rate_pass_kernels
    // (bar.sync orders the CTA's stores before thread 0's system-scope fence: the grid-sync pattern, one fence per CTA)
    __syncthreads();
    if (t == 0) {
      __threadfence_system();
      const uint32_t ticket = atomicAdd(&a.state->ticket2, 1u);
      sh[32] = (ticket == gridDim.x - 1) ? 1.0 : 0.0;
      if (ticket == gridDim.x - 1) a.state->ticket2 = 0;
    }
    __syncthreads();
    if (sh[32] != 0.0 && t < a.peer->world && t != a.peer->me) {  // one lane per rank: the flags leave in parallel
      __threadfence_system();
      st_relaxed_sys(&a.peer->ctrl[t]->halo_flag[a.peer->me], a.halo_seq);
    }
  }
  if (FINAL) {
    const double m = block_max(speed, sh, 0.0);
    if (threadIdx.x == 0) {  // non-negative doubles order like their bit patterns
      atomicMax(reinterpret_cast<unsigned long long*>(&a.state->vmax),
                static_cast<unsigned long long>(__double_as_longlong(m)));
    }
  }
}

constexpr int kRateThreads = 256;

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

// MODE 0: initial pass (state holds gamma_0; g_0 = D^T M D gamma_0, residual only). MODE 1: iteration pass.
// Multi-GPU with peer memory (a.peer != NULL): every CTA first waits until all other ranks have published velocity halo
// a.halo_seq; the last CTA stores this rank's four totals into every rank's control block, publishes a.red_seq, waits for
// the other ranks' totals and reduces them in rank order, so that every rank advances to the same alpha / done without
// a collective call or a host round trip.
template <bool SIX, int MODE>
__global__ void __launch_bounds__(kRateThreads)
    rate_pass_kernel(LcpArgs a, int it, int first, const double2* __restrict__ st_prev, double2* __restrict__ st_cur) {
  if (a.state->done) return;
  __shared__ double sh[33];
  __shared__ bool is_last;
  const double alpha = (MODE == 1) ? a.state->alpha : 0.0;
  double res = -1.0, xx = 0.0, xg = 0.0, gg = 0.0;
  const int t = threadIdx.x;
  if (a.peer) {
    if (t < a.peer->world && t != a.peer->me && !wait_flag(&a.peer->ctrl[a.peer->me]->halo_flag[t], a.halo_seq)) {
      a.state->error = 1;
      a.state->done = 1;
    }
    __syncthreads();
  }
  const int64_t stride = static_cast<int64_t>(gridDim.x) * kRateThreads;
  for (int64_t c = static_cast<int64_t>(blockIdx.x) * kRateThreads + t; c < a.npairs; c += stride) {
    const double2 sp = ldg128(st_prev + c);
    const D4 r0 = ldg256(a.rec + c);
    const V3 n{r0.x, r0.y, r0.z};
    V3 ui, uj, wi{0, 0, 0}, wj{0, 0, 0}, ai{0, 0, 0}, aj{0, 0, 0};
    if (SIX) {
      const D4 r1 = ldg256(a.rec + a.npairs + c), r2 = ldg256(a.rec + 2 * a.npairs + c);
      const int2 p = pair_of(r1);
      const D4 v0 = ldg256(a.vel + 2 * p.x), v1 = ldg256(a.vel + 2 * p.x + 1);
      const D4 v2 = ldg256(a.vel + 2 * p.y), v3 = ldg256(a.vel + 2 * p.y + 1);
      ui = V3{v0.x, v0.y, v0.z}, wi = V3{v1.x, v1.y, v1.z};
      uj = V3{v2.x, v2.y, v2.z}, wj = V3{v3.x, v3.y, v3.z};
      ai = V3{r1.x, r1.y, r1.z}, aj = V3{r2.x, r2.y, r2.z};
    } else {
      const int2 p = __ldg(a.pairs + c);
      const D4 v0 = ldg256(a.vel + p.x), v2 = ldg256(a.vel + p.y);
      ui = V3{v0.x, v0.y, v0.z}, uj = V3{v2.x, v2.y, v2.z};
    }
    const double gam = gamma_of<MODE>(sp, r0.w, alpha, a.dt, first);
    const double g = contact_rate<SIX>(n, ui, uj, wi, wj, ai, aj);
    st_cur[c] = make_double2(gam, g);
    if (a.counted && !a.counted[c]) continue;  // a cut contact is summed by the rank that owns its first body
    const double sep_new = r0.w + a.dt * g;  // KokkosLCP.cpp:500-521
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
    const uint32_t ticket = atomicAdd(&a.state->ticket, 1u);
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
  if (a.peer) {
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
      LcpState* s = a.state;
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
    LcpState* s = a.state;
    s->ticket = 0;
    if (a.red_local) {  // multi-GPU over NCCL: the totals go through an all-gather, advance_kernel finishes
      a.red_local[0] = res, a.red_local[1] = xx, a.red_local[2] = xg, a.red_local[3] = gg;
    } else {
      advance_state(s, res, xx, xg, gg, it, MODE, a.fixed, a.tol, a.it_cap);
    }
  }
}

struct PinnedState {
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

}  // namespace

int resolve_records(cudaStream_t st, const RecordSystem& sys, const Incidence& inc, const mb200_lcp_params* params,
                    const double* U_ext, const double* W_ext, const double* gamma0, double* gamma_out,
                    double* phi0_out, double* F, double* T, double* U, double* W, mb200_collision_result* result_h,
                    const DistSolve* dist) {
  const int64_t n = sys.n, C = sys.npairs;
  const bool six = sys.six;
  const bool fixed = params->fixed_iters >= 0;
  const int it_cap = fixed ? params->fixed_iters : params->max_iters;

  static thread_local PinnedState pinned;
  if (!pinned.h) MB200_CUDA(cudaMallocHost(reinterpret_cast<void**>(&pinned.h), sizeof(LcpState)));

  if (C == 0 && !dist) {  // nothing to solve: zero force / velocity, residual identity of the max-reduction (-1)
    if (F) MB200_CUDA(cudaMemsetAsync(F, 0, 3 * n * sizeof(double), st));
    if (six && T) MB200_CUDA(cudaMemsetAsync(T, 0, 3 * n * sizeof(double), st));
    if (U) MB200_CUDA(cudaMemsetAsync(U, 0, 3 * n * sizeof(double), st));
    if (six && W) MB200_CUDA(cudaMemsetAsync(W, 0, 3 * n * sizeof(double), st));
    MB200_CUDA(cudaStreamSynchronize(st));
    if (result_h) *result_h = mb200_collision_result{-1.0, 0, 1, 0.0};
    return MB200_OK;
  }

  static const int rate_cap6 = resident_grid(rate_pass_kernel<true, 1>, kRateThreads);
  static const int rate_cap3 = resident_grid(rate_pass_kernel<false, 1>, kRateThreads);
  const int rate_grid = grid_for(C, kRateThreads, six ? rate_cap6 : rate_cap3);
  static const int body_cap6 = resident_grid(body_pass_kernel<true, 1, false>, kBodyThreads);
  static const int body_cap3 = resident_grid(body_pass_kernel<false, 1, false>, kBodyThreads);
  const int body_grid = grid_for(n, kBodiesPerCta, six ? body_cap6 : body_cap3);
  const int cgrid = grid_for(C, 256);

  Scratch<double> partials;
  Scratch<double2> s0, s1;
  Scratch<D4> vel, mob;
  Scratch<LcpState> state;
  // The entry mirrors are the largest buffers of the solve (128 B per contact): kept per host thread and only ever
  // grown, so that a step does not carve them out of (and fragment) the stream-ordered pool every time.
  static thread_local DevBuf<D4> ent0, ent1;
  MB200_TRY(ent0.reserve(2 * C > 0 ? 2 * C : 1));
  if (six) MB200_TRY(ent1.reserve(2 * C > 0 ? 2 * C : 1));
  MB200_TRY(s0.alloc(C, st));
  MB200_TRY(s1.alloc(C, st));
  MB200_TRY(partials.alloc(4 * static_cast<size_t>(rate_grid), st));
  const bool peer_mode = dist && dist->peer_dev;
  if (!peer_mode) MB200_TRY(vel.alloc(static_cast<size_t>(six ? 2 : 1) * n, st));
  MB200_TRY(mob.alloc(n, st));
  MB200_TRY(state.alloc(1, st));
  MB200_CUDA(cudaMemsetAsync(state.p, 0, sizeof(LcpState), st));

  LcpArgs a;
  a.n = n, a.npairs = C;
  a.pairs = sys.pairs, a.rec = sys.rec, a.mob = mob.p;
  a.inc_ptr = inc.ptr.p, a.inc_list = inc.list.p;
  a.ent0 = ent0.p, a.ent1 = ent1.p;
  a.vel = peer_mode ? dist->vel_arena : vel.p;
  a.peer = peer_mode ? dist->peer_dev : nullptr;
  a.halo_seq = 0, a.red_seq = 0;
  a.owned = peer_mode ? dist->owned_s : nullptr;
  a.push_ptr = peer_mode ? dist->push_ptr : nullptr;
  a.push_dst = peer_mode ? dist->push_dst : nullptr;
  a.mu = sys.mu, a.dt = sys.dt, a.tol = params->tol;
  a.fixed = fixed ? 1 : 0, a.it_cap = it_cap;
  a.state = state.p, a.partials = partials.p;
  a.counted = dist ? dist->counted : nullptr;
  a.red_local = dist ? dist->red : nullptr;
  const NcclApi* nccl = dist ? nccl_api() : nullptr;
  if (dist && !nccl) return MB200_ERR_COMM;
  const int recs = six ? 2 : 1;
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
  auto launch_body = [&](int mode, bool fin_pass, int first_it, const double2* stp, double* Fo, double* To, double* Uo,
                         double* Wo) {
#define MB200_BODY(KERNEL, SIX)                                                                                      \
  do {                                                                                                               \
    if (fin_pass)                                                                                                    \
      KERNEL<SIX, 0, true><<<body_grid, kBodyThreads, 0, st>>>(a, first_it, stp, Fo, To, Uo, Wo, nullptr, nullptr);  \
    else if (mode == 0)                                                                                              \
      KERNEL<SIX, 0, false><<<body_grid, kBodyThreads, 0, st>>>(a, first_it, stp, Fo, To, Uo, Wo, nullptr, nullptr); \
    else                                                                                                             \
      KERNEL<SIX, 1, false><<<body_grid, kBodyThreads, 0, st>>>(a, first_it, stp, Fo, To, Uo, Wo, nullptr, nullptr); \
  } while (0)
    if (six) MB200_BODY(body_pass_kernel, true); else MB200_BODY(body_pass_kernel, false);
#undef MB200_BODY
  };

  {
    StageScope ps(MB200_STAGE_LCP_SETUP, st);
    mobility_records_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, sys.radius, sys.quat, sys.kind, mob.p);
    MB200_LAUNCHED();
    init_state_kernel<<<cgrid, 256, 0, st>>>(C, gamma0, s0.p);
    MB200_LAUNCHED();
    if (U_ext) {
      if (six)
        ext_correction_kernel<true><<<cgrid, 256, 0, st>>>(C, sys.pairs, sys.rec, U_ext, W_ext, a.dt);
      else
        ext_correction_kernel<false><<<cgrid, 256, 0, st>>>(C, sys.pairs, sys.rec, U_ext, nullptr, a.dt);
      MB200_LAUNCHED();
    }
    if (C > 0) {
      if (six)
        entry_mirror_kernel<true><<<grid_for(2 * C, 256), 256, 0, st>>>(2 * C, C, inc.list.p, sys.rec, ent0.p, ent1.p);
      else
        entry_mirror_kernel<false><<<grid_for(2 * C, 256), 256, 0, st>>>(2 * C, C, inc.list.p, sys.rec, ent0.p, nullptr);
      MB200_LAUNCHED();
    }
  }

  // initial residual: s1 = {gamma_0, g_0 = D^T M D gamma_0}
  if (peer_mode) a.halo_seq = ++(*dist->halo_seq);
  {
    StageScope ps(MB200_STAGE_LCP_BODY, st);
    launch_body(0, false, 0, s0.p, nullptr, nullptr, nullptr, nullptr);
    MB200_LAUNCHED();
  }
  MB200_TRY(halo());
  if (peer_mode) a.red_seq = ++(*dist->red_seq);
  {
    StageScope ps(MB200_STAGE_LCP_RATE, st);
    if (six)
      rate_pass_kernel<true, 0><<<rate_grid, kRateThreads, 0, st>>>(a, 0, 0, s0.p, s1.p);
    else
      rate_pass_kernel<false, 0><<<rate_grid, kRateThreads, 0, st>>>(a, 0, 0, s0.p, s1.p);
    MB200_LAUNCHED();
  }
  MB200_TRY(reduce(0, 0));

  // iteration `it` reads the previous state from s1 when it is odd (s0 when even) and writes the other buffer
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
      const bool odd = it & 1;
      const double2* prev = odd ? s1.p : s0.p;
      double2* cur = odd ? s0.p : s1.p;
      const int first = it == 1 ? 1 : 0;
      if (peer_mode) a.halo_seq = ++(*dist->halo_seq);
      {
        StageScope ps(MB200_STAGE_LCP_BODY, st);
        launch_body(1, false, first, prev, nullptr, nullptr, nullptr, nullptr);
        MB200_LAUNCHED();
      }
      MB200_TRY(halo());
      if (peer_mode) a.red_seq = ++(*dist->red_seq);
      {
        StageScope ps(MB200_STAGE_LCP_RATE, st);
        if (six)
          rate_pass_kernel<true, 1><<<rate_grid, kRateThreads, 0, st>>>(a, it, first, prev, cur);
        else
          rate_pass_kernel<false, 1><<<rate_grid, kRateThreads, 0, st>>>(a, it, first, prev, cur);
        MB200_LAUNCHED();
      }
      MB200_TRY(reduce(it, 1));
    }
    if (batch < 32) batch *= 2;
  }

  // the state of the last executed iteration: written to s0 when it is odd, to s1 when even (or zero)
  const int it_final = hs.it;
  const double2* fin = (it_final & 1) ? s0.p : s1.p;
  {
    StageScope ps(MB200_STAGE_LCP_BODY, st);
    launch_body(0, true, 0, fin, F, six ? T : nullptr, U, six ? W : nullptr);
    MB200_LAUNCHED();
  }
  if (gamma_out || phi0_out) {
    if (six)
      unpack_contacts_kernel<true><<<cgrid, 256, 0, st>>>(C, sys.rec, fin, phi0_out, gamma_out);
    else
      unpack_contacts_kernel<false><<<cgrid, 256, 0, st>>>(C, sys.rec, fin, phi0_out, gamma_out);
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
      pack_contacts_kernel<true><<<grid_for(C, 256), 256, 0, st>>>(C, pairs, op->normal, op->arm_i, op->arm_j, phi0, rec.p);
    else
      pack_contacts_kernel<false><<<grid_for(C, 256), 256, 0, st>>>(C, pairs, op->normal, nullptr, nullptr, phi0, rec.p);
    MB200_LAUNCHED();
  }
  RecordSystem sys{op->n, C, six, pairs, rec.p, op->quat, op->radius, op->kind, op->viscosity, op->dt};
  return resolve_records(st, sys, inc, params, U_ext, W_ext, gamma, gamma, U_ext ? phi0 : nullptr, F, T, U, W, result_h,
                         nullptr);
}

}  // namespace mb200

using namespace mb200;

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
  return resolve_collisions_impl(st, op, inc, params, U_ext, W_ext, phi0, gamma, F, T, U, W, result_h);
}
