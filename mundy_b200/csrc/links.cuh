// The neighbour-link handle: pair list in Morton-order indices + rebuild state + per-step contact arrays.
#pragma once
#include "internal.cuh"

namespace mb200 {
// Where the g-th received ghost (receive order: peers in rank order) sits in the local set of a rank: the ranks' blocks
// are laid out in rank order with the owned block in its place (dist_step.cu).
struct GhostMap {
  int npeers;
  int64_t recv_off[17];   // prefix over peers of the received counts
  int64_t local_beg[16];  // local index of the first ghost of the q-th peer
};
}  // namespace mb200

struct mb200_neighbor_links {
  // builder parameters (GenNeighborLinkers.hpp:401-462)
  double buffer = 0.0;
  int filter = MB200_FILTER_NONE;
  mb200::Periodic periodic = mb200::make_periodic(nullptr);  // orthorhombic periodic box (off by default)
  // state of the last build
  bool built = false;
  bool rods = false;
  int64_t n = 0;
  int64_t npairs = 0;
  mb200::DevBuf<uint32_t> perm;      // Morton-order index -> caller index            [n]
  mb200::DevBuf<double> ref_center;  // centres at build time, caller order           [3n]
  mb200::DevBuf<mb200::D4> ref_b1;   // half-axis records at build time, caller order [n] (shape filter only)
  mb200::DevBuf<int32_t> pairs;      // (i, j) Morton-order indices, sorted           [2 npairs]
  mb200::Incidence inc;              // body-major incidence lists of `pairs`
  // per-step arrays in Morton order (allocated by the step)
  mb200::DevBuf<mb200::D4> b0, b1;
  mb200::DevBuf<mb200::D4> rec;  // contact records {n, phi0} of the current step (lcp_solve.cu)
  mb200::DevBuf<double2> sig;    // rods: axial arm parameters {sigma_i, sigma_j} of the current step
  mb200::DevBuf<double> radius_s, quat_s, vec_a, vec_b, vec_c, vec_d, vec_e, vec_f;
  mb200::DevBuf<int32_t> kind_s;
  mb200::DevBuf<int32_t> flag;  // device flag for the rebuild trigger
  // device mirrors of the host arrays of mb200_collision_step_host
  mb200::DevBuf<double> hm_center, hm_quat, hm_radius, hm_length, hm_fext, hm_text, hm_fcol, hm_tcol, hm_u, hm_w;
  mb200::DevBuf<int32_t> hm_kind;
  // mb200_collision_step_host: the external wrench is uploaded on a side stream behind the body arrays, while the
  // rebuild trigger and the link rebuild (which do not read it) run; the step waits for forces_ev before its first use
  cudaStream_t copy_st = nullptr;
  cudaEvent_t bodies_ev = nullptr, forces_ev = nullptr;
  bool forces_pending = false;
  ~mb200_neighbor_links() {
    if (bodies_ev) cudaEventDestroy(bodies_ev);
    if (forces_ev) cudaEventDestroy(forces_ev);
    if (copy_st) cudaStreamDestroy(copy_st);
  }

  // ---- multi-GPU state (dist_step.cu) ----
  struct Dist {
    bool built = false;
    int64_t n_global = 0, n_local = 0, n_own = 0, o_beg = 0;
    mb200::DevBuf<uint8_t> owned;      // local body owned by this rank                [n_local]
    mb200::DevBuf<double> center, quat, radius, length, fext, text, fcol, uout;   // local body arrays
    mb200::DevBuf<int32_t> kind;
    // ghost bodies: what this rank sends to / receives from each peer at a rebuild (body records) and per step (state)
    int64_t bsend_off[17] = {}, brecv_off[17] = {};
    mb200::DevBuf<uint32_t> bsend_idx;   // owned indices of the bodies ghosted out, peer-major
    mb200::DevBuf<double> bsend_rec, brecv_rec;
    mb200::GhostMap gmap;
    mb200::DevBuf<uint8_t> counted;    // contact counted by this rank                 [npairs]
    mb200::DevBuf<uint32_t> inv_perm;  // local caller index -> Morton index           [n_local]
    mb200::DevBuf<uint32_t> send_idx, recv_idx;
    mb200::DevBuf<uint8_t> owned_s;      // ownership flags in Morton order (peer-memory mode)
    mb200::DevBuf<uint32_t> remote_row;  // [send list] destination rows in the peers' velocity arenas (peer-memory mode)
    mb200::DevBuf<int64_t> push_ptr;     // [n_local + 1] body -> its ghost copies (Morton order)
    mb200::DevBuf<uint32_t> push_dst;    // rank << 28 | row
    mb200::DevBuf<uint32_t> boundary;    // Morton indices of the bodies with ghost copies, ascending
    int64_t n_boundary = 0;
    mb200::DevBuf<mb200::D4> sendbuf, recvbuf;
    mb200::DevBuf<double> red;
    mb200::DevBuf<int32_t> flag_all;
    int npeers = 0;
    int peer[16];
    int64_t send_off[17], recv_off[17];
  } dist;
};

namespace mb200 {
// broadphase.cu: (re)build the link list from the bodies (explicit_S != NULL: search radii given, no buffer/filter).
// owned != NULL (multi-GPU): per-body ownership flags in the caller's order; links between two non-owned bodies are dropped.
int build_links(cudaStream_t st, mb200_neighbor_links* h, const mb200_bodies* b, const double* explicit_S,
                const uint8_t* owned);
// Device flag (h->flag) = 1 when some body of `b` moved more than half the buffer since the last build; body k of `b` is
// body ref_offset + k of the set the links were built from. No host sync.
int links_moved_flag(cudaStream_t st, mb200_neighbor_links* h, const mb200_bodies* b, int64_t ref_offset = 0);
// step.cu
int step_core(mb200_neighbor_links* h, cudaStream_t st, const mb200_bodies* bodies, const mb200_step_params* params,
              const double* F_ext, const double* T_ext, double* F_col, double* T_col, double* U_out, double* W_out,
              int rebuilt, long long launches0, const DistSolve* dist, mb200_step_stats* stats_h);
// pairs in the caller's indices, sorted by (i,j) with i<j, written to out[2*npairs]
int canonical_pairs(cudaStream_t st, mb200_neighbor_links* h, int32_t* out);
int links_generate(cudaStream_t st, mb200_neighbor_links* h, const mb200_bodies* b, int force, int* rebuilt_h);
}  // namespace mb200
