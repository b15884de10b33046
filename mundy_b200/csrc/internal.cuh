// Internal (non-ABI) interfaces shared between the translation units of libmundy_b200.so.
#pragma once
#include "common.cuh"

struct mb200_comm;

namespace mb200 {

// narrow_phase.cu ------------------------------------------------------------------------------------------------
// Caller layout -> 32-byte body records (shapes.cuh). perm == NULL: identity. b1 == NULL: spheres only.
int pack_bodies(cudaStream_t st, const mb200_bodies* b, const int32_t* perm, D4* b0, D4* b1);
int narrow_phase_records(cudaStream_t st, int64_t npairs, const int32_t* pairs, const D4* b0, const D4* b1,
                         double* phi, double* normal, double* arm_i, double* arm_j, const Periodic& P);

// lcp.cu ---------------------------------------------------------------------------------------------------------
// Body-major incidence lists of a pair list: for body b, list[ptr[b] .. ptr[b+1]) holds 2*contact+side in ascending
// order (side 0: b is the first body of the pair, 1: the second).
struct Incidence {
  Scratch<int64_t> ptr;    // n + 1
  Scratch<uint32_t> list;  // 2 * npairs
  Scratch<uint32_t> epos;  // 2 * npairs: epos[2 * contact + side] = position of that entry in `list` (the inverse of it)
  Scratch<int32_t> bad;    // 1: some pair named a body outside [0, n)
};
int build_incidence(cudaStream_t st, int64_t n, int64_t npairs, const int32_t* pairs, Incidence* inc);
// The same lists without a sort, for a pair list sorted by (i, j), i < j, with row pointers rowptr[n + 1] (the link build).
int build_incidence_sorted(cudaStream_t st, int64_t n, int64_t npairs, const int32_t* pairs, const int64_t* rowptr,
                           Incidence* inc);
int check_incidence(cudaStream_t st, const Incidence& inc);  // synchronises; error if `bad` was raised

// The BBPGD solve on contact records (layout: lcp_solve.cu header). rec is modified when U_ext != NULL (phi0 correction).
// F,T,U,W are written in the order of the bodies the records index; U/W hold the collision velocity only.
// Form of the lever arms: none (3-DOF, spheres), axial (arm = sigma * half-axis vector of the body: the step's
// spherocylinders, whose contact points lie on the centre line), general (caller-supplied arm vectors).
enum { ARM_NONE = 0, ARM_AXIAL = 1, ARM_GENERAL = 2 };
struct RecordSystem {
  int64_t n, npairs;
  int arm;
  const int2* pairs;   // [npairs]
  D4* rec;             // R0 [npairs]; ARM_GENERAL: + R1, R2 ([3 npairs])
  const double2* sig;  // ARM_AXIAL: [npairs] {sigma_i, sigma_j}
  const D4* half;      // ARM_AXIAL: [n] {h, kind}, h = (L/2) (q (x) z_hat)
  const double* quat;
  const double* radius;
  const int32_t* kind;
  double mu, dt;
};
// Multi-GPU hooks of the solver (null => one GPU). Body indices are the solver's (the order the records index).
struct DistSolve {
  mb200_comm* comm;
  const uint8_t* counted;      // [npairs] 1 = this rank adds the contact to the global residual / BB sums
  int npeers;
  int peer[16];                // ranks exchanged with
  int64_t send_off[17];        // prefix offsets (in bodies) into send_idx, per peer
  int64_t recv_off[17];
  const uint32_t* send_idx;    // owned bodies whose velocity records go to a peer
  const uint32_t* recv_idx;    // ghost rows filled from a peer (same order as the peer's send list)
  D4* sendbuf;                 // 2 records per body capacity
  D4* recvbuf;
  double* red;                 // device scratch: [4] local totals followed by [4 * world] gathered
  // peer-memory exchange (comm.cuh; peer_dev == NULL: the NCCL calls above are used)
  const struct PeerDev* peer_dev;      // device table of every rank's mapped control block and velocity arena
  D4* vel_arena;                       // this rank's velocity arena (the solver's velocity records live here)
  const int64_t* push_ptr;             // [n + 1] per body: its entries in push_dst (owned boundary bodies only)
  const uint32_t* push_dst;            // rank << 28 | row of the ghost copy in that rank's arena
  const uint8_t* owned_s;              // [n] ownership flags in the solver's (Morton) body order
  const uint32_t* boundary;            // [n_boundary] bodies with ghost copies, ascending: the body pass takes them first
  int64_t n_boundary;
  unsigned long long* halo_seq;        // host counters kept in the communicator (monotone across solves)
  unsigned long long* red_seq;
};
int resolve_records(cudaStream_t st, const RecordSystem& sys, const Incidence& inc, const mb200_lcp_params* params,
                    const double* U_ext, const double* W_ext, const double* gamma0, double* gamma_out,
                    double* phi0_out, double* F, double* T, double* U, double* W, mb200_collision_result* result_h,
                    const DistSolve* dist);
int rate_grid_for(int64_t npairs, int arm);
// Narrow phase straight into contact records (the step's path).
// perm (Morton -> caller index, nullable): pairs are evaluated with the lower caller index first.
// Rods (b1 != NULL): rec = R0 and sig = the axial arm parameters; spheres: rec = R0 only.
int narrow_phase_to_records(cudaStream_t st, int64_t npairs, const int32_t* pairs, const uint32_t* perm, const D4* b0,
                            const D4* b1, D4* rec, double2* sig, const Periodic& P);
// step.cu: the step after the link list is known (dist != NULL on the multi-GPU path)


// The BBPGD solve on caller-layout arrays and an already-built incidence structure (used by mb200_resolve_collisions and the step).
int resolve_collisions_impl(cudaStream_t st, const mb200_contact_op* op, const Incidence& inc,
                            const mb200_lcp_params* params, const double* U_ext, const double* W_ext, double* phi0,
                            double* gamma, double* F, double* T, double* U, double* W,
                            mb200_collision_result* result_h);

}  // namespace mb200

namespace mb200 {
// euler.cu
int euler_update(cudaStream_t st, int64_t n, double* center, double* quat, const int32_t* kind, const double* U,
                 const double* W, double dt);
}  // namespace mb200
