// The neighbour-link handle: pair list in Morton-order indices + rebuild state + per-step contact arrays.
#pragma once
#include "internal.cuh"

struct mb200_neighbor_links {
  // builder parameters (GenNeighborLinkers.hpp:401-462)
  double buffer = 0.0;
  int filter = MB200_FILTER_NONE;
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
  mb200::DevBuf<mb200::D4> rec;  // contact records of the current step (lcp_solve.cu)
  mb200::DevBuf<double> radius_s, quat_s, vec_a, vec_b, vec_c, vec_d, vec_e, vec_f;
  mb200::DevBuf<int32_t> kind_s;
  mb200::DevBuf<int32_t> flag;  // device flag for the rebuild trigger
  // device mirrors of the host arrays of mb200_collision_step_host
  mb200::DevBuf<double> hm_center, hm_quat, hm_radius, hm_length, hm_fext, hm_text, hm_fcol, hm_tcol, hm_u, hm_w;
  mb200::DevBuf<int32_t> hm_kind;
};

namespace mb200 {
// broadphase.cu: (re)build the link list from the bodies (explicit_S != NULL: search radii given, no buffer/filter).
int build_links(cudaStream_t st, mb200_neighbor_links* h, const mb200_bodies* b, const double* explicit_S);
// pairs in the caller's indices, sorted by (i,j) with i<j, written to out[2*npairs]
int canonical_pairs(cudaStream_t st, mb200_neighbor_links* h, int32_t* out);
int links_generate(cudaStream_t st, mb200_neighbor_links* h, const mb200_bodies* b, int force, int* rebuilt_h);
}  // namespace mb200
