// NCCL plumbing for the multi-GPU path (one process per GPU). libnccl.so.2 is opened at run time (the process usually has
// torch's bundled copy loaded already; the same soname resolves to it), so the library itself has no link-time NCCL
// dependency and still loads on a machine without NCCL.
#pragma once
#include <nccl.h>

#include "common.cuh"

namespace mb200 {

constexpr int kMaxRanks = 16;

// Peer-memory exchange (one process per GPU, CUDA IPC over NVLink 5 / NVSwitch): every rank maps every other rank's
// control block and velocity-record arena, so that the solver's kernels exchange their data with plain remote stores
// and system-scope flags instead of NCCL calls (lcp_solve.cu: halo_push_kernel, rate_pass_kernel).
//   control block of a rank (device memory of that rank, zero-initialised, never reallocated):
//     halo_flag[kMaxRanks]        halo_flag[r] = sequence number of the last velocity halo rank r has pushed to this rank
//     red_flag[2][kMaxRanks]      red_flag[s & 1][r] = s once rank r's totals of reduction s are in red[s & 1][r]
//     red[2][kMaxRanks][4]        {residual max, xx, xg, gg} of every rank, double-buffered by sequence parity
struct PeerCtrl {
  unsigned long long halo_flag[kMaxRanks];
  unsigned long long red_flag[2][kMaxRanks];
  double red[2][kMaxRanks][4];
};
// Device-resident table of the mapped addresses, indexed by rank (own rank = local addresses).
struct PeerDev {
  int world, me;
  PeerCtrl* ctrl[kMaxRanks];
  D4* vel[kMaxRanks];
};

}  // namespace mb200

struct mb200_comm {
  ncclComm_t comm = nullptr;
  int world = 1;
  int rank = 0;
  // ---- peer-memory state (comm.cu: peer_setup) ----
  struct Peer {
    bool tried = false;   // setup attempted (collectively) for `rows`
    bool ok = false;      // every rank mapped every other rank's memory
    int64_t rows = 0;     // capacity of the velocity arena in 32-byte records
    mb200::PeerCtrl* ctrl = nullptr;
    mb200::D4* vel = nullptr;
    void* mapped[2 * mb200::kMaxRanks] = {};  // cudaIpcOpenMemHandle results to close
    mb200::PeerDev* dev = nullptr;            // device copy of the address table
    unsigned long long halo_seq = 0, red_seq = 0;
  } peer;
};

namespace mb200 {

struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
// nullptr (and the error string set) when libnccl cannot be loaded
const NcclApi* nccl_api();

// Collective: allocate this rank's control block and a velocity arena of `rows` records, exchange the CUDA IPC handles
// through the communicator and map every other rank's memory. On return comm->peer.ok says whether EVERY rank succeeded
// (identical on all ranks); when it is false the caller keeps the NCCL exchange. Never fails the step for an IPC problem.
int peer_setup(mb200_comm* comm, cudaStream_t st, int64_t rows);
void peer_teardown(mb200_comm* comm);
// Collective: all-gather `bytes` host bytes per rank (src_h -> dst_h[world * bytes]) through the communicator.
int comm_allgather_host(mb200_comm* comm, cudaStream_t st, const void* src_h, void* dst_h, size_t bytes);

#define MB200_NCCL(expr)                                                                                   \
  do {                                                                                                     \
    ncclResult_t _r = (expr);                                                                              \
    if (_r != ncclSuccess) {                                                                               \
      ::mb200::set_error(std::string(#expr) + ": " + ::mb200::nccl_api()->GetErrorString(_r) + " (" +      \
                         __FILE__ + ":" + std::to_string(__LINE__) + ")");                                 \
      return MB200_ERR_COMM;                                                                               \
    }                                                                                                      \
  } while (0)

}  // namespace mb200
