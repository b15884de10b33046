// NCCL plumbing for the multi-GPU path (one process per GPU). libnccl.so.2 is opened at run time (the process usually has
// torch's bundled copy loaded already; the same soname resolves to it), so the library itself has no link-time NCCL
// dependency and still loads on a machine without NCCL.
#pragma once
#include <nccl.h>

#include "common.cuh"

struct mb200_comm {
  ncclComm_t comm = nullptr;
  int world = 1;
  int rank = 0;
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
