// Communicator handle of the multi-GPU path: NCCL over NVLink 5 / NVSwitch, one process per GPU.
// Replaces the reference's stk::ParallelMachine / MPI usage on this path (all_reduce_max of the residual,
// three all_reduce_sum of the BB dot products, ghosting of neighbour bodies).
#include <dlfcn.h>

#include <cstring>
#include <mutex>

#include "comm.cuh"

namespace mb200 {

const NcclApi* nccl_api() {
  static NcclApi api;
  static bool tried = false, ok = false;
  static std::mutex mu;
  std::lock_guard<std::mutex> lk(mu);
  if (tried) return ok ? &api : nullptr;
  tried = true;
  void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) {
    set_error(std::string("cannot load libnccl.so.2: ") + dlerror());
    return nullptr;
  }
  bool all = true;
  auto sym = [&](const char* name) {
    void* p = dlsym(lib, name);
    if (!p) all = false;
    return p;
  };
  api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
  api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
  api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
  api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
  api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
  api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
  api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
  api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
  api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
  api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
  if (!all) {
    set_error("libnccl.so.2 lacks a required symbol");
    return nullptr;
  }
  ok = true;
  return &api;
}

}  // namespace mb200

using namespace mb200;

extern "C" {

int mb200_comm_get_unique_id(char* id_out_h) {
  MB200_REQUIRE(id_out_h, "null id buffer");
  static_assert(sizeof(ncclUniqueId) == MB200_COMM_ID_BYTES, "ncclUniqueId size");
  const NcclApi* api = nccl_api();
  if (!api) return MB200_ERR_COMM;
  ncclUniqueId id;
  MB200_NCCL(api->GetUniqueId(&id));
  std::memcpy(id_out_h, &id, sizeof(id));
  return MB200_OK;
}

int mb200_comm_create(mb200_comm** out, int world, int rank, const char* id_h) {
  MB200_REQUIRE(out && id_h && world >= 1 && rank >= 0 && rank < world, "bad arguments");
  MB200_TRY(require_device());
  const NcclApi* api = nccl_api();
  if (!api) return MB200_ERR_COMM;
  mb200_comm* c = new (std::nothrow) mb200_comm();
  MB200_REQUIRE(c, "allocation failed");
  c->world = world, c->rank = rank;
  ncclUniqueId id;
  std::memcpy(&id, id_h, sizeof(id));
  ncclResult_t r = api->CommInitRank(&c->comm, world, id, rank);
  if (r != ncclSuccess) {
    set_error(std::string("ncclCommInitRank: ") + api->GetErrorString(r));
    delete c;
    return MB200_ERR_COMM;
  }
  *out = c;
  return MB200_OK;
}

int mb200_comm_destroy(mb200_comm* c) {
  if (!c) return MB200_OK;
  const NcclApi* api = nccl_api();
  if (api && c->comm) api->CommDestroy(c->comm);
  delete c;
  return MB200_OK;
}

int mb200_comm_world(const mb200_comm* c) { return c ? c->world : 1; }
int mb200_comm_rank(const mb200_comm* c) { return c ? c->rank : 0; }

}  // extern "C"
