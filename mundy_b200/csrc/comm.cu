// Communicator handle of the multi-GPU path: NCCL over NVLink 5 / NVSwitch, one process per GPU.
// Replaces the reference's stk::ParallelMachine / MPI usage on this path (all_reduce_max of the residual,
// three all_reduce_sum of the BB dot products, ghosting of neighbour bodies).
#include <dlfcn.h>

#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

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

int comm_allgather_host(mb200_comm* comm, cudaStream_t st, const void* src_h, void* dst_h, size_t bytes) {
  const NcclApi* api = nccl_api();
  if (!api) return MB200_ERR_COMM;
  Scratch<unsigned char> buf;
  const size_t W = static_cast<size_t>(comm->world);
  MB200_TRY(buf.alloc((W + 1) * bytes, st));
  MB200_CUDA(cudaMemcpyAsync(buf.p, src_h, bytes, cudaMemcpyHostToDevice, st));
  MB200_NCCL(api->AllGather(buf.p, buf.p + bytes, bytes, ncclUint8, comm->comm, st));
  MB200_CUDA(cudaMemcpyAsync(dst_h, buf.p + bytes, W * bytes, cudaMemcpyDeviceToHost, st));
  MB200_CUDA(cudaStreamSynchronize(st));
  return MB200_OK;
}

void peer_teardown(mb200_comm* comm) {
  auto& P = comm->peer;
  for (void*& m : P.mapped) {
    if (m) cudaIpcCloseMemHandle(m);
    m = nullptr;
  }
  if (P.ctrl) cudaFree(P.ctrl);
  if (P.vel) cudaFree(P.vel);
  if (P.dev) cudaFree(P.dev);
  P.ctrl = nullptr, P.vel = nullptr, P.dev = nullptr;
  P.ok = false, P.tried = false, P.rows = 0;
}

int peer_setup(mb200_comm* comm, cudaStream_t st, int64_t rows) {
  auto& P = comm->peer;
  const int W = comm->world, me = comm->rank;
  if (P.tried && rows <= P.rows) return MB200_OK;
  const char* env = getenv("MB200_DIST_EXCHANGE");
  // (MB200_DIST_EXCHANGE=peer also takes the peer-memory kernels with a single rank: a diagnostic of their own cost)
  const bool want = !(env && std::strcmp(env, "nccl") == 0) && (W > 1 || (env && std::strcmp(env, "peer") == 0)) &&
                    W <= kMaxRanks;
  if (P.tried) {
    // a larger arena is needed: every rank gets here in the same step (rows derives from the global body count);
    // peers must have closed their mappings before the memory is freed
    MB200_CUDA(cudaStreamSynchronize(st));
    char z = 0;
    std::vector<char> all(W);
    for (void*& m : P.mapped) {
      if (m) cudaIpcCloseMemHandle(m);
      m = nullptr;
    }
    MB200_TRY(comm_allgather_host(comm, st, &z, all.data(), 1));  // barrier: all mappings closed
    peer_teardown(comm);
  }
  P.tried = true, P.ok = false, P.rows = rows;
  P.halo_seq = 0, P.red_seq = 0;
  if (!want) return MB200_OK;

  struct Info {
    cudaIpcMemHandle_t ctrl, vel;
    int ok;
  } mine{};
  mine.ok = 1;
  if (cudaMalloc(reinterpret_cast<void**>(&P.ctrl), sizeof(PeerCtrl)) != cudaSuccess ||
      cudaMalloc(reinterpret_cast<void**>(&P.vel), static_cast<size_t>(rows) * sizeof(D4)) != cudaSuccess ||
      cudaMalloc(reinterpret_cast<void**>(&P.dev), sizeof(PeerDev)) != cudaSuccess ||
      cudaMemset(P.ctrl, 0, sizeof(PeerCtrl)) != cudaSuccess ||
      cudaIpcGetMemHandle(&mine.ctrl, P.ctrl) != cudaSuccess || cudaIpcGetMemHandle(&mine.vel, P.vel) != cudaSuccess) {
    mine.ok = 0;
    cudaGetLastError();
  }
  std::vector<Info> all(W);
  MB200_TRY(comm_allgather_host(comm, st, &mine, all.data(), sizeof(Info)));
  bool every = true;
  for (int r = 0; r < W; ++r) every = every && all[r].ok;
  PeerDev table{};
  table.world = W, table.me = me;
  int opened = every ? 1 : 0;
  if (every) {
    for (int r = 0; r < W && opened; ++r) {
      if (r == me) {
        table.ctrl[r] = P.ctrl, table.vel[r] = P.vel;
        continue;
      }
      void *c = nullptr, *v = nullptr;
      if (cudaIpcOpenMemHandle(&c, all[r].ctrl, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess ||
          cudaIpcOpenMemHandle(&v, all[r].vel, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        opened = 0;
        cudaGetLastError();
      }
      P.mapped[2 * r] = c, P.mapped[2 * r + 1] = v;
      table.ctrl[r] = static_cast<PeerCtrl*>(c), table.vel[r] = static_cast<D4*>(v);
    }
  }
  std::vector<int> oks(W);
  MB200_TRY(comm_allgather_host(comm, st, &opened, oks.data(), sizeof(int)));
  bool ok = true;
  for (int r = 0; r < W; ++r) ok = ok && oks[r];
  if (ok) {
    MB200_CUDA(cudaMemcpy(P.dev, &table, sizeof(PeerDev), cudaMemcpyHostToDevice));
    P.ok = true;
  } else {  // keep the NCCL exchange; release what was allocated (after everyone has stopped touching it: nobody has)
    for (void*& m : P.mapped) {
      if (m) cudaIpcCloseMemHandle(m);
      m = nullptr;
    }
    char z = 0;
    std::vector<char> b(W);
    MB200_TRY(comm_allgather_host(comm, st, &z, b.data(), 1));
    if (P.ctrl) cudaFree(P.ctrl);
    if (P.vel) cudaFree(P.vel);
    if (P.dev) cudaFree(P.dev);
    P.ctrl = nullptr, P.vel = nullptr, P.dev = nullptr;
  }
  return MB200_OK;
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
  peer_teardown(c);
  if (api && c->comm) api->CommDestroy(c->comm);
  delete c;
  return MB200_OK;
}

int mb200_comm_world(const mb200_comm* c) { return c ? c->world : 1; }
int mb200_comm_rank(const mb200_comm* c) { return c ? c->rank : 0; }

}  // extern "C"
