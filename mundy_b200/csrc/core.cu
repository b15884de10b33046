// Library-wide plumbing: error string, launch counter, device check.
#include <mutex>
#include <vector>

#include "common.cuh"

namespace mb200 {

// ---- stage timers ------------------------------------------------------------------------------------------------
std::atomic<int> g_profile_on{0};
namespace {
struct ProfRegion {
  int id;
  int stage;
  cudaEvent_t a, b;
  bool closed;
};
int g_prof_next_id = 0;
std::mutex g_prof_mu;
std::vector<ProfRegion> g_prof_open;       // recorded, not yet folded into the totals
std::vector<cudaEvent_t> g_prof_free;      // recycled events
double g_prof_ms[MB200_STAGE_COUNT];
long long g_prof_regions[MB200_STAGE_COUNT];
long long g_prof_calls[MB200_STAGE_COUNT];  // regions entered, bracketed or not

cudaEvent_t prof_event() {
  if (!g_prof_free.empty()) {
    cudaEvent_t e = g_prof_free.back();
    g_prof_free.pop_back();
    return e;
  }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}
void prof_fold() {  // g_prof_mu held. Regions still open (begin recorded, end not yet) stay in the list.
  std::vector<ProfRegion> still_open;
  for (ProfRegion& r : g_prof_open) {
    if (!r.closed) {
      still_open.push_back(r);
      continue;
    }
    float ms = 0.f;
    if (cudaEventSynchronize(r.b) == cudaSuccess && cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) {
      g_prof_ms[r.stage] += ms;
      g_prof_regions[r.stage] += 1;
    }
    g_prof_free.push_back(r.a);
    g_prof_free.push_back(r.b);
  }
  g_prof_open.swap(still_open);
}
}  // namespace

int prof_begin(int stage, cudaStream_t st) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  // mb200_profile_enable(k > 1): of the per-launch stages (two or three regions per BBPGD iteration) only every k-th
  // region is bracketed: an event pair between back-to-back kernels costs about as much as the launch gap it replaces
  const int stride = g_profile_on.load(std::memory_order_relaxed);
  const long long call = g_prof_calls[stage]++;
  const bool per_launch = stage == MB200_STAGE_LCP_BODY || stage == MB200_STAGE_LCP_RATE || stage == MB200_STAGE_LCP_COMM;
  if (per_launch && stride > 1 && call % stride != 0) return -1;
  ProfRegion r{g_prof_next_id, stage, prof_event(), prof_event(), false};
  if (!r.a || !r.b) return -1;
  g_prof_next_id = (g_prof_next_id + 1) & 0x3fffffff;
  cudaEventRecord(r.a, st);
  g_prof_open.push_back(r);
  return r.id;
}
void prof_end(int id, cudaStream_t st) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (size_t k = g_prof_open.size(); k-- > 0;) {  // the region just opened is at or near the back
    ProfRegion& r = g_prof_open[k];
    if (r.id == id && !r.closed) {
      cudaEventRecord(r.b, st);
      r.closed = true;
      return;
    }
  }
}

// NVTX range names of the stages (the reference brackets its phases with named regions,
// alens/tests/performance_tests/NgpHP1.cpp:1503,1643)
const char* const kStageNames[MB200_STAGE_COUNT] = {"mb200:trigger",   "mb200:broadphase", "mb200:pack",     "mb200:narrow",
                                                    "mb200:lcp_setup", "mb200:lcp_body",   "mb200:lcp_rate", "mb200:euler",
                                                    "mb200:lcp_comm",  "mb200:state_comm"};

static thread_local std::string t_error;
std::atomic<long long> g_launches{0};

void set_error(const std::string& msg) { t_error = msg; }

namespace {
struct PinnedScratch {
  void* p = nullptr;
  size_t cap = 0;
  ~PinnedScratch() {
    if (p) cudaFreeHost(p);
  }
};
}  // namespace
void* pinned_scratch(size_t bytes) {
  static thread_local PinnedScratch s;
  if (bytes > s.cap) {
    if (s.p) cudaFreeHost(s.p);
    s.p = nullptr, s.cap = 0;
    size_t want = 4096;
    while (want < bytes) want <<= 1;
    if (cudaHostAlloc(&s.p, want, cudaHostAllocPortable) != cudaSuccess) {
      (void)cudaGetLastError();
      s.p = nullptr;
      set_error("cudaHostAlloc of " + std::to_string(want) + " bytes of page-locked host memory failed");
      return nullptr;
    }
    s.cap = want;
  }
  return s.p;
}

int require_device() {
  static std::atomic<int> ok{0};
  if (ok.load(std::memory_order_acquire)) return MB200_OK;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count <= 0) {
    (void)cudaGetLastError();
    set_error(std::string("no CUDA device available (") + (e != cudaSuccess ? cudaGetErrorString(e) : "count = 0") +
              "); mundy_b200 has no CPU fallback");
    return MB200_ERR_CUDA;
  }
  int dev = 0;
  MB200_CUDA(cudaGetDevice(&dev));
  // Keep stream-ordered scratch memory cached in the pool between calls (a step allocates the same sizes again).
  cudaMemPool_t pool;
  MB200_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
  uint64_t threshold = UINT64_MAX;
  MB200_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold));
  ok.store(1, std::memory_order_release);
  return MB200_OK;
}

}  // namespace mb200

extern "C" {

int mb200_abi_version(void) { return MB200_ABI_VERSION; }

const char* mb200_last_error_string(void) { return mb200::t_error.c_str(); }

int mb200_device_count(void) {
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess) {
    (void)cudaGetLastError();
    return 0;
  }
  return count;
}

int mb200_device_alloc(void** out, int64_t bytes) {
  MB200_REQUIRE(out && bytes >= 0, "bad arguments");
  MB200_TRY(mb200::require_device());
  *out = nullptr;
  if (bytes > 0) MB200_CUDA(cudaMalloc(out, static_cast<size_t>(bytes)));
  return MB200_OK;
}
int mb200_device_free(void* ptr) {
  if (ptr) MB200_CUDA(cudaFree(ptr));
  return MB200_OK;
}
int mb200_copy_to_device(mb200_stream stream, void* dst, const void* src_h, int64_t bytes) {
  MB200_REQUIRE(bytes >= 0 && (bytes == 0 || (dst && src_h)), "bad arguments");
  MB200_TRY(mb200::require_device());
  if (bytes > 0) MB200_CUDA(cudaMemcpyAsync(dst, src_h, static_cast<size_t>(bytes), cudaMemcpyHostToDevice, mb200::as_stream(stream)));
  return MB200_OK;
}
int mb200_copy_to_host(mb200_stream stream, void* dst_h, const void* src, int64_t bytes) {
  MB200_REQUIRE(bytes >= 0 && (bytes == 0 || (dst_h && src)), "bad arguments");
  MB200_TRY(mb200::require_device());
  if (bytes > 0) MB200_CUDA(cudaMemcpyAsync(dst_h, src, static_cast<size_t>(bytes), cudaMemcpyDeviceToHost, mb200::as_stream(stream)));
  return MB200_OK;
}
int mb200_stream_synchronize(mb200_stream stream) {
  MB200_TRY(mb200::require_device());
  MB200_CUDA(cudaStreamSynchronize(mb200::as_stream(stream)));
  return MB200_OK;
}

int64_t mb200_kernel_launch_count(void) { return mb200::g_launches.load(std::memory_order_relaxed); }

int mb200_profile_enable(int on) {
  std::lock_guard<std::mutex> lk(mb200::g_prof_mu);
  mb200::prof_fold();
  if (on) {
    for (int k = 0; k < MB200_STAGE_COUNT; ++k)
      mb200::g_prof_ms[k] = 0.0, mb200::g_prof_regions[k] = 0, mb200::g_prof_calls[k] = 0;
  }
  mb200::g_profile_on.store(on > 0 ? on : 0);
  return MB200_OK;
}

int mb200_profile_read(int stage, double* ms_total_h, int64_t* regions_h) {
  MB200_REQUIRE(stage >= 0 && stage < MB200_STAGE_COUNT && ms_total_h && regions_h, "bad stage or null output");
  std::lock_guard<std::mutex> lk(mb200::g_prof_mu);
  mb200::prof_fold();
  *ms_total_h = mb200::g_prof_ms[stage];
  *regions_h = mb200::g_prof_regions[stage];
  return MB200_OK;
}

int mb200_profile_calls(int stage, int64_t* calls_h) {
  MB200_REQUIRE(stage >= 0 && stage < MB200_STAGE_COUNT && calls_h, "bad stage or null output");
  std::lock_guard<std::mutex> lk(mb200::g_prof_mu);
  *calls_h = mb200::g_prof_calls[stage];
  return MB200_OK;
}

}  // extern "C"
