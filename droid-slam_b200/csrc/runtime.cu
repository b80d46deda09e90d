// Process-wide runtime state of libdroid_b200: error slot, launch counter, per-phase timing and the PER-DEVICE
// contexts (function attributes, side streams, events, pinned scratch).  CUDA function attributes, streams and events
// belong to one device / context, so nothing of the kind is kept in function-local statics: every entry point asks
// dev_ctx() for the context of the CURRENT device, created once under a mutex.
#include <mutex>
#include <vector>
#include "common.cuh"

namespace droid {
cudaError_t g_last_error = cudaSuccess;
std::atomic<unsigned long long> g_launch_count{0};

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_ctx_mutex;
DevCtx* g_ctx[kMaxDevices] = {};
}  // namespace

DevCtx& dev_ctx() {
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= kMaxDevices) dev = 0;
  std::lock_guard<std::mutex> lk(g_ctx_mutex);
  DevCtx* c = g_ctx[dev];
  if (c == nullptr) {
    c = new DevCtx();
    c->device = dev;
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
    c->sm_count = sms;
    g_ctx[dev] = c;
  }
  return *c;
}

bool DevCtx::first(int what) {
  std::lock_guard<std::mutex> lk(g_ctx_mutex);
  if (done[what]) return false;
  done[what] = true;
  return true;
}

int* DevCtx::pinned_ints() {
  std::lock_guard<std::mutex> lk(g_ctx_mutex);
  if (h_ints == nullptr && cudaHostAlloc((void**)&h_ints, 64 * sizeof(int), cudaHostAllocDefault) != cudaSuccess)
    h_ints = nullptr;
  return h_ints;
}

// ---- per-phase timing -----------------------------------------------------------------------
namespace {
struct ProfRec { int phase; cudaEvent_t a, b; };
bool g_prof_on = false;
std::mutex g_prof_mutex;
std::vector<ProfRec> g_prof_recs;
std::vector<cudaEvent_t> g_prof_pool;
cudaEvent_t prof_event() {
  if (!g_prof_pool.empty()) { cudaEvent_t e = g_prof_pool.back(); g_prof_pool.pop_back(); return e; }
  cudaEvent_t e; cudaEventCreate(&e); return e;
}
}  // namespace
void prof_mark(int phase, bool begin, cudaStream_t st) {
  if (!g_prof_on) return;
  std::lock_guard<std::mutex> lk(g_prof_mutex);
  if (begin) {
    ProfRec r{phase, prof_event(), prof_event()};
    cudaEventRecord(r.a, st);
    g_prof_recs.push_back(r);
  } else {
    for (size_t i = g_prof_recs.size(); i-- > 0;)
      if (g_prof_recs[i].phase == phase) { cudaEventRecord(g_prof_recs[i].b, st); break; }
  }
}

}  // namespace droid

using namespace droid;

extern "C" const char* droid_version(void) { return "droid_b200 0.2 sm_100a"; }
extern "C" unsigned long long droid_launch_count(void) { return g_launch_count.load(); }
extern "C" int droid_prof_enable(int on) {
  g_prof_on = on != 0;
  return DROID_OK;
}
extern "C" int droid_prof_read(double* ms, int* counts, int nphase) {
  if (!ms || !counts || nphase <= 0) return DROID_ERR_INVALID_ARGUMENT;
  for (int i = 0; i < nphase; i++) { ms[i] = 0.0; counts[i] = 0; }
  std::lock_guard<std::mutex> lk(g_prof_mutex);
  for (auto& r : g_prof_recs) {
    float t = 0.f;
    cudaEventSynchronize(r.b);
    if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess && r.phase >= 0 && r.phase < nphase) {
      ms[r.phase] += (double)t;
      counts[r.phase] += 1;
    }
    g_prof_pool.push_back(r.a);
    g_prof_pool.push_back(r.b);
  }
  g_prof_recs.clear();
  cudaGetLastError();
  return DROID_OK;
}
extern "C" const char* droid_last_cuda_error(void) { return cudaGetErrorString(g_last_error); }
