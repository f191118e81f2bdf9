// Library-level entry points of libpn_b200.so.
#include "pn_common.cuh"

thread_local char g_pn_err[512] = "";
std::atomic<int64_t> g_pn_launches{0};

extern "C" int pn_version(void) { return PN_VERSION; }

extern "C" const char* pn_last_error(void) { return g_pn_err; }

extern "C" int64_t pn_launch_count(void) { return g_pn_launches.load(); }

extern "C" int pn_device_check(void) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess)
    return pn_fail(PN_E_ARCH, "pn_device_check: no CUDA device: %s", cudaGetErrorString(e));
  int major = 0, minor = 0;
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
  if (major != 10)
    return pn_fail(PN_E_ARCH,
                   "pn_device_check: device %d is sm_%d%d; this library is built for sm_100a only",
                   dev, major, minor);
  return PN_OK;
}

// ---------------------------------------------------------------- per-kernel-class event timing
// A ring of event pairs; a pair older than the ring length has certainly completed (the launch queue
// is far shorter), so harvesting it on wrap-around does not stall the stream.
#include <mutex>
#include <vector>

std::atomic<int> g_pn_prof_on{0};

namespace {
struct ProfPair { cudaEvent_t a = nullptr, b = nullptr; int kind = -1; double work = 0; bool live = false; };
constexpr int kProfRing = 16384;
std::vector<ProfPair> g_ring;
int g_head = 0;
std::mutex g_prof_mu;
double g_prof_ms[PN_PROF_KINDS];
double g_prof_work[PN_PROF_KINDS];
int64_t g_prof_n[PN_PROF_KINDS];

void harvest(ProfPair& p) {
  if (!p.live) return;
  float ms = 0.0f;
  if (cudaEventSynchronize(p.b) == cudaSuccess && cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) {
    g_prof_ms[p.kind] += ms;
    g_prof_work[p.kind] += p.work;
    g_prof_n[p.kind] += 1;
  }
  p.live = false;
}
}  // namespace

void pn_prof_begin_impl(int kind, double work, cudaStream_t st) {
  std::lock_guard<std::mutex> g(g_prof_mu);
  if (g_ring.empty()) g_ring.resize(kProfRing);
  ProfPair& p = g_ring[g_head];
  harvest(p);
  if (!p.a) { cudaEventCreate(&p.a); cudaEventCreate(&p.b); }
  p.kind = kind; p.work = work; p.live = true;
  cudaEventRecord(p.a, st);
}

void pn_prof_end_impl(cudaStream_t st) {
  std::lock_guard<std::mutex> g(g_prof_mu);
  ProfPair& p = g_ring[g_head];
  cudaEventRecord(p.b, st);
  g_head = (g_head + 1) % kProfRing;
}

extern "C" int pn_prof_enable(int on) {
  std::lock_guard<std::mutex> g(g_prof_mu);
  if (on) {
    for (auto& p : g_ring) p.live = false;
    for (int k = 0; k < PN_PROF_KINDS; ++k) { g_prof_ms[k] = 0; g_prof_work[k] = 0; g_prof_n[k] = 0; }
  }
  g_pn_prof_on.store(on ? 1 : 0);
  return PN_OK;
}

extern "C" int pn_prof_read(int kind, double* ms_sum, int64_t* launches, double* work_sum) {
  PN_CHECK_ARG(kind >= 0 && kind < PN_PROF_KINDS);
  std::lock_guard<std::mutex> g(g_prof_mu);
  for (auto& p : g_ring) harvest(p);
  if (ms_sum) *ms_sum = g_prof_ms[kind];
  if (launches) *launches = g_prof_n[kind];
  if (work_sum) *work_sum = g_prof_work[kind];
  return PN_OK;
}
