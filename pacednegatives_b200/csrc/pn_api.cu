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
