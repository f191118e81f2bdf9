// Shared host/device helpers for libpn_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <atomic>

#include "../../include/pn_b200.h"

// SMs of a B200 (the only target: sm_100a).  A compile-time constant on purpose: persistent grids, CTAs per head and
// k-split counts are sized from it in constant expressions; pn_device_check() (pn_api.cu) refuses other devices.
#define PN_SM_COUNT 148

extern thread_local char g_pn_err[512];
extern std::atomic<int64_t> g_pn_launches;

static inline int pn_fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_pn_err, sizeof(g_pn_err), fmt, ap);
  va_end(ap);
  return code;
}

#define PN_CHECK_ARG(cond)                                                             \
  do {                                                                                 \
    if (!(cond)) return pn_fail(PN_E_BADARG, "%s: bad argument: %s", __func__, #cond); \
  } while (0)

#define PN_CUDA(call)                                                                              \
  do {                                                                                             \
    cudaError_t e__ = (call);                                                                      \
    if (e__ != cudaSuccess)                                                                        \
      return pn_fail(PN_E_LAUNCH, "%s: %s failed: %s", __func__, #call, cudaGetErrorString(e__)); \
  } while (0)

// after a <<<>>> launch
static inline void pn_count_launch() { g_pn_launches.fetch_add(1, std::memory_order_relaxed); }
#define PN_LAUNCHED()                                                                         \
  do {                                                                                        \
    g_pn_launches.fetch_add(1, std::memory_order_relaxed);                                    \
    cudaError_t e__ = cudaGetLastError();                                                     \
    if (e__ != cudaSuccess)                                                                   \
      return pn_fail(PN_E_LAUNCH, "%s: launch failed: %s", __func__, cudaGetErrorString(e__)); \
  } while (0)

// ---- per-kernel-class device timing (pn_prof_*, pn_api.cu): CUDA events recorded on the launching
// stream around a launch when profiling is enabled; bench.py reads the sums for its roofline object.
enum { PN_PROF_GEMM = 0, PN_PROF_ATTN = 1, PN_PROF_ELEMWISE = 2, PN_PROF_WGRAD = 3, PN_PROF_ATTN_BWD = 4,
       PN_PROF_GEMM_HBM = 5,   // GEMM launches whose operands + output outweigh their FLOPs (work = bytes)
       PN_PROF_KINDS = 8 };
extern std::atomic<int> g_pn_prof_on;
void pn_prof_begin_impl(int kind, double work, cudaStream_t st);
void pn_prof_end_impl(cudaStream_t st);
struct PnProfScope {
  cudaStream_t st;
  bool on;
  PnProfScope(int kind, double work, cudaStream_t s) : st(s), on(g_pn_prof_on.load(std::memory_order_relaxed) != 0) {
    if (on) pn_prof_begin_impl(kind, work, st);
  }
  ~PnProfScope() {
    if (on) pn_prof_end_impl(st);
  }
};

// grid for a grid-stride bandwidth kernel: a multiple of the SM count, capped by the work
static inline int pn_grid(int64_t work_items, int threads, int items_per_thread, int ctas_per_sm) {
  int64_t need = (work_items + (int64_t)threads * items_per_thread - 1) /
                 ((int64_t)threads * items_per_thread);
  int64_t cap = (int64_t)PN_SM_COUNT * ctas_per_sm;
  if (need < 1) need = 1;
  if (need > cap) need = cap;
  return (int)need;
}

#ifdef __CUDACC__
__device__ __forceinline__ float pn_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double pn_warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float pn_warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// streaming 128-bit load that does not allocate in L1 (inputs are touched once)
__device__ __forceinline__ float4 pn_ldg_stream(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
// request the 128-byte line holding p into L2 (no register, no wait)
__device__ __forceinline__ void pn_prefetch_l2(const void* p) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}
// 256-bit streaming accesses (sm_100: LDG.256 / STG.256): 32-byte aligned, eight floats per thread per access
struct __align__(32) pn_float8 { float v[8]; };
__device__ __forceinline__ pn_float8 pn_ldg_stream8(const pn_float8* p) {
  pn_float8 r;
  asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]),
                 "=f"(r.v[7])
               : "l"(p));
  return r;
}
__device__ __forceinline__ void pn_stg_stream8(pn_float8* p, const pn_float8& r) {
  asm volatile("st.global.L1::no_allocate.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(r.v[0]), "f"(r.v[1]),
               "f"(r.v[2]), "f"(r.v[3]), "f"(r.v[4]), "f"(r.v[5]), "f"(r.v[6]), "f"(r.v[7])
               : "memory");
}
__device__ __forceinline__ void pn_stg_stream(float4* p, float4 v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
#endif
