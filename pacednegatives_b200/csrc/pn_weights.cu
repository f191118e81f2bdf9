// Self-paced weighting kernels.
//   pn_weights_fwd           lceT5.py:91-99, weights.py:26-45,59-78
//   pn_weights_table_update  pairwrapper.py:219-239, weights.py:87-88
// HBM-bound: 4 B read + 4 B write (+1 B keep mask) per element; the six reductions ride along
// in registers -> warp shuffle -> one fp64 atomic per CTA per statistic.
#include <stdlib.h>
#include "pn_common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kUnroll = 4;  // float4 loads in flight per thread

// Raw per-thread sums: everything the six statistics need, as plain sums of l and l^2 split by the keep mask,
// so the per-element work is one IEEE divide, two compares and four fp32 adds; fp32 partials are flushed to
// fp64 every 16 elements.  The statistics are assembled from the raw sums by the last CTA to finish:
//   kept (l <= eta):  w = l/eta (LCE) | 1 - l/eta (ETA)      over:  w = l*mask*eta
//   sum w, sum w*l and d(sum w*l)/d eta are linear in  A1 = sum_kept l, A2 = sum_kept l^2, B1 = sum_over l,
//   B2 = sum_over l^2  (they differ from the sums of the ROUNDED fp32 weights by ~1e-7 relative).
struct Raw {
  float a1, a2, b1, b2;
  int nkeep, nlt;
};

template <int VARIANT>
__device__ __forceinline__ float weight_one(float l, float eta, float mask, Raw& r) {
  const bool over = l > eta;  // the reference's `if loss[i] > self.eta`
  const float l2 = l * l;
  float w;
  if (over) {
    w = __fmul_rn(__fmul_rn(l, mask), eta);
    r.b1 += l;
    r.b2 += l2;
  } else {
    const float q = __fdiv_rn(l, eta);
    w = (VARIANT == PN_W_LCE) ? q : __fadd_rn(-q, 1.0f);
    r.a1 += l;
    r.a2 += l2;
    r.nkeep += 1;
  }
  r.nlt += (l < eta) ? 1 : 0;
  return w;
}

template <int VARIANT, int VEC>   // VEC: 0 scalar, 1 = 128-bit accesses, 2 = 256-bit accesses (32-byte aligned)
__global__ void __launch_bounds__(kThreads)
weights_fwd_kernel(const float* __restrict__ loss, int64_t n, const float* __restrict__ eta_p,
                   float mask, float* __restrict__ w_out, uint8_t* __restrict__ keep_out,
                   double* __restrict__ stats) {
  const float eta = __ldg(eta_p);
  double acc[6] = {0, 0, 0, 0, 0, 0};   // A1 A2 B1 B2 nkeep nlt
  auto flush = [&](Raw& r) {
    acc[0] += (double)r.a1; acc[1] += (double)r.a2; acc[2] += (double)r.b1; acc[3] += (double)r.b2;
    acc[4] += (double)r.nkeep; acc[5] += (double)r.nlt;
    r = Raw{0.f, 0.f, 0.f, 0.f, 0, 0};
  };
  Raw r = {0.f, 0.f, 0.f, 0.f, 0, 0};
  const int64_t tid = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  const int64_t nthreads = (int64_t)gridDim.x * kThreads;

  if (VEC == 2) {
    const int64_t n8 = n >> 3;
    const pn_float8* in8 = reinterpret_cast<const pn_float8*>(loss);
    pn_float8* out8 = reinterpret_cast<pn_float8*>(w_out);
    // a CTA works on one contiguous 32 KB tile at a time (kUnroll x 256 x 32 B): its accesses stay inside a few DRAM
    // pages instead of being spread kUnroll grid-widths apart
    const int64_t tile_elems = (int64_t)kUnroll * kThreads;
    for (int64_t tile = blockIdx.x; tile * tile_elems < n8; tile += gridDim.x) {
      const int64_t base = tile * tile_elems + threadIdx.x;
      pn_float8 v[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t i = base + (int64_t)u * kThreads;
        if (i < n8) v[u] = pn_ldg_stream8(in8 + i);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t i = base + (int64_t)u * kThreads;
        if (i < n8) {
          pn_float8 w;
#pragma unroll
          for (int j = 0; j < 8; ++j) w.v[j] = weight_one<VARIANT>(v[u].v[j], eta, mask, r);
          pn_stg_stream8(out8 + i, w);
          if (keep_out) {
            uint32_t lo = 0, hi = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              lo |= (uint32_t)(!(v[u].v[j] > eta)) << (8 * j);
              hi |= (uint32_t)(!(v[u].v[4 + j] > eta)) << (8 * j);
            }
            reinterpret_cast<uint2*>(keep_out)[i] = make_uint2(lo, hi);
          }
        }
        flush(r);
      }
    }
    for (int64_t i = (n8 << 3) + tid; i < n; i += nthreads) {   // scalar tail
      const float l = loss[i];
      w_out[i] = weight_one<VARIANT>(l, eta, mask, r);
      if (keep_out) keep_out[i] = !(l > eta);
    }
    flush(r);
  } else if (VEC) {
    const int64_t n4 = n >> 2;
    const float4* in4 = reinterpret_cast<const float4*>(loss);
    float4* out4 = reinterpret_cast<float4*>(w_out);
    const int64_t tile_elems = (int64_t)kUnroll * kThreads;   // contiguous 16 KB tile per CTA, as above
    for (int64_t tile = blockIdx.x; tile * tile_elems < n4; tile += gridDim.x) {
      const int64_t base = tile * tile_elems + threadIdx.x;
      float4 v[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t i = base + (int64_t)u * kThreads;
        if (i < n4) v[u] = pn_ldg_stream(in4 + i);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t i = base + (int64_t)u * kThreads;
        if (i < n4) {
          float4 w;
          w.x = weight_one<VARIANT>(v[u].x, eta, mask, r);
          w.y = weight_one<VARIANT>(v[u].y, eta, mask, r);
          w.z = weight_one<VARIANT>(v[u].z, eta, mask, r);
          w.w = weight_one<VARIANT>(v[u].w, eta, mask, r);
          pn_stg_stream(out4 + i, w);
          if (keep_out) {
            uchar4 k;
            k.x = !(v[u].x > eta);
            k.y = !(v[u].y > eta);
            k.z = !(v[u].z > eta);
            k.w = !(v[u].w > eta);
            reinterpret_cast<uchar4*>(keep_out)[i] = k;
          }
        }
      }
      flush(r);
    }
    // scalar tail
    for (int64_t i = (n4 << 2) + tid; i < n; i += nthreads) {
      const float l = loss[i];
      w_out[i] = weight_one<VARIANT>(l, eta, mask, r);
      if (keep_out) keep_out[i] = !(l > eta);
    }
    flush(r);
  } else {
    int cnt = 0;
    for (int64_t i = tid; i < n; i += nthreads) {
      const float l = loss[i];
      w_out[i] = weight_one<VARIANT>(l, eta, mask, r);
      if (keep_out) keep_out[i] = !(l > eta);
      if (++cnt == 16) { flush(r); cnt = 0; }
    }
    flush(r);
  }

  if (stats == nullptr) return;
  __shared__ double red[6][kThreads / 32];
  __shared__ bool last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    const double t = pn_warp_sum(acc[k]);
    if (lane == 0) red[k][warp] = t;
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    double t = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) t += red[threadIdx.x][w];
    atomicAdd(stats + 6 + threadIdx.x, t);   // raw sums live in stats[6..11] until the last CTA folds them
    __threadfence();
  }
  __syncthreads();
  unsigned int* done = reinterpret_cast<unsigned int*>(stats + 12);   // zeroed with the scratch by the host
  if (threadIdx.x == 0) last = atomicAdd(done, 1u) == gridDim.x - 1;
  __syncthreads();
  if (last && threadIdx.x == 0) {
    __threadfence();
    volatile double* raw = stats + 6;
    const double A1 = raw[0], A2 = raw[1], B1 = raw[2], B2 = raw[3], nkeep = raw[4], nlt = raw[5];
    const double e = (double)eta, mk = (double)mask;
    double sw, swl, dswl;
    if (VARIANT == PN_W_LCE) {
      sw = A1 / e + mk * e * B1;
      swl = A2 / e + mk * e * B2;
      dswl = -A2 / (e * e) + mk * B2;
    } else {
      sw = nkeep - A1 / e + mk * e * B1;
      swl = A1 - A2 / e + mk * e * B2;
      dswl = A2 / (e * e) + mk * B2;
    }
    stats[0] = sw; stats[1] = swl; stats[2] = dswl; stats[3] = nlt; stats[4] = A1 + B1; stats[5] = nkeep;
  }
}

__device__ __forceinline__ float table_update_one(float v, float cp, float cn, bool has_p, bool has_n, float inv_len,
                                                  float lr) {
  float acc = 0.0f;
  if (has_p) acc = __fmul_rn(cp, inv_len);
  if (has_n) acc = has_p ? __fadd_rn(acc, __fmul_rn(cn, inv_len)) : __fmul_rn(cn, inv_len);
  const float g = __fadd_rn(acc, -1.0f);
  const float x = __fsub_rn(v, __fmul_rn(lr, g));
  return 1.0f / (1.0f + expf(-x));
}

// VEC: the three arrays are 16-byte aligned -> float4 accesses, one contiguous 16 KB tile per CTA at a time
template <bool VEC>
__global__ void __launch_bounds__(kThreads)
table_update_kernel(float* __restrict__ row, const float* __restrict__ ce_p,
                    const float* __restrict__ ce_n, int64_t n, float inv_len, float lr) {
  const bool hp = ce_p != nullptr, hn = ce_n != nullptr;
  int64_t done = 0;
  if (VEC) {
    const int64_t n4 = n >> 2;
    const int64_t tile_elems = (int64_t)kUnroll * kThreads;
    float4* r4 = reinterpret_cast<float4*>(row);
    const float4* p4 = reinterpret_cast<const float4*>(ce_p);
    const float4* q4 = reinterpret_cast<const float4*>(ce_n);
    for (int64_t tile = blockIdx.x; tile * tile_elems < n4; tile += gridDim.x) {
      const int64_t base = tile * tile_elems + threadIdx.x;
      float4 v[kUnroll], a[kUnroll], b[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t i = base + (int64_t)u * kThreads;
        if (i < n4) {
          v[u] = r4[i];
          a[u] = hp ? pn_ldg_stream(p4 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
          b[u] = hn ? pn_ldg_stream(q4 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t i = base + (int64_t)u * kThreads;
        if (i < n4) {
          float4 o;
          o.x = table_update_one(v[u].x, a[u].x, b[u].x, hp, hn, inv_len, lr);
          o.y = table_update_one(v[u].y, a[u].y, b[u].y, hp, hn, inv_len, lr);
          o.z = table_update_one(v[u].z, a[u].z, b[u].z, hp, hn, inv_len, lr);
          o.w = table_update_one(v[u].w, a[u].w, b[u].w, hp, hn, inv_len, lr);
          r4[i] = o;
        }
      }
    }
    done = n4 << 2;
  }
  for (int64_t i = done + (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads)
    row[i] = table_update_one(row[i], hp ? ce_p[i] : 0.0f, hn ? ce_n[i] : 0.0f, hp, hn, inv_len, lr);
}

template <int VARIANT>
__global__ void __launch_bounds__(kThreads)
weights_bwd_kernel(const float* __restrict__ loss, int64_t n, const float* __restrict__ eta_p,
                   float mask, const float* __restrict__ g, float* __restrict__ dloss,
                   double* __restrict__ deta) {
  const float eta = __ldg(eta_p);
  const double ed = (double)eta;
  double acc = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * kThreads) {
    const float l = loss[i], gi = g[i];
    double dwde, dwdl;
    if (l > eta) {
      dwde = (double)l * (double)mask;
      dwdl = (double)mask * ed;
    } else {
      const double sgn = (VARIANT == PN_W_LCE) ? 1.0 : -1.0;
      dwde = -sgn * (double)l / (ed * ed);
      dwdl = sgn / ed;
    }
    acc += (double)gi * dwde;
    if (dloss) dloss[i] = (float)((double)gi * dwdl);
  }
  acc = pn_warp_sum(acc);
  __shared__ double red[kThreads / 32];
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double r = 0;
    for (int w = 0; w < kThreads / 32; ++w) r += red[w];
    atomicAdd(deta, r);
  }
}

}  // namespace

extern "C" int pn_weights_bwd(const float* loss, int64_t n, const float* eta, float mask,
                              int variant, const float* g, float* dloss_out, double* deta_out,
                              void* stream) {
  PN_CHECK_ARG(n >= 0);
  PN_CHECK_ARG(variant == PN_W_LCE || variant == PN_W_ETA);
  PN_CHECK_ARG(eta != nullptr && deta_out != nullptr);
  cudaStream_t st = (cudaStream_t)stream;
  PN_CUDA(cudaMemsetAsync(deta_out, 0, sizeof(double), st));
  if (n == 0) return PN_OK;
  PN_CHECK_ARG(loss != nullptr && g != nullptr);
  const int grid = pn_grid(n, kThreads, 4, 8);
  if (variant == PN_W_LCE)
    weights_bwd_kernel<PN_W_LCE><<<grid, kThreads, 0, st>>>(loss, n, eta, mask, g, dloss_out, deta_out);
  else
    weights_bwd_kernel<PN_W_ETA><<<grid, kThreads, 0, st>>>(loss, n, eta, mask, g, dloss_out, deta_out);
  PN_LAUNCHED();
  return PN_OK;
}

extern "C" int pn_weights_fwd(const float* loss, int64_t n, const float* eta, float mask,
                              int variant, float* w_out, uint8_t* keep_out, double* stats_out,
                              void* stream) {
  PN_CHECK_ARG(n >= 0);
  PN_CHECK_ARG(variant == PN_W_LCE || variant == PN_W_ETA);
  PN_CHECK_ARG(eta != nullptr);
  cudaStream_t st = (cudaStream_t)stream;
  // the kernel accumulates raw sums (and its done-counter) in stats_out[6..12] — the CALLER's buffer, so calls on
  // different streams or from different host threads share nothing — and the last CTA folds them into [0..5]
  double* stats_dev = stats_out;
  if (stats_out) PN_CUDA(cudaMemsetAsync(stats_out, 0, 13 * sizeof(double), st));
  if (n == 0) return PN_OK;
  PN_CHECK_ARG(loss != nullptr && w_out != nullptr);
  const bool vec = (((uintptr_t)loss | (uintptr_t)w_out) & 15) == 0 &&
                   (keep_out == nullptr || ((uintptr_t)keep_out & 3) == 0);
  static int want8 = -1;        // PN_W_VEC8=0 falls back to 128-bit accesses
  if (want8 < 0) {
    const char* e = getenv("PN_W_VEC8");
    want8 = e ? atoi(e) : 1;
  }
  const bool vec8 = want8 && (((uintptr_t)loss | (uintptr_t)w_out) & 31) == 0 &&
                    (keep_out == nullptr || ((uintptr_t)keep_out & 7) == 0);
  static int ctas_per_sm = 0;   // PN_W_CTAS: tuning knob for the grid size (resident CTAs per SM)
  if (!ctas_per_sm) {
    const char* e = getenv("PN_W_CTAS");
    ctas_per_sm = e && atoi(e) > 0 ? atoi(e) : 8;
  }
  const int grid = pn_grid(n, kThreads, vec8 ? 8 * kUnroll : (vec ? 4 * kUnroll : 4), ctas_per_sm);
#define LAUNCH(V, VEC)                                                                       \
  weights_fwd_kernel<V, VEC><<<grid, kThreads, 0, st>>>(loss, n, eta, mask, w_out, keep_out, \
                                                        stats_dev)
  if (variant == PN_W_LCE) {
    if (vec8) LAUNCH(PN_W_LCE, 2); else if (vec) LAUNCH(PN_W_LCE, 1); else LAUNCH(PN_W_LCE, 0);
  } else {
    if (vec8) LAUNCH(PN_W_ETA, 2); else if (vec) LAUNCH(PN_W_ETA, 1); else LAUNCH(PN_W_ETA, 0);
  }
#undef LAUNCH
  PN_LAUNCHED();
  return PN_OK;
}

extern "C" int pn_weights_table_update(float* row, const float* ce_p, const float* ce_n, int64_t n,
                                       float inv_len, float lr, void* stream) {
  PN_CHECK_ARG(n >= 0);
  if (n == 0) return PN_OK;
  PN_CHECK_ARG(row != nullptr);
  PN_CHECK_ARG(ce_p != nullptr || ce_n != nullptr);
  const bool vec = (((uintptr_t)row | (uintptr_t)ce_p | (uintptr_t)ce_n) & 15) == 0;
  const int grid = pn_grid(n, kThreads, vec ? 4 * kUnroll : 4, 8);
  if (vec) table_update_kernel<true><<<grid, kThreads, 0, (cudaStream_t)stream>>>(row, ce_p, ce_n, n, inv_len, lr);
  else table_update_kernel<false><<<grid, kThreads, 0, (cudaStream_t)stream>>>(row, ce_p, ce_n, n, inv_len, lr);
  PN_LAUNCHED();
  return PN_OK;
}
