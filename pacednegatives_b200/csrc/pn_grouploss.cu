// LCE loss kernels.
//   pn_token_ce_{fwd,bwd}        nn.CrossEntropyLoss(reduction='none') at lceT5.py:110,117-118
//   pn_group_loss_{fwd,bwd}      view(-1,k) -> sum|mean -> pce + nce, lceT5.py:119-121
//   pn_lce_restricted_{fwd,bwd}  two-row head score (distill/wrapper.py:34-38) + group log-softmax
// All HBM-bound; token CE streams each logits row exactly once (online max/sum), 128-bit loads.
#include "pn_common.cuh"

namespace {

constexpr int kThreads = 256;

template <typename T> struct Vec;
template <> struct Vec<float> {
  static constexpr int N = 4;
  __device__ static void load(const float* p, float (&o)[4]) {
    const float4 v = pn_ldg_stream(reinterpret_cast<const float4*>(p));
    o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w;
  }
};
template <> struct Vec<__nv_bfloat16> {
  static constexpr int N = 8;
  __device__ static void load(const __nv_bfloat16* p, float (&o)[8]) {
    const float4 raw = pn_ldg_stream(reinterpret_cast<const float4*>(p));
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __bfloat1622float2(h[i]);
      o[2 * i] = f.x; o[2 * i + 1] = f.y;
    }
  }
};
__device__ __forceinline__ float to_f(float x) { return x; }
__device__ __forceinline__ float to_f(__nv_bfloat16 x) { return __bfloat162float(x); }

__device__ __forceinline__ void online_add(float& m, float& s, float x) {
  if (x > m) {
    s = s * __expf(m - x) + 1.0f;
    m = x;
  } else {
    s += __expf(x - m);
  }
}
__device__ __forceinline__ void online_merge(float& m, float& s, float m2, float s2) {
  const float mm = fmaxf(m, m2);
  if (mm == -INFINITY) return;
  s = s * __expf(m - mm) + s2 * __expf(m2 - mm);
  m = mm;
}

// one CTA per row
template <typename T>
__global__ void __launch_bounds__(kThreads)
token_ce_fwd_kernel(const T* __restrict__ logits, int64_t rows, int32_t V,
                    const int64_t* __restrict__ target, int64_t ignore_index,
                    float* __restrict__ ce_out, float* __restrict__ lse_out, bool vec_ok) {
  constexpr int N = Vec<T>::N;
  __shared__ float sm[kThreads / 32], ss[kThreads / 32];
  for (int64_t r = blockIdx.x; r < rows; r += gridDim.x) {
    const T* row = logits + r * (int64_t)V;
    float m = -INFINITY, s = 0.0f;
    int v0 = 0;
    if (vec_ok) {
      const int nv = V / N;
      for (int i = threadIdx.x; i < nv; i += kThreads) {
        float x[N];
        Vec<T>::load(row + (int64_t)i * N, x);
        float lm = x[0];
#pragma unroll
        for (int j = 1; j < N; ++j) lm = fmaxf(lm, x[j]);
        float ls = 0.0f;
#pragma unroll
        for (int j = 0; j < N; ++j) ls += __expf(x[j] - lm);
        online_merge(m, s, lm, ls);
      }
      v0 = nv * N;
    }
    for (int i = v0 + threadIdx.x; i < V; i += kThreads) online_add(m, s, to_f(row[i]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float m2 = __shfl_xor_sync(0xffffffffu, m, o);
      const float s2 = __shfl_xor_sync(0xffffffffu, s, o);
      online_merge(m, s, m2, s2);
    }
    if ((threadIdx.x & 31) == 0) { sm[threadIdx.x >> 5] = m; ss[threadIdx.x >> 5] = s; }
    __syncthreads();
    if (threadIdx.x == 0) {
      float M = sm[0], S = ss[0];
      for (int w = 1; w < kThreads / 32; ++w) online_merge(M, S, sm[w], ss[w]);
      const float lse = M + logf(S);
      if (lse_out) lse_out[r] = lse;
      const int64_t t = target[r];
      float ce = 0.0f;
      if (t != ignore_index && t >= 0 && t < V) ce = lse - to_f(row[t]);
      ce_out[r] = ce;
    }
    __syncthreads();
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
token_ce_bwd_kernel(const T* __restrict__ logits, int64_t rows, int32_t V,
                    const int64_t* __restrict__ target, int64_t ignore_index,
                    const float* __restrict__ lse, const float* __restrict__ g,
                    float* __restrict__ dlogits) {
  for (int64_t r = blockIdx.x; r < rows; r += gridDim.x) {
    const T* row = logits + r * (int64_t)V;
    float* drow = dlogits + r * (int64_t)V;
    const int64_t t = target[r];
    const bool live = (t != ignore_index && t >= 0 && t < V);
    const float gr = live ? g[r] : 0.0f;
    const float L = lse[r];
    for (int i = threadIdx.x; i < V; i += kThreads) {
      float p = __expf(to_f(row[i]) - L);
      if (i == t) p -= 1.0f;
      drow[i] = gr * p;
    }
  }
}

__global__ void __launch_bounds__(kThreads)
group_loss_fwd_kernel(const float* __restrict__ pce, const float* __restrict__ nce, int64_t m,
                      int32_t k, int use_mean, float* __restrict__ ce_out,
                      double* __restrict__ stats) {
  double local = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < m;
       i += (int64_t)gridDim.x * kThreads) {
    const float* p = nce + i * k;
    float acc = 0.0f;
    for (int j = 0; j < k; ++j) acc += p[j];
    if (use_mean) acc = acc / (float)k;
    const float ce = pce[i] + acc;
    ce_out[i] = ce;
    local += (double)ce;
  }
  if (stats) {
    local = pn_warp_sum(local);
    if ((threadIdx.x & 31) == 0) atomicAdd(stats, local);
  }
}

__global__ void __launch_bounds__(kThreads)
group_loss_bwd_kernel(const float* __restrict__ g, int64_t m, int32_t k, int use_mean,
                      float* __restrict__ dpce, float* __restrict__ dnce) {
  const float scale = use_mean ? 1.0f / (float)k : 1.0f;
  const int64_t total = m * (int64_t)k;
  for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < total;
       e += (int64_t)gridDim.x * kThreads) {
    const int64_t i = e / k;
    const float gi = g[i];
    dnce[e] = gi * scale;
    if (e - i * k == 0) dpce[i] = gi;
  }
}

// s = log_softmax([zt, zf])[0] = -softplus(zf - zt)
__device__ __forceinline__ float rel_score(float zt, float zf) {
  const float d = zf - zt;
  return d > 0.0f ? -(d + log1pf(expf(-d))) : -log1pf(expf(d));
}

// LANES (a power of two >= min(G, 32)) lanes per query: groups of up to 16 documents share a warp, so
// no lane idles and the reduction is log2(LANES) shuffle rounds
template <int LANES>
__global__ void __launch_bounds__(kThreads)
lce_restricted_fwd_kernel(const float* __restrict__ z2, int64_t Q, int32_t G,
                          float* __restrict__ loss_out, float* __restrict__ score_out) {
  // A CTA owns kU x (queries per CTA pass) CONSECUTIVE queries at a time: the first document of each lane for all kU
  // of them is requested before any arithmetic (kU loads in flight per thread instead of one), and the CTA's reads
  // and writes stay inside one contiguous span of z2 / score_out.
  constexpr int kU = 4;
  constexpr int kPerWarp = 32 / LANES;
  constexpr int kPerCta = (kThreads / 32) * kPerWarp;
  const int lane = threadIdx.x & 31;
  const int sub = lane % LANES;
  const int local = (threadIdx.x >> 5) * kPerWarp + lane / LANES;
  const float2* zall = reinterpret_cast<const float2*>(z2);
  for (int64_t tile = blockIdx.x; tile * (kU * kPerCta) < Q; tile += gridDim.x) {   // uniform trip count per CTA
    const int64_t q0 = tile * (kU * kPerCta) + local;
    float2 zf[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t q = q0 + (int64_t)u * kPerCta;
      zf[u] = (q < Q && sub < G) ? __ldg(zall + q * G + sub) : make_float2(0.0f, 0.0f);
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t q = q0 + (int64_t)u * kPerCta;
      const bool live = q < Q;
      float m = -INFINITY, s = 0.0f, s0 = 0.0f;
      if (live) {
        for (int j = sub; j < G; j += LANES) {
          const float2 z = j == sub ? zf[u] : __ldg(zall + q * G + j);
          const float sc = rel_score(z.x, z.y);
          if (score_out) score_out[q * G + j] = sc;
          if (j == 0) s0 = sc;
          online_add(m, s, sc);
        }
      }
      // group max first, then ONE exponential per lane to bring its partial sum to that reference (a pairwise
      // online merge costs two exponentials per lane per shuffle round)
      float gm = m;
#pragma unroll
      for (int o = LANES / 2; o > 0; o >>= 1) gm = fmaxf(gm, __shfl_xor_sync(0xffffffffu, gm, o));
      float gs = (m == -INFINITY) ? 0.0f : s * __expf(m - gm);
#pragma unroll
      for (int o = LANES / 2; o > 0; o >>= 1) gs += __shfl_xor_sync(0xffffffffu, gs, o);
      s0 = __shfl_sync(0xffffffffu, s0, lane - sub);
      if (live && sub == 0) loss_out[q] = (gm + logf(gs)) - s0;
    }
  }
}

__global__ void __launch_bounds__(kThreads)
lce_restricted_bwd_kernel(const float* __restrict__ z2, int64_t Q, int32_t G,
                          const float* __restrict__ g, float* __restrict__ dz2) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * kThreads) >> 5;
  for (int64_t q = warp; q < Q; q += nwarps) {
    const float2* zz = reinterpret_cast<const float2*>(z2) + q * G;
    float2* dz = reinterpret_cast<float2*>(dz2) + q * G;
    float m = -INFINITY, s = 0.0f;
    for (int j = lane; j < G; j += 32) {
      const float2 z = zz[j];
      online_add(m, s, rel_score(z.x, z.y));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float m2 = __shfl_xor_sync(0xffffffffu, m, o);
      const float s2 = __shfl_xor_sync(0xffffffffu, s, o);
      online_merge(m, s, m2, s2);
    }
    const float lse = m + logf(s);
    const float gq = g[q];
    for (int j = lane; j < G; j += 32) {
      const float2 z = zz[j];
      const float sc = rel_score(z.x, z.y);
      // dloss/ds_j = softmax_j - [j==0];  ds/dzt = sigmoid(zf - zt) = 1 - exp(s);  ds/dzf = -that
      const float ds = gq * (__expf(sc - lse) - (j == 0 ? 1.0f : 0.0f));
      const float pf = 1.0f - expf(sc);
      dz[j] = make_float2(ds * pf, -ds * pf);
    }
  }
}

}  // namespace

extern "C" int pn_token_ce_fwd(const void* logits, int dtype, int64_t rows, int32_t V,
                               const int64_t* target, int64_t ignore_index, float* ce_out,
                               float* lse_out, void* stream) {
  PN_CHECK_ARG(rows >= 0 && V > 0);
  PN_CHECK_ARG(dtype == 0 || dtype == 1);
  if (rows == 0) return PN_OK;
  PN_CHECK_ARG(logits && target && ce_out);
  const int grid = (int)(rows < PN_SM_COUNT * 8 ? rows : PN_SM_COUNT * 8);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == 0) {
    const bool vec = ((uintptr_t)logits & 15) == 0 && (V % 4) == 0;
    token_ce_fwd_kernel<float><<<grid, kThreads, 0, st>>>((const float*)logits, rows, V, target,
                                                          ignore_index, ce_out, lse_out, vec);
  } else {
    const bool vec = ((uintptr_t)logits & 15) == 0 && (V % 8) == 0;
    token_ce_fwd_kernel<__nv_bfloat16><<<grid, kThreads, 0, st>>>(
        (const __nv_bfloat16*)logits, rows, V, target, ignore_index, ce_out, lse_out, vec);
  }
  PN_LAUNCHED();
  return PN_OK;
}

extern "C" int pn_token_ce_bwd(const void* logits, int dtype, int64_t rows, int32_t V,
                               const int64_t* target, int64_t ignore_index, const float* lse,
                               const float* g, float* dlogits, void* stream) {
  PN_CHECK_ARG(rows >= 0 && V > 0);
  PN_CHECK_ARG(dtype == 0 || dtype == 1);
  if (rows == 0) return PN_OK;
  PN_CHECK_ARG(logits && target && lse && g && dlogits);
  const int grid = (int)(rows < PN_SM_COUNT * 8 ? rows : PN_SM_COUNT * 8);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == 0)
    token_ce_bwd_kernel<float><<<grid, kThreads, 0, st>>>((const float*)logits, rows, V, target,
                                                          ignore_index, lse, g, dlogits);
  else
    token_ce_bwd_kernel<__nv_bfloat16><<<grid, kThreads, 0, st>>>(
        (const __nv_bfloat16*)logits, rows, V, target, ignore_index, lse, g, dlogits);
  PN_LAUNCHED();
  return PN_OK;
}

extern "C" int pn_group_loss_fwd(const float* pce, const float* nce, int64_t m, int32_t k,
                                 int use_mean, float* ce_out, double* stats_out, void* stream) {
  PN_CHECK_ARG(m >= 0 && k > 0);
  cudaStream_t st = (cudaStream_t)stream;
  if (stats_out) PN_CUDA(cudaMemsetAsync(stats_out, 0, sizeof(double), st));
  if (m == 0) return PN_OK;
  PN_CHECK_ARG(pce && nce && ce_out);
  group_loss_fwd_kernel<<<pn_grid(m, kThreads, 1, 8), kThreads, 0, st>>>(pce, nce, m, k, use_mean,
                                                                         ce_out, stats_out);
  PN_LAUNCHED();
  return PN_OK;
}

extern "C" int pn_group_loss_bwd(const float* g, int64_t m, int32_t k, int use_mean, float* dpce,
                                 float* dnce, void* stream) {
  PN_CHECK_ARG(m >= 0 && k > 0);
  if (m == 0) return PN_OK;
  PN_CHECK_ARG(g && dpce && dnce);
  group_loss_bwd_kernel<<<pn_grid(m * k, kThreads, 1, 8), kThreads, 0, (cudaStream_t)stream>>>(
      g, m, k, use_mean, dpce, dnce);
  PN_LAUNCHED();
  return PN_OK;
}

extern "C" int pn_lce_restricted_fwd(const float* z2, int64_t Q, int32_t G, float* loss_out,
                                     float* score_out, void* stream) {
  PN_CHECK_ARG(Q >= 0 && G > 0);
  if (Q == 0) return PN_OK;
  PN_CHECK_ARG(z2 && loss_out);
  cudaStream_t st = (cudaStream_t)stream;
  if (G <= 4)
    lce_restricted_fwd_kernel<4><<<pn_grid(Q * 4, kThreads, 4, 8), kThreads, 0, st>>>(z2, Q, G, loss_out, score_out);
  else if (G <= 8)
    lce_restricted_fwd_kernel<8><<<pn_grid(Q * 8, kThreads, 4, 8), kThreads, 0, st>>>(z2, Q, G, loss_out, score_out);
  else if (G <= 16)
    lce_restricted_fwd_kernel<16><<<pn_grid(Q * 16, kThreads, 4, 8), kThreads, 0, st>>>(z2, Q, G, loss_out, score_out);
  else
    lce_restricted_fwd_kernel<32><<<pn_grid(Q * 32, kThreads, 4, 8), kThreads, 0, st>>>(z2, Q, G, loss_out, score_out);
  PN_LAUNCHED();
  return PN_OK;
}

extern "C" int pn_lce_restricted_bwd(const float* z2, int64_t Q, int32_t G, const float* g,
                                     float* dz2, void* stream) {
  PN_CHECK_ARG(Q >= 0 && G > 0);
  if (Q == 0) return PN_OK;
  PN_CHECK_ARG(z2 && g && dz2);
  lce_restricted_bwd_kernel<<<pn_grid(Q * 32, kThreads, 1, 8), kThreads, 0,
                              (cudaStream_t)stream>>>(z2, Q, G, g, dz2);
  PN_LAUNCHED();
  return PN_OK;
}

// ---- check_success / check_probs (old/levels.py:35-39, old/eta_bound.py:51-55): pair i succeeds when
// softmax(pos[i, 0, (REL, NREL)])[0] > softmax(neg[i, 0, (REL, NREL)])[0]; only those four logits of each pair
// are read from the [N, T, V] logits (row_stride = T*V elements between documents).
__device__ __forceinline__ float pn_p_rel(float zt, float zf) {
  // torch's fp32 softmax over two entries: exp(x - max) / sum
  const float m = fmaxf(zt, zf);
  const float et = expf(zt - m), ef = expf(zf - m);
  return et / (et + ef);
}

__global__ void __launch_bounds__(kThreads) check_success_kernel(const float* __restrict__ pos,
                                                                 const float* __restrict__ neg, int64_t n,
                                                                 int64_t pos_stride, int64_t neg_stride, int rel,
                                                                 int nrel, int32_t* __restrict__ count) {
  int local = 0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float* p = pos + i * pos_stride;
    const float* q = neg + i * neg_stride;
    local += pn_p_rel(__ldg(p + rel), __ldg(p + nrel)) > pn_p_rel(__ldg(q + rel), __ldg(q + nrel)) ? 1 : 0;
  }
  local = __reduce_add_sync(0xffffffffu, local);
  if ((threadIdx.x & 31) == 0 && local) atomicAdd(count, local);
}

extern "C" int pn_check_success(const float* pos_logits, const float* neg_logits, int64_t n, int64_t pos_stride,
                                int64_t neg_stride, int32_t rel_id, int32_t nrel_id, int32_t* count_out,
                                void* stream) {
  PN_CHECK_ARG(n >= 0 && rel_id >= 0 && nrel_id >= 0 && count_out);
  PN_CHECK_ARG(pos_stride > rel_id && pos_stride > nrel_id && neg_stride > rel_id && neg_stride > nrel_id);
  cudaStream_t st = (cudaStream_t)stream;
  PN_CUDA(cudaMemsetAsync(count_out, 0, sizeof(int32_t), st));
  if (n == 0) return PN_OK;
  PN_CHECK_ARG(pos_logits && neg_logits);
  check_success_kernel<<<pn_grid(n, kThreads, 1, 8), kThreads, 0, st>>>(pos_logits, neg_logits, n, pos_stride,
                                                                        neg_stride, rel_id, nrel_id, count_out);
  PN_LAUNCHED();
  return PN_OK;
}
