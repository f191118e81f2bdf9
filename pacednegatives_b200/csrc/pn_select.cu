// Paced negative selection.
//   level mode     dataloader.py:29-33   r = floor|ceil(w * n_neg) in fp64, doc = cand[q][r]
//   binomial mode  dataloader.py:134-151 scipy binom.pmf + numpy legacy choice(replace=False, p)
// Candidate ids are int32, row-major [Q,C] (the reference's per-query list) or rank-major [C,Q]
// (one contiguous row per difficulty rank, so a scalar level is a coalesced copy).
// fp64 arithmetic below avoids FMA contraction on purpose: the index decisions must be bit-identical
// to numpy's.  This file is compiled with -fmad=false as well.
#include "pn_common.cuh"
#include <math.h>

namespace {

constexpr int kThreads = 256;

// ---------------------------------------------------------------- level mode
__global__ void __launch_bounds__(kThreads)
level_scalar_kernel(const int32_t* __restrict__ cand, int64_t Q, int32_t C, int layout, int32_t r,
                    int32_t* __restrict__ idx_out, int32_t* __restrict__ doc_out,
                    int32_t* __restrict__ status) {
  const int64_t tid = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  const int64_t nth = (int64_t)gridDim.x * kThreads;
  const bool ok = (r >= 0 && r < C);
  if (tid == 0) {
    if (idx_out) idx_out[0] = ok ? r : -1;
    if (!ok && status) atomicAdd(status, (int32_t)(Q > 0x7fffffff ? 0x7fffffff : Q));
  }
  if (!ok) {
    for (int64_t q = tid; q < Q; q += nth) doc_out[q] = -1;
    return;
  }
  if (layout == PN_LAYOUT_RANK) {
    const int32_t* src = cand + (int64_t)r * Q;
    if ((((uintptr_t)src | (uintptr_t)doc_out) & 15) == 0) {
      const int64_t q4 = Q >> 2;
      const int4* s4 = reinterpret_cast<const int4*>(src);
      int4* d4 = reinterpret_cast<int4*>(doc_out);
      for (int64_t i = tid; i < q4; i += nth) {
        int4 v;
        asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(s4 + i));
        d4[i] = v;
      }
      for (int64_t q = (q4 << 2) + tid; q < Q; q += nth) doc_out[q] = src[q];
    } else {
      for (int64_t q = tid; q < Q; q += nth) doc_out[q] = src[q];
    }
  } else {
    for (int64_t q = tid; q < Q; q += nth) doc_out[q] = __ldg(cand + q * (int64_t)C + r);
  }
}

__global__ void __launch_bounds__(kThreads)
level_items_kernel(const int32_t* __restrict__ cand, int64_t Q, int32_t C, int layout,
                   const float* __restrict__ w, int use_ceil, int32_t* __restrict__ idx_out,
                   int32_t* __restrict__ doc_out, int32_t* __restrict__ status) {
  const double n_neg = (double)(C - 1);
  int bad = 0;
  for (int64_t q = (int64_t)blockIdx.x * kThreads + threadIdx.x; q < Q;
       q += (int64_t)gridDim.x * kThreads) {
    // weights[i].item() -> Python float (fp64); weight * n_neg; floor | ceil
    const double x = __dmul_rn((double)w[q], n_neg);
    const double rr = use_ceil ? ceil(x) : floor(x);
    int32_t r = -1, d = -1;
    if (rr >= 0.0 && rr < (double)C) {
      r = (int32_t)rr;
      d = (layout == PN_LAYOUT_RANK) ? __ldg(cand + (int64_t)r * Q + q)
                                     : __ldg(cand + q * (int64_t)C + r);
    } else {
      bad++;
    }
    if (idx_out) idx_out[q] = r;
    doc_out[q] = d;
  }
  if (status && bad) atomicAdd(status, bad);
}

// ---------------------------------------------------------------- binomial table
// numpy's pairwise summation for a contiguous fp64 reduce (n <= 8192 is a single inner-loop call).
__device__ double np_pairwise_sum(const double* a, int n) {
  if (n < 8) {
    double res = 0.0;
    for (int i = 0; i < n; ++i) res = __dadd_rn(res, a[i]);
    return res;
  } else if (n <= 128) {
    double r[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = a[j];
    int i;
    for (i = 8; i < n - (n % 8); i += 8) {
#pragma unroll
      for (int j = 0; j < 8; ++j) r[j] = __dadd_rn(r[j], a[i + j]);
    }
    double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                           __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
    for (; i < n; ++i) res = __dadd_rn(res, a[i]);
    return res;
  } else {
    int n2 = n / 2;
    n2 -= n2 % 8;
    return __dadd_rn(np_pairwise_sum(a, n2), np_pairwise_sum(a + n2, n - n2));
  }
}

// table_out[0..C) = p, table_out[C..2C) = first-round cdf.  One CTA; thread 0 does the ordered parts.
__global__ void __launch_bounds__(kThreads)
binomial_table_kernel(double mean, double var, int32_t C, const double* __restrict__ pmf_in,
                      double* __restrict__ table, double* __restrict__ scratch,
                      int32_t* __restrict__ status) {
  double* p = table;
  double* cdf = table + C;
  const int n = C - 1;
  if (pmf_in) {
    for (int i = threadIdx.x; i < C; i += kThreads) p[i] = pmf_in[i];
  } else {
    const double lp = log(mean), lq = log1p(-mean), lgn = lgamma((double)n + 1.0);
    for (int i = threadIdx.x; i < C; i += kThreads) {
      const double k = (double)i;
      const double lg = lgn - lgamma(k + 1.0) - lgamma((double)n - k + 1.0) + k * lp +
                        ((double)n - k) * lq;
      p[i] = exp(lg);
    }
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  // adjusted_probabilities = probabilities / probabilities.sum()
  const double s = np_pairwise_sum(p, C);
  for (int i = 0; i < C; ++i) p[i] = __ddiv_rn(p[i], s);
  // np.var: mean = sum/N ; x = a - mean ; x = x*x ; var = sum(x)/N
  const double mu = __ddiv_rn(np_pairwise_sum(p, C), (double)C);
  for (int i = 0; i < C; ++i) {
    const double d = __dsub_rn(p[i], mu);
    scratch[i] = __dmul_rn(d, d);
  }
  const double v = __ddiv_rn(np_pairwise_sum(scratch, C), (double)C);
  const double scale = sqrt(__ddiv_rn(var, v));
  for (int i = 0; i < C; ++i) p[i] = __dmul_rn(p[i], scale);
  const double s2 = np_pairwise_sum(p, C);
  int nonzero = 0, bad = 0;
  for (int i = 0; i < C; ++i) {
    p[i] = __ddiv_rn(p[i], s2);
    if (p[i] > 0.0) nonzero++;
    if (!(p[i] >= 0.0) || isinf(p[i])) bad = 1;  // NaN / negative / inf -> numpy raises ValueError
  }
  // cdf = cumsum(p); cdf /= cdf[-1]
  double c = 0.0;
  for (int i = 0; i < C; ++i) {
    c = (i == 0) ? p[0] : __dadd_rn(c, p[i]);
    cdf[i] = c;
  }
  const double tot = cdf[C - 1];
  for (int i = 0; i < C; ++i) cdf[i] = __ddiv_rn(cdf[i], tot);
  status[0] = nonzero;
  status[1] = bad;
}

// ---------------------------------------------------------------- binomial select
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                              uint32_t k0, uint32_t k1, uint32_t (&out)[4]) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
// numpy legacy double: (a >> 5, b >> 6) -> (a * 2^26 + b) / 2^53
__device__ __forceinline__ double res53(uint32_t a, uint32_t b) {
  return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) / 9007199254740992.0;
}

struct UniformSrc {
  const double* u;  // nullptr -> philox
  int64_t avail;
  uint64_t seed;
  uint64_t q;
  __device__ double get(int64_t i) const {
    if (u) return u[i];
    uint32_t o[4];
    philox4x32_10((uint32_t)(i >> 1), (uint32_t)((uint64_t)i >> 33), (uint32_t)q,
                  (uint32_t)(q >> 32), (uint32_t)seed, (uint32_t)(seed >> 32), o);
    return (i & 1) ? res53(o[2], o[3]) : res53(o[0], o[1]);
  }
};

constexpr int kMaxK = 64;

// searchsorted(cdf, x, side='right') on the shared first-round cdf
__device__ __forceinline__ int upper_bound(const double* cdf, int C, double x) {
  int lo = 0, hi = C;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (cdf[mid] <= x) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// One query, numpy RandomState.choice(replace=False, p) verbatim.  `found` is the output row.
// `mask` = per-thread bitmask of found ranks in shared memory (C bits).
// Returns uniforms consumed, or -1 on failure (ran out of uniforms).
__device__ int choice_one(const double* sp, const double* scdf, int C, int k, const UniformSrc& src,
                          int64_t pos0, int32_t* found, uint32_t* mask, int mask_words) {
  for (int w = 0; w < mask_words; ++w) mask[w] = 0u;
  double xs[kMaxK];
  int n_uniq = 0;
  int64_t pos = pos0;
  int rounds = 0;
  while (n_uniq < k) {
    // numpy raises before sampling when fewer than k entries of p are non-zero; the host checks the
    // table status for that.  The round cap only keeps a bad table from spinning the GPU.
    if (++rounds > 4 * kMaxK) return -1;
    const int m = k - n_uniq;
    if (src.u && pos + m > src.avail) return -1;
    for (int j = 0; j < m; ++j) xs[j] = src.get(pos + j);
    pos += m;
    int32_t* nw = found + n_uniq;  // new[] lives in the tail of the output row
    if (n_uniq == 0) {
      for (int j = 0; j < m; ++j) nw[j] = upper_bound(scdf, C, xs[j]);
    } else {
      // p[found] = 0 ; cdf = cumsum(p) ; cdf /= cdf[-1] ; searchsorted(x, 'right')
      double tot = 0.0;
      for (int i = 0; i < C; ++i) {
        const double pi = (mask[i >> 5] >> (i & 31)) & 1u ? 0.0 : sp[i];
        tot = (i == 0) ? pi : __dadd_rn(tot, pi);
      }
      // searchsorted(cdf / tot, x, 'right') = index of the first entry with fl(c_i / tot) > x (the cdf is
      // non-decreasing).  The draws are visited in ascending order, so each entry is compared with ONE draw; the
      // correctly rounded quotient is only formed when c_i lies within 2^-48 (relative) of x * tot — outside that
      // band the comparison of fl(c_i / tot) with x is decided by c_i against x * tot alone.
      int ord[kMaxK];
      for (int j = 0; j < m; ++j) {
        int t = j;
        while (t > 0 && xs[ord[t - 1]] > xs[j]) { ord[t] = ord[t - 1]; --t; }
        ord[t] = j;
      }
      int t = 0;
      double lo_t = 0.0, hi_t = 0.0, x_t = 0.0;
      auto next_draw = [&]() {
        if (t < m) {
          x_t = xs[ord[t]];
          const double xt = __dmul_rn(x_t, tot);
          lo_t = __dmul_rn(xt, 1.0 - 0x1p-48);
          hi_t = __dmul_rn(xt, 1.0 + 0x1p-48);
        }
      };
      next_draw();
      double c = 0.0;
      for (int i = 0; i < C && t < m; ++i) {
        const double pi = (mask[i >> 5] >> (i & 31)) & 1u ? 0.0 : sp[i];
        c = (i == 0) ? pi : __dadd_rn(c, pi);
        while (t < m) {
          bool le;
          if (c <= lo_t) le = true;
          else if (c > hi_t) le = false;
          else le = __ddiv_rn(c, tot) <= x_t;
          if (le) break;
          nw[ord[t]] = i;
          ++t;
          next_draw();
        }
      }
      for (; t < m; ++t) nw[ord[t]] = C;
    }
    // first-occurrence unique, original order kept
    int kept = 0;
    for (int j = 0; j < m; ++j) {
      const int32_t v = nw[j];
      bool dup = false;
      for (int t = 0; t < kept; ++t) dup |= (nw[t] == v);
      if (!dup) {
        nw[kept++] = v;
        if (v >= 0 && v < C) mask[v >> 5] |= 1u << (v & 31);
      }
    }
    n_uniq += kept;
  }
  return (int)(pos - pos0);
}

__device__ __forceinline__ void gather_docs(const int32_t* cand, int64_t Q, int32_t C, int layout,
                                            int64_t q, int k, const int32_t* idx, int32_t* doc) {
  for (int j = 0; j < k; ++j) {
    const int32_t r = idx[j];
    doc[j] = (r < 0 || r >= C) ? -1
             : (layout == PN_LAYOUT_RANK) ? __ldg(cand + (int64_t)r * Q + q)
                                          : __ldg(cand + q * (int64_t)C + r);
  }
}

constexpr int kSelThreads = 128;

// stream_mode 0: one thread per query, independent uniform substreams.
__global__ void __launch_bounds__(kSelThreads)
select_binomial_kernel(const int32_t* __restrict__ cand, int64_t Q, int32_t C, int layout,
                       const double* __restrict__ table, int32_t k,
                       const double* __restrict__ uniforms, int64_t U, uint64_t seed,
                       int32_t* __restrict__ idx_out, int32_t* __restrict__ doc_out,
                       int32_t* __restrict__ consumed_out, int32_t* __restrict__ status) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* sp = reinterpret_cast<double*>(smem_raw);
  double* scdf = sp + C;
  const int mask_words = (C + 31) >> 5;
  uint32_t* smask = reinterpret_cast<uint32_t*>(scdf + C) + threadIdx.x * mask_words;
  for (int i = threadIdx.x; i < 2 * C; i += kSelThreads) sp[i] = table[i];
  __syncthreads();
  for (int64_t q = (int64_t)blockIdx.x * kSelThreads + threadIdx.x; q < Q;
       q += (int64_t)gridDim.x * kSelThreads) {
    UniformSrc src;
    src.u = uniforms ? uniforms + q * U : nullptr;
    src.avail = U; src.seed = seed; src.q = (uint64_t)q;
    int32_t* row = idx_out + q * k;
    const int used = choice_one(sp, scdf, C, k, src, 0, row, smask, mask_words);
    if (used < 0) {
      for (int j = 0; j < k; ++j) row[j] = -1;
      if (status) atomicAdd(status, 1);
    }
    if (consumed_out) consumed_out[q] = used;
    if (doc_out) gather_docs(cand, Q, C, layout, q, k, row, doc_out + q * k);
  }
}

// stream_mode 1: the reference's single global RNG stream, consumed by the queries in order.
__global__ void __launch_bounds__(32)
select_binomial_seq_kernel(const int32_t* __restrict__ cand, int64_t Q, int32_t C, int layout,
                           const double* __restrict__ table, int32_t k,
                           const double* __restrict__ uniforms, int64_t U, uint64_t seed,
                           int32_t* __restrict__ idx_out, int32_t* __restrict__ doc_out,
                           int32_t* __restrict__ consumed_out, int32_t* __restrict__ status) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* sp = reinterpret_cast<double*>(smem_raw);
  double* scdf = sp + C;
  const int mask_words = (C + 31) >> 5;
  uint32_t* smask = reinterpret_cast<uint32_t*>(scdf + C);
  for (int i = threadIdx.x; i < 2 * C; i += 32) sp[i] = table[i];
  __syncwarp();
  if (threadIdx.x != 0) return;
  int64_t pos = 0;
  bool dead = false;
  for (int64_t q = 0; q < Q; ++q) {
    UniformSrc src;
    src.u = uniforms; src.avail = U; src.seed = seed; src.q = 0;
    int32_t* row = idx_out + q * k;
    int used = dead ? -1 : choice_one(sp, scdf, C, k, src, pos, row, smask, mask_words);
    if (used < 0) {
      dead = true;  // the stream position is unknown from here on
      for (int j = 0; j < k; ++j) row[j] = -1;
      if (status) atomicAdd(status, 1);
    } else {
      pos += used;
    }
    if (consumed_out) consumed_out[q] = used;
    if (doc_out) gather_docs(cand, Q, C, layout, q, k, row, doc_out + q * k);
  }
}

}  // namespace

extern "C" int pn_select_level_scalar(const int32_t* cand, int64_t Q, int32_t C, int layout,
                                      double w, int use_ceil, int32_t* idx_out, int32_t* doc_out,
                                      int32_t* status_out, void* stream) {
  PN_CHECK_ARG(Q >= 0 && C >= 1);
  PN_CHECK_ARG(layout == PN_LAYOUT_ROW || layout == PN_LAYOUT_RANK);
  cudaStream_t st = (cudaStream_t)stream;
  if (status_out) PN_CUDA(cudaMemsetAsync(status_out, 0, sizeof(int32_t), st));
  PN_CHECK_ARG(Q == 0 || (cand && doc_out));
  // Python: self.round(weight * self.n_neg) with float64 operands (dataloader.py:32)
  const double x = w * (double)(C - 1);
  const double rr = use_ceil ? ceil(x) : floor(x);
  int32_t r = -1;
  if (rr >= 0.0 && rr < (double)C) r = (int32_t)rr;
  const int grid = pn_grid(Q > 0 ? Q : 1, kThreads, 4, 8);
  level_scalar_kernel<<<grid, kThreads, 0, st>>>(cand, Q, C, layout, r, idx_out, doc_out,
                                                 status_out);
  PN_LAUNCHED();
  return PN_OK;
}

extern "C" int pn_select_level_items(const int32_t* cand, int64_t Q, int32_t C, int layout,
                                     const float* w, int use_ceil, int32_t* idx_out,
                                     int32_t* doc_out, int32_t* status_out, void* stream) {
  PN_CHECK_ARG(Q >= 0 && C >= 1);
  PN_CHECK_ARG(layout == PN_LAYOUT_ROW || layout == PN_LAYOUT_RANK);
  cudaStream_t st = (cudaStream_t)stream;
  if (status_out) PN_CUDA(cudaMemsetAsync(status_out, 0, sizeof(int32_t), st));
  if (Q == 0) return PN_OK;
  PN_CHECK_ARG(cand && w && doc_out);
  level_items_kernel<<<pn_grid(Q, kThreads, 1, 8), kThreads, 0, st>>>(cand, Q, C, layout, w,
                                                                      use_ceil, idx_out, doc_out,
                                                                      status_out);
  PN_LAUNCHED();
  return PN_OK;
}

extern "C" int pn_binomial_table(double mean, double clip_min, double clip_max, double var,
                                 int32_t C, const double* pmf_in, double* table_out,
                                 int32_t* status_out, void* stream) {
  PN_CHECK_ARG(C >= 2 && C <= 8192);
  PN_CHECK_ARG(table_out && status_out);
  cudaStream_t st = (cudaStream_t)stream;
  // np.clip(mean, self.min, self.max)  (dataloader.py:135)
  double m = mean;
  if (m < clip_min) m = clip_min;
  if (m > clip_max) m = clip_max;
  double* scratch = nullptr;
  PN_CUDA(cudaMallocAsync((void**)&scratch, sizeof(double) * C, st));
  binomial_table_kernel<<<1, kThreads, 0, st>>>(m, var, C, pmf_in, table_out, scratch, status_out);
  g_pn_launches.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  cudaFreeAsync(scratch, st);
  if (e != cudaSuccess)
    return pn_fail(PN_E_LAUNCH, "pn_binomial_table: launch failed: %s", cudaGetErrorString(e));
  return PN_OK;
}

extern "C" int pn_select_binomial(const int32_t* cand, int64_t Q, int32_t C, int layout,
                                  const double* table, int32_t k, const double* uniforms, int64_t U,
                                  int stream_mode, uint64_t philox_seed, int32_t* idx_out,
                                  int32_t* doc_out, int32_t* consumed_out, int32_t* status_out,
                                  void* stream) {
  PN_CHECK_ARG(Q >= 0 && C >= 2 && C <= 8192);
  PN_CHECK_ARG(k >= 1 && k <= kMaxK && k <= C);
  PN_CHECK_ARG(layout == PN_LAYOUT_ROW || layout == PN_LAYOUT_RANK);
  PN_CHECK_ARG(stream_mode == 0 || stream_mode == 1);
  PN_CHECK_ARG(doc_out == nullptr || cand != nullptr);
  PN_CHECK_ARG(uniforms == nullptr || U >= k);
  cudaStream_t st = (cudaStream_t)stream;
  if (status_out) PN_CUDA(cudaMemsetAsync(status_out, 0, sizeof(int32_t), st));
  if (Q == 0) return PN_OK;
  PN_CHECK_ARG(table && idx_out);
  const int mask_words = (C + 31) >> 5;
  if (stream_mode == 0) {
    const size_t smem = sizeof(double) * 2 * C + sizeof(uint32_t) * mask_words * kSelThreads;
    PN_CUDA(cudaFuncSetAttribute(select_binomial_kernel,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = pn_grid(Q, kSelThreads, 1, 4);
    select_binomial_kernel<<<grid, kSelThreads, smem, st>>>(cand, Q, C, layout, table, k, uniforms,
                                                            U, philox_seed, idx_out, doc_out,
                                                            consumed_out, status_out);
  } else {
    const size_t smem = sizeof(double) * 2 * C + sizeof(uint32_t) * mask_words;
    PN_CUDA(cudaFuncSetAttribute(select_binomial_seq_kernel,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    select_binomial_seq_kernel<<<1, 32, smem, st>>>(cand, Q, C, layout, table, k, uniforms, U,
                                                    philox_seed, idx_out, doc_out, consumed_out,
                                                    status_out);
  }
  PN_LAUNCHED();
  return PN_OK;
}
