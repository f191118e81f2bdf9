// Per-query ranking + trec_eval measures (eval.py:26,41-43 -> pt.Experiment with
// eval_metrics=["map","ndcg_cut_10","recip_rank"]).  PyTerrier / pytrec_eval are third-party and not
// vendored in the reference, so this follows trec_eval's published definitions:
//   ranking   : score desc, ties by docno desc (trec_eval sorts docno as a string; ids here are int32)
//   ndcg_cut_K: gain = rel (linear), discount log2(rank+1), ideal = the query's rels sorted desc
//   recip_rank: 1 / rank of the first doc with rel >= 1 (whole list); rr@K the same cut at K
//   map       : mean over relevant docs of precision at their rank
// One CTA per query: C candidates (<= 4096) are packed to 64-bit keys in shared memory and sorted with
// a bitonic network; HBM traffic is the C*(4+4+1) input bytes once.
#include "pn_common.cuh"

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ uint32_t orderable(float f) {
  // -0.0f ties with +0.0f (trec_eval compares the floats), so both map to the key of +0.0f
  const uint32_t u = (f == 0.0f) ? 0u : __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float unorderable(uint32_t o) {
  const uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
  return __uint_as_float(u);
}

__global__ void __launch_bounds__(kThreads)
topk_metrics_kernel(const float* __restrict__ scores, const int32_t* __restrict__ docids,
                    const uint8_t* __restrict__ rels, int64_t Q, int32_t C, int32_t P, int32_t K,
                    int32_t* __restrict__ top_ids, float* __restrict__ top_scores,
                    float* __restrict__ metrics) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t* key = reinterpret_cast<uint64_t*>(smem_raw);       // [P]
  uint16_t* pos = reinterpret_cast<uint16_t*>(key + P);        // [P]
  uint8_t* srel = reinterpret_cast<uint8_t*>(pos + P);         // [P] rel by original position
  __shared__ int hist[256];
  __shared__ double s_part[kThreads / 32][2];
  __shared__ int s_first[kThreads / 32];

  for (int64_t q = blockIdx.x; q < Q; q += gridDim.x) {
    const float* sc = scores + q * (int64_t)C;
    const int32_t* ids = docids ? docids + q * (int64_t)C : nullptr;
    const uint8_t* rl = rels ? rels + q * (int64_t)C : nullptr;
    if (threadIdx.x < 256) hist[threadIdx.x] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < P; i += kThreads) {
      if (i < C) {
        const int32_t id = ids ? ids[i] : i;
        key[i] = ((uint64_t)orderable(sc[i]) << 32) | (uint32_t)(id ^ 0x80000000);
        pos[i] = (uint16_t)i;
        const uint8_t r = rl ? rl[i] : 0;
        srel[i] = r;
        if (r) atomicAdd(&hist[r], 1);
      } else {
        key[i] = 0ull;  // below every real key (real keys have a non-zero high word)
        pos[i] = 0xFFFFu;
        srel[i] = 0;
      }
    }
    __syncthreads();
    // bitonic sort, descending
    for (int size = 2; size <= P; size <<= 1) {
      for (int stride = size >> 1; stride > 0; stride >>= 1) {
        for (int t = threadIdx.x; t < (P >> 1); t += kThreads) {
          const int lo = 2 * t - (t & (stride - 1));
          const int hi = lo + stride;
          const bool desc = ((lo & size) == 0);
          const uint64_t a = key[lo], b = key[hi];
          if ((a < b) == desc) {
            key[lo] = b; key[hi] = a;
            const uint16_t pa = pos[lo]; pos[lo] = pos[hi]; pos[hi] = pa;
          }
        }
        __syncthreads();
      }
    }
    // top-K
    for (int i = threadIdx.x; i < K; i += kThreads) {
      const bool live = i < C;
      if (top_ids) top_ids[q * K + i] = live ? (int32_t)((uint32_t)key[i] ^ 0x80000000u) : -1;
      if (top_scores) top_scores[q * K + i] = live ? unorderable((uint32_t)(key[i] >> 32)) : 0.0f;
    }
    if (metrics) {
      // DCG@K and first relevant rank: parallel over ranks; AP needs a running count -> warp scan
      const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
      double dcg = 0.0;
      int first = 0x7fffffff;
      for (int i = threadIdx.x; i < C; i += kThreads) {
        const int r = srel[pos[i]];
        if (r) {
          if (i < K) dcg += (double)r / log2((double)(i + 2));
          first = min(first, i);
        }
      }
      dcg = pn_warp_sum(dcg);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) first = min(first, __shfl_xor_sync(0xffffffffu, first, o));
      if (lane == 0) { s_part[warp][0] = dcg; s_first[warp] = first; }
      __syncthreads();
      if (warp == 0) {
        // average precision: ranks in chunks of 32, ballot prefix
        int seen = 0;
        double ap = 0.0;
        for (int base = 0; base < C; base += 32) {
          const int i = base + lane;
          const bool rel = (i < C) && srel[pos[i]] > 0;
          const unsigned b = __ballot_sync(0xffffffffu, rel);
          if (rel) {
            const int upto = seen + __popc(b & (0xffffffffu >> (31 - lane)));
            ap += (double)upto / (double)(i + 1);
          }
          seen += __popc(b);
        }
        ap = pn_warp_sum(ap);
        if (lane == 0) {
          double d = 0.0;
          int f = 0x7fffffff;
          for (int w = 0; w < kThreads / 32; ++w) { d += s_part[w][0]; f = min(f, s_first[w]); }
          // ideal DCG@K from the histogram of grades, high to low
          double idcg = 0.0;
          int slot = 0, nrel = 0;
          for (int g = 255; g >= 1; --g) {
            const int c = hist[g];
            nrel += c;
            for (int j = 0; j < c && slot < K; ++j, ++slot) idcg += (double)g / log2((double)(slot + 2));
          }
          float* m = metrics + q * 4;
          m[0] = idcg > 0.0 ? (float)(d / idcg) : 0.0f;
          m[1] = f < C ? (float)(1.0 / (double)(f + 1)) : 0.0f;
          m[2] = f < K ? (float)(1.0 / (double)(f + 1)) : 0.0f;
          m[3] = nrel > 0 ? (float)(ap / (double)nrel) : 0.0f;
        }
      }
    }
    __syncthreads();
  }
}

}  // namespace

extern "C" int pn_topk_metrics(const float* scores, const int32_t* docids, const uint8_t* rels,
                               int64_t Q, int32_t C, int32_t K, int32_t* top_ids, float* top_scores,
                               float* metrics_out, void* stream) {
  PN_CHECK_ARG(Q >= 0 && C >= 1 && C <= 4096 && K >= 1);
  if (Q == 0) return PN_OK;
  PN_CHECK_ARG(scores != nullptr);
  int P = 2;
  while (P < C) P <<= 1;
  const size_t smem = (size_t)P * (8 + 2 + 1);
  PN_CUDA(cudaFuncSetAttribute(topk_metrics_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               (int)smem));
  const int grid = (int)(Q < 148 * 4 ? Q : 148 * 4);
  topk_metrics_kernel<<<grid, kThreads, smem, (cudaStream_t)stream>>>(
      scores, docids, rels, Q, C, P, K, top_ids, top_scores, metrics_out);
  PN_LAUNCHED();
  return PN_OK;
}
