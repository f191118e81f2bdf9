// On-device prompt packing: the monoT5 prompt
//   'Query: ' + q + ' Document: ' + d + ' Relevant:'            (dataloader.py:43-44, 66-67, 153-155)
// assembled from pre-tokenised pieces that already live in HBM — the three literal pieces, a CSR store of query
// tokens and a CSR store of document tokens — for the (query, document) pairs the selection kernels picked, written
// straight into the int32 [N, S] id matrix the scorer reads, padded to S like the tokenizer's padding=True
// (lceT5.py:46-47, pairwrapper.py:79-80).  HBM-bound: one warp per prompt row, lanes over consecutive positions
// (128-byte coalesced stores; the CSR segments are contiguous, so the loads coalesce too).
#include "pn_common.cuh"

namespace {

constexpr int kThreads = 256;

struct Pieces {
  const int32_t* tok[3];   // prefix 'Query:', middle 'Document:', suffix 'Relevant:' + </s>
  int32_t len[3];
};

// row length after the optional document truncation: the document loses its tail so that the prompt still ends in
// ' Relevant:</s>' (max_len = 0: no truncation, the reference's behaviour)
__device__ __forceinline__ void prompt_layout(const Pieces& P, int lq, int ld, int max_len, int* ld_used, int* total) {
  const int fixed = P.len[0] + P.len[1] + P.len[2] + lq;
  int d = ld;
  if (max_len > 0 && fixed + d > max_len) d = max_len > fixed ? max_len - fixed : 0;
  *ld_used = d;
  *total = fixed + d;
}

__global__ void __launch_bounds__(kThreads)
pack_prompts_kernel(Pieces P, const int32_t* __restrict__ q_tok, const int64_t* __restrict__ q_off,
                    const int32_t* __restrict__ d_tok, const int64_t* __restrict__ d_off,
                    const int32_t* __restrict__ q_idx, const int32_t* __restrict__ d_idx, int64_t n, int64_t nq,
                    int64_t nd, int S, int max_len, int pad_id, int32_t* __restrict__ ids_out,
                    int32_t* __restrict__ len_out, int32_t* __restrict__ stat /* [max length, #bad rows] */) {
  const int lane = threadIdx.x & 31;
  const int64_t wid = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t)gridDim.x * kThreads) >> 5;
  int wmax = 0, wbad = 0;
  for (int64_t r = wid; r < n; r += nw) {
    const int64_t qi = q_idx[r], di = d_idx[r];
    const bool ok = qi >= 0 && qi < nq && di >= 0 && di < nd;
    int64_t qs = 0, ds = 0;
    int lq = 0, ld = 0;
    if (ok) {
      qs = q_off[qi]; lq = (int)(q_off[qi + 1] - qs);
      ds = d_off[di]; ld = (int)(d_off[di + 1] - ds);
    }
    int ldu, total;
    prompt_layout(P, lq, ld, max_len, &ldu, &total);
    const bool fits = ok && (ids_out == nullptr || total <= S);
    if (lane == 0) {
      if (len_out) len_out[r] = ok ? total : -1;
      wmax = max(wmax, ok ? total : 0);
      wbad += fits ? 0 : 1;
    }
    if (ids_out == nullptr) continue;
    const int b0 = P.len[0], b1 = b0 + lq, b2 = b1 + P.len[1], b3 = b2 + ldu;
    int32_t* row = ids_out + r * (int64_t)S;
    for (int c = lane; c < S; c += 32) {
      int32_t v = pad_id;
      if (fits && c < total) {
        if (c < b0) v = __ldg(P.tok[0] + c);
        else if (c < b1) v = __ldg(q_tok + qs + (c - b0));
        else if (c < b2) v = __ldg(P.tok[1] + (c - b1));
        else if (c < b3) v = __ldg(d_tok + ds + (c - b2));
        else v = __ldg(P.tok[2] + (c - b3));
      }
      row[c] = v;
    }
  }
  if (lane == 0 && stat) {
    if (wmax) atomicMax(stat, wmax);
    if (wbad) atomicAdd(stat + 1, wbad);
  }
}

}  // namespace

extern "C" int pn_pack_prompts(const int32_t* prefix, int32_t n_prefix, const int32_t* middle, int32_t n_middle,
                               const int32_t* suffix, int32_t n_suffix, const int32_t* q_tok, const int64_t* q_off,
                               int64_t n_queries, const int32_t* d_tok, const int64_t* d_off, int64_t n_docs,
                               const int32_t* q_idx, const int32_t* d_idx, int64_t n, int32_t S, int32_t max_len,
                               int32_t pad_id, int32_t* ids_out, int32_t* len_out, int32_t* status_out,
                               void* stream) {
  PN_CHECK_ARG(n >= 0 && n_prefix >= 0 && n_middle >= 0 && n_suffix >= 0 && n_queries >= 0 && n_docs >= 0);
  PN_CHECK_ARG(max_len >= 0 && status_out);
  PN_CHECK_ARG(ids_out == nullptr || S >= 1);
  cudaStream_t st = (cudaStream_t)stream;
  PN_CUDA(cudaMemsetAsync(status_out, 0, 2 * sizeof(int32_t), st));
  if (n == 0) return PN_OK;
  PN_CHECK_ARG(q_off && d_off && q_idx && d_idx);
  PN_CHECK_ARG((prefix || !n_prefix) && (middle || !n_middle) && (suffix || !n_suffix));
  Pieces P;
  P.tok[0] = prefix; P.tok[1] = middle; P.tok[2] = suffix;
  P.len[0] = n_prefix; P.len[1] = n_middle; P.len[2] = n_suffix;
  // one warp per row: 8 rows per CTA
  pack_prompts_kernel<<<pn_grid(n, kThreads / 32, 1, 8), kThreads, 0, st>>>(
      P, q_tok, q_off, d_tok, d_off, q_idx, d_idx, n, n_queries, n_docs, S, max_len, pad_id, ids_out, len_out,
      status_out);
  PN_LAUNCHED();
  return PN_OK;
}
