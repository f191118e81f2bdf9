// Counter-based dropout masks for the training pass (HF T5 in train mode, as Lightning runs the reference's
// LCEModel, lceT5.py:131-133,149-150: modeling_t5.py applies nn.Dropout(dropout_rate) to the embedding output, every
// sub-layer output before its residual add, the FFN hidden activation, the attention probabilities and the two
// final norms; dropout_rate 0.1 in t5-* / monot5-* configs).
//
// A mask is a pure function of (seed of the forward call, site, element index), so the backward pass REGENERATES it
// instead of storing it.  Element idx keeps its value iff the 16-bit half (idx & 1) of
//     h = mix32(lo32(idx >> 1) ^ key_lo ^ mix32(hi32(idx >> 1) ^ key_hi))
// is below thresh = round((1 - p) * 65536); kept values are multiplied by 1 / (1 - p) like torch.nn.functional.dropout.
// key = splitmix64(seed ^ (site + 1) * golden).  `site` names the dropout call (pn_drop_site below) and idx is the flat
// index of the element in the tensor HF passes to that call, document axis first (documents are numbered across the
// whole step: doc_offset when a batch is walked in pieces); attention-probability rows use an even pitch
// (S + (S & 1)).  tests/golden/make_golden.py restates this in numpy and replays the same masks inside HF itself.
#pragma once
#include <stdint.h>

struct PnDrop {
  uint32_t key_lo, key_hi;
  uint32_t thresh;       // keep iff 16 random bits < thresh; 0 = dropout off (every element kept, scale 1)
  float scale;           // 1 / (1 - p)
  long long idx_offset;  // added to every element index (doc_offset * elements per document at this site)
};

// site numbering: stack (0 encoder, 1 decoder) * 1000 + layer * 10 + kind
enum { PN_DROP_EMBED = 900, PN_DROP_FINAL = 901,            // + stack * 1000
       PN_DROP_SELF_P = 0, PN_DROP_SELF_OUT = 1, PN_DROP_CROSS_P = 2, PN_DROP_CROSS_OUT = 3, PN_DROP_FF_HID = 4,
       PN_DROP_FF_OUT = 5 };
static inline uint32_t pn_drop_site(int stack, int layer, int kind) {
  return (uint32_t)(stack * 1000 + (kind >= 900 ? kind : layer * 10 + kind));
}

static inline uint64_t pn_splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

// p <= 0 gives the disabled mask; doc_offset * elems_per_doc positions the call inside the whole step's batch
static inline PnDrop pn_drop_make(uint64_t seed, uint32_t site, float p, long long doc_offset, long long elems_per_doc) {
  PnDrop d;
  const uint64_t k = pn_splitmix64(seed ^ ((uint64_t)(site + 1) * 0x9E3779B97F4A7C15ull));
  d.key_lo = (uint32_t)k;
  d.key_hi = (uint32_t)(k >> 32);
  d.thresh = p > 0.0f ? (uint32_t)((1.0 - (double)p) * 65536.0 + 0.5) : 0u;
  d.scale = p > 0.0f ? 1.0f / (1.0f - p) : 1.0f;
  d.idx_offset = doc_offset * elems_per_doc;
  return d;
}
static inline PnDrop pn_drop_off() { return pn_drop_make(0, 0, 0.0f, 0, 0); }

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t pn_mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du;
  x ^= x >> 15; x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}
// 32 random bits shared by elements 2*pair and 2*pair + 1 (indices already include idx_offset)
__device__ __forceinline__ uint32_t pn_drop_bits(const PnDrop& d, unsigned long long pair) {
  const uint32_t hh = pn_mix32((uint32_t)(pair >> 32) ^ d.key_hi);
  return pn_mix32((uint32_t)pair ^ d.key_lo ^ hh);
}
// multiplier (0 or scale) of element idx
__device__ __forceinline__ float pn_drop_mul(const PnDrop& d, unsigned long long idx) {
  if (!d.thresh) return 1.0f;
  const uint32_t h = pn_drop_bits(d, idx >> 1);
  const uint32_t u = (idx & 1ull) ? (h >> 16) : (h & 0xffffu);
  return u < d.thresh ? d.scale : 0.0f;
}
// multipliers of the aligned pair (idx_even, idx_even + 1); the caller has checked d.thresh != 0
__device__ __forceinline__ void pn_drop_mul2(const PnDrop& d, unsigned long long idx_even, float& m0, float& m1) {
  const uint32_t h = pn_drop_bits(d, idx_even >> 1);
  m0 = (h & 0xffffu) < d.thresh ? d.scale : 0.0f;
  m1 = (h >> 16) < d.thresh ? d.scale : 0.0f;
}
// Row form for kernels that walk a run of consecutive elements (an attention row chunk): when the run's pair indices
// [pair0, pair0 + n_pairs) share their high 32 bits — always, except for one run in 2^32 pairs — the inner hash of
// pn_drop_bits is the same for the whole run and is folded into the key once.  Same bits as pn_drop_mul2.
struct PnDropRun {
  uint32_t k;        // key_lo ^ mix32(hi32(pair0) ^ key_hi)
  uint32_t lo0;      // lo32(pair0)
  bool fast;         // false: the run crosses a 2^32 pair boundary, use pn_drop_mul2 on the full index
};
__device__ __forceinline__ PnDropRun pn_drop_run(const PnDrop& d, unsigned long long idx_even0, uint32_t n_pairs) {
  const unsigned long long pair0 = idx_even0 >> 1;
  PnDropRun r;
  r.lo0 = (uint32_t)pair0;
  r.k = d.key_lo ^ pn_mix32((uint32_t)(pair0 >> 32) ^ d.key_hi);
  r.fast = r.lo0 <= 0xffffffffu - n_pairs;
  return r;
}
// multipliers of elements (idx_even0 + 2 t, idx_even0 + 2 t + 1) of a fast run
__device__ __forceinline__ void pn_drop_run_mul2(const PnDrop& d, const PnDropRun& r, uint32_t t, float& m0, float& m1) {
  const uint32_t h = pn_mix32((r.lo0 + t) ^ r.k);
  m0 = (h & 0xffffu) < d.thresh ? d.scale : 0.0f;
  m1 = (h >> 16) < d.thresh ? d.scale : 0.0f;
}
// run form with the rare slow path folded in: multipliers of elements (idx_even0 + 2 t, + 1)
__device__ __forceinline__ void pn_drop_run_get(const PnDrop& d, const PnDropRun& r, unsigned long long idx_even0,
                                                uint32_t t, float& m0, float& m1) {
  if (r.fast) pn_drop_run_mul2(d, r, t, m0, m1);
  else pn_drop_mul2(d, idx_even0 + 2ull * t, m0, m1);
}
#endif
