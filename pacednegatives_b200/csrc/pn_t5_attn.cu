// T5 encoder self-attention (HF modeling_t5.py:308-336): softmax(Q K^T + bias[h, j-i]) V, no 1/sqrt(d)
// scaling, additive learned relative-position bias, fp32 softmax.  Flash-style on tcgen05:
//   one CTA per (document, head); K and V (S x 64 bf16 each) stay in shared memory, the query rows are
//   walked in blocks of 128; per 128-key chunk  S = Q K^T  (UMMA 128x128x64, accumulator in TMEM),
//   online softmax by 128 threads (thread r owns query row r = TMEM lane r), P written to shared memory
//   as bf16 in the swizzled K-major layout, O_chunk = P V (UMMA 128x64x128, V consumed MN-major straight
//   from its row-major tile), O accumulated in registers.  The [N,h,S,S] score tensor never exists.
// Warp roles: warp 0 lane 0 = TMA + MMA issue, warps 4..7 = softmax (TMEM lane quarter = warp % 4).
#include "pn_common.cuh"
#include "pn_tc.cuh"
#include "pn_t5_internal.cuh"

using namespace pntc;

namespace {

constexpr int kAttnThreads = 256;
constexpr int DK = 64;            // d_kv of every T5 size
constexpr int QB = 128;           // query rows per block
constexpr int KC = 128;           // keys per chunk
constexpr int TILE_BYTES = 128 * DK * 2;  // one 128-row x 64-col bf16 tile = 16 KB
constexpr int MAX_S = 512;
constexpr float kLog2e = 1.4426950408889634f;

struct AttnArgs {
  int S;            // sequence length (<= 512)
  int n_heads;
  int ld_qkv;       // elements per row of the fused qkv matrix
  int q_col, k_col, v_col;  // first column of the q / k / v blocks
  const float* bias;        // [n_heads, 2*MAX_S-1] additive bias by (j - i + MAX_S - 1)
  __nv_bfloat16* out;       // [rows, ld_out]
  int ld_out;
};

// smem: K tiles [4][16KB] | V tiles [4][16KB] | Q [2][16KB] | P [2 sub-tiles][16KB] | bias 4KB | barriers
constexpr int SM_K = 0;
constexpr int SM_V = SM_K + 4 * TILE_BYTES;
constexpr int SM_Q = SM_V + 4 * TILE_BYTES;
constexpr int SM_P = SM_Q + 2 * TILE_BYTES;
constexpr int SM_BIAS = SM_P + 2 * TILE_BYTES;
constexpr int SM_BAR = SM_BIAS + (2 * MAX_S) * 4;
constexpr int ATTN_SMEM = SM_BAR + 256 + 1024;

// TMEM columns: S chunk buffers 0/1 at 0 and 128, O chunk buffers at 256 and 320
constexpr uint32_t TM_S = 0, TM_O = 256, TM_COLS = 512;

__global__ void __launch_bounds__(kAttnThreads, 1)
t5_self_attn_kernel(const __grid_constant__ CUtensorMap tmQKV, const AttnArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  float* sbias = reinterpret_cast<float*>(gbase + SM_BIAS);
  const uint32_t bar = base + SM_BAR;
  // barriers: kv_full, q_full[2], s_full[2], o_full[2], p_ready, q_empty[2]
  const uint32_t b_kv = bar, b_qfull = bar + 8, b_sfull = bar + 24, b_ofull = bar + 40,
                 b_pready = bar + 56, b_qempty = bar + 64;
  const uint32_t tmem_slot = bar + 96;
  uint32_t* tmem_slot_gen = reinterpret_cast<uint32_t*>(gbase + SM_BAR + 96);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int doc = blockIdx.x / a.n_heads, head = blockIdx.x % a.n_heads;
  const int S = a.S;
  const int n_qb = (S + QB - 1) / QB, n_kc = (S + KC - 1) / KC;
  const int steps = n_qb * n_kc;
  const int row0 = doc * S;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmQKV);
    mbar_init(b_kv, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(b_qfull + 8 * i, 1);
      mbar_init(b_sfull + 8 * i, 1);
      mbar_init(b_ofull + 8 * i, 1);
      mbar_init(b_qempty + 8 * i, 1);
    }
    mbar_init(b_pready, 4);  // one arrive per softmax warp
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc<1>(tmem_slot, TM_COLS);
    tmem_relinquish<1>();
  }
  // relative-position bias row of this head: index (j - i) + MAX_S - 1
  for (int i = threadIdx.x; i < 2 * MAX_S - 1; i += kAttnThreads)
    sbias[i] = a.bias[(size_t)head * (2 * MAX_S - 1) + i];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot_gen;

  if (warp == 0 && lane == 0) {
    // ============ TMA loads + MMA issue (one thread) ============
    mbar_arrive_expect_tx(b_kv, 2 * n_kc * TILE_BYTES);
    for (int c = 0; c < n_kc; ++c) {
      tma_load_2d(base + SM_K + c * TILE_BYTES, &tmQKV, b_kv, a.k_col + head * DK, row0 + c * KC);
      tma_load_2d(base + SM_V + c * TILE_BYTES, &tmQKV, b_kv, a.v_col + head * DK, row0 + c * KC);
    }
    auto load_q = [&](int qb) {
      const int buf = qb & 1;
      mbar_wait(b_qempty + 8 * buf, ((qb >> 1) & 1) ^ 1);
      mbar_arrive_expect_tx(b_qfull + 8 * buf, TILE_BYTES);
      tma_load_2d(base + SM_Q + buf * TILE_BYTES, &tmQKV, b_qfull + 8 * buf, a.q_col + head * DK,
                  row0 + qb * QB);
    };
    load_q(0);
    if (n_qb > 1) load_q(1);
    mbar_wait(b_kv, 0);

    constexpr uint32_t IDESC_QK = make_idesc_bf16(128, KC, 0, 0);  // A=Q K-major, B=K K-major
    constexpr uint32_t IDESC_PV = make_idesc_bf16(128, DK, 0, 1);  // A=P K-major, B=V MN-major
    auto issue_qk = [&](int t) {
      const int qb = t / n_kc, c = t % n_kc, buf = qb & 1;
      if (c == 0) {
        mbar_wait(b_qfull + 8 * buf, (qb >> 1) & 1);
        tc_fence_after();
      }
      const uint64_t adesc = make_desc_kmajor_sw128(base + SM_Q + buf * TILE_BYTES);
      const uint64_t bdesc = make_desc_kmajor_sw128(base + SM_K + c * TILE_BYTES);
#pragma unroll
      for (int k = 0; k < DK / 16; ++k)
        umma_bf16<1>(tmem + TM_S + (t & 1) * KC, adesc + 2u * k, bdesc + 2u * k, IDESC_QK, k ? 1u : 0u);
      umma_commit<1>(b_sfull + 8 * (t & 1));
      if (c == n_kc - 1) {
        umma_commit<1>(b_qempty + 8 * buf);        // Q buffer reusable once these MMAs retire
        if (qb + 2 < n_qb) load_q(qb + 2);
      }
    };
    issue_qk(0);
    if (steps > 1) issue_qk(1);
    for (int u = 0; u < steps; ++u) {
      mbar_wait(b_pready, u & 1);
      tc_fence_after();
      const int c = u % n_kc;
      // O_chunk[128 x 64] = P[128 x 128] . V_chunk[128 x 64]
      const uint32_t pbase = base + SM_P;
      const uint32_t vbase = base + SM_V + c * TILE_BYTES;
#pragma unroll
      for (int k = 0; k < KC / 16; ++k) {
        const uint64_t adesc = make_desc_kmajor_sw128(pbase + (k >> 2) * TILE_BYTES) + 2u * (k & 3);
        const uint64_t bdesc = make_desc_mnmajor_sw128(vbase + k * 16 * 128, 1024);
        umma_bf16<1>(tmem + TM_O + (u & 1) * DK, adesc, bdesc, IDESC_PV, k ? 1u : 0u);
      }
      umma_commit<1>(b_ofull + 8 * (u & 1));
      if (u + 2 < steps) issue_qk(u + 2);
    }
  } else if (warp >= 4) {
    // ============ softmax: thread r owns query row r of the block ============
    const int r = threadIdx.x - 128;          // 0..127 = TMEM lane
    const uint32_t lane_addr = (uint32_t)((warp - 4) * 32) << 16;
    float o[DK];
    float m = -INFINITY, l = 0.0f;
#pragma unroll
    for (int d = 0; d < DK; ++d) o[d] = 0.0f;

    auto add_o = [&](int u) {  // o += O_chunk(u)
      mbar_wait(b_ofull + 8 * (u & 1), (u >> 1) & 1);
      tc_fence_after();
#pragma unroll
      for (int h2 = 0; h2 < 2; ++h2) {
        uint32_t v[32];
        tmem_ld_32x32(tmem + lane_addr + TM_O + (u & 1) * DK + h2 * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int d = 0; d < 32; ++d) o[h2 * 32 + d] += __uint_as_float(v[d]);
      }
    };
    auto finalize = [&](int qb) {  // ctx[row, head*64 ..] = o / l
      const int qi = qb * QB + r;
      if (qi < S) {
        const float inv = 1.0f / l;
        __nv_bfloat16* dst = a.out + (size_t)(row0 + qi) * a.ld_out + head * DK;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          uint32_t pk[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            __nv_bfloat162 h = __floats2bfloat162_rn(o[j * 8 + 2 * e] * inv, o[j * 8 + 2 * e + 1] * inv);
            pk[e] = *reinterpret_cast<uint32_t*>(&h);
          }
          *reinterpret_cast<uint4*>(dst + j * 8) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        }
      }
      m = -INFINITY; l = 0.0f;
#pragma unroll
      for (int d = 0; d < DK; ++d) o[d] = 0.0f;
    };

    for (int u = 0; u < steps; ++u) {
      const int qb = u / n_kc, c = u % n_kc;
      const int qi = qb * QB + r;                 // query position of this thread
      const int key0 = c * KC;
      const float* brow = sbias + (key0 - qi + MAX_S - 1);  // + jj
      const uint32_t s_addr = tmem + lane_addr + TM_S + (u & 1) * KC;
      mbar_wait(b_sfull + 8 * (u & 1), (u >> 1) & 1);
      tc_fence_after();
      // pass 1: chunk max of (s + bias), keys beyond S masked out
      float cmax = -INFINITY;
#pragma unroll 1
      for (int g = 0; g < KC / 32; ++g) {
        uint32_t v[32];
        tmem_ld_32x32(s_addr + g * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int jj = 0; jj < 32; ++jj) {
          const int j = g * 32 + jj;
          const float s = __uint_as_float(v[jj]) + brow[j];
          if (key0 + j < S) cmax = fmaxf(cmax, s);
        }
      }
      if (u > 0) {
        add_o(u - 1);
        if (c == 0) finalize(qb - 1);
      }
      const float m_new = fmaxf(m, cmax);
      const float alpha = exp2f((m - m_new) * kLog2e);   // m = -inf on a fresh block -> 0
      l *= alpha;
#pragma unroll
      for (int d = 0; d < DK; ++d) o[d] *= alpha;
      m = m_new;
      const float mb = m * kLog2e;
      // pass 2: p = exp(s - m), row sum, P -> smem (bf16, 128B-swizzled K-major, two 64-key sub-tiles)
      float lsum = 0.0f;
#pragma unroll 1
      for (int g = 0; g < KC / 32; ++g) {
        uint32_t v[32];
        tmem_ld_32x32(s_addr + g * 32, v);
        tmem_ld_wait();
        uint32_t pk[16];
#pragma unroll
        for (int jj = 0; jj < 32; jj += 2) {
          const int j = g * 32 + jj;
          float p0 = exp2f((__uint_as_float(v[jj]) + brow[j]) * kLog2e - mb);
          float p1 = exp2f((__uint_as_float(v[jj + 1]) + brow[j + 1]) * kLog2e - mb);
          if (key0 + j >= S) p0 = 0.0f;
          if (key0 + j + 1 >= S) p1 = 0.0f;
          lsum += p0 + p1;
          __nv_bfloat162 h = __floats2bfloat162_rn(p0, p1);
          pk[jj >> 1] = *reinterpret_cast<uint32_t*>(&h);
        }
        // 32 keys = 4 chunks of 16 B; key j lives in sub-tile j/64, 16B-chunk ((j%64)/8) ^ (r%8)
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          const int j = g * 32 + q4 * 8;
          const uint32_t addr = base + SM_P + (j >> 6) * TILE_BYTES + r * 128 +
                                ((((j & 63) >> 3) ^ (r & 7)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(pk[q4 * 4]),
                       "r"(pk[q4 * 4 + 1]), "r"(pk[q4 * 4 + 2]), "r"(pk[q4 * 4 + 3])
                       : "memory");
        }
      }
      l += lsum;
      fence_proxy_async();   // generic-proxy smem writes -> visible to the tensor core's async proxy
      tc_fence_before();     // our TMEM reads of S(u) are ordered before the MMA that reuses the buffer
      __syncwarp();
      if (lane == 0) mbar_arrive(b_pready);
    }
    add_o(steps - 1);
    finalize(n_qb - 1);
  }

  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<1>(tmem, TM_COLS);
}

// ---------------------------------------------------------------------------------------------
// Decoder attention for T <= 2 query tokens per document (one decoder step of monoT5, or the two
// label positions of the LCE path).  Memory-bound CUDA-core kernel: one warp per (doc, head, t).
//   self-attention (causal, own bias table)  : keys = the <= T decoder tokens
//   cross-attention (no bias, HF:318-323)    : keys = the S encoder states' K/V projections
struct DecAttnArgs {
  const __nv_bfloat16* q; int ld_q;       // [docs*T, ld_q], head h at column q_col + h*64
  int q_col;
  const __nv_bfloat16* k; const __nv_bfloat16* v; int ld_kv;  // rows: doc*Sk + j
  int k_col, v_col;
  int Sk;                 // keys per doc
  int T;                  // query tokens per doc
  int causal;             // 1: key j visible to query t iff j <= t
  const float* bias;      // nullable [n_heads, T, Sk] additive
  __nv_bfloat16* out; int ld_out;         // [docs*T, ld_out]
  int n_heads; int docs;
  const int32_t* klen;                     // optional [docs]: cross-attention key-padding mask
  PnDrop drop;                             // dropout of the attention probabilities (training), index [doc,h,t,j]
};

__global__ void __launch_bounds__(128)
dec_attn_kernel(const DecAttnArgs a) {
  __shared__ float sp[4][MAX_S];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t gw = (int64_t)blockIdx.x * 4 + w;
  const int64_t total = (int64_t)a.docs * a.n_heads * a.T;
  if (gw >= total) return;
  const int t = (int)(gw % a.T);
  const int head = (int)((gw / a.T) % a.n_heads);
  const int doc = (int)(gw / ((int64_t)a.T * a.n_heads));
  const __nv_bfloat16* q = a.q + ((size_t)doc * a.T + t) * a.ld_q + a.q_col + head * DK;
  float qf[DK];
#pragma unroll
  for (int d = 0; d < DK; d += 2) {
    const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(q + d));
    qf[d] = f.x; qf[d + 1] = f.y;
  }
  const int nk = a.causal ? (t + 1) : (a.klen ? min(a.Sk, max(1, a.klen[doc])) : a.Sk);
  float mx = -INFINITY;
  for (int j = lane; j < nk; j += 32) {
    const __nv_bfloat16* kr = a.k + ((size_t)doc * a.Sk + j) * a.ld_kv + a.k_col + head * DK;
    float s = 0.0f;
#pragma unroll
    for (int d = 0; d < DK; d += 8) {
      const uint4 raw = *reinterpret_cast<const uint4*>(kr + d);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = __bfloat1622float2(h[e]);
        s += qf[d + 2 * e] * f.x + qf[d + 2 * e + 1] * f.y;
      }
    }
    if (a.bias) s += a.bias[((size_t)head * a.T + t) * a.Sk + j];
    sp[w][j] = s;
    mx = fmaxf(mx, s);
  }
  mx = pn_warp_max(mx);
  float sum = 0.0f;
  const int skp = a.Sk + (a.Sk & 1);
  const unsigned long long dbase = (unsigned long long)a.drop.idx_offset +
                                   ((unsigned long long)(doc * a.n_heads + head) * a.T + t) * skp;
  for (int j = lane; j < nk; j += 32) {
    const float p = __expf(sp[w][j] - mx);
    sp[w][j] = p * pn_drop_mul(a.drop, dbase + j);   // the row sum stays that of the un-dropped softmax
    sum += p;
  }
  sum = pn_warp_sum(sum);
  __syncwarp();
  // out[d] = sum_j p_j V[j][d]; lane owns dims 2*lane, 2*lane+1
  float o0 = 0.0f, o1 = 0.0f;
  for (int j = 0; j < nk; ++j) {
    const __nv_bfloat16* vr = a.v + ((size_t)doc * a.Sk + j) * a.ld_kv + a.v_col + head * DK;
    const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(vr + 2 * lane));
    const float p = sp[w][j];
    o0 += p * f.x; o1 += p * f.y;
  }
  const float inv = 1.0f / sum;
  __nv_bfloat16* dst = a.out + ((size_t)doc * a.T + t) * a.ld_out + head * DK + 2 * lane;
  *reinterpret_cast<__nv_bfloat162*>(dst) = __floats2bfloat162_rn(o0 * inv, o1 * inv);
}

}  // namespace

int pn_t5_self_attn_v1(const __nv_bfloat16* qkv, int ld_qkv, int q_col, int k_col, int v_col, int docs,
                       int S, int n_heads, const float* bias, __nv_bfloat16* out, int ld_out,
                       cudaStream_t st) {
  if (S < 1 || S > MAX_S) return pn_fail(PN_E_UNSUPPORTED, "self-attention: S=%d not in [1,512]", S);
  CUtensorMap tm;
  int rc = pn_make_tmap_bf16_2d(&tm, qkv, (uint64_t)docs * S, (uint64_t)ld_qkv, ld_qkv, 128, DK);
  if (rc) return rc;
  AttnArgs a;
  a.S = S; a.n_heads = n_heads; a.ld_qkv = ld_qkv; a.q_col = q_col; a.k_col = k_col; a.v_col = v_col;
  a.bias = bias; a.out = out; a.ld_out = ld_out;
  static bool attr_done = false;
  if (!attr_done) {
    PN_CUDA(cudaFuncSetAttribute(t5_self_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 ATTN_SMEM));
    attr_done = true;
  }
  PnProfScope prof(PN_PROF_ATTN, 4.0 * docs * n_heads * (double)S * S * DK, st);
  t5_self_attn_kernel<<<docs * n_heads, kAttnThreads, ATTN_SMEM, st>>>(tm, a);
  PN_LAUNCHED();
  return PN_OK;
}

int pn_t5_dec_attn(const __nv_bfloat16* q, int ld_q, int q_col, const __nv_bfloat16* k,
                   const __nv_bfloat16* v, int ld_kv, int k_col, int v_col, int docs, int T, int Sk,
                   int n_heads, int causal, const float* bias, __nv_bfloat16* out, int ld_out,
                   cudaStream_t st, const int32_t* klen, PnDrop drop) {
  if (Sk < 1 || Sk > MAX_S) return pn_fail(PN_E_UNSUPPORTED, "decoder attention: Sk=%d not in [1,512]", Sk);
  DecAttnArgs a;
  a.q = q; a.ld_q = ld_q; a.q_col = q_col; a.k = k; a.v = v; a.ld_kv = ld_kv; a.k_col = k_col;
  a.v_col = v_col; a.Sk = Sk; a.T = T; a.causal = causal; a.bias = bias; a.out = out;
  a.ld_out = ld_out; a.n_heads = n_heads; a.docs = docs;
  a.klen = klen;
  a.drop = drop;
  const int64_t warps = (int64_t)docs * n_heads * T;
  dec_attn_kernel<<<(unsigned)((warps + 3) / 4), 128, 0, st>>>(a);
  PN_LAUNCHED();
  return PN_OK;
}
