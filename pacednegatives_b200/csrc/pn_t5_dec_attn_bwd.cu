// Backward of the decoder attentions (pn_t5_dec_attn; autograd of HF modeling_t5.py:308-336 for the decoder's
// self-attention over T <= 2 label tokens and its cross-attention over the S encoder states).  With at most two
// query rows per (document, head) this is bandwidth work over K, V and their gradients, not MMA work:
//   p = softmax(q K^T + bias),  dp = do V^T,  ds = p (.) (dp - sum_j p dp),
//   dq = ds K,  dK = ds^T q,  dV = p^T do,  d rel_bias[dist(t - j), h] += ds[t, j]
// One CTA per (document, head); scores live in shared memory, K / V are re-read from L2.
#include "pn_common.cuh"
#include "pn_t5_internal.cuh"

namespace {

constexpr int DK = 64;
constexpr int MAX_S = 512;
constexpr int MAX_T = 2;
constexpr int kThreads = 256;

struct DecBwdArgs {
  const __nv_bfloat16 *q, *k, *v, *dout;
  int ld_q, q_col, ld_kv, k_col, v_col, ld_do;
  int docs, T, Sk, n_heads, causal;
  const float* bias;       // [H, T, T] (self-attention) or null
  const float *rs_q, *rs_kv;
  __nv_bfloat16 *dq, *dkv;
  int ld_dq, dq_col, ld_dkv, dk_col, dv_col;
  float* grad_rel;         // [buckets, H] gradient of the decoder's relative_attention_bias (self) or null
  PnDrop drop;             // dropout of the attention probabilities in the forward (regenerated here)
};

__device__ __forceinline__ void unpack8(const uint4& raw, float (&f)[8]) {
  const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    f[2 * e] = __uint_as_float(w[e] << 16);
    f[2 * e + 1] = __uint_as_float(w[e] & 0xffff0000u);
  }
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

__global__ void __launch_bounds__(kThreads)
dec_attn_bwd_kernel(const DecBwdArgs a) {
  __shared__ float sq[MAX_T][DK], sdo[MAX_T][DK];
  __shared__ float sp[MAX_T][MAX_S], sds[MAX_T][MAX_S];
  __shared__ float sred[4][MAX_T][DK];
  const int head = blockIdx.x % a.n_heads, doc = blockIdx.x / a.n_heads;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int T = a.T, Sk = a.Sk;
  const size_t qrow0 = (size_t)doc * T, krow0 = (size_t)doc * Sk;
  for (int i = tid; i < T * DK; i += kThreads) {
    const int t = i / DK, c = i % DK;
    sq[t][c] = __bfloat162float(a.q[(qrow0 + t) * a.ld_q + a.q_col + head * DK + c]);
    sdo[t][c] = __bfloat162float(a.dout[(qrow0 + t) * a.ld_do + head * DK + c]);
  }
  __syncthreads();
  // scores and dp: one key per thread
  for (int j = tid; j < Sk; j += kThreads) {
    const __nv_bfloat16* kr = a.k + (krow0 + j) * a.ld_kv + a.k_col + head * DK;
    const __nv_bfloat16* vr = a.v + (krow0 + j) * a.ld_kv + a.v_col + head * DK;
    float s[MAX_T] = {0.0f, 0.0f}, dp[MAX_T] = {0.0f, 0.0f};
#pragma unroll
    for (int c = 0; c < DK; c += 8) {
      float kf[8], vf[8];
      unpack8(*reinterpret_cast<const uint4*>(kr + c), kf);
      unpack8(*reinterpret_cast<const uint4*>(vr + c), vf);
#pragma unroll
      for (int t = 0; t < MAX_T; ++t)
        if (t < T) {
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            s[t] = fmaf(sq[t][c + e], kf[e], s[t]);
            dp[t] = fmaf(sdo[t][c + e], vf[e], dp[t]);
          }
        }
    }
    for (int t = 0; t < T; ++t) {
      if (a.bias) s[t] += a.bias[((size_t)head * T + t) * Sk + j];
      if (a.causal && j > t) s[t] = -INFINITY;
      sp[t][j] = s[t];
      sds[t][j] = dp[t];
    }
  }
  __syncthreads();
  // softmax and ds: warp t handles query row t
  if (warp < T) {
    const int t = warp;
    float mx = -INFINITY;
    for (int j = lane; j < Sk; j += 32) mx = fmaxf(mx, sp[t][j]);
    mx = pn_warp_max(mx);
    float sum = 0.0f;
    for (int j = lane; j < Sk; j += 32) {
      const float p = __expf(sp[t][j] - mx);
      sp[t][j] = p;
      sum += p;
    }
    sum = pn_warp_sum(sum);
    const float inv = 1.0f / sum;
    // with dropout O = sum_j p_j m_j v_j: dV sees p m, dp is masked, and D = sum_j p_j m_j dp_j
    const int skp = Sk + (Sk & 1);
    const unsigned long long dbase = (unsigned long long)a.drop.idx_offset +
                                     ((unsigned long long)(doc * a.n_heads + head) * T + t) * skp;
    float dsum = 0.0f;
    for (int j = lane; j < Sk; j += 32) {
      const float p = sp[t][j] * inv;
      sp[t][j] = p;
      sds[t][j] *= pn_drop_mul(a.drop, dbase + j);
      dsum += p * sds[t][j];
    }
    dsum = pn_warp_sum(dsum);
    for (int j = lane; j < Sk; j += 32) {
      const float ds = sp[t][j] * (sds[t][j] - dsum);
      sds[t][j] = ds;
      sp[t][j] *= pn_drop_mul(a.drop, dbase + j);    // what dV = (p m)^T do reads below
      // decoder buckets are the distance itself for distance < 16 (unidirectional, HF:218-234)
      if (a.grad_rel && j <= t) atomicAdd(a.grad_rel + (size_t)(t - j) * a.n_heads + head, ds);
    }
  }
  __syncthreads();
  // dK, dV: eight columns of one key row per thread
  for (int i = tid; i < Sk * (DK / 8); i += kThreads) {
    const int j = i >> 3, c = (i & 7) * 8;
    const float sc = a.rs_kv ? a.rs_kv[krow0 + j] : 1.0f;
    float dk[8], dv[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) dk[e] = dv[e] = 0.0f;
    for (int t = 0; t < T; ++t) {
      const float ds = sds[t][j], p = sp[t][j];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        dk[e] = fmaf(ds, sq[t][c + e], dk[e]);
        dv[e] = fmaf(p, sdo[t][c + e], dv[e]);
      }
    }
    uint4 uk, uv;
    uk.x = pack2(dk[0] * sc, dk[1] * sc); uk.y = pack2(dk[2] * sc, dk[3] * sc);
    uk.z = pack2(dk[4] * sc, dk[5] * sc); uk.w = pack2(dk[6] * sc, dk[7] * sc);
    uv.x = pack2(dv[0] * sc, dv[1] * sc); uv.y = pack2(dv[2] * sc, dv[3] * sc);
    uv.z = pack2(dv[4] * sc, dv[5] * sc); uv.w = pack2(dv[6] * sc, dv[7] * sc);
    __nv_bfloat16* drow = a.dkv + (krow0 + j) * a.ld_dkv + head * DK + c;
    *reinterpret_cast<uint4*>(drow + a.dk_col) = uk;
    *reinterpret_cast<uint4*>(drow + a.dv_col) = uv;
  }
  // dq[t, c] = sum_j ds[t, j] K[j, c]: four key groups x 64 columns, reduced through shared memory
  {
    const int c = tid & 63, grp = tid >> 6;
    float acc[MAX_T] = {0.0f, 0.0f};
    for (int j = grp; j < Sk; j += 4) {
      const float kv = __bfloat162float(a.k[(krow0 + j) * a.ld_kv + a.k_col + head * DK + c]);
#pragma unroll
      for (int t = 0; t < MAX_T; ++t)
        if (t < T) acc[t] = fmaf(sds[t][j], kv, acc[t]);
    }
#pragma unroll
    for (int t = 0; t < MAX_T; ++t) sred[grp][t][c] = acc[t];
  }
  __syncthreads();
  for (int i = tid; i < T * DK; i += kThreads) {
    const int t = i / DK, c = i % DK;
    const float v = (sred[0][t][c] + sred[1][t][c]) + (sred[2][t][c] + sred[3][t][c]);
    const float sc = a.rs_q ? a.rs_q[qrow0 + t] : 1.0f;
    a.dq[(qrow0 + t) * a.ld_dq + a.dq_col + head * DK + c] = __float2bfloat16_rn(v * sc);
  }
}

}  // namespace

int pn_t5_dec_attn_bwd(const __nv_bfloat16* q, int ld_q, int q_col, const __nv_bfloat16* k,
                       const __nv_bfloat16* v, int ld_kv, int k_col, int v_col, const __nv_bfloat16* dout,
                       int ld_do, int docs, int T, int Sk, int n_heads, int causal, const float* bias,
                       const float* rs_q, const float* rs_kv, __nv_bfloat16* dq, int ld_dq, int dq_col,
                       __nv_bfloat16* dkv, int ld_dkv, int dk_col, int dv_col, float* grad_rel,
                       cudaStream_t st, PnDrop drop) {
  if (Sk < 1 || Sk > MAX_S) return pn_fail(PN_E_UNSUPPORTED, "decoder attention backward: Sk=%d not in [1,512]", Sk);
  if (T < 1 || T > MAX_T) return pn_fail(PN_E_UNSUPPORTED, "decoder attention backward: T=%d not in [1,2]", T);
  if (docs <= 0) return PN_OK;
  DecBwdArgs a;
  a.q = q; a.k = k; a.v = v; a.dout = dout;
  a.ld_q = ld_q; a.q_col = q_col; a.ld_kv = ld_kv; a.k_col = k_col; a.v_col = v_col; a.ld_do = ld_do;
  a.docs = docs; a.T = T; a.Sk = Sk; a.n_heads = n_heads; a.causal = causal;
  a.bias = bias; a.rs_q = rs_q; a.rs_kv = rs_kv;
  a.dq = dq; a.dkv = dkv; a.ld_dq = ld_dq; a.dq_col = dq_col; a.ld_dkv = ld_dkv; a.dk_col = dk_col; a.dv_col = dv_col;
  a.grad_rel = grad_rel;
  a.drop = drop;
  dec_attn_bwd_kernel<<<docs * n_heads, kThreads, 0, st>>>(a);
  PN_LAUNCHED();
  return PN_OK;
}
