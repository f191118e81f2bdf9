// Bandwidth-bound pieces of the T5 forward: embedding gather, T5 RMSNorm (HF:55-69),
// weight conversion, and the two-row monoT5 head (HF:1107-1110 + distill/wrapper.py:34-38).
#include "pn_common.cuh"
#include "pn_t5_internal.cuh"

namespace {

constexpr int kThreads = 256;

// hidden[row, :] = table[ids[row], :]   (fp32 table -> fp32 residual stream), one warp per row; optionally also
// the bf16 copy of the row and its sum of squares (inputs of the RMSNorm fused into the next GEMM)
__global__ void __launch_bounds__(kThreads)
embed_kernel(const int32_t* __restrict__ ids, const float* __restrict__ table, int64_t rows, int d,
             int vocab, float* __restrict__ hidden, __nv_bfloat16* __restrict__ copy,
             float* __restrict__ sumsq, const PnDrop drop) {
  const int lane = threadIdx.x & 31;
  const int64_t w = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t)gridDim.x * kThreads) >> 5;
  for (int64_t r = w; r < rows; r += nw) {
    int id = ids[r];
    id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
    const float4* src = reinterpret_cast<const float4*>(table + (size_t)id * d);
    float4* dst = reinterpret_cast<float4*>(hidden + (size_t)r * d);
    float ss = 0.0f;
    for (int i = lane; i < d / 4; i += 32) {
      float4 v = src[i];
      if (drop.thresh) {   // dropout of the embedding output (T5Stack: self.dropout(inputs_embeds))
        const unsigned long long e0 = (unsigned long long)drop.idx_offset + (unsigned long long)r * d + 4 * i;
        float m0, m1, m2, m3;
        pn_drop_mul2(drop, e0, m0, m1);
        pn_drop_mul2(drop, e0 + 2, m2, m3);
        v.x *= m0; v.y *= m1; v.z *= m2; v.w *= m3;
      }
      dst[i] = v;
      ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
      if (copy) {
        __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
        reinterpret_cast<uint2*>(copy + (size_t)r * d)[i] =
            make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
      }
    }
    if (sumsq) {
      ss = pn_warp_sum(ss);
      if (lane == 0) sumsq[r] = ss;
    }
  }
}

// packed[r, c] = bf16( master[r, c] * (colscale ? colscale[c] : 1) ): folds an RMSNorm weight into the
// [out, in] matrix of the linear that consumes the normalised activations; packedT[c, r] (ld = ldT) is the
// transposed copy the data-gradient GEMM reads.  32x32 tiles through shared memory: both stores coalesced.
__global__ void __launch_bounds__(256)
pack_weight_kernel(const float* __restrict__ master, const float* __restrict__ colscale, int64_t rows,
                   int cols, __nv_bfloat16* __restrict__ packed, __nv_bfloat16* __restrict__ packedT, int ldT) {
  __shared__ float tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
  const int64_t tiles_c = (cols + 31) / 32, tiles_r = (rows + 31) / 32;
  for (int64_t t = blockIdx.x; t < tiles_c * tiles_r; t += gridDim.x) {
    const int64_t r0 = (t / tiles_c) * 32;
    const int c0 = (int)(t % tiles_c) * 32;
    const int c = c0 + tx;
    const float g = (colscale && c < cols) ? colscale[c] : 1.0f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int64_t r = r0 + ty + 8 * k;
      float v = 0.0f;
      if (r < rows && c < cols) {
        v = master[r * cols + c] * g;
        packed[r * cols + c] = __float2bfloat16_rn(v);
      }
      tile[ty + 8 * k][tx] = v;
    }
    __syncthreads();
    if (packedT) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int cc = c0 + ty + 8 * k;
        const int64_t rr = r0 + tx;
        if (cc < cols && rr < rows) packedT[(int64_t)cc * ldT + rr] = __float2bfloat16_rn(tile[tx][ty + 8 * k]);
      }
    }
    __syncthreads();
  }
}

// out[row, :] = bf16( x * rsqrt(mean(x^2) + eps) * w ), one warp per row, row held in registers
template <int MAXV>  // float4 per lane
__global__ void __launch_bounds__(kThreads)
rmsnorm_kernel(const float* __restrict__ x, const float* __restrict__ w, int64_t rows, int d,
               float eps, __nv_bfloat16* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t wid = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t)gridDim.x * kThreads) >> 5;
  const int nv = d / 4;
  for (int64_t r = wid; r < rows; r += nw) {
    const float4* src = reinterpret_cast<const float4*>(x + (size_t)r * d);
    float4 v[MAXV];
    float ss = 0.0f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int c = lane + i * 32;
      if (c < nv) {
        v[i] = src[c];
        ss += v[i].x * v[i].x + v[i].y * v[i].y + v[i].z * v[i].z + v[i].w * v[i].w;
      }
    }
    ss = pn_warp_sum(ss);
    const float rs = rsqrtf(ss / (float)d + eps);
    uint2* dst = reinterpret_cast<uint2*>(out + (size_t)r * d);
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int c = lane + i * 32;
      if (c < nv) {
        const float4 g = reinterpret_cast<const float4*>(w)[c];
        __nv_bfloat162 a = __floats2bfloat162_rn(v[i].x * rs * g.x, v[i].y * rs * g.y);
        __nv_bfloat162 b = __floats2bfloat162_rn(v[i].z * rs * g.z, v[i].w * rs * g.w);
        dst[c] = make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
      }
    }
  }
}

__global__ void __launch_bounds__(kThreads)
f32_to_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * kThreads)
    dst[i] = __float2bfloat16_rn(src[i]);
}

// One warp per (doc, t): y = rmsnorm(x) * w * d^-0.5 ; z = y . E[rel], y . E[nrel]
// score = log_softmax([z_true, z_false])[0]
__global__ void __launch_bounds__(kThreads)
head2_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ emb,
             int64_t rows, int d, float eps, int rel_id, int nrel_id, float* __restrict__ score,
             float* __restrict__ z2, const PnDrop drop) {
  const int lane = threadIdx.x & 31;
  const int64_t r = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5;
  if (r >= rows) return;
  const float* xr = x + (size_t)r * d;
  const float* et = emb + (size_t)rel_id * d;
  const float* ef = emb + (size_t)nrel_id * d;
  float ss = 0.0f, zt = 0.0f, zf = 0.0f;
  for (int i = lane; i < d; i += 32) {
    const float v = xr[i];
    ss += v * v;
    const float y = v * w[i] * pn_drop_mul(drop, (unsigned long long)drop.idx_offset + (unsigned long long)r * d + i);
    zt += y * et[i];
    zf += y * ef[i];
  }
  ss = pn_warp_sum(ss);
  zt = pn_warp_sum(zt);
  zf = pn_warp_sum(zf);
  if (lane == 0) {
    const float k = rsqrtf(ss / (float)d + eps) * rsqrtf((float)d);
    zt *= k; zf *= k;
    const float dlt = zf - zt;
    const float s = dlt > 0.0f ? -(dlt + log1pf(expf(-dlt))) : -log1pf(expf(dlt));
    if (score) score[r] = s;
    if (z2) { z2[2 * r] = zt; z2[2 * r + 1] = zf; }
  }
}

// Backward of head2_kernel's two dot products (restricted-head training): with head_in = x * w * rs * d^-0.5,
// z[r, c] = head_in[r, :] . E[id_c, :]  ->  dhead_in[r, :] = dz[r, 0] E[rel] + dz[r, 1] E[nrel] (what
// final_norm_bwd consumes) and grad_E[id_c, :] += sum_r dz[r, c] head_in[r, :].  One warp per decoder row.
__global__ void __launch_bounds__(kThreads)
head2_bwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ emb,
                 const float* __restrict__ dz2, int64_t rows, int d, float eps, int rel_id, int nrel_id,
                 float* __restrict__ dhead_in, float* __restrict__ grad_emb, const PnDrop drop) {
  const int lane = threadIdx.x & 31;
  const int64_t r = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5;
  if (r >= rows) return;
  const float* xr = x + (size_t)r * d;
  const float* et = emb + (size_t)rel_id * d;
  const float* ef = emb + (size_t)nrel_id * d;
  float ss = 0.0f;
  for (int i = lane; i < d; i += 32) ss += xr[i] * xr[i];
  ss = pn_warp_sum(ss);
  const float k = rsqrtf(ss / (float)d + eps) * rsqrtf((float)d);
  const float g0 = dz2[2 * r], g1 = dz2[2 * r + 1];
  for (int i = lane; i < d; i += 32) {
    dhead_in[(size_t)r * d + i] = g0 * et[i] + g1 * ef[i];
    const float hi = xr[i] * w[i] * k * pn_drop_mul(drop, (unsigned long long)drop.idx_offset + (unsigned long long)r * d + i);
    atomicAdd(grad_emb + (size_t)rel_id * d + i, g0 * hi);
    atomicAdd(grad_emb + (size_t)nrel_id * d + i, g1 * hi);
  }
}

// bf16 y = rmsnorm(x) * w * d^-0.5 for the full-vocabulary head GEMM
__global__ void __launch_bounds__(kThreads)
head_prep_kernel(const float* __restrict__ x, const float* __restrict__ w, int64_t rows, int d,
                 float eps, __nv_bfloat16* __restrict__ out, const PnDrop drop) {
  const int lane = threadIdx.x & 31;
  const int64_t r = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5;
  if (r >= rows) return;
  const float* xr = x + (size_t)r * d;
  float ss = 0.0f;
  for (int i = lane; i < d; i += 32) ss += xr[i] * xr[i];
  ss = pn_warp_sum(ss);
  const float k = rsqrtf(ss / (float)d + eps) * rsqrtf((float)d);
  for (int i = lane; i < d; i += 32)   // dropout of the final decoder norm's output (T5Stack: dropout(final_layer_norm))
    out[(size_t)r * d + i] = __float2bfloat16_rn(
        xr[i] * w[i] * k * pn_drop_mul(drop, (unsigned long long)drop.idx_offset + (unsigned long long)r * d + i));
}

// dst = bf16(src (.) mask): the dropped copy of the encoder output the cross K/V projections read.  The fused norm
// scales rows in the GEMM epilogue and dropout is elementwise, so dropout(rs h g) W^T = rs ((h (.) M) (g W)^T).
__global__ void __launch_bounds__(kThreads)
mask_copy_kernel(const __nv_bfloat16* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t n, const PnDrop drop) {
  for (int64_t i = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * 2; i < n; i += (int64_t)gridDim.x * kThreads * 2) {
    float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(src + i));
    float m0, m1;
    pn_drop_mul2(drop, (unsigned long long)drop.idx_offset + (unsigned long long)i, m0, m1);
    *reinterpret_cast<__nv_bfloat162*>(dst + i) = __floats2bfloat162_rn(f.x * m0, f.y * m1);
  }
}

// hidden[row, :] = table[ids[row], :] for the decoder inputs (ids int32 [rows])
}  // namespace

int pn_t5_embed(const int32_t* ids, const float* table, int64_t rows, int d, int vocab, float* hidden,
                __nv_bfloat16* copy, float* sumsq, cudaStream_t st, PnDrop drop) {
  if (rows == 0) return PN_OK;
  if (d % 4) return pn_fail(PN_E_UNSUPPORTED, "embed: d_model %% 4 != 0");
  embed_kernel<<<pn_grid(rows * 32, kThreads, 1, 8), kThreads, 0, st>>>(ids, table, rows, d, vocab,
                                                                        hidden, copy, sumsq, drop);
  PN_LAUNCHED();
  return PN_OK;
}

int pn_t5_pack_weight(const float* master, const float* colscale, int64_t rows, int cols,
                      __nv_bfloat16* packed, __nv_bfloat16* packedT, int ldT, cudaStream_t st) {
  if (rows * cols == 0) return PN_OK;
  const int64_t tiles = ((rows + 31) / 32) * ((cols + 31) / 32);
  const int grid = (int)(tiles < (int64_t)PN_SM_COUNT * 16 ? tiles : (int64_t)PN_SM_COUNT * 16);
  pack_weight_kernel<<<grid, 256, 0, st>>>(master, colscale, rows, cols, packed, packedT, ldT);
  PN_LAUNCHED();
  return PN_OK;
}

int pn_t5_rmsnorm(const float* x, const float* w, int64_t rows, int d, float eps, __nv_bfloat16* out,
                  cudaStream_t st) {
  if (rows == 0) return PN_OK;
  if (d % 4 || d > 32 * 4 * 16) return pn_fail(PN_E_UNSUPPORTED, "rmsnorm: d_model=%d unsupported", d);
  const int grid = pn_grid(rows * 32, kThreads, 1, 8);
  if (d <= 32 * 4 * 4)
    rmsnorm_kernel<4><<<grid, kThreads, 0, st>>>(x, w, rows, d, eps, out);
  else if (d <= 32 * 4 * 8)
    rmsnorm_kernel<8><<<grid, kThreads, 0, st>>>(x, w, rows, d, eps, out);
  else
    rmsnorm_kernel<16><<<grid, kThreads, 0, st>>>(x, w, rows, d, eps, out);
  PN_LAUNCHED();
  return PN_OK;
}

int pn_t5_f32_to_bf16(const float* src, __nv_bfloat16* dst, int64_t n, cudaStream_t st) {
  if (n == 0) return PN_OK;
  f32_to_bf16_kernel<<<pn_grid(n, kThreads, 4, 8), kThreads, 0, st>>>(src, dst, n);
  PN_LAUNCHED();
  return PN_OK;
}

int pn_t5_head2(const float* x, const float* w, const float* emb, int64_t rows, int d, float eps,
                int rel_id, int nrel_id, float* score, float* z2, cudaStream_t st, PnDrop drop) {
  if (rows == 0) return PN_OK;
  head2_kernel<<<(unsigned)((rows * 32 + kThreads - 1) / kThreads), kThreads, 0, st>>>(
      x, w, emb, rows, d, eps, rel_id, nrel_id, score, z2, drop);
  PN_LAUNCHED();
  return PN_OK;
}

int pn_t5_head2_bwd(const float* x, const float* w, const float* emb, const float* dz2, int64_t rows, int d, float eps,
                    int rel_id, int nrel_id, float* dhead_in, float* grad_emb, cudaStream_t st, PnDrop drop) {
  if (rows == 0) return PN_OK;
  head2_bwd_kernel<<<(unsigned)((rows * 32 + kThreads - 1) / kThreads), kThreads, 0, st>>>(
      x, w, emb, dz2, rows, d, eps, rel_id, nrel_id, dhead_in, grad_emb, drop);
  PN_LAUNCHED();
  return PN_OK;
}

int pn_t5_head_prep(const float* x, const float* w, int64_t rows, int d, float eps,
                    __nv_bfloat16* out, cudaStream_t st, PnDrop drop) {
  if (rows == 0) return PN_OK;
  head_prep_kernel<<<(unsigned)((rows * 32 + kThreads - 1) / kThreads), kThreads, 0, st>>>(
      x, w, rows, d, eps, out, drop);
  PN_LAUNCHED();
  return PN_OK;
}

int pn_t5_mask_copy(const __nv_bfloat16* src, __nv_bfloat16* dst, int64_t n, PnDrop drop, cudaStream_t st) {
  if (n == 0) return PN_OK;
  if ((n & 1) || !drop.thresh) return pn_fail(PN_E_BADARG, "mask_copy: even element count and an enabled mask");
  mask_copy_kernel<<<pn_grid(n / 2, kThreads, 4, 8), kThreads, 0, st>>>(src, dst, n, drop);
  PN_LAUNCHED();
  return PN_OK;
}

// ---------------------------------------------------------------------------------------------------------
// Cross-attention without the K/V projection of the encoder states (decoder has T <= 2 tokens per document):
//   s[t,h,j] = q[t,h,:] . (Wk_h x_j) = x_j . (Wk_h^T q[t,h,:])            -> scores = X . QK^T, QK = Qexp . Wk
//   o[t,h,:] = sum_j p[t,h,j] Wv_h x_j = Wv_h (sum_j p[t,h,j] x_j)        -> C' = P^T . X,     O = C' . Wv^T
// which replaces a [rows, d] x [d, 2*d_inner] GEMM per decoder layer (13 % of the forward FLOPs) by two skinny
// products over the same rows (T*H columns) and three tiny ones over docs*T*H rows.
namespace {

// Qexp[doc*Mp + t*H + h, :] = q[doc*T + t, head block h] placed at columns h*64.., zero elsewhere (and in pad rows)
__global__ void __launch_bounds__(256)
cross_expand_q_kernel(const __nv_bfloat16* __restrict__ q, int ld_q, int64_t docs, int T, int H, int Mp,
                      __nv_bfloat16* __restrict__ qexp) {
  const int c8n = H * 8;   // 16-byte chunks per row (d_inner / 8)
  const int64_t n = docs * Mp * c8n;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
    const int c8 = (int)(i % c8n);
    const int64_t row = i / c8n;
    const int th = (int)(row % Mp);
    const int64_t doc = row / Mp;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (th < T * H && (c8 >> 3) == th % H)
      v = *reinterpret_cast<const uint4*>(q + (size_t)(doc * T + th / H) * ld_q + c8 * 8);
    reinterpret_cast<uint4*>(qexp)[i] = v;
  }
}

// P[row, th] = softmax over the document's S rows of scores[row, th], times rs[row] (the RMS scale of x_row, so that
// P^T . X sums the normalised states); pad columns th >= TH are written as zero.  One CTA per document.
__global__ void __launch_bounds__(256)
cross_softmax_kernel(const float* __restrict__ scores, int lds, const float* __restrict__ rs, int S, int TH, int Mp,
                     __nv_bfloat16* __restrict__ P, const int32_t* __restrict__ klen, const PnDrop drop, int H,
                     __nv_bfloat16* __restrict__ Pc) {
  __shared__ float red[8][32];
  __shared__ float stat[32];
  const int64_t row0 = (int64_t)blockIdx.x * S;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // thread layout: th = tid % 32 (columns), row phase = tid / 32 (8 rows in flight per pass)
  const int th = lane;
  const bool act = th < TH;
  float v[64];   // this lane's column, rows warp, warp + 8, ... (S <= 512)
  float mx = -INFINITY;
  const int nk = klen ? min(S, max(1, klen[blockIdx.x])) : S;   // encoder rows >= klen[doc] are padding: P = 0 there
#pragma unroll
  for (int i = 0; i < 64; ++i) {
    const int j = warp + 8 * i;
    v[i] = (act && j < nk) ? scores[(size_t)(row0 + j) * lds + th] : -INFINITY;
    mx = fmaxf(mx, v[i]);
  }
  red[warp][lane] = mx;
  __syncthreads();
  if (warp == 0) {
    float m = red[0][lane];
#pragma unroll
    for (int k = 1; k < 8; ++k) m = fmaxf(m, red[k][lane]);
    stat[lane] = m;
  }
  __syncthreads();
  mx = stat[lane];
  __syncthreads();
  float sum = 0.0f;
#pragma unroll
  for (int i = 0; i < 64; ++i) {
    v[i] = act ? __expf(v[i] - mx) : 0.0f;   // exp(-inf) = 0 for rows beyond S
    sum += v[i];
  }
  red[warp][lane] = sum;
  __syncthreads();
  if (warp == 0) {
    float t = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += red[k][lane];
    stat[lane] = 1.0f / t;
  }
  __syncthreads();
  const float inv = stat[lane];
  if (th < Mp) {
    // training with dropout (HF:331 on the cross-attention probabilities [doc, h, t, j], even row pitch): P carries the
    // mask (it feeds P^T X), Pc keeps the clean probabilities for the softmax backward
    const int T = TH / H;
    const unsigned long long dbase =
        (unsigned long long)drop.idx_offset +
        (((unsigned long long)blockIdx.x * H + (th % H)) * T + (th / H)) * (unsigned long long)(S + (S & 1));
#pragma unroll
    for (int i = 0; i < 64; ++i) {
      const int j = warp + 8 * i;
      if (j < S) {
        const float pv = act ? v[i] * inv * rs[row0 + j] : 0.0f;
        if (Pc) Pc[(size_t)(row0 + j) * Mp + th] = __float2bfloat16_rn(pv);
        const float mk = (act && drop.thresh) ? pn_drop_mul(drop, dbase + j) : 1.0f;
        P[(size_t)(row0 + j) * Mp + th] = __float2bfloat16_rn(pv * mk);
      }
    }
  }
}

// ctx[doc*T + t, h*64 + c] = full[doc*Mp + t*H + h, h*64 + c]
__global__ void __launch_bounds__(256)
cross_gather_kernel(const __nv_bfloat16* __restrict__ full, int Mp, int64_t docs, int T, int H,
                    __nv_bfloat16* __restrict__ ctx, int ld_ctx) {
  const int c8n = H * 8;
  const int64_t n = docs * T * c8n;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
    const int c8 = (int)(i % c8n);
    const int64_t dt = i / c8n;
    const int64_t doc = dt / T;
    const int t = (int)(dt % T), h = c8 >> 3;
    const uint4 v = *reinterpret_cast<const uint4*>(full + (size_t)(doc * Mp + t * H + h) * (H * 64) + c8 * 8);
    *reinterpret_cast<uint4*>(ctx + (size_t)dt * ld_ctx + c8 * 8) = v;
  }
}

}  // namespace

int pn_t5_cross_expand_q(const __nv_bfloat16* q, int ld_q, int64_t docs, int T, int H, int Mp, __nv_bfloat16* qexp,
                         cudaStream_t st) {
  if (docs == 0) return PN_OK;
  cross_expand_q_kernel<<<pn_grid(docs * Mp * H * 8, 256, 1, 8), 256, 0, st>>>(q, ld_q, docs, T, H, Mp, qexp);
  PN_LAUNCHED();
  return PN_OK;
}
int pn_t5_cross_softmax(const float* scores, int lds, const float* rs, int64_t docs, int S, int TH, int Mp,
                        __nv_bfloat16* P, cudaStream_t st, const int32_t* klen, PnDrop drop, int H,
                        __nv_bfloat16* Pc) {
  if (docs == 0) return PN_OK;
  if (Mp > 32 || TH > Mp) return pn_fail(PN_E_UNSUPPORTED, "cross softmax: T*H=%d > 32", TH);
  if (drop.thresh && (H <= 0 || TH % H || !Pc))
    return pn_fail(PN_E_BADARG, "cross softmax: dropout needs the head count and a buffer for the clean probabilities");
  if (H <= 0) H = TH;
  cross_softmax_kernel<<<(unsigned)docs, 256, 0, st>>>(scores, lds, rs, S, TH, Mp, P, klen, drop, H, Pc);
  PN_LAUNCHED();
  return PN_OK;
}
int pn_t5_cross_gather(const __nv_bfloat16* full, int Mp, int64_t docs, int T, int H, __nv_bfloat16* ctx, int ld_ctx,
                       cudaStream_t st) {
  if (docs == 0) return PN_OK;
  cross_gather_kernel<<<pn_grid(docs * T * H * 8, 256, 1, 8), 256, 0, st>>>(full, Mp, docs, T, H, ctx, ld_ctx);
  PN_LAUNCHED();
  return PN_OK;
}

// ---- backward pieces of the projection-free cross-attention (training, pn_t5_train.cu)
namespace {

// Per document: P = Pp / rs (Pp is what cross_softmax_kernel stored), dSc = P (.) (dP - sum_j P dP).
//   U[row, u_col + th]      = P     (value path:  d x_hat += P  . dC')
//   U[row, u_col + Mp + th] = dSc   (score path:  d x_hat += dSc . QK),   columns up to 64 zero
//   V2[row, th]             = rs * dSc            (d QK = V2^T X)
// One CTA per document, one thread per encoder row: the Mp values of a row are contiguous in Pp (bf16) and dP (fp32),
// so every access is a 16-byte vector and a row's 128-byte slice of U is written by one thread.
template <int MP8>
__global__ void __launch_bounds__(512)
cross_softmax_bwd_kernel(const __nv_bfloat16* __restrict__ Pp, const float* __restrict__ dP, int ldd,
                         const float* __restrict__ rs, int S, int TH, __nv_bfloat16* __restrict__ U, int ldu,
                         int u_col, __nv_bfloat16* __restrict__ V2, const PnDrop drop, int H) {
  constexpr int Mp = MP8 * 8;
  __shared__ float red[16][Mp];
  __shared__ float stat[Mp];
  const int64_t row0 = (int64_t)blockIdx.x * S;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // With dropout (Pp = the CLEAN probabilities then, times rs): out = (P (.) M)^T x, so the value path carries P (.) M,
  // dP arrives for the dropped probabilities and dSc = P (.) (M (.) dP - sum_j P M dP).  `mm` holds the regenerated M.
  float p[Mp], g[Mp], dot[Mp], mm[Mp];
  float r = 1.0f;
  const int T = TH / H;
  const unsigned long long Sp = (unsigned long long)(S + (S & 1));
#pragma unroll
  for (int c = 0; c < Mp; ++c) mm[c] = 1.0f;
  auto load = [&](int j) {
    if (j < S) {
      r = rs[row0 + j];
      const float inv = 1.0f / r;
#pragma unroll
      for (int c = 0; c < MP8; ++c) {
        const uint4 raw = reinterpret_cast<const uint4*>(Pp + (size_t)(row0 + j) * Mp)[c];
        const uint32_t w4[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w4[k]));
          p[c * 8 + 2 * k] = f.x * inv;
          p[c * 8 + 2 * k + 1] = f.y * inv;
        }
        const float4 g0 = reinterpret_cast<const float4*>(dP + (size_t)(row0 + j) * ldd)[2 * c];
        const float4 g1 = reinterpret_cast<const float4*>(dP + (size_t)(row0 + j) * ldd)[2 * c + 1];
        g[c * 8 + 0] = g0.x; g[c * 8 + 1] = g0.y; g[c * 8 + 2] = g0.z; g[c * 8 + 3] = g0.w;
        g[c * 8 + 4] = g1.x; g[c * 8 + 5] = g1.y; g[c * 8 + 6] = g1.z; g[c * 8 + 7] = g1.w;
      }
#pragma unroll
      for (int c = 0; c < Mp; ++c)
        if (c >= TH) { p[c] = 0.0f; g[c] = 0.0f; }
      if (drop.thresh) {
#pragma unroll
        for (int c = 0; c < Mp; ++c)
          if (c < TH) {
            mm[c] = pn_drop_mul(drop, (unsigned long long)drop.idx_offset +
                                          (((unsigned long long)blockIdx.x * H + (c % H)) * T + (c / H)) * Sp + j);
            g[c] *= mm[c];
          }
      }
    } else {
#pragma unroll
      for (int c = 0; c < Mp; ++c) { p[c] = 0.0f; g[c] = 0.0f; }
    }
  };
#pragma unroll
  for (int c = 0; c < Mp; ++c) dot[c] = 0.0f;
  for (int j0 = 0; j0 < S; j0 += 512) {
    load(j0 + tid);
#pragma unroll
    for (int c = 0; c < Mp; ++c) dot[c] += p[c] * g[c];
  }
#pragma unroll
  for (int c = 0; c < Mp; ++c) {
    float t = pn_warp_sum(dot[c]);
    if (lane == 0) red[warp][c] = t;
  }
  __syncthreads();
  if (tid < Mp) {
    float t = 0.0f;
#pragma unroll
    for (int k = 0; k < 16; ++k) t += red[k][tid];
    stat[tid] = t;
  }
  __syncthreads();
  for (int j0 = 0; j0 < S; j0 += 512) {
    const int j = j0 + tid;
    if (S > 512) load(j);   // a document of at most 512 rows still holds its row in registers
    if (j >= S) continue;
    uint4* urow = reinterpret_cast<uint4*>(U + (size_t)(row0 + j) * ldu + u_col);
    uint4* vrow = reinterpret_cast<uint4*>(V2 + (size_t)(row0 + j) * Mp);
    uint32_t up[Mp / 2], ud[Mp / 2], uv[Mp / 2];
#pragma unroll
    for (int c = 0; c < Mp; c += 2) {
      const float d0 = p[c] * (g[c] - stat[c]), d1 = p[c + 1] * (g[c + 1] - stat[c + 1]);
      __nv_bfloat162 a = __floats2bfloat162_rn(p[c] * mm[c], p[c + 1] * mm[c + 1]);
      __nv_bfloat162 b = __floats2bfloat162_rn(d0, d1);
      __nv_bfloat162 e = __floats2bfloat162_rn(d0 * r, d1 * r);
      up[c / 2] = *reinterpret_cast<uint32_t*>(&a);
      ud[c / 2] = *reinterpret_cast<uint32_t*>(&b);
      uv[c / 2] = *reinterpret_cast<uint32_t*>(&e);
    }
#pragma unroll
    for (int c = 0; c < MP8; ++c) {
      urow[c] = make_uint4(up[4 * c], up[4 * c + 1], up[4 * c + 2], up[4 * c + 3]);
      urow[MP8 + c] = make_uint4(ud[4 * c], ud[4 * c + 1], ud[4 * c + 2], ud[4 * c + 3]);
      vrow[c] = make_uint4(uv[4 * c], uv[4 * c + 1], uv[4 * c + 2], uv[4 * c + 3]);
    }
#pragma unroll
    for (int c = 2 * MP8; c < 8; ++c) urow[c] = make_uint4(0u, 0u, 0u, 0u);
  }
}

// BT[doc*d + n, k_col + k] = k < Mp ? dC[doc*Mp + k, n] : (k < 2*Mp ? QK[doc*Mp + k - Mp, n] : 0),  k in [0, 64)
__global__ void __launch_bounds__(256)
cross_bt_pack_kernel(const __nv_bfloat16* __restrict__ dC, const __nv_bfloat16* __restrict__ QK, int64_t docs, int d,
                     int Mp, __nv_bfloat16* __restrict__ BT, int ldb, int k_col) {
  const int64_t n_items = docs * d;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n_items; i += (int64_t)gridDim.x * 256) {
    const int n = (int)(i % d);
    const int64_t doc = i / d;
    __nv_bfloat16 v[64];
#pragma unroll
    for (int k = 0; k < 64; ++k) {
      __nv_bfloat16 x = __float2bfloat16_rn(0.0f);
      if (k < Mp) x = dC[(size_t)(doc * Mp + k) * d + n];
      else if (k < 2 * Mp) x = QK[(size_t)(doc * Mp + k - Mp) * d + n];
      v[k] = x;
    }
    uint4* dst = reinterpret_cast<uint4*>(BT + (size_t)i * ldb + k_col);
#pragma unroll
    for (int c = 0; c < 8; ++c) dst[c] = reinterpret_cast<const uint4*>(v)[c];
  }
}

}  // namespace

int pn_t5_cross_softmax_bwd(const __nv_bfloat16* Pp, const float* dP, int ldd, const float* rs, int64_t docs, int S,
                            int TH, int Mp, __nv_bfloat16* U, int ldu, int u_col, __nv_bfloat16* V2, cudaStream_t st,
                            PnDrop drop, int H) {
  if (docs == 0) return PN_OK;
  if (drop.thresh && (H <= 0 || TH % H)) return pn_fail(PN_E_BADARG, "cross softmax backward: dropout needs the head count");
  if (H <= 0) H = TH;
  if (2 * Mp > 64 || TH > Mp || (Mp % 8)) return pn_fail(PN_E_UNSUPPORTED, "cross softmax backward: T*H=%d > 32", TH);
  if ((ldd % 4) || (ldu % 8) || (u_col % 8))
    return pn_fail(PN_E_BADARG, "cross softmax backward: ldd %% 4, ldu %% 8, u_col %% 8 must be 0");
#define PN_CSB(N) cross_softmax_bwd_kernel<N><<<(unsigned)docs, 512, 0, st>>>(Pp, dP, ldd, rs, S, TH, U, ldu, u_col, V2, drop, H)
  switch (Mp / 8) {
    case 1: PN_CSB(1); break;
    case 2: PN_CSB(2); break;
    case 3: PN_CSB(3); break;
    default: PN_CSB(4); break;
  }
#undef PN_CSB
  PN_LAUNCHED();
  return PN_OK;
}
int pn_t5_cross_bt_pack(const __nv_bfloat16* dC, const __nv_bfloat16* QK, int64_t docs, int d, int Mp,
                        __nv_bfloat16* BT, int ldb, int k_col, cudaStream_t st) {
  if (docs == 0) return PN_OK;
  if ((ldb % 8) || (k_col % 8)) return pn_fail(PN_E_BADARG, "cross bt pack: ldb / k_col must be multiples of 8");
  cross_bt_pack_kernel<<<pn_grid(docs * d, 256, 1, 8), 256, 0, st>>>(dC, QK, docs, d, Mp, BT, ldb, k_col);
  PN_LAUNCHED();
  return PN_OK;
}
