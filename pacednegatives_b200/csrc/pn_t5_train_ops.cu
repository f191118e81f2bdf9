// Bandwidth-bound pieces of the T5 backward pass (autograd of HF modeling_t5.py under lceT5.py:154-157):
// RMSNorm backward in the "fused norm" formulation, ReLU mask, folded-weight gradient unfold, embedding
// scatter, final-norm/head backward, AdamW.  All one pass over their operands, 128-bit accesses.
//
// Fused-norm algebra (forward: y = rs * (hb @ W'^T), rs = rsqrt(mean(h^2)+eps), W' = W diag(g), hb = bf16(h)):
//   Ys = rs (.) dY  (row-scaled upstream gradient, what the producers of dY write)
//   dW' = Ys^T hb            (weight-gradient GEMM on the saved bf16 residual copy)
//   G   = Ys W'              (data-gradient GEMM)      = rs * d x_hat
//   dh  = G - rs^2 * hb * mean_k(G_k hb_k)             (the rs of x_hat and of d x_hat cancel inside the mean)
//   dW  = dW' diag(g),  dg_k = sum_n dW'[n,k] W[n,k]
#include "pn_t5_model.cuh"

namespace {

constexpr int kThreads = 256;

// rs[row] = rsqrt(sum_p ss[p*rows + row] / d + eps)
__global__ void __launch_bounds__(kThreads)
rs_kernel(const float* __restrict__ ss, int parts, int64_t rows, int d, float eps, float* __restrict__ rs) {
  for (int64_t r = (int64_t)blockIdx.x * kThreads + threadIdx.x; r < rows; r += (int64_t)gridDim.x * kThreads) {
    float s = 0.0f;
    for (int p = 0; p < parts; ++p) s += ss[(size_t)p * rows + r];
    rs[r] = rsqrtf(s / (float)d + eps);
  }
}

// y[r, c] = (act[r, c] > 0 ? y[r, c] : 0) * rs[r]   in place, bf16 [rows, cols]
__global__ void __launch_bounds__(kThreads)
relu_mask_scale_kernel(__nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ act,
                       const float* __restrict__ rs, int64_t rows, int cols) {
  const int64_t n8 = rows * cols / 8;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n8; i += (int64_t)gridDim.x * kThreads) {
    const int64_t r = (i * 8) / cols;
    const float s = rs ? rs[r] : 1.0f;
    uint4 yv = reinterpret_cast<const uint4*>(y)[i];
    const uint4 av = reinterpret_cast<const uint4*>(act)[i];
    __nv_bfloat162* yh = reinterpret_cast<__nv_bfloat162*>(&yv);
    const __nv_bfloat162* ah = reinterpret_cast<const __nv_bfloat162*>(&av);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float2 f = __bfloat1622float2(yh[e]);
      const float2 a = __bfloat1622float2(ah[e]);
      f.x = a.x > 0.0f ? f.x * s : 0.0f;
      f.y = a.y > 0.0f ? f.y * s : 0.0f;
      yh[e] = __floats2bfloat162_rn(f.x, f.y);
    }
    reinterpret_cast<uint4*>(y)[i] = yv;
  }
}

// y[r, :] *= rs[r]   in place, bf16 [rows, cols] with row stride ld (a column block of a wider matrix)
__global__ void __launch_bounds__(kThreads)
row_scale_kernel(__nv_bfloat16* __restrict__ y, int ld, const float* __restrict__ rs, int64_t rows, int cols) {
  const int c8 = cols / 8;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < rows * c8; i += (int64_t)gridDim.x * kThreads) {
    const int64_t r = i / c8;
    const int c = (int)(i % c8);
    const float s = rs[r];
    uint4* p = reinterpret_cast<uint4*>(y + (size_t)r * ld) + c;
    uint4 v = *p;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float2 f = __bfloat1622float2(h[e]);
      h[e] = __floats2bfloat162_rn(f.x * s, f.y * s);
    }
    *p = v;
  }
}

// Fused-norm backward, one warp per row:  dh[r,:] (+)= G[r,:] - rs[r]^2 * hb[r,:] * mean_k(G[r,k] hb[r,k]);
// also refreshes the bf16 copy of dh that the next GEMMs read.  accumulate = 0 overwrites dh.
template <int MAXV, bool kG16>
__global__ void __launch_bounds__(kThreads)
norm_bwd_kernel(const void* __restrict__ Gv, const __nv_bfloat16* __restrict__ hb, const float* __restrict__ rs,
                int64_t rows, int d, int accumulate, float* __restrict__ dh, __nv_bfloat16* __restrict__ dhb,
                const PnDrop gdrop, const PnDrop odrop) {
  const float* G = reinterpret_cast<const float*>(Gv);
  const __nv_bfloat16* G16 = reinterpret_cast<const __nv_bfloat16*>(Gv);
  const int lane = threadIdx.x & 31;
  const int64_t wid = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t)gridDim.x * kThreads) >> 5;
  const int nv = d / 4;
  for (int64_t r = wid; r < rows; r += nw) {
    float4 g[MAXV], h[MAXV], pv[MAXV];   // the read half of the dh update is issued with the other loads
    float dot = 0.0f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int c = lane + i * 32;
      if (c < nv) {
        pv[i] = accumulate ? reinterpret_cast<const float4*>(dh + (size_t)r * d)[c] : make_float4(0.f, 0.f, 0.f, 0.f);
        if (kG16) {
          const uint2 gr = reinterpret_cast<const uint2*>(G16 + (size_t)r * d)[c];
          const float2 ga = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gr.x));
          const float2 gb = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gr.y));
          g[i] = make_float4(ga.x, ga.y, gb.x, gb.y);
        } else {
          g[i] = reinterpret_cast<const float4*>(G + (size_t)r * d)[c];
        }
        if (gdrop.thresh) {
          const unsigned long long e0 = (unsigned long long)gdrop.idx_offset + (unsigned long long)r * d + 4 * c;
          float m0, m1, m2, m3;
          pn_drop_mul2(gdrop, e0, m0, m1);
          pn_drop_mul2(gdrop, e0 + 2, m2, m3);
          g[i].x *= m0; g[i].y *= m1; g[i].z *= m2; g[i].w *= m3;
        }
        const uint2 raw = reinterpret_cast<const uint2*>(hb + (size_t)r * d)[c];
        const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.x));
        const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.y));
        h[i] = make_float4(a.x, a.y, b.x, b.y);
        dot += g[i].x * h[i].x + g[i].y * h[i].y + g[i].z * h[i].z + g[i].w * h[i].w;
      }
    }
    dot = pn_warp_sum(dot);
    const float s = rs[r];
    const float k = s * s * dot / (float)d;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int c = lane + i * 32;
      if (c < nv) {
        float4 o;
        o.x = g[i].x - k * h[i].x; o.y = g[i].y - k * h[i].y; o.z = g[i].z - k * h[i].z; o.w = g[i].w - k * h[i].w;
        float4* dst = reinterpret_cast<float4*>(dh + (size_t)r * d) + c;
        o.x += pv[i].x; o.y += pv[i].y; o.z += pv[i].z; o.w += pv[i].w;
        *dst = o;
        if (dhb) {
          if (odrop.thresh) {
            const unsigned long long e0 = (unsigned long long)odrop.idx_offset + (unsigned long long)r * d + 4 * c;
            float m0, m1, m2, m3;
            pn_drop_mul2(odrop, e0, m0, m1);
            pn_drop_mul2(odrop, e0 + 2, m2, m3);
            o.x *= m0; o.y *= m1; o.z *= m2; o.w *= m3;
          }
          __nv_bfloat162 lo = __floats2bfloat162_rn(o.x, o.y), hi = __floats2bfloat162_rn(o.z, o.w);
          reinterpret_cast<uint2*>(dhb + (size_t)r * d)[c] =
              make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
        }
      }
    }
  }
}

// coefficient of the RMSNorm backward folded into a data-gradient GEMM: coef[r] = rs[r] * sum_p part[p][r] / d
__global__ void __launch_bounds__(kThreads)
norm_coef_kernel(const float* __restrict__ part, int parts, const float* __restrict__ rs, int64_t rows, float inv_d,
                 float* __restrict__ coef) {
  for (int64_t r = (int64_t)blockIdx.x * kThreads + threadIdx.x; r < rows; r += (int64_t)gridDim.x * kThreads) {
    float s = 0.0f;
    for (int p = 0; p < parts; ++p) s += part[(size_t)p * rows + r];
    coef[r] = rs[r] * s * inv_d;
  }
}

// gradW[n,k] += dWp[n,k] * g[k];  gradg[k] += sum_n dWp[n,k] * W[n,k]   (W = fp32 master, dWp = d W')
// A CTA owns 128 columns (one float4 per lane) x 32 rows; every load of a thread is independent and issued up front,
// one atomic per column per CTA.
__global__ void __launch_bounds__(kThreads)
unfold_kernel(const float* __restrict__ dWp, const float* __restrict__ W, const float* __restrict__ g,
              int64_t rows, int cols, float* __restrict__ gradW, float* __restrict__ gradg) {
  __shared__ float4 red[8][32];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = (blockIdx.x * 32 + tx) * 4;
  const int64_t r0 = (int64_t)blockIdx.y * 32 + ty;
  float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
  if (c < cols) {
    const float4 gc = *reinterpret_cast<const float4*>(g + c);
    float4 v[4], w[4], o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t r = r0 + 8 * i;
      if (r < rows) {
        v[i] = *reinterpret_cast<const float4*>(dWp + r * cols + c);
        w[i] = *reinterpret_cast<const float4*>(W + r * cols + c);
        o[i] = *reinterpret_cast<const float4*>(gradW + r * cols + c);
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t r = r0 + 8 * i;
      if (r < rows) {
        o[i].x += v[i].x * gc.x; o[i].y += v[i].y * gc.y; o[i].z += v[i].z * gc.z; o[i].w += v[i].w * gc.w;
        *reinterpret_cast<float4*>(gradW + r * cols + c) = o[i];
        acc.x += v[i].x * w[i].x; acc.y += v[i].y * w[i].y; acc.z += v[i].z * w[i].z; acc.w += v[i].w * w[i].w;
      }
    }
  }
  red[ty][tx] = acc;
  __syncthreads();
  if (ty < 4 && c < cols) {   // warp ty sums component ty of the 8 partials
    float t = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += reinterpret_cast<const float*>(&red[k][tx])[ty];
    atomicAdd(gradg + c + ty, t);
  }
}

// grad_table[ids[r], :] += dh[r, :]  (tied input embeddings), one warp per row, vector atomics
__global__ void __launch_bounds__(kThreads)
embed_bwd_kernel(const int32_t* __restrict__ ids, const float* __restrict__ dh, int64_t rows, int d, int vocab,
                 float* __restrict__ grad_table, const PnDrop drop) {
  const int lane = threadIdx.x & 31;
  const int64_t w = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t)gridDim.x * kThreads) >> 5;
  for (int64_t r = w; r < rows; r += nw) {
    int id = ids[r];
    id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
    const float4* src = reinterpret_cast<const float4*>(dh + (size_t)r * d);
    float4* dst = reinterpret_cast<float4*>(grad_table + (size_t)id * d);
    for (int i = lane; i < d / 4; i += 32) {
      float4 v = src[i];
      if (drop.thresh) {   // the embedding output was dropped: its gradient passes through the same mask
        const unsigned long long e0 = (unsigned long long)drop.idx_offset + (unsigned long long)r * d + 4 * i;
        float m0, m1, m2, m3;
        pn_drop_mul2(drop, e0, m0, m1);
        pn_drop_mul2(drop, e0 + 2, m2, m3);
        v.x *= m0; v.y *= m1; v.z *= m2; v.w *= m3;
      }
      atomicAdd(dst + i, v);
    }
  }
}

// Final decoder norm + head scale, y = x * rsqrt(mean(x^2)+eps) * w * d^-0.5 (HF:55-69,1107-1108):
//   dx = k * (w (.) dy - x_hat * mean(x_hat (.) w (.) dy)),  dw += k' * sum_rows dy (.) x_hat,   k = rs * d^-0.5
__global__ void __launch_bounds__(kThreads)
final_norm_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ w,
                      int64_t rows, int d, float eps, float* __restrict__ dx, __nv_bfloat16* __restrict__ dxb,
                      float* __restrict__ dw, const PnDrop ydrop, const PnDrop odrop) {
  const int lane = threadIdx.x & 31;
  const int64_t r = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5;
  if (r >= rows) return;
  const float* xr = x + (size_t)r * d;
  const float* gr = dy + (size_t)r * d;
  float ss = 0.0f, dot = 0.0f;
  const unsigned long long yb = (unsigned long long)ydrop.idx_offset + (unsigned long long)r * d;
  for (int i = lane; i < d; i += 32) {
    ss += xr[i] * xr[i];
    dot += xr[i] * w[i] * gr[i] * pn_drop_mul(ydrop, yb + i);
  }
  ss = pn_warp_sum(ss);
  dot = pn_warp_sum(dot);
  const float rs = rsqrtf(ss / (float)d + eps);
  const float sc = rsqrtf((float)d);
  const float m = dot * rs * rs / (float)d;     // mean(x_hat w dy) * rs, folded for the x (not x_hat) product
  for (int i = lane; i < d; i += 32) {
    const float gi = gr[i] * pn_drop_mul(ydrop, yb + i);
    const float v = sc * rs * (w[i] * gi - xr[i] * m);
    dx[(size_t)r * d + i] = v;
    if (dxb)
      dxb[(size_t)r * d + i] = __float2bfloat16_rn(
          v * pn_drop_mul(odrop, (unsigned long long)odrop.idx_offset + (unsigned long long)r * d + i));
    atomicAdd(dw + i, sc * gi * xr[i] * rs);
  }
}

// transformers.AdamW (the class the reference constructs, lceT5.py:171-175): step_size = lr*sqrt(bc2)/bc1,
// p -= step_size * m / (sqrt(v) + eps); decoupled decay p -= lr * wd * p
__device__ __forceinline__ void adamw_one(float& p, float g, float& m, float& v, float step, float lr, float b1,
                                          float b2, float eps, float wd, float gscale) {
  const float gi = g * gscale;
  const float mi = b1 * m + (1.0f - b1) * gi;
  const float vi = b2 * v + (1.0f - b2) * gi * gi;
  m = mi;
  v = vi;
  float x = p;
  x -= step * (mi / (sqrtf(vi) + eps));
  if (wd != 0.0f) x -= lr * wd * x;
  p = x;
}

// 28 B per parameter (p, m, v read + written, g read).  VEC: all four arrays 16-byte aligned -> float4 accesses, and
// a CTA walks one contiguous 8 KB tile of each array at a time (2 x 256 float4) so that the seven streams stay
// inside a few DRAM pages.
template <bool VEC>
__global__ void __launch_bounds__(kThreads)
adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
             int64_t n, float lr, float b1, float b2, float eps, float wd, float bc1, float bc2, float gscale) {
  const float step = lr * sqrtf(bc2) / bc1;
  int64_t done = 0;
  if (VEC) {
    constexpr int kU = 2;
    const int64_t n4 = n >> 2;
    const int64_t tile_elems = (int64_t)kU * kThreads;
    float4* p4 = reinterpret_cast<float4*>(p);
    const float4* g4 = reinterpret_cast<const float4*>(g);
    float4* m4 = reinterpret_cast<float4*>(m);
    float4* v4 = reinterpret_cast<float4*>(v);
    for (int64_t tile = blockIdx.x; tile * tile_elems < n4; tile += gridDim.x) {
      const int64_t base = tile * tile_elems + threadIdx.x;
      float4 pp[kU], gg[kU], mm[kU], vv[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const int64_t i = base + (int64_t)u * kThreads;
        if (i < n4) { pp[u] = p4[i]; gg[u] = g4[i]; mm[u] = m4[i]; vv[u] = v4[i]; }
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const int64_t i = base + (int64_t)u * kThreads;
        if (i < n4) {
          adamw_one(pp[u].x, gg[u].x, mm[u].x, vv[u].x, step, lr, b1, b2, eps, wd, gscale);
          adamw_one(pp[u].y, gg[u].y, mm[u].y, vv[u].y, step, lr, b1, b2, eps, wd, gscale);
          adamw_one(pp[u].z, gg[u].z, mm[u].z, vv[u].z, step, lr, b1, b2, eps, wd, gscale);
          adamw_one(pp[u].w, gg[u].w, mm[u].w, vv[u].w, step, lr, b1, b2, eps, wd, gscale);
          p4[i] = pp[u]; m4[i] = mm[u]; v4[i] = vv[u];
        }
      }
    }
    done = n4 << 2;
  }
  for (int64_t i = done + (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads)
    adamw_one(p[i], g[i], m[i], v[i], step, lr, b1, b2, eps, wd, gscale);
}

}  // namespace

int pn_t5_rs(const float* ss, int parts, int64_t rows, int d, float eps, float* rs, cudaStream_t st) {
  if (rows == 0) return PN_OK;
  rs_kernel<<<pn_grid(rows, kThreads, 1, 8), kThreads, 0, st>>>(ss, parts, rows, d, eps, rs);
  PN_LAUNCHED();
  return PN_OK;
}

int pn_t5_relu_mask_scale(__nv_bfloat16* y, const __nv_bfloat16* act, const float* rs, int64_t rows, int cols,
                          cudaStream_t st) {
  if (rows == 0) return PN_OK;
  if (cols % 8) return pn_fail(PN_E_UNSUPPORTED, "relu_mask_scale: cols %% 8 != 0");
  relu_mask_scale_kernel<<<pn_grid(rows * cols / 8, kThreads, 2, 8), kThreads, 0, st>>>(y, act, rs, rows, cols);
  PN_LAUNCHED();
  return PN_OK;
}

int pn_t5_row_scale(__nv_bfloat16* y, int ld, const float* rs, int64_t rows, int cols, cudaStream_t st) {
  if (rows == 0) return PN_OK;
  if ((cols % 8) || (ld % 8)) return pn_fail(PN_E_UNSUPPORTED, "row_scale: cols/ld %% 8 != 0");
  row_scale_kernel<<<pn_grid(rows * (cols / 8), kThreads, 2, 8), kThreads, 0, st>>>(y, ld, rs, rows, cols);
  PN_LAUNCHED();
  return PN_OK;
}

int pn_t5_norm_bwd(const void* G, int g_bf16, const __nv_bfloat16* hb, const float* rs, int64_t rows, int d,
                   int accumulate, float* dh, __nv_bfloat16* dhb, cudaStream_t st, PnDrop gdrop, PnDrop odrop) {
  if (rows == 0) return PN_OK;
  if (d % 4 || d > 32 * 4 * 16) return pn_fail(PN_E_UNSUPPORTED, "norm_bwd: d_model=%d unsupported", d);
  const int grid = pn_grid(rows * 32, kThreads, 1, 8);
#define PN_NB(V, B) norm_bwd_kernel<V, B><<<grid, kThreads, 0, st>>>(G, hb, rs, rows, d, accumulate, dh, dhb, gdrop, odrop)
  if (d <= 32 * 4 * 4) { if (g_bf16) PN_NB(4, true); else PN_NB(4, false); }
  else if (d <= 32 * 4 * 6) { if (g_bf16) PN_NB(6, true); else PN_NB(6, false); }
  else if (d <= 32 * 4 * 8) { if (g_bf16) PN_NB(8, true); else PN_NB(8, false); }
  else { if (g_bf16) PN_NB(16, true); else PN_NB(16, false); }
#undef PN_NB
  PN_LAUNCHED();
  return PN_OK;
}

int pn_t5_unfold(const float* dWp, const float* W, const float* g, int64_t rows, int cols, float* gradW,
                 float* gradg, cudaStream_t st) {
  if (rows * cols == 0) return PN_OK;
  if (cols % 4 || ((reinterpret_cast<uintptr_t>(dWp) | reinterpret_cast<uintptr_t>(W) | reinterpret_cast<uintptr_t>(g) |
                    reinterpret_cast<uintptr_t>(gradW)) & 15))
    return pn_fail(PN_E_UNSUPPORTED, "unfold: cols %% 4 != 0 or operands not 16-byte aligned");
  dim3 grid((cols / 4 + 31) / 32, (unsigned)((rows + 31) / 32));
  unfold_kernel<<<grid, kThreads, 0, st>>>(dWp, W, g, rows, cols, gradW, gradg);
  PN_LAUNCHED();
  return PN_OK;
}

int pn_t5_embed_bwd(const int32_t* ids, const float* dh, int64_t rows, int d, int vocab, float* grad_table,
                    cudaStream_t st, PnDrop drop) {
  if (rows == 0) return PN_OK;
  embed_bwd_kernel<<<pn_grid(rows * 32, kThreads, 1, 8), kThreads, 0, st>>>(ids, dh, rows, d, vocab, grad_table, drop);
  PN_LAUNCHED();
  return PN_OK;
}

extern "C" int pn_t5_norm_coef(const float* part, int32_t parts, const float* rs, int64_t rows, int32_t d,
                               float* coef, void* stream) {
  PN_CHECK_ARG(part && rs && coef && parts > 0 && d > 0);
  return pn_t5_norm_coef(part, (int)parts, rs, rows, (int)d, coef, (cudaStream_t)stream, 1.0f);
}

int pn_t5_norm_coef(const float* part, int parts, const float* rs, int64_t rows, int d, float* coef, cudaStream_t st,
                    float scale) {
  if (rows <= 0) return PN_OK;
  norm_coef_kernel<<<(unsigned)std::min<int64_t>((rows + kThreads - 1) / kThreads, 4 * PN_SM_COUNT), kThreads, 0, st>>>(
      part, parts, rs, rows, scale / (float)d, coef);
  PN_LAUNCHED();
  return PN_OK;
}

int pn_t5_final_norm_bwd(const float* dy, const float* x, const float* w, int64_t rows, int d, float eps,
                         float* dx, __nv_bfloat16* dxb, float* dw, cudaStream_t st, PnDrop ydrop, PnDrop odrop) {
  if (rows == 0) return PN_OK;
  final_norm_bwd_kernel<<<(unsigned)((rows * 32 + kThreads - 1) / kThreads), kThreads, 0, st>>>(dy, x, w, rows, d,
                                                                                               eps, dx, dxb, dw,
                                                                                               ydrop, odrop);
  PN_LAUNCHED();
  return PN_OK;
}

namespace {
__global__ void __launch_bounds__(kThreads)
sgd_kernel(float* __restrict__ p, const float* __restrict__ g, int64_t n, float lr) {
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads)
    p[i] -= lr * g[i];
}
}  // namespace

// p -= lr * g: the functional one-step copy of StdWrapper.update_params (pairwrapper.py:68-73), 12 B per parameter
int pn_t5_sgd(float* p, const float* g, int64_t n, float lr, cudaStream_t st) {
  if (n <= 0) return PN_OK;
  sgd_kernel<<<pn_grid(n, kThreads, 4, 8), kThreads, 0, st>>>(p, g, n, lr);
  PN_LAUNCHED();
  return PN_OK;
}

int pn_t5_adamw(float* p, const float* g, float* m, float* v, int64_t n, float lr, float b1, float b2, float eps,
                float wd, int64_t step, float gscale, cudaStream_t st) {
  if (n == 0) return PN_OK;
  const float bc1 = 1.0f - powf(b1, (float)step), bc2 = 1.0f - powf(b2, (float)step);
  const bool vec = (((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0;
  if (vec)
    adamw_kernel<true><<<pn_grid(n, kThreads, 8, 8), kThreads, 0, st>>>(p, g, m, v, n, lr, b1, b2, eps, wd, bc1, bc2, gscale);
  else
    adamw_kernel<false><<<pn_grid(n, kThreads, 4, 8), kThreads, 0, st>>>(p, g, m, v, n, lr, b1, b2, eps, wd, bc1, bc2, gscale);
  PN_LAUNCHED();
  return PN_OK;
}
