// monoT5 scorer handle: weights by HuggingFace state_dict name -> packed bf16 device layouts, and the
// forward pass that strings the kernels together (call sites lceT5.py:132-133,
// pairwrapper.py:223-247, distill/wrapper.py:34-38; arithmetic HF modeling_t5.py:992-1137).
//
// Data layout in HBM for a chunk of `docs` documents, rows = docs*S:
//   hidden  fp32 [rows, d]        residual stream (fp32 like the oracle; the bf16 copies feed the GEMMs)
//   xn      bf16 [rows, d]        RMSNorm output / final encoder states
//   ctx     bf16 [rows, d_inner]  attention output
//   big     bf16 [rows, max(3*d_inner, d_ff)]   fused q|k|v, FFN hidden, or cross k|v
//   decoder buffers are the same at rows = docs*T.
#include "pn_t5_model.cuh"

#include <math.h>
#include <stdlib.h>
#include <string.h>

namespace {

// HF:188-234 for the bidirectional (encoder) case in exact integer arithmetic; equals HF's fp32
// expression for every |rel| <= 700 (tests/test_oracle_golden.py::test_relative_bucket_vs_hf).
int rel_bucket_bidir(int rel, int num_buckets, int max_distance) {
  int nb = num_buckets / 2;
  int ret = rel > 0 ? nb : 0;
  int n = rel < 0 ? -rel : rel;
  const int max_exact = nb / 2;
  if (n < max_exact) return ret + n;
  // max_exact + floor( log(n/max_exact) / log(max_distance/max_exact) * (nb - max_exact) )
  int add;
  if (nb == 16 && max_distance == 128) {
    // T5's constants: 8 * log16(n/8) = log2(n^2/64), exact in integers
    const unsigned v = (unsigned)(n * n) / 64u;
    add = 31 - __builtin_clz(v);
  } else {
    const double v = log((double)n / max_exact) / log((double)max_distance / max_exact) * (nb - max_exact);
    add = (int)floor(v + 1e-9);
  }
  int val = max_exact + add;
  if (val > nb - 1) val = nb - 1;
  return ret + val;
}

// y = RMSNorm(h) . W^T with the norm fused: A = bf16 copy of h, W has the norm weight folded in, and the
// epilogue scales each row by rsqrt(mean(h^2) + eps) from the partial sums of squares `ss` ([parts][M])
int gemm_normed(const PnT5* m, const __nv_bfloat16* hb, const float* ss, int parts, const PnLin& W, void* out,
                int ldc, int M, int relu, cudaStream_t st) {
  PnGemmExtra ex = {};
  ex.sumsq_in = ss; ex.sumsq_parts = parts; ex.norm_dim = m->cfg.d_model; ex.eps = m->cfg.eps;
  return pn_gemm_bf16_ex(hb, W.cols, W.packed, W.cols, out, ldc, M, (int)W.rows, W.cols, PN_EPI_BF16, relu, &ex,
                         m->cta_group, st);
}
// h += x . W^T, also refreshing the bf16 copy of h and its per-n-tile sums of squares for the next fused norm
int gemm_resid(const PnT5* m, const __nv_bfloat16* x, const PnLin& W, float* h, __nv_bfloat16* hb, float* ss,
               int M, cudaStream_t st) {
  PnGemmExtra ex = {};
  ex.copy_out = hb; ex.ld_copy = (int)W.rows; ex.sumsq_out = ss;
  return pn_gemm_bf16_ex(x, W.cols, W.packed, W.cols, h, (int)W.rows, M, (int)W.rows, W.cols, PN_EPI_RESID_F32, 0,
                         &ex, m->cta_group, st);
}

struct Workspace {
  float* hidden; __nv_bfloat16 *xn, *ctx, *big; float* ss;
  float* dh; __nv_bfloat16 *dxn, *dctx, *dbig; float* dss;
  int32_t* dec_ids;
  __nv_bfloat16* head_in; float* logits;
  // cross-attention without the K/V projection (see pn_t5_ops.cu): R = docs * Mp rows, Mp = T*H rounded up to 8
  float *rs_enc, *scores, *cprime;
  __nv_bfloat16 *qexp, *qk, *P, *cprime_b, *ctxfull;
};

size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }

size_t carve(const PnT5* m, int64_t docs, int S, int T, int want_logits, uint8_t* base, Workspace* w) {
  const int d = m->cfg.d_model, di = m->d_inner, ff = m->cfg.d_ff;
  const size_t rows = (size_t)docs * S, drows = (size_t)docs * T;
  const size_t bigc = (size_t)(3 * di > ff ? 3 * di : ff);
  const size_t parts = (size_t)pn_gemm_sumsq_parts(d);
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off += align_up(bytes); return base ? base + o : nullptr; };
  w->hidden = (float*)take(rows * d * 4);
  w->xn = (__nv_bfloat16*)take(rows * d * 2);
  w->ctx = (__nv_bfloat16*)take(rows * di * 2);
  w->big = (__nv_bfloat16*)take(rows * bigc * 2);
  w->ss = (float*)take(rows * parts * 4);
  w->dh = (float*)take(drows * d * 4);
  w->dxn = (__nv_bfloat16*)take(drows * d * 2);
  w->dctx = (__nv_bfloat16*)take(drows * di * 2);
  w->dbig = (__nv_bfloat16*)take(drows * bigc * 2);
  w->dss = (float*)take(drows * parts * 4);
  w->dec_ids = (int32_t*)take(drows * 4);
  w->head_in = (__nv_bfloat16*)take(drows * d * 2);
  w->logits = (float*)take(want_logits ? drows * (size_t)m->vocab_pad * 4 : 0);
  const size_t Mp = (size_t)((T * m->cfg.n_heads + 7) & ~7), R = (size_t)docs * Mp;
  w->rs_enc = (float*)take(rows * 4);
  w->scores = (float*)take(rows * 128 * 4);
  w->cprime = (float*)take(R * d * 4);
  w->qexp = (__nv_bfloat16*)take(R * di * 2);
  w->qk = (__nv_bfloat16*)take(R * d * 2);
  w->P = (__nv_bfloat16*)take(rows * Mp * 2);
  w->cprime_b = (__nv_bfloat16*)take(R * d * 2);
  w->ctxfull = (__nv_bfloat16*)take(R * di * 2);
  return off;
}

// Encoder: xn always holds the bf16 copy of the residual stream and ss its row sums of squares, so every
// RMSNorm is fused into the GEMM that consumes it (no standalone norm kernel, no extra pass over hidden).
// On return xn/ss describe the PRE-final-norm encoder output; the final norm is folded into the cross K/V
// projections of the decoder.
int encoder_forward(const PnT5* m, const int32_t* ids, int64_t docs, int S, const Workspace& w,
                    cudaStream_t st, const int32_t* lens = nullptr) {
  const int d = m->cfg.d_model, di = m->d_inner, ff = m->cfg.d_ff, H = m->cfg.n_heads;
  const int rows = (int)(docs * S);
  const int parts = pn_gemm_sumsq_parts(d);
  RC(pn_t5_embed(ids, m->emb, rows, d, m->cfg.vocab, w.hidden, w.xn, w.ss, st));
  int cur_parts = 1;
  for (size_t l = 0; l < m->enc.size(); ++l) {
    const PnEncLayer& L = m->enc[l];
    RC(gemm_normed(m, w.xn, w.ss, cur_parts, L.wqkv, w.big, 3 * di, rows, 0, st));
    RC(pn_t5_self_attn_masked(w.big, 3 * di, 0, di, 2 * di, (int)docs, S, H, m->enc_bias, lens, w.ctx, di, nullptr,
                              st));
    RC(gemm_resid(m, w.ctx, L.wo, w.hidden, w.xn, w.ss, rows, st));
    cur_parts = parts;
    RC(gemm_normed(m, w.xn, w.ss, cur_parts, L.wi, w.big, ff, rows, 1, st));
    RC(gemm_resid(m, w.big, L.wo2, w.hidden, w.xn, w.ss, rows, st));
  }
  return PN_OK;
}

// decoder over T tokens per doc (dec_ids [docs*T]); leaves the pre-final-norm hidden in w.dh
int decoder_forward(const PnT5* m, int64_t docs, int S, int T, const Workspace& w, cudaStream_t st,
                    const int32_t* lens = nullptr) {
  const int d = m->cfg.d_model, di = m->d_inner, ff = m->cfg.d_ff, H = m->cfg.n_heads;
  const int rows = (int)(docs * S), drows = (int)(docs * T);
  const int parts = pn_gemm_sumsq_parts(d);
  const int enc_parts = m->enc.empty() ? 1 : parts;
  RC(pn_t5_embed(w.dec_ids, m->emb, drows, d, m->cfg.vocab, w.dh, w.dxn, w.dss, st));
  // Cross-attention: with T <= 2 decoder tokens the K/V projection of all S encoder rows is replaced by
  // scores = X (Wk^T q) and out = Wv (P^T X) (pn_t5_ops.cu); needs whole GEMM tiles per document.
  const int TH = T * H, Mp = (TH + 7) & ~7;
  const bool xattn = m->xattn && (S % 128) == 0 && Mp <= 32;
  const int R = (int)docs * Mp;
  if (xattn) RC(pn_t5_rs(w.ss, enc_parts, rows, d, m->cfg.eps, w.rs_enc, st));
  int cur_parts = 1;
  for (size_t l = 0; l < m->dec.size(); ++l) {
    const PnDecLayer& L = m->dec[l];
    // self-attention over the <= T decoder tokens (causal + decoder bias)
    RC(gemm_normed(m, w.dxn, w.dss, cur_parts, L.wqkv, w.dbig, 3 * di, drows, 0, st));
    RC(pn_t5_dec_attn(w.dbig, 3 * di, 0, w.dbig, w.dbig, 3 * di, di, 2 * di, (int)docs, T, T, H, 1,
                      m->dec_bias, w.dctx, di, st));
    RC(gemm_resid(m, w.dctx, L.wo, w.dh, w.dxn, w.dss, drows, st));
    cur_parts = parts;
    // cross-attention: K/V projection of all S encoder states (the 13% of FLOPs), no bias; the encoder's
    // final RMSNorm is fused into this GEMM
    RC(gemm_normed(m, w.dxn, w.dss, cur_parts, L.cq, w.dbig, di, drows, 0, st));
    if (xattn) {
      // QK[doc, (t,h), :] = Wk_h^T q[doc, t, h]  (B = the K half of the transposed packed weight, [d, d_inner])
      RC(pn_t5_cross_expand_q(w.dbig, di, docs, T, H, Mp, w.qexp, st));
      RC(pn_gemm_bf16_internal(w.qexp, di, L.ckv.packedT, 2 * di, w.qk, d, R, d, di, PN_EPI_BF16, 0, nullptr,
                               m->cta_group, st));
      // scores[row, (t,h)] = rs[row] * (x_row . QK[doc(row), (t,h)]): block-diagonal B, one QK per document
      PnGemmExtra ex = {};
      ex.row_scale = w.rs_enc; ex.group_rows = S; ex.group_b_rows = Mp;
      RC(pn_gemm_bf16_ex(w.xn, d, w.qk, d, w.scores, 128, rows, 128, d, PN_EPI_F32, 0, &ex, (S % 256) == 0 ? 2 : 1,
                         st));
      RC(pn_t5_cross_softmax(w.scores, 128, w.rs_enc, docs, S, TH, Mp, w.P, st, lens));
      // C'[doc] = P[doc]^T . X[doc]  (batched MN-major GEMM), then out = C' . Wv^T and the per-head diagonal blocks
      PN_CUDA(cudaMemsetAsync(w.cprime, 0, (size_t)R * d * 4, st));
      RC(pn_gemm_tn_batched(w.P, Mp, w.xn, d, w.cprime, d, Mp, d, S, (int)docs, st));
      RC(pn_t5_f32_to_bf16(w.cprime, w.cprime_b, (int64_t)R * d, st));
      RC(pn_gemm_bf16_internal(w.cprime_b, d, L.ckv.packed + (size_t)di * d, d, w.ctxfull, di, R, di, d, PN_EPI_BF16,
                               0, nullptr, m->cta_group, st));
      RC(pn_t5_cross_gather(w.ctxfull, Mp, docs, T, H, w.dctx, di, st));
    } else {
      RC(gemm_normed(m, w.xn, w.ss, enc_parts, L.ckv, w.big, 2 * di, rows, 0, st));
      RC(pn_t5_dec_attn(w.dbig, di, 0, w.big, w.big, 2 * di, 0, di, (int)docs, T, S, H, 0, nullptr,
                        w.dctx, di, st, lens));
    }
    RC(gemm_resid(m, w.dctx, L.co, w.dh, w.dxn, w.dss, drows, st));
    // FFN
    RC(gemm_normed(m, w.dxn, w.dss, cur_parts, L.wi, w.dbig, ff, drows, 1, st));
    RC(gemm_resid(m, w.dbig, L.wo2, w.dh, w.dxn, w.dss, drows, st));
  }
  return PN_OK;
}

__global__ void fill_i32_kernel(int32_t* p, int64_t n, int32_t v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    p[i] = v;
}
// decoder_input_ids = shift_right(labels): [start, labels[:, :-1]] with -100 -> pad (HF:595-614)
__global__ void shift_right_kernel(const int64_t* labels, int64_t docs, int T, int32_t start,
                                   int32_t pad, int32_t* out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < docs * T;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int t = (int)(i % T);
    int64_t v = t == 0 ? start : labels[i - 1];
    if (v == -100) v = pad;
    out[i] = (int32_t)v;
  }
}

// (re)build every packed bf16 operand from the flat fp32 parameters, with each RMSNorm weight folded into the
// linear that consumes the normalised activations (encoder final norm -> every decoder cross K/V projection),
// and re-expand the relative-position bias tables
__global__ void expand_bias_kernel(const float* __restrict__ enc_rel, const float* __restrict__ dec_rel,
                                   const int* __restrict__ bucket_of, int H, float* __restrict__ enc_bias,
                                   float* __restrict__ dec_bias) {
  const int W = 1023;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < H * W; i += gridDim.x * blockDim.x) {
    const int h = i / W, k = i % W;
    enc_bias[i] = enc_rel[bucket_of[k] * H + h];
  }
  // decoder (unidirectional): bucket = distance for distance < 16; T <= 2 here
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < H * 4; i += gridDim.x * blockDim.x) {
    const int h = i / 4, t = (i / 2) & 1, j = i & 1;
    const int dist = t - j > 0 ? t - j : 0;
    dec_bias[i] = dec_rel[dist * H + h];
  }
}

}  // namespace

int pn_t5_repack(PnT5* m, cudaStream_t st) {
  auto pk = [&](const PnLin& L) {
    return pn_t5_pack_weight(L.master, L.fold, L.rows, L.cols, L.packed, L.packedT, (int)L.rows, st);
  };
  for (auto& L : m->enc) { RC(pk(L.wqkv)); RC(pk(L.wo)); RC(pk(L.wi)); RC(pk(L.wo2)); }
  for (auto& L : m->dec) {
    RC(pk(L.wqkv)); RC(pk(L.wo)); RC(pk(L.cq)); RC(pk(L.ckv)); RC(pk(L.co)); RC(pk(L.wi)); RC(pk(L.wo2));
  }
  RC(pn_t5_pack_weight(m->emb, nullptr, m->cfg.vocab, m->cfg.d_model, m->emb_bf16, m->embT_bf16, m->vocab_pad, st));
  expand_bias_kernel<<<8, 256, 0, st>>>(m->enc_rel, m->dec_rel, (const int*)m->allocs[0], m->cfg.n_heads,
                                        m->enc_bias, m->dec_bias);
  PN_LAUNCHED();
  return PN_OK;
}

extern "C" int pn_t5_create(const PnT5Config* cfg, PnT5** out) {
  PN_CHECK_ARG(cfg && out);
  PN_CHECK_ARG(cfg->d_kv == 64);
  PN_CHECK_ARG(cfg->d_model % 128 == 0 && cfg->d_ff % 128 == 0);
  PN_CHECK_ARG((cfg->n_heads * cfg->d_kv) % 128 == 0);
  PN_CHECK_ARG(cfg->n_layers > 0 && cfg->n_dec_layers > 0 && cfg->vocab > 0);
  PN_CHECK_ARG(cfg->rel_buckets > 0 && cfg->rel_buckets <= 64);
  PnT5* m = new PnT5();
  m->cfg = *cfg;
  const int d = cfg->d_model, di = cfg->n_heads * cfg->d_kv, ff = cfg->d_ff, H = cfg->n_heads;
  m->d_inner = di;
  m->vocab_pad = (cfg->vocab + 127) / 128 * 128;
  m->enc.resize(cfg->n_layers);
  m->dec.resize(cfg->n_dec_layers);
  // ---- parameter table in HuggingFace state_dict order
  int64_t off = 0;
  auto add = [&](const std::string& name, int64_t numel) {
    m->index[name] = (int)m->params.size();
    m->params.push_back({name, off, numel});
    const int64_t o = off;
    off += numel;
    return o;
  };
  char b[160];
  m->emb_off = add("shared.weight", (int64_t)cfg->vocab * d);
  const int64_t dd = (int64_t)di * d;
  for (int i = 0; i < cfg->n_layers; ++i) {
    PnEncLayer& L = m->enc[i];
    auto nm = [&](const char* s) { snprintf(b, sizeof(b), "encoder.block.%d.layer.%s.weight", i, s); return std::string(b); };
    L.wqkv.off = add(nm("0.SelfAttention.q"), dd); add(nm("0.SelfAttention.k"), dd); add(nm("0.SelfAttention.v"), dd);
    L.wqkv.rows = 3 * di; L.wqkv.cols = d;
    L.wo.off = add(nm("0.SelfAttention.o"), dd); L.wo.rows = d; L.wo.cols = di;
    if (i == 0) m->enc_rel_off = add(nm("0.SelfAttention.relative_attention_bias"), (int64_t)cfg->rel_buckets * H);
    L.ln0_off = add(nm("0.layer_norm"), d);
    L.wi.off = add(nm("1.DenseReluDense.wi"), (int64_t)ff * d); L.wi.rows = ff; L.wi.cols = d;
    L.wo2.off = add(nm("1.DenseReluDense.wo"), (int64_t)ff * d); L.wo2.rows = d; L.wo2.cols = ff;
    L.ln1_off = add(nm("1.layer_norm"), d);
    L.wqkv.fold_off = L.ln0_off; L.wi.fold_off = L.ln1_off;
  }
  m->enc_final_off = add("encoder.final_layer_norm.weight", d);
  for (int i = 0; i < cfg->n_dec_layers; ++i) {
    PnDecLayer& L = m->dec[i];
    auto nm = [&](const char* s) { snprintf(b, sizeof(b), "decoder.block.%d.layer.%s.weight", i, s); return std::string(b); };
    L.wqkv.off = add(nm("0.SelfAttention.q"), dd); add(nm("0.SelfAttention.k"), dd); add(nm("0.SelfAttention.v"), dd);
    L.wqkv.rows = 3 * di; L.wqkv.cols = d;
    L.wo.off = add(nm("0.SelfAttention.o"), dd); L.wo.rows = d; L.wo.cols = di;
    if (i == 0) m->dec_rel_off = add(nm("0.SelfAttention.relative_attention_bias"), (int64_t)cfg->rel_buckets * H);
    L.ln0_off = add(nm("0.layer_norm"), d);
    L.cq.off = add(nm("1.EncDecAttention.q"), dd); L.cq.rows = di; L.cq.cols = d;
    L.ckv.off = add(nm("1.EncDecAttention.k"), dd); add(nm("1.EncDecAttention.v"), dd);
    L.ckv.rows = 2 * di; L.ckv.cols = d;
    L.co.off = add(nm("1.EncDecAttention.o"), dd); L.co.rows = d; L.co.cols = di;
    L.ln1_off = add(nm("1.layer_norm"), d);
    L.wi.off = add(nm("2.DenseReluDense.wi"), (int64_t)ff * d); L.wi.rows = ff; L.wi.cols = d;
    L.wo2.off = add(nm("2.DenseReluDense.wo"), (int64_t)ff * d); L.wo2.rows = d; L.wo2.cols = ff;
    L.ln2_off = add(nm("2.layer_norm"), d);
    L.wqkv.fold_off = L.ln0_off; L.cq.fold_off = L.ln1_off; L.wi.fold_off = L.ln2_off;
    L.ckv.fold_off = m->enc_final_off;
  }
  m->dec_final_off = add("decoder.final_layer_norm.weight", d);
  for (const char* alias : {"encoder.embed_tokens.weight", "decoder.embed_tokens.weight", "lm_head.weight"})
    m->index[alias] = 0;   // tied to shared.weight
  m->n_params = off;
  m->loaded.assign(m->params.size(), 0);
  *out = m;
  return PN_OK;
}

// device storage: the bucket-of-relative-position table (allocs[0]), the flat parameters, the packed operands
static int t5_allocate(PnT5* m) {
  if (m->flat) return PN_OK;
  const int H = m->cfg.n_heads, nb = m->cfg.rel_buckets;
  int* bucket_of = nullptr;
  RC(pn_dev_alloc(m, &bucket_of, 1023));   // must stay allocs[0]
  std::vector<int> bo(1023);
  for (int k = 0; k < 1023; ++k) bo[k] = rel_bucket_bidir(k - 511, nb, m->cfg.rel_max_dist);
  PN_CUDA(cudaMemcpy(bucket_of, bo.data(), 1023 * sizeof(int), cudaMemcpyHostToDevice));
  RC(pn_dev_alloc(m, &m->flat, (size_t)m->n_params));
  float* f = m->flat;
  auto bind = [&](PnLin& L, bool needT = false) -> int {
    L.master = f + L.off;
    L.fold = L.fold_off >= 0 ? f + L.fold_off : nullptr;
    RC(pn_dev_alloc(m, &L.packed, (size_t)L.rows * L.cols));
    if (m->want_train || needT) RC(pn_dev_alloc(m, &L.packedT, (size_t)L.rows * L.cols));
    return PN_OK;
  };
  for (auto& L : m->enc) {
    RC(bind(L.wqkv)); RC(bind(L.wo)); RC(bind(L.wi)); RC(bind(L.wo2));
    L.ln0 = f + L.ln0_off; L.ln1 = f + L.ln1_off;
  }
  for (auto& L : m->dec) {
    RC(bind(L.wqkv)); RC(bind(L.wo)); RC(bind(L.cq)); RC(bind(L.ckv, true)); RC(bind(L.co)); RC(bind(L.wi)); RC(bind(L.wo2));
    L.ln0 = f + L.ln0_off; L.ln1 = f + L.ln1_off; L.ln2 = f + L.ln2_off;
  }
  m->emb = f + m->emb_off;
  m->enc_final = f + m->enc_final_off; m->dec_final = f + m->dec_final_off;
  m->enc_rel = f + m->enc_rel_off; m->dec_rel = f + m->dec_rel_off;
  const size_t ve = (size_t)m->vocab_pad * m->cfg.d_model;
  RC(pn_dev_alloc(m, &m->emb_bf16, ve));
  PN_CUDA(cudaMemset(m->emb_bf16, 0, ve * 2));
  if (m->want_train) {
    RC(pn_dev_alloc(m, &m->embT_bf16, ve));
    PN_CUDA(cudaMemset(m->embT_bf16, 0, ve * 2));
  }
  RC(pn_dev_alloc(m, &m->enc_bias, (size_t)H * 1023));
  RC(pn_dev_alloc(m, &m->dec_bias, (size_t)H * 4));
  return PN_OK;
}

extern "C" int pn_t5_destroy(PnT5* m) {
  if (!m) return PN_OK;
  for (void* p : m->allocs) cudaFree(p);
  delete m;
  return PN_OK;
}

extern "C" int pn_t5_set_option(PnT5* m, const char* name, int64_t value) {
  PN_CHECK_ARG(m && name);
  if (!strcmp(name, "cta_group")) {
    PN_CHECK_ARG(value == 1 || value == 2);
    m->cta_group = (int)value;
    return PN_OK;
  }
  if (!strcmp(name, "xattn")) {   // 0: always project the encoder states to K/V for the cross-attention
    m->xattn = value != 0;
    return PN_OK;
  }
  if (!strcmp(name, "dgrad_bf16")) {   // 0: the encoder's data gradients reach norm_bwd as fp32
    m->dgrad_bf16 = value != 0;
    return PN_OK;
  }
  if (!strcmp(name, "fuse_norm_bwd")) {   // 0: the encoder's RMSNorm backward runs as its own bandwidth pass
    m->fuse_norm_bwd = value != 0;
    return PN_OK;
  }
  if (!strcmp(name, "grad_ranges")) {   // encoder layer groups handed to the gradient hook separately
    PN_CHECK_ARG(value >= 1 && value <= 64);
    m->grad_ranges = (int)value;
    return PN_OK;
  }
  if (!strcmp(name, "train")) {   // must precede the first pn_t5_load_weight: also keeps transposed operands
    if (m->flat) return pn_fail(PN_E_BADARG, "pn_t5_set_option: \"train\" must be set before weights are loaded");
    m->want_train = value != 0;
    return PN_OK;
  }
  return pn_fail(PN_E_BADARG, "pn_t5_set_option: unknown option %s", name);
}

// src: fp32 device tensor, contiguous, in HF's [out, in] layout.
extern "C" int pn_t5_load_weight(PnT5* m, const char* name, const float* src, int64_t numel,
                                 void* stream) {
  PN_CHECK_ARG(m && name && src);
  auto it = m->index.find(name);
  if (it == m->index.end())
    return pn_fail(PN_E_BADARG, "pn_t5_load_weight: unrecognised parameter name %s", name);
  const PnParam& p = m->params[it->second];
  if (numel != p.numel)
    return pn_fail(PN_E_BADARG, "pn_t5_load_weight: %s has %lld elements, expected %lld", name,
                   (long long)numel, (long long)p.numel);
  RC(t5_allocate(m));
  if (it->second == 0 && m->loaded[0] && strcmp(name, "shared.weight")) return PN_OK;  // tied copies
  PN_CUDA(cudaMemcpyAsync(m->flat + p.off, src, numel * sizeof(float), cudaMemcpyDeviceToDevice,
                          (cudaStream_t)stream));
  m->loaded[it->second] = 1;
  m->finalized = false;
  return PN_OK;
}

extern "C" int pn_t5_finalize(PnT5* m, void* stream) {
  PN_CHECK_ARG(m);
  // every parameter of a tied-embedding T5 must have been supplied
  for (size_t i = 0; i < m->params.size(); ++i)
    if (!m->loaded[i]) return pn_fail(PN_E_BADARG, "pn_t5_finalize: missing weight %s", m->params[i].name.c_str());
  RC(pn_t5_repack(m, (cudaStream_t)stream));
  m->finalized = true;
  return PN_OK;
}

/* flat parameter access (training, checkpoint export): HF order, one contiguous fp32 buffer */
extern "C" int64_t pn_t5_num_params(const PnT5* m) { return m ? m->n_params : -1; }
extern "C" int32_t pn_t5_num_tensors(const PnT5* m) { return m ? (int32_t)m->params.size() : -1; }
extern "C" int pn_t5_tensor_info(const PnT5* m, int32_t i, char* name_out, int32_t name_cap,
                                 int64_t* offset_out, int64_t* numel_out) {
  PN_CHECK_ARG(m && i >= 0 && i < (int32_t)m->params.size());
  const PnParam& p = m->params[i];
  if (name_out && name_cap > 0) snprintf(name_out, name_cap, "%s", p.name.c_str());
  if (offset_out) *offset_out = p.off;
  if (numel_out) *numel_out = p.numel;
  return PN_OK;
}
extern "C" float* pn_t5_params(PnT5* m) {
  if (!m || t5_allocate(m)) return nullptr;
  return m->flat;
}

extern "C" int64_t pn_t5_workspace_bytes(const PnT5* m, int64_t docs, int32_t S, int32_t T,
                                         int want_logits) {
  if (!m || docs < 0 || S < 1 || T < 1) return -1;
  Workspace w;
  return (int64_t)carve(m, docs, S, T, want_logits, nullptr, &w);
}

extern "C" int pn_t5_score(PnT5* m, const int32_t* ids, int64_t docs, int32_t S, int32_t rel_id,
                           int32_t nrel_id, float* score_out, float* z2_out, void* workspace,
                           int64_t workspace_bytes, void* stream) {
  return pn_t5_score_lens(m, ids, nullptr, docs, S, rel_id, nrel_id, score_out, z2_out, workspace, workspace_bytes,
                          stream);
}

// monoT5 re-rank score: encoder + ONE decoder step + two-row head (distill/wrapper.py:34-38); lens (nullable) =
// the tokenizer's attention_mask of right-padded prompts as per-document lengths
extern "C" int pn_t5_score_lens(PnT5* m, const int32_t* ids, const int32_t* lens, int64_t docs, int32_t S,
                                int32_t rel_id, int32_t nrel_id, float* score_out, float* z2_out, void* workspace,
                                int64_t workspace_bytes, void* stream) {
  PN_CHECK_ARG(m && m->finalized);
  PN_CHECK_ARG(docs >= 0 && S >= 1 && S <= 512);
  PN_CHECK_ARG(rel_id >= 0 && rel_id < m->cfg.vocab && nrel_id >= 0 && nrel_id < m->cfg.vocab);
  if (docs == 0) return PN_OK;
  PN_CHECK_ARG(ids && workspace && (score_out || z2_out));
  PN_CHECK_ARG(docs * (int64_t)S < (1ll << 31));
  Workspace w;
  const size_t need = carve(m, docs, S, 1, 0, (uint8_t*)workspace, &w);
  if ((int64_t)need > workspace_bytes)
    return pn_fail(PN_E_BADARG, "pn_t5_score: workspace %lld < %zu bytes", (long long)workspace_bytes, need);
  cudaStream_t st = (cudaStream_t)stream;
  RC(encoder_forward(m, ids, docs, S, w, st, lens));
  fill_i32_kernel<<<pn_grid(docs, 256, 1, 4), 256, 0, st>>>(w.dec_ids, docs, m->cfg.pad_id);
  PN_LAUNCHED();
  RC(decoder_forward(m, docs, S, 1, w, st, lens));
  RC(pn_t5_head2(w.dh, m->dec_final, m->emb, docs, m->cfg.d_model, m->cfg.eps, rel_id, nrel_id,
                 score_out, z2_out, st));
  return PN_OK;
}

// model(input_ids=, labels=).logits for T <= 2 label positions (lceT5.py:132-133): fp32 [docs*T, vocab_pad]
// (columns >= vocab are zero-weight padding; ld = vocab_pad).  logits_out may be NULL when only the
// two-row z2 [docs*T, 2] is wanted.
extern "C" int pn_t5_logits(PnT5* m, const int32_t* ids, const int64_t* labels, int64_t docs, int32_t S,
                            int32_t T, int32_t rel_id, int32_t nrel_id, float* logits_out,
                            float* z2_out, void* workspace, int64_t workspace_bytes, void* stream) {
  PN_CHECK_ARG(m && m->finalized);
  PN_CHECK_ARG(docs >= 0 && S >= 1 && S <= 512 && (T == 1 || T == 2));
  if (docs == 0) return PN_OK;
  PN_CHECK_ARG(ids && labels && workspace);
  Workspace w;
  const size_t need = carve(m, docs, S, T, 0, (uint8_t*)workspace, &w);
  if ((int64_t)need > workspace_bytes)
    return pn_fail(PN_E_BADARG, "pn_t5_logits: workspace %lld < %zu bytes", (long long)workspace_bytes, need);
  cudaStream_t st = (cudaStream_t)stream;
  RC(encoder_forward(m, ids, docs, S, w, st));
  shift_right_kernel<<<pn_grid(docs * T, 256, 1, 4), 256, 0, st>>>(labels, docs, T, m->cfg.pad_id,
                                                                  m->cfg.pad_id, w.dec_ids);
  PN_LAUNCHED();
  RC(decoder_forward(m, docs, S, T, w, st));
  const int drows = (int)(docs * T), d = m->cfg.d_model;
  if (z2_out)
    RC(pn_t5_head2(w.dh, m->dec_final, m->emb, drows, d, m->cfg.eps, rel_id, nrel_id, nullptr, z2_out, st));
  if (logits_out) {
    RC(pn_t5_head_prep(w.dh, m->dec_final, drows, d, m->cfg.eps, w.head_in, st));
    RC(pn_gemm_bf16_internal(w.head_in, d, m->emb_bf16, d, logits_out, m->vocab_pad, drows,
                             m->vocab_pad, d, PN_EPI_F32, 0, nullptr, m->cta_group, st));
  }
  return PN_OK;
}

extern "C" int32_t pn_t5_vocab_pad(const PnT5* m) { return m ? m->vocab_pad : -1; }

extern "C" int pn_t5_self_attention_lse(const void* qkv, int32_t docs, int32_t S, int32_t n_heads,
                                        const float* bias, void* out, float* lse2, void* stream) {
  PN_CHECK_ARG(qkv && bias && out && lse2 && docs >= 0 && n_heads > 0);
  if (docs == 0) return PN_OK;
  const int di = n_heads * 64;
  return pn_t5_self_attn((const __nv_bfloat16*)qkv, 3 * di, 0, di, 2 * di, docs, S, n_heads, bias,
                         (__nv_bfloat16*)out, di, lse2, (cudaStream_t)stream);
}

extern "C" int pn_t5_self_attention(const void* qkv, int32_t docs, int32_t S, int32_t n_heads,
                                    const float* bias, void* out, void* stream) {
  PN_CHECK_ARG(qkv && bias && out && docs >= 0 && n_heads > 0);
  if (docs == 0) return PN_OK;
  const int di = n_heads * 64;
  return pn_t5_self_attn((const __nv_bfloat16*)qkv, 3 * di, 0, di, 2 * di, docs, S, n_heads, bias,
                         (__nv_bfloat16*)out, di, nullptr, (cudaStream_t)stream);
}
