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
#include "pn_common.cuh"
#include "pn_t5_internal.cuh"

#include <map>
#include <string>
#include <vector>
#include <math.h>
#include <string.h>

int pn_t5_embed(const int32_t* ids, const float* table, int64_t rows, int d, int vocab, float* hidden,
                __nv_bfloat16* copy, float* sumsq, cudaStream_t st);
int pn_t5_pack_weight(const float* master, const float* colscale, int64_t rows, int cols,
                      __nv_bfloat16* packed, cudaStream_t st);
int pn_t5_rmsnorm(const float* x, const float* w, int64_t rows, int d, float eps, __nv_bfloat16* out,
                  cudaStream_t st);
int pn_t5_f32_to_bf16(const float* src, __nv_bfloat16* dst, int64_t n, cudaStream_t st);
int pn_t5_head2(const float* x, const float* w, const float* emb, int64_t rows, int d, float eps,
                int rel_id, int nrel_id, float* score, float* z2, cudaStream_t st);
int pn_t5_head_prep(const float* x, const float* w, int64_t rows, int d, float eps,
                    __nv_bfloat16* out, cudaStream_t st);

#define RC(x) do { int rc__ = (x); if (rc__) return rc__; } while (0)

namespace {

// One linear: the fp32 master in HF's [out, in] layout (what the optimiser updates) and the packed bf16 copy
// the tcgen05 GEMM reads, with the RMSNorm weight of its input folded into the columns (pn_t5_finalize).
struct Lin {
  float* master = nullptr;
  __nv_bfloat16* packed = nullptr;
  int64_t rows = 0;
  int cols = 0;
};
struct EncLayer {
  Lin wqkv, wo, wi, wo2;
  float *ln0 = nullptr, *ln1 = nullptr;
};
struct DecLayer {
  Lin wqkv, wo, cq, ckv, co, wi, wo2;
  float *ln0 = nullptr, *ln1 = nullptr, *ln2 = nullptr;
};

}  // namespace

struct PnT5 {
  PnT5Config cfg;
  int d_inner;
  std::vector<void*> allocs;
  float* emb = nullptr;              // fp32 [vocab, d]
  __nv_bfloat16* emb_bf16 = nullptr; // bf16 [vocab_pad, d] for the full-vocabulary head
  int vocab_pad = 0;
  std::vector<EncLayer> enc;
  std::vector<DecLayer> dec;
  float *enc_final = nullptr, *dec_final = nullptr;
  float *enc_rel = nullptr, *dec_rel = nullptr;  // [buckets, heads] as in HF
  float *enc_bias = nullptr;                     // [heads, 1023] expanded by (j - i + 511)
  float *dec_bias = nullptr;                     // [heads, 2, 2] for T <= 2
  std::map<std::string, int> loaded;
  bool finalized = false;
  int cta_group = 2;
};

namespace {

template <typename T>
int dev_alloc(PnT5* m, T** p, size_t n) {
  void* q = nullptr;
  cudaError_t e = cudaMalloc(&q, n * sizeof(T));
  if (e != cudaSuccess) return pn_fail(PN_E_OOM, "pn_t5: cudaMalloc(%zu) failed: %s", n * sizeof(T),
                                       cudaGetErrorString(e));
  m->allocs.push_back(q);
  *p = (T*)q;
  return PN_OK;
}

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
int gemm_normed(const PnT5* m, const __nv_bfloat16* hb, const float* ss, int parts, const Lin& W, void* out,
                int ldc, int M, int relu, cudaStream_t st) {
  PnGemmExtra ex = {};
  ex.sumsq_in = ss; ex.sumsq_parts = parts; ex.norm_dim = m->cfg.d_model; ex.eps = m->cfg.eps;
  return pn_gemm_bf16_ex(hb, W.cols, W.packed, W.cols, out, ldc, M, (int)W.rows, W.cols, PN_EPI_BF16, relu, &ex,
                         m->cta_group, st);
}
// h += x . W^T, also refreshing the bf16 copy of h and its per-n-tile sums of squares for the next fused norm
int gemm_resid(const PnT5* m, const __nv_bfloat16* x, const Lin& W, float* h, __nv_bfloat16* hb, float* ss,
               int M, cudaStream_t st) {
  PnGemmExtra ex = {};
  ex.copy_out = hb; ex.ld_copy = (int)W.rows; ex.sumsq_out = ss;
  return pn_gemm_bf16_ex(x, W.cols, W.packed, W.cols, h, (int)W.rows, M, (int)W.rows, W.cols, PN_EPI_RESID_F32, 0,
                         &ex, m->cta_group, st);
}

// bf16 GEMM operands from the fp32 masters, with each RMSNorm weight folded into the linear that consumes the
// normalised activations (encoder final norm -> every decoder layer's cross K/V projection)
int pack_all(PnT5* m, cudaStream_t st) {
  auto pk = [&](const Lin& L, const float* ln) { return pn_t5_pack_weight(L.master, ln, L.rows, L.cols, L.packed, st); };
  for (auto& L : m->enc) {
    RC(pk(L.wqkv, L.ln0)); RC(pk(L.wo, nullptr)); RC(pk(L.wi, L.ln1)); RC(pk(L.wo2, nullptr));
  }
  for (auto& L : m->dec) {
    RC(pk(L.wqkv, L.ln0)); RC(pk(L.wo, nullptr)); RC(pk(L.cq, L.ln1)); RC(pk(L.ckv, m->enc_final));
    RC(pk(L.co, nullptr)); RC(pk(L.wi, L.ln2)); RC(pk(L.wo2, nullptr));
  }
  return PN_OK;
}

struct Workspace {
  float* hidden; __nv_bfloat16 *xn, *ctx, *big; float* ss;
  float* dh; __nv_bfloat16 *dxn, *dctx, *dbig; float* dss;
  int32_t* dec_ids;
  __nv_bfloat16* head_in; float* logits;
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
  return off;
}

// Encoder: xn always holds the bf16 copy of the residual stream and ss its row sums of squares, so every
// RMSNorm is fused into the GEMM that consumes it (no standalone norm kernel, no extra pass over hidden).
// On return xn/ss describe the PRE-final-norm encoder output; the final norm is folded into the cross K/V
// projections of the decoder.
int encoder_forward(const PnT5* m, const int32_t* ids, int64_t docs, int S, const Workspace& w,
                    cudaStream_t st) {
  const int d = m->cfg.d_model, di = m->d_inner, ff = m->cfg.d_ff, H = m->cfg.n_heads;
  const int rows = (int)(docs * S);
  const int parts = pn_gemm_sumsq_parts(d);
  RC(pn_t5_embed(ids, m->emb, rows, d, m->cfg.vocab, w.hidden, w.xn, w.ss, st));
  int cur_parts = 1;
  for (size_t l = 0; l < m->enc.size(); ++l) {
    const EncLayer& L = m->enc[l];
    RC(gemm_normed(m, w.xn, w.ss, cur_parts, L.wqkv, w.big, 3 * di, rows, 0, st));
    RC(pn_t5_self_attn(w.big, 3 * di, 0, di, 2 * di, (int)docs, S, H, m->enc_bias, w.ctx, di, st));
    RC(gemm_resid(m, w.ctx, L.wo, w.hidden, w.xn, w.ss, rows, st));
    cur_parts = parts;
    RC(gemm_normed(m, w.xn, w.ss, cur_parts, L.wi, w.big, ff, rows, 1, st));
    RC(gemm_resid(m, w.big, L.wo2, w.hidden, w.xn, w.ss, rows, st));
  }
  return PN_OK;
}

// decoder over T tokens per doc (dec_ids [docs*T]); leaves the pre-final-norm hidden in w.dh
int decoder_forward(const PnT5* m, int64_t docs, int S, int T, const Workspace& w, cudaStream_t st) {
  const int d = m->cfg.d_model, di = m->d_inner, ff = m->cfg.d_ff, H = m->cfg.n_heads;
  const int rows = (int)(docs * S), drows = (int)(docs * T);
  const int parts = pn_gemm_sumsq_parts(d);
  const int enc_parts = m->enc.empty() ? 1 : parts;
  RC(pn_t5_embed(w.dec_ids, m->emb, drows, d, m->cfg.vocab, w.dh, w.dxn, w.dss, st));
  int cur_parts = 1;
  for (size_t l = 0; l < m->dec.size(); ++l) {
    const DecLayer& L = m->dec[l];
    // self-attention over the <= T decoder tokens (causal + decoder bias)
    RC(gemm_normed(m, w.dxn, w.dss, cur_parts, L.wqkv, w.dbig, 3 * di, drows, 0, st));
    RC(pn_t5_dec_attn(w.dbig, 3 * di, 0, w.dbig, w.dbig, 3 * di, di, 2 * di, (int)docs, T, T, H, 1,
                      m->dec_bias, w.dctx, di, st));
    RC(gemm_resid(m, w.dctx, L.wo, w.dh, w.dxn, w.dss, drows, st));
    cur_parts = parts;
    // cross-attention: K/V projection of all S encoder states (the 13% of FLOPs), no bias; the encoder's
    // final RMSNorm is fused into this GEMM
    RC(gemm_normed(m, w.dxn, w.dss, cur_parts, L.cq, w.dbig, di, drows, 0, st));
    RC(gemm_normed(m, w.xn, w.ss, enc_parts, L.ckv, w.big, 2 * di, rows, 0, st));
    RC(pn_t5_dec_attn(w.dbig, di, 0, w.big, w.big, 2 * di, 0, di, (int)docs, T, S, H, 0, nullptr,
                      w.dctx, di, st));
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

}  // namespace

extern "C" int pn_t5_create(const PnT5Config* cfg, PnT5** out) {
  PN_CHECK_ARG(cfg && out);
  PN_CHECK_ARG(cfg->d_kv == 64);
  PN_CHECK_ARG(cfg->d_model % 64 == 0 && cfg->d_ff % 128 == 0);
  PN_CHECK_ARG((cfg->n_heads * cfg->d_kv) % 128 == 0);
  PN_CHECK_ARG(cfg->n_layers > 0 && cfg->n_dec_layers > 0 && cfg->vocab > 0);
  PN_CHECK_ARG(cfg->rel_buckets > 0 && cfg->rel_buckets <= 64);
  PnT5* m = new PnT5();
  m->cfg = *cfg;
  m->d_inner = cfg->n_heads * cfg->d_kv;
  m->vocab_pad = (cfg->vocab + 127) / 128 * 128;
  m->enc.resize(cfg->n_layers);
  m->dec.resize(cfg->n_dec_layers);
  *out = m;
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
  return pn_fail(PN_E_BADARG, "pn_t5_set_option: unknown option %s", name);
}

// src: fp32 device tensor, contiguous, in HF's [out, in] layout.
extern "C" int pn_t5_load_weight(PnT5* m, const char* name, const float* src, int64_t numel,
                                 void* stream) {
  PN_CHECK_ARG(m && name && src);
  cudaStream_t st = (cudaStream_t)stream;
  const int d = m->cfg.d_model, di = m->d_inner, ff = m->cfg.d_ff, H = m->cfg.n_heads;
  std::string n(name);
  auto expect = [&](int64_t want) -> int {
    if (numel != want)
      return pn_fail(PN_E_BADARG, "pn_t5_load_weight: %s has %lld elements, expected %lld", name,
                     (long long)numel, (long long)want);
    return PN_OK;
  };
  // fp32 master [rows_total, cols]; this parameter fills rows [row_off, row_off + cnt / cols)
  auto put_lin = [&](Lin* L, int64_t rows_total, int cols, int64_t row_off, int64_t cnt) -> int {
    if (!L->master) {
      RC(dev_alloc(m, &L->master, (size_t)rows_total * cols));
      RC(dev_alloc(m, &L->packed, (size_t)rows_total * cols));
      L->rows = rows_total;
      L->cols = cols;
    }
    PN_CUDA(cudaMemcpyAsync(L->master + row_off * cols, src, cnt * sizeof(float), cudaMemcpyDeviceToDevice, st));
    m->finalized = false;
    return PN_OK;
  };
  auto put_f32 = [&](float** slot, int64_t cnt) -> int {
    if (!*slot) RC(dev_alloc(m, slot, (size_t)cnt));
    PN_CUDA(cudaMemcpyAsync(*slot, src, cnt * sizeof(float), cudaMemcpyDeviceToDevice, st));
    return PN_OK;
  };
  m->loaded[n] = 1;
  if (n == "shared.weight" || n == "encoder.embed_tokens.weight" ||
      n == "decoder.embed_tokens.weight" || n == "lm_head.weight") {
    RC(expect((int64_t)m->cfg.vocab * d));
    if (m->emb && n != "shared.weight") return PN_OK;  // tied copies
    RC(put_f32(&m->emb, numel));
    if (!m->emb_bf16) {
      RC(dev_alloc(m, &m->emb_bf16, (size_t)m->vocab_pad * d));
      PN_CUDA(cudaMemsetAsync(m->emb_bf16, 0, (size_t)m->vocab_pad * d * 2, st));
    }
    m->loaded["shared.weight"] = 1;
    return pn_t5_f32_to_bf16(src, m->emb_bf16, numel, st);
  }
  if (n == "encoder.final_layer_norm.weight") { RC(expect(d)); return put_f32(&m->enc_final, d); }
  if (n == "decoder.final_layer_norm.weight") { RC(expect(d)); return put_f32(&m->dec_final, d); }
  char stack[16], rest[96];
  int blk = -1, sub = -1;
  if (sscanf(name, "%7[a-z].block.%d.layer.%d.%95s", stack, &blk, &sub, rest) != 4)
    return pn_fail(PN_E_BADARG, "pn_t5_load_weight: unrecognised parameter name %s", name);
  const bool is_dec = !strcmp(stack, "decoder");
  if (!is_dec && strcmp(stack, "encoder"))
    return pn_fail(PN_E_BADARG, "pn_t5_load_weight: unrecognised parameter name %s", name);
  if (blk < 0 || blk >= (int)(is_dec ? m->dec.size() : m->enc.size()))
    return pn_fail(PN_E_BADARG, "pn_t5_load_weight: block index out of range in %s", name);
  std::string r(rest);
  const int64_t dd = (int64_t)di * d;
  if (r == "SelfAttention.relative_attention_bias.weight") {
    RC(expect((int64_t)m->cfg.rel_buckets * H));
    return put_f32(is_dec ? &m->dec_rel : &m->enc_rel, numel);
  }
  if (!is_dec) {
    EncLayer& L = m->enc[blk];
    if (sub == 0) {
      if (r == "SelfAttention.q.weight") { RC(expect(dd)); return put_lin(&L.wqkv, 3 * di, d, 0, dd); }
      if (r == "SelfAttention.k.weight") { RC(expect(dd)); return put_lin(&L.wqkv, 3 * di, d, di, dd); }
      if (r == "SelfAttention.v.weight") { RC(expect(dd)); return put_lin(&L.wqkv, 3 * di, d, 2 * di, dd); }
      if (r == "SelfAttention.o.weight") { RC(expect(dd)); return put_lin(&L.wo, d, di, 0, dd); }
      if (r == "layer_norm.weight") { RC(expect(d)); return put_f32(&L.ln0, d); }
    } else if (sub == 1) {
      if (r == "DenseReluDense.wi.weight") { RC(expect((int64_t)ff * d)); return put_lin(&L.wi, ff, d, 0, (int64_t)ff * d); }
      if (r == "DenseReluDense.wo.weight") { RC(expect((int64_t)ff * d)); return put_lin(&L.wo2, d, ff, 0, (int64_t)ff * d); }
      if (r == "layer_norm.weight") { RC(expect(d)); return put_f32(&L.ln1, d); }
    }
  } else {
    DecLayer& L = m->dec[blk];
    if (sub == 0) {
      if (r == "SelfAttention.q.weight") { RC(expect(dd)); return put_lin(&L.wqkv, 3 * di, d, 0, dd); }
      if (r == "SelfAttention.k.weight") { RC(expect(dd)); return put_lin(&L.wqkv, 3 * di, d, di, dd); }
      if (r == "SelfAttention.v.weight") { RC(expect(dd)); return put_lin(&L.wqkv, 3 * di, d, 2 * di, dd); }
      if (r == "SelfAttention.o.weight") { RC(expect(dd)); return put_lin(&L.wo, d, di, 0, dd); }
      if (r == "layer_norm.weight") { RC(expect(d)); return put_f32(&L.ln0, d); }
    } else if (sub == 1) {
      if (r == "EncDecAttention.q.weight") { RC(expect(dd)); return put_lin(&L.cq, di, d, 0, dd); }
      if (r == "EncDecAttention.k.weight") { RC(expect(dd)); return put_lin(&L.ckv, 2 * di, d, 0, dd); }
      if (r == "EncDecAttention.v.weight") { RC(expect(dd)); return put_lin(&L.ckv, 2 * di, d, di, dd); }
      if (r == "EncDecAttention.o.weight") { RC(expect(dd)); return put_lin(&L.co, d, di, 0, dd); }
      if (r == "layer_norm.weight") { RC(expect(d)); return put_f32(&L.ln1, d); }
    } else if (sub == 2) {
      if (r == "DenseReluDense.wi.weight") { RC(expect((int64_t)ff * d)); return put_lin(&L.wi, ff, d, 0, (int64_t)ff * d); }
      if (r == "DenseReluDense.wo.weight") { RC(expect((int64_t)ff * d)); return put_lin(&L.wo2, d, ff, 0, (int64_t)ff * d); }
      if (r == "layer_norm.weight") { RC(expect(d)); return put_f32(&L.ln2, d); }
    }
  }
  return pn_fail(PN_E_BADARG, "pn_t5_load_weight: unrecognised parameter name %s", name);
}

extern "C" int pn_t5_finalize(PnT5* m, void* stream) {
  PN_CHECK_ARG(m);
  cudaStream_t st = (cudaStream_t)stream;
  // every parameter of a tied-embedding T5 must have been supplied
  std::vector<std::string> need = {"shared.weight", "encoder.final_layer_norm.weight",
                                   "decoder.final_layer_norm.weight",
                                   "encoder.block.0.layer.0.SelfAttention.relative_attention_bias.weight",
                                   "decoder.block.0.layer.0.SelfAttention.relative_attention_bias.weight"};
  char buf[128];
  for (size_t i = 0; i < m->enc.size(); ++i) {
    for (const char* s : {"0.SelfAttention.q", "0.SelfAttention.k", "0.SelfAttention.v",
                          "0.SelfAttention.o", "0.layer_norm", "1.DenseReluDense.wi",
                          "1.DenseReluDense.wo", "1.layer_norm"}) {
      snprintf(buf, sizeof(buf), "encoder.block.%zu.layer.%s.weight", i, s);
      need.push_back(buf);
    }
  }
  for (size_t i = 0; i < m->dec.size(); ++i) {
    for (const char* s : {"0.SelfAttention.q", "0.SelfAttention.k", "0.SelfAttention.v",
                          "0.SelfAttention.o", "0.layer_norm", "1.EncDecAttention.q",
                          "1.EncDecAttention.k", "1.EncDecAttention.v", "1.EncDecAttention.o",
                          "1.layer_norm", "2.DenseReluDense.wi", "2.DenseReluDense.wo", "2.layer_norm"}) {
      snprintf(buf, sizeof(buf), "decoder.block.%zu.layer.%s.weight", i, s);
      need.push_back(buf);
    }
  }
  for (const auto& s : need)
    if (!m->loaded.count(s)) return pn_fail(PN_E_BADARG, "pn_t5_finalize: missing weight %s", s.c_str());
  // expand the bias tables on the host (tiny) from the [buckets, heads] embeddings
  const int H = m->cfg.n_heads, nb = m->cfg.rel_buckets;
  std::vector<float> rel((size_t)nb * H), drel((size_t)nb * H);
  PN_CUDA(cudaStreamSynchronize(st));
  PN_CUDA(cudaMemcpy(rel.data(), m->enc_rel, rel.size() * 4, cudaMemcpyDeviceToHost));
  PN_CUDA(cudaMemcpy(drel.data(), m->dec_rel, drel.size() * 4, cudaMemcpyDeviceToHost));
  const int W = 1023;
  std::vector<float> eb((size_t)H * W), db((size_t)H * 4);
  for (int h = 0; h < H; ++h) {
    for (int i = 0; i < W; ++i) {
      const int relpos = i - 511;  // j - i
      eb[(size_t)h * W + i] = rel[(size_t)rel_bucket_bidir(relpos, nb, m->cfg.rel_max_dist) * H + h];
    }
    // decoder (unidirectional): bucket = min(i - j, ...) = distance for distance < 16; T <= 2 here
    for (int t = 0; t < 2; ++t)
      for (int j = 0; j < 2; ++j) {
        const int dist = t - j > 0 ? t - j : 0;
        db[((size_t)h * 2 + t) * 2 + j] = drel[(size_t)dist * H + h];
      }
  }
  if (!m->enc_bias) RC(dev_alloc(m, &m->enc_bias, eb.size()));
  if (!m->dec_bias) RC(dev_alloc(m, &m->dec_bias, db.size()));
  PN_CUDA(cudaMemcpy(m->enc_bias, eb.data(), eb.size() * 4, cudaMemcpyHostToDevice));
  PN_CUDA(cudaMemcpy(m->dec_bias, db.data(), db.size() * 4, cudaMemcpyHostToDevice));
  RC(pack_all(m, st));
  m->finalized = true;
  return PN_OK;
}

extern "C" int64_t pn_t5_workspace_bytes(const PnT5* m, int64_t docs, int32_t S, int32_t T,
                                         int want_logits) {
  if (!m || docs < 0 || S < 1 || T < 1) return -1;
  Workspace w;
  return (int64_t)carve(m, docs, S, T, want_logits, nullptr, &w);
}

// monoT5 re-rank score: encoder + ONE decoder step + two-row head (distill/wrapper.py:34-38)
extern "C" int pn_t5_score(PnT5* m, const int32_t* ids, int64_t docs, int32_t S, int32_t rel_id,
                           int32_t nrel_id, float* score_out, float* z2_out, void* workspace,
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
  RC(encoder_forward(m, ids, docs, S, w, st));
  fill_i32_kernel<<<pn_grid(docs, 256, 1, 4), 256, 0, st>>>(w.dec_ids, docs, m->cfg.pad_id);
  PN_LAUNCHED();
  RC(decoder_forward(m, docs, S, 1, w, st));
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

extern "C" int pn_t5_self_attention(const void* qkv, int32_t docs, int32_t S, int32_t n_heads,
                                    const float* bias, void* out, void* stream) {
  PN_CHECK_ARG(qkv && bias && out && docs >= 0 && n_heads > 0);
  if (docs == 0) return PN_OK;
  const int di = n_heads * 64;
  return pn_t5_self_attn((const __nv_bfloat16*)qkv, 3 * di, 0, di, 2 * di, docs, S, n_heads, bias,
                         (__nv_bfloat16*)out, di, (cudaStream_t)stream);
}
