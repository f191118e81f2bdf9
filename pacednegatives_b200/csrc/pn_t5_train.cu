// Training pass of the monoT5 model (LCEModel.training_step, lceT5.py:124-167): the
// gradient-tracking forward `self.model(**p).logits` / `self.model(**n).logits` (lceT5.py:149-150), the
// autograd of `self.manual_backward(main_loss.mean())` (lceT5.py:154-155) and `opt.step()` (lceT5.py:156) with
// the optimiser of configure_optimizers (transformers.AdamW, lceT5.py:171-172).
//
// Same kernels and fused-norm formulation as the scorer (pn_t5.cu); what differs is that every GEMM input the
// backward pass needs is kept.  Saved per encoder layer, rows = docs*S:
//   hb[2l], hb[2l+1] bf16 [rows, d]  residual stream entering the attention / FFN sub-layer
//   rs[2l], rs[2l+1] fp32 [rows]     its RMSNorm scale rsqrt(mean(h^2) + eps)
//   qkv[l] bf16 [rows, 3*di],  lse[l] fp32 [rows, H] (base-2 logsumexp),  ctx[l] bf16 [rows, di],
//   act[l] bf16 [rows, ff] (post-ReLU)
// plus hb[2L], rs[2L] (encoder output before its final norm, which is folded into the cross K/V projections) and,
// per decoder layer, the cross-attention operands (QK, P, C' of the projection-free form, or ckv[l] bf16
// [rows, 2*di] when the shapes rule it out) and the decoder-row (docs*T) analogues.  Nothing [S, S]-sized is
// kept: the attention backward recomputes the probabilities from lse.
//
// Backward per linear y = rs * (hb W'^T) (see pn_t5_train_ops.cu for the algebra): one weight-gradient GEMM
// (MN-major tcgen05, split-K fp32 atomics), one data-gradient GEMM against the transposed packed weight, and a
// single bandwidth pass (norm_bwd) that folds the RMSNorm Jacobian into the residual-stream gradient.
#include "pn_t5_model.cuh"

#include <string.h>
#include <algorithm>

namespace {

size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }

struct TrainWs {
  // encoder rows
  float* hidden;                       // residual stream (forward); sum of cross K/V data gradients (backward)
  std::vector<__nv_bfloat16*> hb;      // 2L+1
  std::vector<float*> rs;              // 2L+1
  float* ss;
  std::vector<__nv_bfloat16*> qkv, ctx, act, ckv;
  std::vector<float*> lse;
  float *dH, *gbuf, *dvec;
  __nv_bfloat16 *dHb, *dbig, *dctx;
  float *dWtmp, *diag;
  // decoder rows
  float* dec_hidden;
  std::vector<__nv_bfloat16*> dhb;     // 3Ld+1
  std::vector<float*> drs;
  float* dss;
  std::vector<__nv_bfloat16*> dqkv, dctx_s, cq, dctx_c, dact;
  int32_t* dec_ids;
  __nv_bfloat16* head_in;
  float *ddH, *dgbuf;
  __nv_bfloat16 *ddHb, *ddbig, *ddctx, *dlb;
  // projection-free cross-attention (pn_t5_ops.cu): R = docs*Mp rows, Mp = T*H rounded up to 8
  bool xattn;
  int Mp;
  std::vector<__nv_bfloat16*> qk, Pp, cpb;   // per decoder layer: QK [R, d], P*rs [rows, Mp], C' [R, d]
  std::vector<__nv_bfloat16*> Pc;            // with dropout: the clean P*rs (Pp then carries the mask)
  float *scores, *cprime;                   // [rows, 128], [R, d]
  __nv_bfloat16 *qexp, *ctxfull, *dOfull, *dCp, *V2, *dQKb, *A_all, *B_all;
  __nv_bfloat16* xdrop;   // dropped bf16 copy of the encoder output (dropout of the encoder's final norm), or null
  int64_t* header;   // reserved
};

size_t carve_train(const PnT5* m, int64_t docs, int S, int T, uint8_t* base, TrainWs* w, bool dropping) {
  const int d = m->cfg.d_model, di = m->d_inner, ff = m->cfg.d_ff, H = m->cfg.n_heads;
  const size_t L = m->enc.size(), Ld = m->dec.size();
  const size_t rows = (size_t)docs * S, drows = (size_t)docs * T;
  const size_t bigc = (size_t)(3 * di > ff ? 3 * di : ff);
  const size_t parts = (size_t)pn_gemm_sumsq_parts(d);
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off += align_up(bytes); return base ? base + o : nullptr; };
  w->header = (int64_t*)take(64);
  w->hidden = (float*)take(rows * d * 4);
  w->hb.resize(2 * L + 1); w->rs.resize(2 * L + 1);
  for (auto& p : w->hb) p = (__nv_bfloat16*)take(rows * d * 2);
  for (auto& p : w->rs) p = (float*)take(rows * 4);
  w->ss = (float*)take(rows * parts * 4);
  w->qkv.resize(L); w->ctx.resize(L); w->act.resize(L); w->lse.resize(L); w->ckv.resize(Ld);
  for (auto& p : w->qkv) p = (__nv_bfloat16*)take(rows * 3 * di * 2);
  for (auto& p : w->ctx) p = (__nv_bfloat16*)take(rows * di * 2);
  for (auto& p : w->act) p = (__nv_bfloat16*)take(rows * ff * 2);
  for (auto& p : w->lse) p = (float*)take(rows * H * 4);
  w->Mp = (T * H + 7) & ~7;
  w->xattn = m->xattn && (S % 128) == 0 && w->Mp <= 32;
  w->xdrop = dropping ? (__nv_bfloat16*)take(rows * d * 2) : nullptr;
  const size_t Mp = (size_t)w->Mp, R = (size_t)docs * Mp;
  if (!w->xattn) {
    for (auto& p : w->ckv) p = (__nv_bfloat16*)take(rows * 2 * di * 2);
  } else {
    w->qk.resize(Ld); w->Pp.resize(Ld); w->cpb.resize(Ld);
    for (auto& p : w->qk) p = (__nv_bfloat16*)take(R * d * 2);
    for (auto& p : w->Pp) p = (__nv_bfloat16*)take(rows * Mp * 2);
    w->Pc.clear();
    if (dropping) {
      w->Pc.resize(Ld);
      for (auto& p : w->Pc) p = (__nv_bfloat16*)take(rows * Mp * 2);
    }
    for (auto& p : w->cpb) p = (__nv_bfloat16*)take(R * d * 2);
    w->scores = (float*)take(rows * 128 * 4);
    w->cprime = (float*)take(R * d * 4);
    w->qexp = (__nv_bfloat16*)take(R * di * 2);
    w->ctxfull = (__nv_bfloat16*)take(R * di * 2);
    w->dOfull = (__nv_bfloat16*)take(R * di * 2);
    w->dCp = (__nv_bfloat16*)take(R * d * 2);
    w->V2 = (__nv_bfloat16*)take(rows * Mp * 2);
    w->dQKb = (__nv_bfloat16*)take(R * d * 2);
    w->A_all = (__nv_bfloat16*)take(rows * 64 * Ld * 2);
    w->B_all = (__nv_bfloat16*)take((size_t)docs * d * 64 * Ld * 2);
  }
  w->dH = (float*)take(rows * d * 4);
  w->gbuf = (float*)take(rows * d * 4);
  w->dvec = (float*)take(rows * H * 4);
  w->dHb = (__nv_bfloat16*)take(rows * d * 2);
  w->dbig = (__nv_bfloat16*)take(rows * bigc * 2);
  w->dctx = (__nv_bfloat16*)take(rows * di * 2);
  w->dWtmp = (float*)take(bigc * d * 4);
  w->diag = (float*)take((size_t)H * 1023 * 4);
  w->dec_hidden = (float*)take(drows * d * 4);
  w->dhb.resize(3 * Ld + 1); w->drs.resize(3 * Ld + 1);
  for (auto& p : w->dhb) p = (__nv_bfloat16*)take(drows * d * 2);
  for (auto& p : w->drs) p = (float*)take(drows * 4);
  w->dss = (float*)take(drows * parts * 4);
  w->dqkv.resize(Ld); w->dctx_s.resize(Ld); w->cq.resize(Ld); w->dctx_c.resize(Ld); w->dact.resize(Ld);
  for (auto& p : w->dqkv) p = (__nv_bfloat16*)take(drows * 3 * di * 2);
  for (auto& p : w->dctx_s) p = (__nv_bfloat16*)take(drows * di * 2);
  for (auto& p : w->cq) p = (__nv_bfloat16*)take(drows * di * 2);
  for (auto& p : w->dctx_c) p = (__nv_bfloat16*)take(drows * di * 2);
  for (auto& p : w->dact) p = (__nv_bfloat16*)take(drows * ff * 2);
  w->dec_ids = (int32_t*)take(drows * 4);
  w->head_in = (__nv_bfloat16*)take(drows * d * 2);
  w->ddH = (float*)take(drows * d * 4);
  w->dgbuf = (float*)take(drows * d * 4);
  w->ddHb = (__nv_bfloat16*)take(drows * d * 2);
  w->ddbig = (__nv_bfloat16*)take(drows * bigc * 2);
  w->ddctx = (__nv_bfloat16*)take(drows * di * 2);
  w->dlb = (__nv_bfloat16*)take(drows * (size_t)m->vocab_pad * 2);
  return off;
}

// decoder_input_ids = shift_right(labels): [start, labels[:, :-1]] with -100 -> pad (HF:595-614)
__global__ void shift_right_train_kernel(const int64_t* labels, int64_t docs, int T, int32_t start, int32_t pad,
                                         int32_t* out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < docs * T;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int t = (int)(i % T);
    int64_t v = t == 0 ? start : labels[i - 1];
    if (v == -100) v = pad;
    out[i] = (int32_t)v;
  }
}

// y = rs (.) (hb W'^T), bf16 out
int lin_fwd(const PnT5* m, const __nv_bfloat16* hb, const float* rs, const PnLin& W, __nv_bfloat16* out, int M,
            int relu, cudaStream_t st, PnDrop drop = PnDrop()) {
  PnGemmExtra ex = {};
  ex.row_scale = rs; ex.drop = drop;
  return pn_gemm_bf16_ex(hb, W.cols, W.packed, W.cols, out, (int)W.rows, M, (int)W.rows, W.cols, PN_EPI_BF16, relu,
                         &ex, m->cta_group, st);
}
// h += x W^T; hb_next = bf16(h); rs_next from the epilogue's row sums of squares
int lin_resid(const PnT5* m, const __nv_bfloat16* x, const PnLin& W, float* h, __nv_bfloat16* hb_next, float* ss,
              float* rs_next, int M, cudaStream_t st, PnDrop drop = PnDrop()) {
  PnGemmExtra ex = {};
  ex.drop = drop;
  ex.copy_out = hb_next; ex.ld_copy = (int)W.rows; ex.sumsq_out = ss;
  RC(pn_gemm_bf16_ex(x, W.cols, W.packed, W.cols, h, (int)W.rows, M, (int)W.rows, W.cols, PN_EPI_RESID_F32, 0, &ex,
                     m->cta_group, st));
  return pn_t5_rs(ss, pn_gemm_sumsq_parts((int)W.rows), M, (int)W.rows, m->cfg.eps, rs_next, st);
}

// data gradient: out[M, in] = dY[M, out] . W  (B operand = the transposed packed weight [in, out])
int dgrad(const PnT5* m, const __nv_bfloat16* dY, int ld_dy, const PnLin& W, void* out, int mode, int M,
          cudaStream_t st) {
  return pn_gemm_bf16_ex(dY, ld_dy, W.packedT, (int)W.rows, out, W.cols, M, W.cols, (int)W.rows, mode, 0, nullptr,
                         m->cta_group, st);
}
// data gradient through Wo of the FFN with the ReLU backward and the fused-norm row scale in the epilogue:
// Ys = rs (.) (dY Wo (.) [act > 0])
// (with dropout of the hidden activation the saved act is m s relu(u): [act > 0] = m [u > 0], times the scale s)
int dgrad_relu(const PnT5* m, const __nv_bfloat16* dY, const PnLin& W, const __nv_bfloat16* act, const float* rs,
               __nv_bfloat16* out, int M, cudaStream_t st, float act_scale = 1.0f, float* dot_out = nullptr) {
  PnGemmExtra ex = {};
  ex.row_scale = rs; ex.relu_mask = act; ex.ld_mask = W.cols; ex.out_scale = act_scale; ex.dot_out = dot_out;
  return pn_gemm_bf16_ex(dY, (int)W.rows, W.packedT, (int)W.rows, out, W.cols, M, W.cols, (int)W.rows, PN_EPI_BF16, 0,
                         &ex, m->cta_group, st);
}
// data gradient of a normed linear with the RMSNorm backward folded into the epilogue:
//   dh += Ys W' - coef (.) hb,  dhb = bf16(dh)        (coef from pn_t5_norm_coef)
int dgrad_norm(const PnT5* m, const __nv_bfloat16* Ys, int ld_ys, const PnLin& W, const __nv_bfloat16* hb,
               const float* coef, float* dh, __nv_bfloat16* dhb, int M, cudaStream_t st, PnDrop copy_drop = PnDrop()) {
  PnGemmExtra ex = {};
  ex.drop = copy_drop;
  ex.nb_coef = coef; ex.nb_hb = hb; ex.ld_hb = W.cols; ex.copy_out = dhb; ex.ld_copy = W.cols;
  return pn_gemm_bf16_ex(Ys, ld_ys, W.packedT, (int)W.rows, dh, W.cols, M, W.cols, (int)W.rows, PN_EPI_RESID_F32, 0, &ex,
                         m->cta_group, st);
}
// weight gradient of an un-normed linear: grads[W] += dY^T X
int wgrad_plain(const PnT5* m, const __nv_bfloat16* dY, int ld_dy, const __nv_bfloat16* X, const PnLin& W, int K,
                cudaStream_t st) {
  return pn_gemm_wgrad(dY, ld_dy, X, W.cols, m->grads + W.off, W.cols, (int)W.rows, W.cols, K, m->cta_group, st);
}
// weight gradient of a linear behind a fused RMSNorm: dW' = Ys^T hb, then unfold into grads[W] and grads[g]
int wgrad_folded(const PnT5* m, const __nv_bfloat16* Ys, int ld_ys, const __nv_bfloat16* hb, const PnLin& W,
                 float* tmp, int K, cudaStream_t st) {
  PN_CUDA(cudaMemsetAsync(tmp, 0, (size_t)W.rows * W.cols * 4, st));
  RC(pn_gemm_wgrad(Ys, ld_ys, hb, W.cols, tmp, W.cols, (int)W.rows, W.cols, K, m->cta_group, st));
  return pn_t5_unfold(tmp, W.master, W.fold, W.rows, W.cols, m->grads + W.off, m->grads + W.fold_off, st);
}

}  // namespace

extern "C" int64_t pn_t5_train_workspace_bytes(const PnT5* m, int64_t docs, int32_t S, int32_t T) {
  if (!m || docs < 0 || S < 1 || T < 1) return -1;
  TrainWs w;
  return (int64_t)carve_train(m, docs, S, T, nullptr, &w, m->drop_p > 0.0f);
}

extern "C" int pn_t5_set_train_buffers(PnT5* m, float* grads, float* adam_m, float* adam_v) {
  PN_CHECK_ARG(m && grads);
  if (!m->want_train) return pn_fail(PN_E_BADARG, "pn_t5_set_train_buffers: model was not created with option \"train\"");
  m->grads = grads; m->adam_m = adam_m; m->adam_v = adam_v;
  return PN_OK;
}

// logits_out fp32 [docs*T, vocab_pad] = self.model(input_ids=ids, labels=labels).logits with every activation the
// backward needs left in `workspace`
static int train_forward_impl(PnT5* m, const int32_t* ids, const int64_t* labels, int64_t docs, int32_t S,
                              int32_t T, float* logits_out, float* z2_out, int32_t rel_id, int32_t nrel_id,
                              void* workspace, int64_t workspace_bytes, void* stream) {
  PN_CHECK_ARG(m && m->finalized);
  if (!m->want_train) return pn_fail(PN_E_BADARG, "pn_t5_train_forward: model was not created with option \"train\"");
  PN_CHECK_ARG(docs >= 0 && S >= 1 && S <= 512 && (T == 1 || T == 2));
  if (docs == 0) return PN_OK;
  PN_CHECK_ARG(ids && labels && (logits_out || z2_out) && workspace);
  PN_CHECK_ARG(docs * (int64_t)S < (1ll << 31));
  if (z2_out) PN_CHECK_ARG(rel_id >= 0 && rel_id < m->cfg.vocab && nrel_id >= 0 && nrel_id < m->cfg.vocab);
  const float dp = m->drop_p;
  const uint64_t dseed = m->drop_seed;
  const long long doff = m->drop_doc_offset;
  TrainWs w;
  const size_t need = carve_train(m, docs, S, T, (uint8_t*)workspace, &w, dp > 0.0f);
  if ((int64_t)need > workspace_bytes)
    return pn_fail(PN_E_BADARG, "pn_t5_train_forward: workspace %lld < %zu bytes", (long long)workspace_bytes, need);
  cudaStream_t st = (cudaStream_t)stream;
  const int d = m->cfg.d_model, di = m->d_inner, H = m->cfg.n_heads;
  const int rows = (int)(docs * S), drows = (int)(docs * T);
  const size_t L = m->enc.size(), Ld = m->dec.size();
  int32_t dp_bits;
  memcpy(&dp_bits, &dp, 4);
  // host-side record: backward checks it without a sync and regenerates the same dropout masks
  m->fwd_ws[workspace] = {docs, (int64_t)S, (int64_t)T, (int64_t)dseed, (int64_t)dp_bits, (int64_t)doff};
  const int ff = m->cfg.d_ff;
  const long long Sp = S + (S & 1), Tp = T + (T & 1);
  // mask of dropout call (stack, layer, kind) over `per_doc` elements per document (pn_dropout.cuh)
  auto D = [&](int stack, int layer, int kind, long long per_doc) {
    return dp > 0.0f ? pn_drop_make(dseed, pn_drop_site(stack, layer, kind), dp, doff, per_doc) : PnDrop();
  };
  auto DR = D;   // the decoder loops name their layer `D`

  // ---- encoder
  RC(pn_t5_embed(ids, m->emb, rows, d, m->cfg.vocab, w.hidden, w.hb[0], w.ss, st, D(0, 0, PN_DROP_EMBED, (long long)S * d)));
  RC(pn_t5_rs(w.ss, 1, rows, d, m->cfg.eps, w.rs[0], st));
  for (size_t l = 0; l < L; ++l) {
    const PnEncLayer& E = m->enc[l];
    const int li = (int)l;
    RC(lin_fwd(m, w.hb[2 * l], w.rs[2 * l], E.wqkv, w.qkv[l], rows, 0, st));
    RC(pn_t5_self_attn_drop(w.qkv[l], 3 * di, 0, di, 2 * di, (int)docs, S, H, m->enc_bias, w.ctx[l], di, w.lse[l],
                            D(0, li, PN_DROP_SELF_P, (long long)H * S * Sp), st));
    RC(lin_resid(m, w.ctx[l], E.wo, w.hidden, w.hb[2 * l + 1], w.ss, w.rs[2 * l + 1], rows, st,
                 D(0, li, PN_DROP_SELF_OUT, (long long)S * d)));
    RC(lin_fwd(m, w.hb[2 * l + 1], w.rs[2 * l + 1], E.wi, w.act[l], rows, 1, st,
               D(0, li, PN_DROP_FF_HID, (long long)S * ff)));
    RC(lin_resid(m, w.act[l], E.wo2, w.hidden, w.hb[2 * l + 2], w.ss, w.rs[2 * l + 2], rows, st,
                 D(0, li, PN_DROP_FF_OUT, (long long)S * d)));
  }
  // dropout of the encoder's final norm: a dropped copy of the bf16 encoder output is what the cross K/V projections
  // read (the norm's row scale and weight stay in the GEMM, see mask_copy_kernel)
  const __nv_bfloat16* xsrc = w.hb[2 * L];
  if (dp > 0.0f) {
    RC(pn_t5_mask_copy(w.hb[2 * L], w.xdrop, (int64_t)rows * d, D(0, 0, PN_DROP_FINAL, (long long)S * d), st));
    xsrc = w.xdrop;
  }
  // ---- decoder over shift_right(labels)
  shift_right_train_kernel<<<pn_grid(docs * T, 256, 1, 4), 256, 0, st>>>(labels, docs, T, m->cfg.pad_id,
                                                                        m->cfg.pad_id, w.dec_ids);
  PN_LAUNCHED();
  RC(pn_t5_embed(w.dec_ids, m->emb, drows, d, m->cfg.vocab, w.dec_hidden, w.dhb[0], w.dss, st,
                 D(1, 0, PN_DROP_EMBED, (long long)T * d)));
  RC(pn_t5_rs(w.dss, 1, drows, d, m->cfg.eps, w.drs[0], st));
  for (size_t l = 0; l < Ld; ++l) {
    const PnDecLayer& D = m->dec[l];
    const int li = (int)l;
    RC(lin_fwd(m, w.dhb[3 * l], w.drs[3 * l], D.wqkv, w.dqkv[l], drows, 0, st));
    RC(pn_t5_dec_attn(w.dqkv[l], 3 * di, 0, w.dqkv[l], w.dqkv[l], 3 * di, di, 2 * di, (int)docs, T, T, H, 1,
                      m->dec_bias, w.dctx_s[l], di, st, nullptr, DR(1, li, PN_DROP_SELF_P, (long long)H * T * Tp)));
    RC(lin_resid(m, w.dctx_s[l], D.wo, w.dec_hidden, w.dhb[3 * l + 1], w.dss, w.drs[3 * l + 1], drows, st,
                 DR(1, li, PN_DROP_SELF_OUT, (long long)T * d)));
    RC(lin_fwd(m, w.dhb[3 * l + 1], w.drs[3 * l + 1], D.cq, w.cq[l], drows, 0, st));
    if (w.xattn) {
      // scores = X (Wk^T q), out = Wv (P^T X): no K/V projection of the encoder rows (pn_t5_ops.cu)
      const int Mp = w.Mp, R = (int)docs * Mp;
      RC(pn_t5_cross_expand_q(w.cq[l], di, docs, T, H, Mp, w.qexp, st));
      RC(pn_gemm_bf16_internal(w.qexp, di, D.ckv.packedT, 2 * di, w.qk[l], d, R, d, di, PN_EPI_BF16, 0, nullptr,
                               m->cta_group, st));
      PnGemmExtra ex = {};
      ex.row_scale = w.rs[2 * L]; ex.group_rows = S; ex.group_b_rows = Mp;
      RC(pn_gemm_bf16_ex(xsrc, d, w.qk[l], d, w.scores, 128, rows, 128, d, PN_EPI_F32, 0, &ex,
                         (S % 256) == 0 ? 2 : 1, st));
      RC(pn_t5_cross_softmax(w.scores, 128, w.rs[2 * L], docs, S, T * H, Mp, w.Pp[l], st, nullptr,
                             DR(1, li, PN_DROP_CROSS_P, (long long)H * T * Sp), H, w.Pc.empty() ? nullptr : w.Pc[l]));
      PN_CUDA(cudaMemsetAsync(w.cprime, 0, (size_t)R * d * 4, st));
      RC(pn_gemm_tn_batched(w.Pp[l], Mp, xsrc, d, w.cprime, d, Mp, d, S, (int)docs, st));
      RC(pn_t5_f32_to_bf16(w.cprime, w.cpb[l], (int64_t)R * d, st));
      RC(pn_gemm_bf16_internal(w.cpb[l], d, D.ckv.packed + (size_t)di * d, d, w.ctxfull, di, R, di, d, PN_EPI_BF16, 0,
                               nullptr, m->cta_group, st));
      RC(pn_t5_cross_gather(w.ctxfull, Mp, docs, T, H, w.dctx_c[l], di, st));
    } else {
      RC(lin_fwd(m, xsrc, w.rs[2 * L], D.ckv, w.ckv[l], rows, 0, st));
      RC(pn_t5_dec_attn(w.cq[l], di, 0, w.ckv[l], w.ckv[l], 2 * di, 0, di, (int)docs, T, S, H, 0, nullptr,
                        w.dctx_c[l], di, st, nullptr, DR(1, li, PN_DROP_CROSS_P, (long long)H * T * Sp)));
    }
    RC(lin_resid(m, w.dctx_c[l], D.co, w.dec_hidden, w.dhb[3 * l + 2], w.dss, w.drs[3 * l + 2], drows, st,
                 DR(1, li, PN_DROP_CROSS_OUT, (long long)T * d)));
    RC(lin_fwd(m, w.dhb[3 * l + 2], w.drs[3 * l + 2], D.wi, w.dact[l], drows, 1, st,
               DR(1, li, PN_DROP_FF_HID, (long long)T * ff)));
    RC(lin_resid(m, w.dact[l], D.wo2, w.dec_hidden, w.dhb[3 * l + 3], w.dss, w.drs[3 * l + 3], drows, st,
                 DR(1, li, PN_DROP_FF_OUT, (long long)T * d)));
  }
  // ---- tied head: logits = (RMSNorm(h) * d^-0.5) E^T  (HF:1107-1110); restricted mode: only the two rows of E
  // the re-rank score reads (distill/wrapper.py:34-38), in fp32
  if (z2_out)
    RC(pn_t5_head2(w.dec_hidden, m->dec_final, m->emb, drows, d, m->cfg.eps, rel_id, nrel_id, nullptr, z2_out, st,
                   DR(1, 0, PN_DROP_FINAL, (long long)T * d)));
  if (logits_out) {
    RC(pn_t5_head_prep(w.dec_hidden, m->dec_final, drows, d, m->cfg.eps, w.head_in, st,
                       DR(1, 0, PN_DROP_FINAL, (long long)T * d)));
    RC(pn_gemm_bf16_internal(w.head_in, d, m->emb_bf16, d, logits_out, m->vocab_pad, drows, m->vocab_pad, d,
                             PN_EPI_F32, 0, nullptr, m->cta_group, st));
  }
  return PN_OK;
}

extern "C" int pn_t5_train_forward(PnT5* m, const int32_t* ids, const int64_t* labels, int64_t docs, int32_t S,
                                   int32_t T, float* logits_out, void* workspace, int64_t workspace_bytes,
                                   void* stream) {
  PN_CHECK_ARG(logits_out);
  return train_forward_impl(m, ids, labels, docs, S, T, logits_out, nullptr, 0, 0, workspace, workspace_bytes, stream);
}

extern "C" int pn_t5_train_forward_z2(PnT5* m, const int32_t* ids, const int64_t* labels, int64_t docs, int32_t S,
                                      int32_t T, int32_t rel_id, int32_t nrel_id, float* z2_out, void* workspace,
                                      int64_t workspace_bytes, void* stream) {
  PN_CHECK_ARG(z2_out);
  return train_forward_impl(m, ids, labels, docs, S, T, nullptr, z2_out, rel_id, nrel_id, workspace, workspace_bytes,
                            stream);
}

// Dropout of the NEXT pn_t5_train_forward (and of the backward that consumes its workspace): p = the checkpoint's
// dropout_rate (HF T5 in train mode, as Lightning runs LCEModel), seed names the forward call, doc_offset = index of
// this call's first document inside the whole step's batch (masks are a function of the global document index)
extern "C" int pn_t5_set_dropout(PnT5* m, float p, uint64_t seed, int64_t doc_offset) {
  PN_CHECK_ARG(m && p >= 0.0f && p < 1.0f && doc_offset >= 0);
  m->drop_p = p;
  m->drop_seed = seed;
  m->drop_doc_offset = doc_offset;
  return PN_OK;
}

extern "C" int pn_t5_set_grad_hook(PnT5* m, void (*fn)(void*, int64_t, int64_t), void* user) {
  PN_CHECK_ARG(m);
  m->grad_hook = fn;
  m->grad_hook_user = user;
  return PN_OK;
}

namespace {
// first element of the first parameter whose HF name starts with `prefix` (n_params when there is none)
int64_t first_offset(const PnT5* m, const std::string& prefix) {
  for (const PnParam& p : m->params)
    if (p.name.compare(0, prefix.size(), prefix) == 0) return p.off;
  return m->n_params;
}
}  // namespace

// Accumulates d loss / d parameters into the gradient buffer, given dlogits = d loss / d logits (fp32
// [docs*T, vocab_pad]) for the forward pass whose activations `workspace` holds.
static int train_backward_impl(PnT5* m, const int32_t* ids, const float* dlogits, const float* dz2, int32_t rel_id,
                               int32_t nrel_id, int64_t docs, int32_t S, int32_t T, void* workspace,
                               int64_t workspace_bytes, void* stream) {
  PN_CHECK_ARG(m && m->finalized && m->want_train);
  if (!m->grads) return pn_fail(PN_E_BADARG, "pn_t5_train_backward: no gradient buffer (pn_t5_set_train_buffers)");
  PN_CHECK_ARG(docs >= 0 && S >= 1 && S <= 512 && (T == 1 || T == 2));
  if (docs == 0) return PN_OK;
  PN_CHECK_ARG(ids && (dlogits || dz2) && workspace);
  if (dz2) PN_CHECK_ARG(rel_id >= 0 && rel_id < m->cfg.vocab && nrel_id >= 0 && nrel_id < m->cfg.vocab);
  if (m->cfg.vocab % 8)
    return pn_fail(PN_E_UNSUPPORTED, "pn_t5_train_backward: vocab=%d must be a multiple of 8", m->cfg.vocab);
  cudaStream_t st = (cudaStream_t)stream;
  // the forward this workspace holds: its shape and the dropout (seed, p, document offset) it ran with
  uint64_t dseed = 0;
  float dp = 0.0f;
  long long doff = 0;
  {
    auto it = m->fwd_ws.find(workspace);
    if (it == m->fwd_ws.end() || it->second[0] != docs || it->second[1] != S || it->second[2] != T)
      return pn_fail(PN_E_BADARG, "pn_t5_train_backward: workspace does not hold a forward pass of this shape");
    dseed = (uint64_t)it->second[3];
    const int32_t bits = (int32_t)it->second[4];
    memcpy(&dp, &bits, 4);
    doff = it->second[5];
    m->fwd_ws.erase(it);
  }
  TrainWs w;
  const size_t need = carve_train(m, docs, S, T, (uint8_t*)workspace, &w, dp > 0.0f);
  if ((int64_t)need > workspace_bytes)
    return pn_fail(PN_E_BADARG, "pn_t5_train_backward: workspace %lld < %zu bytes", (long long)workspace_bytes, need);
  const long long Sp = S + (S & 1), Tp = T + (T & 1);
  auto DR = [&](int stack, int layer, int kind, long long per_doc) {
    return dp > 0.0f ? pn_drop_make(dseed, pn_drop_site(stack, layer, kind), dp, doff, per_doc) : PnDrop();
  };
  const float act_scale = dp > 0.0f ? 1.0f / (1.0f - dp) : 1.0f;
  const int d = m->cfg.d_model, di = m->d_inner, ff = m->cfg.d_ff, H = m->cfg.n_heads;
  const int rows = (int)(docs * S), drows = (int)(docs * T);
  const int L = (int)m->enc.size(), Ld = (int)m->dec.size();
  const int Vp = m->vocab_pad;
  float* G = m->grads;
  const int* bucket_of = (const int*)m->allocs[0];

  // ---- head: d head_in = dlogits E,  dE += dlogits^T head_in,  then the final decoder norm
  if (dz2) {
    // restricted head: z[r, c] = head_in[r, :] . E[id_c, :]  ->  d head_in = dz E[(rel, nrel)], dE[id_c] += dz^T head_in
    RC(pn_t5_head2_bwd(w.dec_hidden, m->dec_final, m->emb, dz2, drows, d, m->cfg.eps, rel_id, nrel_id, w.dgbuf,
                       G + m->emb_off, st, DR(1, 0, PN_DROP_FINAL, (long long)T * d)));
  } else {
    RC(pn_t5_f32_to_bf16(dlogits, w.dlb, (int64_t)drows * Vp, st));
    RC(pn_gemm_wgrad(w.dlb, Vp, w.head_in, d, G + m->emb_off, d, m->cfg.vocab, d, drows, m->cta_group, st));
    RC(pn_gemm_bf16_internal(w.dlb, Vp, m->embT_bf16, Vp, w.dgbuf, d, drows, d, Vp, PN_EPI_F32, 0, nullptr,
                             m->cta_group, st));
  }
  // every bf16 copy of the residual-stream gradient is written for ONE consumer — the output projection of the next
  // sub-layer down — whose forward output was dropped: the copy carries that sub-layer's mask (`odrop`)
  RC(pn_t5_final_norm_bwd(w.dgbuf, w.dec_hidden, m->dec_final, drows, d, m->cfg.eps, w.ddH, w.ddHb,
                          G + m->dec_final_off, st, DR(1, 0, PN_DROP_FINAL, (long long)T * d),
                          DR(1, Ld - 1, PN_DROP_FF_OUT, (long long)T * d)));

  // ---- decoder layers, last to first.  ddH / ddHb = gradient of the decoder residual stream
  for (int l = Ld - 1; l >= 0; --l) {
    const PnDecLayer& D = m->dec[l];
    const __nv_bfloat16 *hb0 = w.dhb[3 * l], *hb1 = w.dhb[3 * l + 1], *hb2 = w.dhb[3 * l + 2];
    const float *rs0 = w.drs[3 * l], *rs1 = w.drs[3 * l + 1], *rs2 = w.drs[3 * l + 2];
    // FFN: h3 = h2 + relu(norm(h2) Wi^T) Wo^T
    RC(wgrad_plain(m, w.ddHb, d, w.dact[l], D.wo2, drows, st));
    RC(dgrad_relu(m, w.ddHb, D.wo2, w.dact[l], rs2, w.ddbig, drows, st, act_scale));
    RC(wgrad_folded(m, w.ddbig, ff, hb2, D.wi, w.dWtmp, drows, st));
    RC(dgrad(m, w.ddbig, ff, D.wi, w.dgbuf, PN_EPI_F32, drows, st));
    RC(pn_t5_norm_bwd(w.dgbuf, 0, hb2, rs2, drows, d, 1, w.ddH, w.ddHb, st, PnDrop(),
                      DR(1, l, PN_DROP_CROSS_OUT, (long long)T * d)));
    // cross-attention: h2 = h1 + attn(norm(h1) Cq^T, norm_f(enc) Ckv^T) Co^T
    RC(wgrad_plain(m, w.ddHb, d, w.dctx_c[l], D.co, drows, st));
    RC(dgrad(m, w.ddHb, d, D.co, w.ddctx, PN_EPI_BF16, drows, st));
    if (w.xattn) {
      const int Mp = w.Mp, R = (int)docs * Mp, ldall = 64 * Ld;
      const float* rsE = w.rs[2 * L];
      const __nv_bfloat16* X = w.xdrop ? w.xdrop : w.hb[2 * L];   // the (dropped) encoder output the forward read
      // d out (head blocks back on the diagonal) -> dWv' = dO^T C',  dC' = dO Wv'
      RC(pn_t5_cross_expand_q(w.ddctx, di, docs, T, H, Mp, w.dOfull, st));
      PN_CUDA(cudaMemsetAsync(w.dWtmp, 0, (size_t)2 * di * d * 4, st));
      RC(pn_gemm_wgrad(w.dOfull, di, w.cpb[l], d, w.dWtmp + (size_t)di * d, d, di, d, R, m->cta_group, st));
      RC(pn_gemm_bf16_internal(w.dOfull, di, D.ckv.packedT + di, 2 * di, w.dCp, d, R, d, di, PN_EPI_BF16, 0, nullptr,
                               m->cta_group, st));
      // dP[row, (t,h)] = rs[row] * (x_row . dC'[doc, (t,h)])  (same block-diagonal product as the scores)
      PnGemmExtra ex = {};
      ex.row_scale = rsE; ex.group_rows = S; ex.group_b_rows = Mp;
      RC(pn_gemm_bf16_ex(X, d, w.dCp, d, w.scores, 128, rows, 128, d, PN_EPI_F32, 0, &ex, (S % 256) == 0 ? 2 : 1, st));
      // softmax backward; its outputs and [dC'; QK] become column block l of the operands of ONE product over
      // all layers that yields the gradient of the normalised encoder states (after the loop)
      RC(pn_t5_cross_softmax_bwd(w.Pc.empty() ? w.Pp[l] : w.Pc[l], w.scores, 128, rsE, docs, S, T * H, Mp, w.A_all, ldall,
                                 64 * l, w.V2, st, DR(1, l, PN_DROP_CROSS_P, (long long)H * T * Sp), H));
      RC(pn_t5_cross_bt_pack(w.dCp, w.qk[l], docs, d, Mp, w.B_all, ldall, 64 * l, st));
      // dQK = (rs dSc)^T X -> dWk' = Qexp^T dQK,  dQexp = dQK Wk'^T -> dq (head blocks), scaled for the fused norm
      PN_CUDA(cudaMemsetAsync(w.cprime, 0, (size_t)R * d * 4, st));
      RC(pn_gemm_tn_batched(w.V2, Mp, X, d, w.cprime, d, Mp, d, S, (int)docs, st));
      RC(pn_t5_f32_to_bf16(w.cprime, w.dQKb, (int64_t)R * d, st));
      RC(pn_t5_cross_expand_q(w.cq[l], di, docs, T, H, Mp, w.qexp, st));
      RC(pn_gemm_wgrad(w.qexp, di, w.dQKb, d, w.dWtmp, d, di, d, R, m->cta_group, st));
      RC(pn_t5_unfold(w.dWtmp, D.ckv.master, D.ckv.fold, D.ckv.rows, D.ckv.cols, G + D.ckv.off, G + D.ckv.fold_off,
                      st));
      RC(pn_gemm_bf16_internal(w.dQKb, d, D.ckv.packed, d, w.ctxfull, di, R, di, d, PN_EPI_BF16, 0, nullptr,
                               m->cta_group, st));
      RC(pn_t5_cross_gather(w.ctxfull, Mp, docs, T, H, w.ddbig, di, st));
      RC(pn_t5_row_scale(w.ddbig, di, rs1, drows, di, st));
    } else {
      RC(pn_t5_dec_attn_bwd(w.cq[l], di, 0, w.ckv[l], w.ckv[l], 2 * di, 0, di, w.ddctx, di, (int)docs, T, S, H, 0,
                            nullptr, rs1, w.rs[2 * L], w.ddbig, di, 0, w.dbig, 2 * di, 0, di, nullptr, st,
                            DR(1, l, PN_DROP_CROSS_P, (long long)H * T * Sp)));
      RC(wgrad_folded(m, w.dbig, 2 * di, w.xdrop ? w.xdrop : w.hb[2 * L], D.ckv, w.dWtmp, rows, st));
      // every decoder layer reads the same normalised encoder output: their data gradients add up in `hidden`
      RC(dgrad(m, w.dbig, 2 * di, D.ckv, w.hidden, l == Ld - 1 ? PN_EPI_F32 : PN_EPI_RESID_F32, rows, st));
    }
    RC(wgrad_folded(m, w.ddbig, di, hb1, D.cq, w.dWtmp, drows, st));
    RC(dgrad(m, w.ddbig, di, D.cq, w.dgbuf, PN_EPI_F32, drows, st));
    RC(pn_t5_norm_bwd(w.dgbuf, 0, hb1, rs1, drows, d, 1, w.ddH, w.ddHb, st, PnDrop(),
                      DR(1, l, PN_DROP_SELF_OUT, (long long)T * d)));
    // self-attention: h1 = h0 + attn(norm(h0) Wqkv^T) Wo^T
    RC(wgrad_plain(m, w.ddHb, d, w.dctx_s[l], D.wo, drows, st));
    RC(dgrad(m, w.ddHb, d, D.wo, w.ddctx, PN_EPI_BF16, drows, st));
    RC(pn_t5_dec_attn_bwd(w.dqkv[l], 3 * di, 0, w.dqkv[l], w.dqkv[l], 3 * di, di, 2 * di, w.ddctx, di, (int)docs, T,
                          T, H, 1, m->dec_bias, rs0, rs0, w.ddbig, 3 * di, 0, w.ddbig, 3 * di, di, 2 * di,
                          G + m->dec_rel_off, st, DR(1, l, PN_DROP_SELF_P, (long long)H * T * Tp)));
    RC(wgrad_folded(m, w.ddbig, 3 * di, hb0, D.wqkv, w.dWtmp, drows, st));
    RC(dgrad(m, w.ddbig, 3 * di, D.wqkv, w.dgbuf, PN_EPI_F32, drows, st));
    RC(pn_t5_norm_bwd(w.dgbuf, 0, hb0, rs0, drows, d, 1, w.ddH, w.ddHb, st, PnDrop(),
                      l > 0 ? DR(1, l - 1, PN_DROP_FF_OUT, (long long)T * d) : PnDrop()));
    if (m->grad_hook && Ld >= 2 && l == Ld / 2) {   // decoder layers [Ld/2, Ld) and the final decoder norm are done
      const int64_t dmid = first_offset(m, "decoder.block." + std::to_string(l) + ".");
      m->grad_hook(m->grad_hook_user, dmid, m->n_params - dmid);
    }
  }
  RC(pn_t5_embed_bwd(w.dec_ids, w.ddH, drows, d, m->cfg.vocab, G + m->emb_off, st,
                     DR(1, 0, PN_DROP_EMBED, (long long)T * d)));
  // Gradient ranges in the order they become final (flat buffer = HF order: shared | encoder.block.* |
  // encoder.final_layer_norm | decoder.block.* | decoder.final_layer_norm): the upper decoder half (+ its final
  // norm) left above when decoder layer Ld/2 was done, the lower half now (decoder.block.0 holds the relative bias
  // every decoder layer adds to), then the encoder in `grad_ranges` groups of layers from the top (the first one
  // carries encoder.final_layer_norm, which every cross K/V projection adds to); the bottom group ends with the
  // embedding and the encoder's relative bias, final only after the last kernel.
  const int64_t dec_off = first_offset(m, "decoder.");
  if (m->grad_hook) {
    const int64_t dmid = Ld >= 2 ? first_offset(m, "decoder.block." + std::to_string(Ld / 2) + ".") : m->n_params;
    m->grad_hook(m->grad_hook_user, dec_off, dmid - dec_off);
  }
  const int groups = std::max(1, std::min(m->grad_ranges, L));
  // group g (0 = bottom) covers encoder layers [L*g/groups, L*(g+1)/groups)
  auto group_lo = [&](int g) { return (int)((int64_t)L * g / groups); };
  auto enc_layer_off = [&](int l) { return first_offset(m, "encoder.block." + std::to_string(l) + "."); };
  if (w.xattn) {
    // G = rs (.) ([P | dSc]_all-layers . [dC'; QK]_all-layers): every layer's cross-attention gradient of the
    // normalised encoder states in one block-diagonal GEMM (K = 64 per layer)
    PnGemmExtra ex = {};
    ex.row_scale = w.rs[2 * L]; ex.group_rows = S; ex.group_b_rows = d;
    RC(pn_gemm_bf16_ex(w.A_all, 64 * Ld, w.B_all, 64 * Ld, w.hidden, d, rows, d, 64 * Ld, PN_EPI_F32, 0, &ex,
                       (S % 256) == 0 ? 2 : 1, st));
  }

  // ---- encoder: the final norm (folded into every Ckv), then the layers last to first
  RC(pn_t5_norm_bwd(w.hidden, 0, w.hb[2 * L], w.rs[2 * L], rows, d, 0, w.dH, w.dHb, st,
                    DR(0, 0, PN_DROP_FINAL, (long long)S * d), DR(0, L - 1, PN_DROP_FF_OUT, (long long)S * d)));
  PN_CUDA(cudaMemsetAsync(w.diag, 0, (size_t)H * 1023 * 4, st));
  // G = Ys W' of the encoder's two normed linears goes to norm_bwd as bf16 ("dgrad_bf16", default): the bf16 epilogue
  // leaves by TMA store (the fp32 one through registers) and norm_bwd reads 2 B instead of 4 B per element; dh itself
  // stays fp32.  Parity at the bench widths: tests/test_gpu_train.py::test_training_step_at_bench_widths.
  const int g16 = m->dgrad_bf16 ? 1 : 0;
  const int gmode = g16 ? PN_EPI_BF16 : PN_EPI_F32;
  // "fuse_norm_bwd" (default; 2-CTA tiles, widths in 256s): the row dot <G, h> of each norm's backward is
  // <Ys, y> / rs with y the forward output of the normed linear (y = rs h W'^T, so <Ys W', h> = <Ys, h W'^T>), and the
  // kernels that PRODUCE Ys hold y already: the masked FFN data gradient has the saved activation as its mask tile, the
  // attention backward reads q | k | v.  They leave per-tile partial sums, pn_t5_norm_coef turns them into
  // coef = rs^2 <G, h> / d, and the data-gradient GEMM finishes dh += G - coef h in its epilogue: no G round trip
  // through HBM and no separate bandwidth pass (norm_bwd, 7 % of the step in profiles/r4n_train_launches_summary.txt).
  const bool fuse = m->fuse_norm_bwd && m->cta_group == 2 && (d % 256) == 0 && (ff % 256) == 0 &&
                    ((3 * di) % 256) == 0 && ff / 256 <= 64 && 4 * H <= 64 && d > 65;
  float* part = w.gbuf;                          // [<= 64][rows] partial dots (G itself is never stored when fusing)
  float* coef = w.gbuf + (size_t)64 * rows;      // [rows]
  for (int l = L - 1; l >= 0; --l) {
    const PnEncLayer& E = m->enc[l];
    const __nv_bfloat16 *hb0 = w.hb[2 * l], *hb1 = w.hb[2 * l + 1];
    const float *rs0 = w.rs[2 * l], *rs1 = w.rs[2 * l + 1];
    RC(wgrad_plain(m, w.dHb, d, w.act[l], E.wo2, rows, st));
    RC(dgrad_relu(m, w.dHb, E.wo2, w.act[l], rs1, w.dbig, rows, st, act_scale, fuse ? part : nullptr));
    RC(wgrad_folded(m, w.dbig, ff, hb1, E.wi, w.dWtmp, rows, st));
    if (fuse) {
      // with dropout the saved activation is m s relu(u), s = 1 / (1 - p): <Ys, u> = <Ys, act> / s
      RC(pn_t5_norm_coef(part, ff / 256, rs1, rows, d, coef, st, 1.0f / act_scale));
      RC(dgrad_norm(m, w.dbig, ff, E.wi, hb1, coef, w.dH, w.dHb, rows, st,
                    DR(0, l, PN_DROP_SELF_OUT, (long long)S * d)));
    } else {
      RC(dgrad(m, w.dbig, ff, E.wi, w.gbuf, gmode, rows, st));
      RC(pn_t5_norm_bwd(w.gbuf, g16, hb1, rs1, rows, d, 1, w.dH, w.dHb, st, PnDrop(),
                        DR(0, l, PN_DROP_SELF_OUT, (long long)S * d)));
    }
    RC(wgrad_plain(m, w.dHb, d, w.ctx[l], E.wo, rows, st));
    RC(dgrad(m, w.dHb, d, E.wo, w.dctx, PN_EPI_BF16, rows, st));
    RC(pn_t5_self_attn_bwd(w.qkv[l], 3 * di, 0, di, 2 * di, (int)docs, S, H, m->enc_bias, w.ctx[l], w.dctx, di,
                           w.lse[l], w.dvec, rs0, w.dbig, w.diag, st,
                           DR(0, l, PN_DROP_SELF_P, (long long)H * S * Sp), fuse ? part : nullptr));
    RC(wgrad_folded(m, w.dbig, 3 * di, hb0, E.wqkv, w.dWtmp, rows, st));
    if (fuse) {
      RC(pn_t5_norm_coef(part, 4 * H, rs0, rows, d, coef, st));
      RC(dgrad_norm(m, w.dbig, 3 * di, E.wqkv, hb0, coef, w.dH, w.dHb, rows, st,
                    l > 0 ? DR(0, l - 1, PN_DROP_FF_OUT, (long long)S * d) : PnDrop()));
    } else {
      RC(dgrad(m, w.dbig, 3 * di, E.wqkv, w.gbuf, gmode, rows, st));
      RC(pn_t5_norm_bwd(w.gbuf, g16, hb0, rs0, rows, d, 1, w.dH, w.dHb, st, PnDrop(),
                        l > 0 ? DR(0, l - 1, PN_DROP_FF_OUT, (long long)S * d) : PnDrop()));
    }
    if (m->grad_hook && l > 0)
      for (int g = groups - 1; g >= 1; --g)
        if (group_lo(g) == l) {   // layers [l, top of group g) are done
          const int64_t lo = enc_layer_off(l);
          const int64_t hi = g == groups - 1 ? dec_off : enc_layer_off(group_lo(g + 1));
          m->grad_hook(m->grad_hook_user, lo, hi - lo);
        }
  }
  RC(pn_t5_embed_bwd(ids, w.dH, rows, d, m->cfg.vocab, G + m->emb_off, st, DR(0, 0, PN_DROP_EMBED, (long long)S * d)));
  RC(pn_t5_rel_bias_grad(w.diag, bucket_of, H, m->cfg.rel_buckets, G + m->enc_rel_off, st));
  if (m->grad_hook) {
    const int64_t rest = groups > 1 && group_lo(1) > 0 ? enc_layer_off(group_lo(1)) : dec_off;
    m->grad_hook(m->grad_hook_user, 0, rest);
  }
  return PN_OK;
}

extern "C" int pn_t5_train_backward(PnT5* m, const int32_t* ids, const float* dlogits, int64_t docs, int32_t S,
                                    int32_t T, void* workspace, int64_t workspace_bytes, void* stream) {
  PN_CHECK_ARG(dlogits);
  return train_backward_impl(m, ids, dlogits, nullptr, 0, 0, docs, S, T, workspace, workspace_bytes, stream);
}

extern "C" int pn_t5_train_backward_z2(PnT5* m, const int32_t* ids, const float* dz2, int32_t rel_id,
                                       int32_t nrel_id, int64_t docs, int32_t S, int32_t T, void* workspace,
                                       int64_t workspace_bytes, void* stream) {
  PN_CHECK_ARG(dz2);
  return train_backward_impl(m, ids, nullptr, dz2, rel_id, nrel_id, docs, S, T, workspace, workspace_bytes, stream);
}

extern "C" int pn_t5_zero_grad(PnT5* m, void* stream) {
  PN_CHECK_ARG(m && m->grads);
  PN_CUDA(cudaMemsetAsync(m->grads, 0, (size_t)m->n_params * 4, (cudaStream_t)stream));
  return PN_OK;
}

// transformers.AdamW step over the flat parameter buffer (gradients multiplied by gscale first, e.g. 1/world for
// a summed all-reduce), then the bf16 operands are rebuilt from the updated masters
extern "C" int pn_t5_adamw_step(PnT5* m, float lr, float beta1, float beta2, float eps, float weight_decay,
                                int64_t step, float gscale, void* stream) {
  PN_CHECK_ARG(m && m->finalized && m->grads && m->adam_m && m->adam_v && step >= 1);
  cudaStream_t st = (cudaStream_t)stream;
  RC(pn_t5_adamw(m->flat, m->grads, m->adam_m, m->adam_v, m->n_params, lr, beta1, beta2, eps, weight_decay, step,
                 gscale, st));
  return pn_t5_repack(m, st);
}

// One plain SGD step on the parameters from element `first_elem` on (StdWrapper's look-ahead copy,
// pairwrapper.py:57-74,126-128: theta' = theta - lr * grad), then the bf16 operands are rebuilt
extern "C" int pn_t5_sgd_step(PnT5* m, float lr, int64_t first_elem, void* stream) {
  PN_CHECK_ARG(m && m->finalized && m->grads && first_elem >= 0 && first_elem <= m->n_params);
  cudaStream_t st = (cudaStream_t)stream;
  RC(pn_t5_sgd(m->flat + first_elem, m->grads + first_elem, m->n_params - first_elem, lr, st));
  return pn_t5_repack(m, st);
}

extern "C" int pn_adamw(float* p, const float* g, float* mom, float* var, int64_t n, float lr, float beta1,
                        float beta2, float eps, float weight_decay, int64_t step, float gscale, void* stream) {
  PN_CHECK_ARG(p && g && mom && var && n >= 0 && step >= 1);
  return pn_t5_adamw(p, g, mom, var, n, lr, beta1, beta2, eps, weight_decay, step, gscale, (cudaStream_t)stream);
}
