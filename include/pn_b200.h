/*
 * pn_b200.h — C-ABI of libpn_b200.so: the B200 (sm_100a) implementation of the
 * PacedNegatives LCE / self-paced monoT5 hot path.
 *
 * The reference (Parry-Parry/PacedNegatives) is pure Python and has no FFI; the
 * drop-in boundary is its Python call surface.  Every entry point below names the
 * reference function (file:line under the reference tree) whose arithmetic it
 * replaces; the Python shims in pacednegatives_b200/ bind these with ctypes and
 * keep the reference's class / method names.
 *
 * Conventions
 *   - all pointers are DEVICE pointers unless the name ends in _host;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - calls enqueue work and return without a host sync unless stated;
 *   - return 0 on success, a negative PN_E* code otherwise; the message is in
 *     pn_last_error() (thread-local).  Nothing here calls exit()/abort() — the
 *     reference's `exit()` on sampler failure (dataloader.py:149-151) becomes
 *     PN_E_INFEASIBLE reported through a status word;
 *   - the library never owns caller buffers; only PnT5 handles own memory.
 */
#ifndef PN_B200_H
#define PN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PN_VERSION 100

enum {
  PN_OK = 0,
  PN_E_BADARG = -1,      /* null pointer, negative size, unsupported enum value */
  PN_E_ARCH = -2,        /* device is not compute capability 10.x */
  PN_E_LAUNCH = -3,      /* CUDA launch / runtime error */
  PN_E_INFEASIBLE = -4,  /* sampler: fewer non-zero pmf entries than k */
  PN_E_UNSUPPORTED = -5, /* shape outside what the kernels are built for */
  PN_E_OOM = -6
};

int pn_version(void);
const char* pn_last_error(void);
/* 0 when the current device is sm_100 (B200); PN_E_ARCH otherwise. */
int pn_device_check(void);
/* number of kernels this library has launched in this process (bench.py gpu_launches). */
int64_t pn_launch_count(void);

/* Per-kernel-class device timing for bench.py's roofline object: when enabled, every launch of a
 * class is bracketed by CUDA events on its own stream.  kind 0 = tcgen05 GEMM (work = 2MNK flops),
 * 1 = encoder self-attention (work = 4*S*S*64 flops per (doc, head)), 2 = elementwise, 3 = wgrad GEMM,
 * 4 = attention backward, 5 = HBM-bound GEMM launches (< 200 FLOP per byte of operands + output; work = those bytes).  pn_prof_read waits for outstanding events and returns the sums since the last
 * pn_prof_enable(1). */
int pn_prof_enable(int on);
int pn_prof_read(int kind, double* ms_sum, int64_t* launches, double* work_sum);

/* ------------------------------------------------------------------ a4 / a5
 * Self-paced eta-threshold weights.
 *   variant 0 (PN_W_LCE): w = l/eta        if !(l > eta) else (l*mask)*eta
 *       lceT5.py:91-99 (mask == 0), weights.py:59-78
 *   variant 1 (PN_W_ETA): w = (-(l/eta))+1 if !(l > eta) else (l*mask)*eta
 *       weights.py:26-45
 * fp32 IEEE ops in the reference's order, so w and the keep mask are bit-exact.
 * keep_out[i] = !(loss[i] > eta)  (nullable).
 * stats_out (nullable, double[13], overwritten): the six statistics in [0..5], [6..12] is the call's scratch
 * (raw partial sums + the grid's completion counter; the library keeps no state between or across calls):
 *   [0] sum w   [1] sum w*l   [2] d(sum w*l)/d eta   [3] #(l < eta)  (old/eta_bound.py:57-58)
 *   [4] sum l   [5] #(!(l > eta))
 */
enum { PN_W_LCE = 0, PN_W_ETA = 1 };
int pn_weights_fwd(const float* loss, int64_t n, const float* eta, float mask, int variant,
                   float* w_out, uint8_t* keep_out, double* stats_out, void* stream);

/* Backward of pn_weights_fwd for an upstream gradient g[n] (autograd of weights.forward):
 *   deta_out[0] (double) = sum_i g[i] * dw_i/deta,  dloss_out[i] (nullable) = g[i] * dw_i/dloss_i. */
int pn_weights_bwd(const float* loss, int64_t n, const float* eta, float mask, int variant,
                   const float* g, float* dloss_out, double* deta_out, void* stream);

/* ------------------------------------------------------------------ a6
 * Per-sample weight table row update (pairwrapper.py:219-239, weights.py:87-88):
 *   g_i = (ce_p ? ce_p[i]*inv_len : 0) + (ce_n ? ce_n[i]*inv_len : 0) - 1
 *   row[i] = sigmoid(row[i] - lr * g_i)
 * ce_p / ce_n are nullable (NewWrapper only differentiates the negative term).
 */
int pn_weights_table_update(float* row, const float* ce_p, const float* ce_n, int64_t n,
                            float inv_len, float lr, void* stream);

/* ------------------------------------------------------------------ a3
 * Token cross-entropy, reduction='none' (lceT5.py:110,117-118; pairwrapper.py:48):
 *   ce[r] = logsumexp(logits[r,:]) - logits[r,target[r]],  0 when target[r]==ignore_index.
 * logits is [rows, V] row-major, dtype 0 = fp32, 1 = bf16.  lse_out nullable.
 */
int pn_token_ce_fwd(const void* logits, int dtype, int64_t rows, int32_t V, const int64_t* target,
                    int64_t ignore_index, float* ce_out, float* lse_out, void* stream);
/* dlogits[r,v] = g[r] * (softmax(logits[r,:])[v] - [v==target[r]]), fp32 out. */
int pn_token_ce_bwd(const void* logits, int dtype, int64_t rows, int32_t V, const int64_t* target,
                    int64_t ignore_index, const float* lse, const float* g, float* dlogits,
                    void* stream);

/* Group reduce of the reference's LCE loss (lceT5.py:119-121; utilities/loss.py:9-12):
 *   ce[i] = pce[i] + (use_mean ? mean : sum)(nce[i*k : (i+1)*k]),  i in [0, m)
 * m = Q*T, nce has m*k entries in the reference's flat order.  stats_out nullable double[1] = sum ce. */
int pn_group_loss_fwd(const float* pce, const float* nce, int64_t m, int32_t k, int use_mean,
                      float* ce_out, double* stats_out, void* stream);
/* d pce[i] = g[i];  d nce[i*k+j] = g[i] * (use_mean ? 1/k : 1). */
int pn_group_loss_bwd(const float* g, int64_t m, int32_t k, int use_mean, float* dpce, float* dnce,
                      void* stream);

/* Restricted-head LCE (canonical localized contrastive estimation over the two-row head):
 *   s[d] = log_softmax([z_true[d], z_false[d]])[0]  (distill/wrapper.py:34-38)
 *   loss[q] = logsumexp_j s[q*G+j] - s[q*G+0]       (doc 0 of each group is the positive)
 * z2 is [Q*G, 2] fp32 (true, false).  score_out nullable [Q*G]. */
int pn_lce_restricted_fwd(const float* z2, int64_t Q, int32_t G, float* loss_out, float* score_out,
                          void* stream);
int pn_lce_restricted_bwd(const float* z2, int64_t Q, int32_t G, const float* g, float* dz2,
                          void* stream);

/* check_success / check_probs (old/levels.py:35-39, old/eta_bound.py:51-55): count_out[0] (device int32, zeroed
 * by the call) = #{ i < n : softmax(pos[i,0,(rel,nrel)])[0] > softmax(neg[i,0,(rel,nrel)])[0] }; the caller divides
 * by n.  pos/neg are fp32 logits [n, T, V] read at label position 0; *_stride = T*V elements per document. */
int pn_check_success(const float* pos_logits, const float* neg_logits, int64_t n, int64_t pos_stride,
                     int64_t neg_stride, int32_t rel_id, int32_t nrel_id, int32_t* count_out, void* stream);

/* ------------------------------------------------------------------ a2 / 8f-2
 * On-device prompt packing: 'Query: ' + q + ' Document: ' + d + ' Relevant:' (dataloader.py:43-44,66-67,153-155)
 * from pre-tokenised pieces in HBM, padded to S like tokenizer(..., padding=True) (lceT5.py:46-47,
 * pairwrapper.py:79-80).  prefix / middle / suffix: int32 ids of 'Query:', 'Document:', 'Relevant:' + </s>;
 * q_tok/q_off, d_tok/d_off: CSR token stores (off has n+1 int64 entries); q_idx / d_idx [n]: the (query, document)
 * of each output row (d_idx is what pn_select_* wrote as doc ids).  max_len = 0 keeps whole documents (the
 * reference does not truncate); max_len > 0 drops the document's tail so the row ends in ' Relevant:</s>' within
 * max_len tokens.  ids_out [n, S] nullable (NULL = lengths only), len_out [n] nullable (-1 for an out-of-range
 * index), status_out[2] = {longest row, #rows that are out of range or longer than S (those rows are all pad)}. */
int pn_pack_prompts(const int32_t* prefix, int32_t n_prefix, const int32_t* middle, int32_t n_middle,
                    const int32_t* suffix, int32_t n_suffix, const int32_t* q_tok, const int64_t* q_off,
                    int64_t n_queries, const int32_t* d_tok, const int64_t* d_off, int64_t n_docs,
                    const int32_t* q_idx, const int32_t* d_idx, int64_t n, int32_t S, int32_t max_len,
                    int32_t pad_id, int32_t* ids_out, int32_t* len_out, int32_t* status_out, void* stream);

/* ------------------------------------------------------------------ a7
 * Level-paced negative pick (dataloader.py:29-33): r = floor|ceil(w * n_neg), n_neg = C-1,
 * computed in fp64 like the Python float the reference multiplies.
 * cand is int32 doc ids, layout 0 = row-major [Q,C], 1 = rank-major [C,Q].
 * Out-of-range r (reference: IndexError / negative wrap) gives idx=-1, doc=-1 and bumps status_out[0].
 */
enum { PN_LAYOUT_ROW = 0, PN_LAYOUT_RANK = 1 };
int pn_select_level_scalar(const int32_t* cand, int64_t Q, int32_t C, int layout, double w,
                           int use_ceil, int32_t* idx_out /*[1]*/, int32_t* doc_out /*[Q]*/,
                           int32_t* status_out, void* stream);
/* per-item weights (PairLoader.get_batch, dataloader.py:46-56): w[q] is the fp32 the reference
 * reads with tensor.item(). */
int pn_select_level_items(const int32_t* cand, int64_t Q, int32_t C, int layout, const float* w,
                          int use_ceil, int32_t* idx_out /*[Q]*/, int32_t* doc_out /*[Q]*/,
                          int32_t* status_out, void* stream);

/* ------------------------------------------------------------------ a8
 * Binomial-paced sampling of k ranks without replacement (dataloader.py:134-151 =
 * scipy binom.pmf + numpy legacy RandomState.choice(replace=False, p)).
 *
 * pn_binomial_table: p = clip(mean) -> pmf(arange(C); C-1, mean) -> p/sum -> var rescale -> p/sum,
 *   then the first-round cdf = cumsum(p)/cumsum(p)[-1], all fp64 in numpy's summation order.
 *   table_out is double[2*C]: p then cdf.  pmf_in (nullable, double[C]) overrides the device
 *   lgamma pmf with a caller-supplied one (scipy's, for strict parity).  status_out[0] = number of
 *   non-zero p entries, status_out[1] = PN_E_INFEASIBLE flag if the variance step produced non-finite p.
 */
int pn_binomial_table(double mean, double clip_min, double clip_max, double var, int32_t C,
                      const double* pmf_in, double* table_out, int32_t* status_out, void* stream);
/* pn_select_binomial: for each query draw k distinct ranks with numpy's algorithm.
 *   uniforms (nullable): double [Q, U] per-query substreams (stream_mode 0) or one flat stream of
 *     length U consumed by the queries in order (stream_mode 1, the reference's global RNG);
 *     NULL = counter-based Philox4x32-10 stream keyed by (philox_seed, query).
 *   idx_out int32 [Q,k] rank indices in numpy's order; doc_out (nullable) cand[q][idx]; cand nullable
 *   when doc_out is.  consumed_out (nullable) [Q] uniforms used.  status_out[0] = #queries that ran
 *   out of uniforms or were infeasible (their rows are filled with -1).
 */
int pn_select_binomial(const int32_t* cand, int64_t Q, int32_t C, int layout, const double* table,
                       int32_t k, const double* uniforms, int64_t U, int stream_mode,
                       uint64_t philox_seed, int32_t* idx_out, int32_t* doc_out,
                       int32_t* consumed_out, int32_t* status_out, void* stream);

/* ------------------------------------------------------------------ a12
 * Per-query ranking + trec_eval measures (eval.py:26,41-43): sort C candidates by
 * (score desc, docid desc), emit the top K and ndcg_cut_K (linear gain, log2(i+1) discount, ideal from
 * the query's rels), recip_rank (whole list), recip_rank cut at K, and average precision.
 * rels uint8 [Q,C] graded relevance of each candidate.  metrics_out float [Q,4] =
 * (ndcg@K, rr, rr@K, ap).  C <= 4096.  top_* nullable.
 */
int pn_topk_metrics(const float* scores, const int32_t* docids, const uint8_t* rels, int64_t Q,
                    int32_t C, int32_t K, int32_t* top_ids, float* top_scores, float* metrics_out,
                    void* stream);
/* The same with the query's qrel totals (eval.py:26,41-43: pt.Experiment / trec_eval judge against ALL judged
 * documents of a query, including relevant ones the first stage did not retrieve): qrel_num int32 [Q] (nullable) =
 * number of judged documents with rel >= 1, the average-precision denominator; qrel_idcg double [Q] (nullable) =
 * ideal DCG@K of the query's judged grades.  NULL = derive each from the candidates' rels (every judged-relevant
 * document is in the list). */
int pn_topk_metrics_qrels(const float* scores, const int32_t* docids, const uint8_t* rels,
                          const int32_t* qrel_num, const double* qrel_idcg, int64_t Q, int32_t C, int32_t K,
                          int32_t* top_ids, float* top_scores, float* metrics_out, void* stream);

/* ------------------------------------------------------------------ a1 (building block)
 * bf16 GEMM on the tcgen05 tensor path, the T5 linear layers (HF modeling_t5.py:277,298-299,338,92-103):
 *   C[M,N] = A[M,K] . B[N,K]^T, A row-major activations (lda), B an nn.Linear weight [out,in] (ldb).
 *   mode 0: bf16 out[M,ldc] = relu?(acc * row_scale[row]?) ; mode 1: fp32 out[M,ldc] += acc * row_scale?.
 *   K % 64 == 0, N % 128 == 0, 16-byte aligned operands.  cta_group 1 = single-SM tiles 128xN,
 *   2 = CTA-pair tiles 256xN (cta_group::2). */
int pn_gemm_bf16(const void* A, int64_t lda, const void* B, int64_t ldb, void* out, int64_t ldc,
                 int64_t M, int64_t N, int64_t K, int mode, int relu, const float* row_scale,
                 int cta_group, void* stream);

/* Weight gradient of a linear layer (autograd of HF modeling_t5.py:277,298-299,338,92-103 under
 * lceT5.py:155 manual_backward): C[M,N] (fp32, ldc) += A[K,M]^T . B[K,N] with A = dY and B = X, both bf16 row-major
 * with K rows (documents x positions).  Both operands are consumed MN-major straight from their row-major
 * tensors; the K range is split across SMs and accumulated with fp32 vector atomics (order non-deterministic).
 * N % 128 == 0, M % 8 == 0; any K (TMA zero-fills the tail). */
int pn_gemm_bf16_wgrad(const void* A, int64_t lda, const void* B, int64_t ldb, float* out, int64_t ldc,
                       int64_t M, int64_t N, int64_t K, int cta_group, void* stream);

/* Building blocks of the fused RMSNorm backward of the encoder's normed linears (autograd of HF modeling_t5.py:55-69
 * followed by :298-299 / :92-95 under lceT5.py:155 manual_backward; DESIGN.md section 4 "norm backward").
 * With y = rs h W'^T the forward output of a normed linear and Ys its scaled output gradient,
 *   dh = Ys W' - coef h,   coef[row] = rs^2 <Ys W', h> / d = rs <Ys, y> / d :
 * the row dot is known before the data-gradient product is formed, from the kernels that produce Ys.
 *
 * pn_gemm_bf16_relu_bwd: out bf16 [M, ldc] = row_scale[row]? * (dY[M,K] . Wt[N,K]^T) where act[row, col] > 0, else 0
 *   (ReLU backward of the FFN's first linear; act = the saved forward activation, bf16 [M, ld_act]); dot_out
 *   (nullable) fp32 [N / 256][M] receives the per-256-column partial sums of out[row, col] * act[row, col].
 *   2-CTA tiles only: N % 256 == 0, K % 64 == 0, 16-byte aligned operands. */
int pn_gemm_bf16_relu_bwd(const void* dY, int64_t ld_dy, const void* Wt, int64_t ldw, const void* act, int64_t ld_act,
                          const float* row_scale, void* out, int64_t ldc, float* dot_out, int64_t M, int64_t N,
                          int64_t K, void* stream);
/* pn_gemm_bf16_norm_bwd: dh fp32 [M, ld_dh] += Ys[M,K] . Wt[N,K]^T - coef[row] * hb[row, col]; dhb bf16 [M, ld_dhb] =
 *   the updated dh.  hb bf16 [M, ld_hb] is the residual stream the norm read.  Same shape limits. */
int pn_gemm_bf16_norm_bwd(const void* Ys, int64_t ld_ys, const void* Wt, int64_t ldw, const void* hb, int64_t ld_hb,
                          const float* coef, float* dh, int64_t ld_dh, void* dhb, int64_t ld_dhb, int64_t M, int64_t N,
                          int64_t K, void* stream);
/* coef[row] = rs[row] * sum_p part[p * rows + row] / d */
int pn_t5_norm_coef(const float* part, int32_t parts, const float* rs, int64_t rows, int32_t d, float* coef,
                    void* stream);

/* T5 encoder self-attention (HF modeling_t5.py:308-336): out = softmax(Q K^T + bias[h, j-i]) V per
 * (document, head), no 1/sqrt(d) scaling, d_kv = 64, S <= 512, no mask.  qkv bf16 [docs*S, 3*H*64] holds
 * q | k | v column blocks; bias fp32 [H, 1023] indexed by (j - i + 511); out bf16 [docs*S, H*64]. */
int pn_t5_self_attention(const void* qkv, int32_t docs, int32_t S, int32_t n_heads, const float* bias,
                         void* out, void* stream);

/* Same, also returning the base-2 row logsumexp lse2 fp32 [docs*S, H] = log2(sum_j 2^((s_ij + bias) * log2 e))
 * that the backward pass recomputes the probabilities from. */
int pn_t5_self_attention_lse(const void* qkv, int32_t docs, int32_t S, int32_t n_heads, const float* bias,
                             void* out, float* lse2, void* stream);

/* Backward of pn_t5_self_attention (autograd of HF modeling_t5.py:308-336 under lceT5.py:155 manual_backward).
 * Inputs: qkv and bias as in the forward, ctx = the forward's out, dctx = its gradient (bf16 [docs*S, H*64]),
 * lse2 from pn_t5_self_attention_lse.  Outputs: dqkv bf16 [docs*S, 3*H*64] = dq | dk | dv, each row multiplied
 * by row_scale[row] when row_scale is not NULL; dbias_diag fp32 [H, 1023] += sum over documents of dS along
 * the diagonal (j - i + 511) (caller zeroes it); dvec_ws is a [docs*S, H] fp32 workspace. */
int pn_t5_self_attention_bwd(const void* qkv, int32_t docs, int32_t S, int32_t n_heads, const float* bias,
                             const void* ctx, const void* dctx, const float* lse2, const float* row_scale,
                             void* dqkv, float* dbias_diag, float* dvec_ws, void* stream);

/* pn_t5_self_attention_bwd that also leaves, in dot_out fp32 [4 * n_heads][docs * S], partial row sums of
 * dqkv[row, :] * qkv[row, :] (the row dot of the RMSNorm backward behind the qkv projection; sum the planes). */
int pn_t5_self_attention_bwd_dot(const void* qkv, int32_t docs, int32_t S, int32_t n_heads, const float* bias,
                                 const void* ctx, const void* dctx, const float* lse2, const float* row_scale,
                                 void* dqkv, float* dbias_diag, float* dvec_ws, float* dot_out, void* stream);

/* Allocation-free variant of pn_t5_self_attention_bwd.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid allocation).  Same inputs, outputs and arithmetic, but the [docs*S, n_heads]
 * fp32 workspace is a file-scope device array of the library (4 MiB, held by every GPU that loads the library until
 * exit) instead of a caller buffer: the call makes no cudaMalloc / cudaFree / allocating library call.
 * Limit: docs * S * n_heads <= 1 048 576 (e.g. 128 documents x 512 tokens x 16 heads); larger shapes return
 * PN_E_UNSUPPORTED with a message.  Every call on a device shares the one array: nothing in it has to start at zero
 * (the call overwrites every entry it reads), but two calls that overlap on the GPU (different streams, or graphs
 * replayed concurrently) race on it; serialise them on one stream. */
int pn_t5_self_attention_bwd_static(const void* qkv, int32_t docs, int32_t S, int32_t n_heads, const float* bias,
                                    const void* ctx, const void* dctx, const float* lse2, const float* row_scale,
                                    void* dqkv, float* dbias_diag, void* stream);

/* ------------------------------------------------------------------ a1 / a9 / a12: monoT5 scorer
 * Replaces transformers.T5ForConditionalGeneration as the reference calls it (lceT5.py:106,132-133,
 * 149-150; pairwrapper.py:223-247; the re-rank formula distill/wrapper.py:34-38).  ReLU FFN, RMSNorm,
 * 32-bucket relative bias, tied embedding/head scaled by d_model^-0.5, d_kv = 64, no attention mask
 * (the reference passes none), dropout off.
 */
typedef struct PnT5 PnT5;
typedef struct PnT5Config {
  int32_t vocab, d_model, d_kv, d_ff, n_layers, n_dec_layers, n_heads;
  int32_t rel_buckets, rel_max_dist;
  int32_t pad_id; /* also decoder_start_token_id */
  float eps;
} PnT5Config;

int pn_t5_create(const PnT5Config* cfg, PnT5** out);
int pn_t5_destroy(PnT5* m);
/* options: "cta_group" = 1 | 2 (tcgen05 tile mode of the linears); "train" = 1 (before loading weights):
 * also keep the transposed bf16 operands the data-gradient GEMMs read */
int pn_t5_set_option(PnT5* m, const char* name, int64_t value);
/* One parameter by its HuggingFace state_dict name (e.g. "encoder.block.3.layer.0.SelfAttention.q.weight"),
 * fp32 contiguous DEVICE tensor in HF's layout; converted into the handle's packed bf16 buffers. */
int pn_t5_load_weight(PnT5* m, const char* hf_name, const float* src, int64_t numel, void* stream);
/* checks every parameter arrived, expands the relative-position bias tables (syncs the stream) */
int pn_t5_finalize(PnT5* m, void* stream);
int64_t pn_t5_workspace_bytes(const PnT5* m, int64_t docs, int32_t S, int32_t T, int want_logits);
/* Flat parameter storage: every parameter lives in one contiguous fp32 DEVICE buffer in state_dict order
 * (what the gradient all-reduce of lceT5.py:155 under DDP, AdamW and save_pretrained export walk).
 * pn_t5_tensor_info gives tensor i's HF name, element offset and size; pn_t5_params the buffer (owned by the
 * handle; after writing into it call pn_t5_finalize again to refresh the packed bf16 operands). */
int64_t pn_t5_num_params(const PnT5* m);
int32_t pn_t5_num_tensors(const PnT5* m);
int pn_t5_tensor_info(const PnT5* m, int32_t i, char* name_out, int32_t name_cap, int64_t* offset_out,
                      int64_t* numel_out);
float* pn_t5_params(PnT5* m);
int32_t pn_t5_vocab_pad(const PnT5* m);
/* Re-rank scores for `docs` prompts of S token ids (int32 [docs,S]): encoder, one decoder step from the
 * start token, two-row head: z2_out[d] = (z_true, z_false), score_out[d] = log_softmax(z2)[0]. */
int pn_t5_score(PnT5* m, const int32_t* ids, int64_t docs, int32_t S, int32_t rel_id, int32_t nrel_id,
                float* score_out, float* z2_out, void* workspace, int64_t workspace_bytes, void* stream);
/* The same with the tokenizer's attention_mask, as pyterrier_t5.MonoT5ReRanker (eval.py:14,23) and
 * distill/wrapper.py:36 call the model (`self.model(**enc)`, padding='longest'): lens int32 [docs] (nullable = no
 * mask) is the number of real tokens of each right-padded prompt; keys at positions >= lens[d] are excluded from the
 * encoder self-attention and the cross-attention (HF's extended attention mask), 1 <= lens[d] <= S. */
int pn_t5_score_lens(PnT5* m, const int32_t* ids, const int32_t* lens, int64_t docs, int32_t S, int32_t rel_id,
                     int32_t nrel_id, float* score_out, float* z2_out, void* workspace, int64_t workspace_bytes,
                     void* stream);
/* model(input_ids=ids, labels=labels).logits for T = 1 or 2 label positions (int64 [docs,T]):
 * logits_out fp32 [docs*T, pn_t5_vocab_pad()] (nullable), z2_out fp32 [docs*T, 2] (nullable). */
int pn_t5_logits(PnT5* m, const int32_t* ids, const int64_t* labels, int64_t docs, int32_t S, int32_t T,
                 int32_t rel_id, int32_t nrel_id, float* logits_out, float* z2_out, void* workspace,
                 int64_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------ a10 / a13: the training step
 * Replaces, for LCEModel.training_step (lceT5.py:124-167), the gradient-tracking model calls
 * `self.model(**p).logits`, `self.model(**n).logits` (lceT5.py:149-150), the autograd of
 * `self.manual_backward(main_loss.mean())` (lceT5.py:154-155) and `opt.step()` of the transformers.AdamW built
 * in configure_optimizers (lceT5.py:171-172).  The handle must have been created with option "train" = 1.
 *
 * pn_t5_set_train_buffers: caller-owned fp32 device buffers of pn_t5_num_params() elements, laid out like
 *   pn_t5_params() (HF state_dict order): gradients (accumulated into), Adam first / second moments (zeroed by
 *   the caller once).  One contiguous gradient buffer = one all-reduce per step in data-parallel runs.
 * pn_t5_train_forward: logits_out fp32 [docs*T, vocab_pad]; every activation the backward needs stays in
 *   `workspace` (pn_t5_train_workspace_bytes), which must be left untouched until pn_t5_train_backward.
 * pn_t5_train_backward: dlogits fp32 [docs*T, vocab_pad] = d loss / d logits; gradients += .  vocab % 8 == 0.
 * pn_t5_adamw_step: p -= lr*sqrt(1-b2^t)/(1-b1^t) * m/(sqrt(v)+eps) (+ decoupled weight decay) on gscale*grad,
 *   then the packed bf16 operands are rebuilt.  pn_adamw: the same update on any fp32 tensor (eta). */
int64_t pn_t5_train_workspace_bytes(const PnT5* m, int64_t docs, int32_t S, int32_t T);
int pn_t5_set_train_buffers(PnT5* m, float* grads, float* adam_m, float* adam_v);
int pn_t5_train_forward(PnT5* m, const int32_t* ids, const int64_t* labels, int64_t docs, int32_t S, int32_t T,
                        float* logits_out, void* workspace, int64_t workspace_bytes, void* stream);
/* Train-mode dropout (the reference's LCEModel is a LightningModule in train() mode, so HF T5 applies
 * nn.Dropout(config.dropout_rate = 0.1) in BOTH forwards of lceT5.py:131-133,149-150: embedding output, every
 * sub-layer output, FFN hidden activation, attention probabilities, the two final norms).  p, seed and doc_offset
 * apply to the NEXT pn_t5_train_forward / _forward_z2 and to the backward that consumes its workspace; p = 0
 * (default) is eval-mode arithmetic.  Masks are counter-based — a pure function of (seed, dropout site, element
 * index with documents numbered from doc_offset) — so the backward regenerates them instead of storing them, and
 * tests replay the same masks inside HF.  pn_t5_train_workspace_bytes depends on whether p > 0 (the projection-free
 * cross-attention is the p = 0 path). */
int pn_t5_set_dropout(PnT5* m, float p, uint64_t seed, int64_t doc_offset);
int pn_t5_train_backward(PnT5* m, const int32_t* ids, const float* dlogits, int64_t docs, int32_t S, int32_t T,
                         void* workspace, int64_t workspace_bytes, void* stream);
/* Restricted-head training (north_star kernels (1) + (2): the LM head restricted to the 'true' / 'false' rows and
 * the group-wise log-softmax LCE of pn_lce_restricted_fwd / _bwd): the same forward with the two-row head of the
 * re-rank score (distill/wrapper.py:34-38) instead of the [docs*T, vocab] logits — z2_out fp32 [docs*T, 2] =
 * (z_true, z_false) — and the backward from dz2 = d loss / d z2 (the [docs*T, vocab_pad] logits and their gradient
 * never exist).  Same workspace contract as pn_t5_train_forward / _backward. */
int pn_t5_train_forward_z2(PnT5* m, const int32_t* ids, const int64_t* labels, int64_t docs, int32_t S, int32_t T,
                           int32_t rel_id, int32_t nrel_id, float* z2_out, void* workspace, int64_t workspace_bytes,
                           void* stream);
int pn_t5_train_backward_z2(PnT5* m, const int32_t* ids, const float* dz2, int32_t rel_id, int32_t nrel_id,
                            int64_t docs, int32_t S, int32_t T, void* workspace, int64_t workspace_bytes,
                            void* stream);
/* Data-parallel overlap (DDP's bucketed all-reduce under `manual_backward`, lceT5.py:155 with strategy='ddp'
 * train/train_lce.py:89-91): fn(user, offset, numel) is called on the host thread while pn_t5_train_backward
 * enqueues work, each time a contiguous range of the gradient buffer has received its last contribution
 * (decoder parameters first, the upper encoder layers half way, the rest at the end; the ranges tile the buffer).
 * The range is final once the work enqueued so far has run: record an event, all-reduce on a side stream.
 * fn == NULL removes the hook.  Only meaningful when one backward call produces the whole step's gradient. */
int pn_t5_set_grad_hook(PnT5* m, void (*fn)(void* user, int64_t offset, int64_t numel), void* user);
int pn_t5_zero_grad(PnT5* m, void* stream);
int pn_t5_adamw_step(PnT5* m, float lr, float beta1, float beta2, float eps, float weight_decay, int64_t step,
                     float gscale, void* stream);
int pn_adamw(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
             float eps, float weight_decay, int64_t step, float gscale, void* stream);
/* StdWrapper's look-ahead copy (pairwrapper.py:57-74 update_params, called at :126-128): one plain SGD step
 * p -= lr * grad on the flat parameters from element `first_elem` on (0 = every parameter; the offset of the first
 * non-embedding tensor reproduces transformers 5.x, where replacing `shared.weight` does not reach the tied
 * embedding / head modules), then the packed bf16 operands are rebuilt. */
int pn_t5_sgd_step(PnT5* m, float lr, int64_t first_elem, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PN_B200_H */
