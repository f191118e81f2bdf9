// The PnT5 handle shared by the scorer (pn_t5.cu) and the trainer (pn_t5_train.cu).
//
// Parameters live in ONE flat fp32 buffer in HuggingFace state_dict order (so the data-parallel gradient
// all-reduce, AdamW and checkpoint export each touch a single contiguous tensor); q|k|v and cross k|v are
// adjacent in that order, which makes the fused [3*d_inner, d] / [2*d_inner, d] projection matrices plain views.
// Every linear also has packed bf16 copies for the tcgen05 GEMMs: `packed` [out, in] (forward) and `packedT`
// [in, out] (data gradient), both with the RMSNorm weight of the linear's input folded into the `in` axis.
#pragma once
#include "pn_common.cuh"
#include "pn_t5_internal.cuh"

#include <array>
#include <map>
#include <string>
#include <vector>

#define RC(x) do { int rc__ = (x); if (rc__) return rc__; } while (0)

struct PnLin {
  float* master = nullptr;           // view into PnT5::flat, [rows, cols] = HF's [out, in]
  int64_t off = 0;                   // element offset of `master` in the flat buffer (same offset in grads)
  __nv_bfloat16* packed = nullptr;   // bf16 [rows, cols], norm weight folded into the columns
  __nv_bfloat16* packedT = nullptr;  // bf16 [cols, rows] (training only)
  int64_t rows = 0;
  int cols = 0;
  const float* fold = nullptr;       // the RMSNorm weight folded in (view into flat), or null
  int64_t fold_off = -1;
};
struct PnEncLayer {
  PnLin wqkv, wo, wi, wo2;
  float *ln0 = nullptr, *ln1 = nullptr;
  int64_t ln0_off = 0, ln1_off = 0;
};
struct PnDecLayer {
  PnLin wqkv, wo, cq, ckv, co, wi, wo2;
  float *ln0 = nullptr, *ln1 = nullptr, *ln2 = nullptr;
  int64_t ln0_off = 0, ln1_off = 0, ln2_off = 0;
};
struct PnParam {
  std::string name;
  int64_t off, numel;
};

struct PnT5 {
  PnT5Config cfg;
  int d_inner = 0;
  std::vector<void*> allocs;
  // flat parameter storage
  float* flat = nullptr;
  float* grads = nullptr;            // allocated by the first training call
  float *adam_m = nullptr, *adam_v = nullptr;
  int64_t n_params = 0;
  std::vector<PnParam> params;       // HF order
  std::map<std::string, int> index;  // name (and tied aliases) -> params[]
  std::vector<char> loaded;
  // views
  float* emb = nullptr;              // fp32 [vocab, d] = shared.weight
  int64_t emb_off = 0;
  __nv_bfloat16* emb_bf16 = nullptr; // bf16 [vocab_pad, d] for the full-vocabulary head
  __nv_bfloat16* embT_bf16 = nullptr;// bf16 [d, vocab_pad] (training: d head_in = dlogits . E)
  int vocab_pad = 0;
  std::vector<PnEncLayer> enc;
  std::vector<PnDecLayer> dec;
  float *enc_final = nullptr, *dec_final = nullptr;
  int64_t enc_final_off = 0, dec_final_off = 0;
  float *enc_rel = nullptr, *dec_rel = nullptr;  // [buckets, heads] as in HF
  int64_t enc_rel_off = 0, dec_rel_off = 0;
  float *enc_bias = nullptr;                     // [heads, 1023] expanded by (j - i + 511)
  float *dec_bias = nullptr;                     // [heads, 2, 2] for T <= 2
  // called on the host while pn_t5_train_backward enqueues work: gradients [offset, offset + numel) are final once
  // the work enqueued so far has run (data-parallel callers start that range's all-reduce on a side stream)
  void (*grad_hook)(void* user, int64_t offset, int64_t numel) = nullptr;
  void* grad_hook_user = nullptr;
  // training workspaces holding a forward pass: docs, S, T, dropout seed, dropout p (float bits), doc offset
  std::map<const void*, std::array<int64_t, 6>> fwd_ws;
  // dropout of the NEXT training forward (pn_t5_set_dropout); 0 = eval-mode arithmetic
  float drop_p = 0.0f;
  uint64_t drop_seed = 0;
  int64_t drop_doc_offset = 0;
  bool finalized = false;
  bool xattn = true;                 // projection-free cross-attention where the shapes allow ("xattn" option)
  bool want_train = false;           // also keep packedT / embT (pn_t5_set_option "train")
  int cta_group = 2;
  bool fuse_norm_bwd = true;         // encoder RMSNorm backward inside the data-gradient GEMM ("fuse_norm_bwd" option)
  bool dgrad_bf16 = true;            // encoder data gradients reach norm_bwd as bf16 ("dgrad_bf16" option)
  int grad_ranges = 4;               // encoder layer groups whose gradients leave separately ("grad_ranges" option)
};

template <typename T>
static inline int pn_dev_alloc(PnT5* m, T** p, size_t n) {
  void* q = nullptr;
  cudaError_t e = cudaMalloc(&q, (n ? n : 1) * sizeof(T));
  if (e != cudaSuccess)
    return pn_fail(PN_E_OOM, "pn_t5: cudaMalloc(%zu) failed: %s", n * sizeof(T), cudaGetErrorString(e));
  m->allocs.push_back(q);
  *p = (T*)q;
  return PN_OK;
}

// pn_t5_ops.cu
int pn_t5_embed(const int32_t* ids, const float* table, int64_t rows, int d, int vocab, float* hidden,
                __nv_bfloat16* copy, float* sumsq, cudaStream_t st, PnDrop drop = PnDrop());
int pn_t5_mask_copy(const __nv_bfloat16* src, __nv_bfloat16* dst, int64_t n, PnDrop drop, cudaStream_t st);
int pn_t5_pack_weight(const float* master, const float* colscale, int64_t rows, int cols,
                      __nv_bfloat16* packed, __nv_bfloat16* packedT, int ldT, cudaStream_t st);
int pn_t5_rmsnorm(const float* x, const float* w, int64_t rows, int d, float eps, __nv_bfloat16* out,
                  cudaStream_t st);
int pn_t5_f32_to_bf16(const float* src, __nv_bfloat16* dst, int64_t n, cudaStream_t st);
int pn_t5_head2(const float* x, const float* w, const float* emb, int64_t rows, int d, float eps,
                int rel_id, int nrel_id, float* score, float* z2, cudaStream_t st, PnDrop drop = PnDrop());
int pn_t5_head2_bwd(const float* x, const float* w, const float* emb, const float* dz2, int64_t rows, int d, float eps,
                    int rel_id, int nrel_id, float* dhead_in, float* grad_emb, cudaStream_t st, PnDrop drop = PnDrop());
int pn_t5_head_prep(const float* x, const float* w, int64_t rows, int d, float eps,
                    __nv_bfloat16* out, cudaStream_t st, PnDrop drop = PnDrop());

// cross-attention without the K/V projection (pn_t5_ops.cu)
int pn_t5_cross_expand_q(const __nv_bfloat16* q, int ld_q, int64_t docs, int T, int H, int Mp, __nv_bfloat16* qexp,
                         cudaStream_t st);
int pn_t5_cross_softmax(const float* scores, int lds, const float* rs, int64_t docs, int S, int TH, int Mp,
                        __nv_bfloat16* P, cudaStream_t st, const int32_t* klen = nullptr, PnDrop drop = PnDrop(),
                        int H = 0, __nv_bfloat16* Pc = nullptr);   // drop: P carries the mask, Pc the clean values
int pn_t5_cross_gather(const __nv_bfloat16* full, int Mp, int64_t docs, int T, int H, __nv_bfloat16* ctx, int ld_ctx,
                       cudaStream_t st);

int pn_t5_cross_softmax_bwd(const __nv_bfloat16* Pp, const float* dP, int ldd, const float* rs, int64_t docs, int S,
                            int TH, int Mp, __nv_bfloat16* U, int ldu, int u_col, __nv_bfloat16* V2, cudaStream_t st,
                            PnDrop drop = PnDrop(), int H = 0);   // drop: Pp = the clean probabilities (Pc above)
int pn_t5_cross_bt_pack(const __nv_bfloat16* dC, const __nv_bfloat16* QK, int64_t docs, int d, int Mp,
                        __nv_bfloat16* BT, int ldb, int k_col, cudaStream_t st);

// pn_t5.cu: (re)build every packed bf16 operand and the expanded bias tables from the flat fp32 parameters
int pn_t5_repack(PnT5* m, cudaStream_t st);

// pn_t5_train_ops.cu
int pn_t5_rs(const float* ss, int parts, int64_t rows, int d, float eps, float* rs, cudaStream_t st);
int pn_t5_relu_mask_scale(__nv_bfloat16* y, const __nv_bfloat16* act, const float* rs, int64_t rows, int cols,
                          cudaStream_t st);
int pn_t5_row_scale(__nv_bfloat16* y, int ld, const float* rs, int64_t rows, int cols, cudaStream_t st);
// gdrop: mask applied to G as it is read (the sub-layer's input was dropped: encoder final norm); odrop: mask applied
// to the bf16 copy dhb only (the NEXT consumer of dhb is the output projection of a sub-layer whose output was dropped)
int pn_t5_norm_bwd(const void* G, int g_bf16, const __nv_bfloat16* hb, const float* rs, int64_t rows, int d,
                   int accumulate, float* dh, __nv_bfloat16* dhb, cudaStream_t st, PnDrop gdrop = PnDrop(),
                   PnDrop odrop = PnDrop());
int pn_t5_unfold(const float* dWp, const float* W, const float* g, int64_t rows, int cols, float* gradW,
                 float* gradg, cudaStream_t st);
int pn_t5_embed_bwd(const int32_t* ids, const float* dh, int64_t rows, int d, int vocab, float* grad_table,
                    cudaStream_t st, PnDrop drop = PnDrop());
// ydrop: the norm's output was dropped (dy is masked on load); odrop: mask on the bf16 copy dxb only (see norm_bwd)
int pn_t5_final_norm_bwd(const float* dy, const float* x, const float* w, int64_t rows, int d, float eps,
                         float* dx, __nv_bfloat16* dxb, float* dw, cudaStream_t st, PnDrop ydrop = PnDrop(),
                         PnDrop odrop = PnDrop());
int pn_t5_sgd(float* p, const float* g, int64_t n, float lr, cudaStream_t st);
int pn_t5_adamw(float* p, const float* g, float* m, float* v, int64_t n, float lr, float b1, float b2, float eps,
                float wd, int64_t step, float gscale, cudaStream_t st);
