"""monoT5 relevance scorer on the B200 — the model side of the reference's hot path
(``T5ForConditionalGeneration`` as called at lceT5.py:132-133,149-150, pairwrapper.py:223-247 and the
re-rank formula of distill/wrapper.py:34-38 / pyterrier_t5.MonoT5ReRanker in eval.py:14,23).

``MonoT5Scorer`` owns a ``PnT5`` handle of libpn_b200.so (tcgen05 GEMMs + fused attention); weights are
handed over by HuggingFace state_dict name, so a ``save_pretrained`` directory's tensors load as they are.
"""
import ctypes as C
from dataclasses import dataclass

import torch

from . import _lib
from .metrics import topk_metrics


class _Cfg(C.Structure):
    _fields_ = [("vocab", C.c_int32), ("d_model", C.c_int32), ("d_kv", C.c_int32), ("d_ff", C.c_int32),
                ("n_layers", C.c_int32), ("n_dec_layers", C.c_int32), ("n_heads", C.c_int32),
                ("rel_buckets", C.c_int32), ("rel_max_dist", C.c_int32), ("pad_id", C.c_int32),
                ("eps", C.c_float)]


@dataclass
class T5Dims:
    vocab: int = 32128
    d_model: int = 512
    d_kv: int = 64
    d_ff: int = 2048
    n_layers: int = 6
    n_dec_layers: int = 6
    n_heads: int = 8
    rel_buckets: int = 32
    rel_max_dist: int = 128
    eps: float = 1e-6
    pad_id: int = 0

    @property
    def d_inner(self):
        return self.n_heads * self.d_kv


T5_SMALL = T5Dims()
T5_BASE = T5Dims(d_model=768, d_ff=3072, n_layers=12, n_dec_layers=12, n_heads=12)
T5_LARGE = T5Dims(d_model=1024, d_ff=4096, n_layers=24, n_dec_layers=24, n_heads=16)

# monoT5 convention (castorini/monot5-*): first sentencepiece ids of 'true' / 'false'
REL_ID, NREL_ID = 1176, 6136


def flops_per_doc(dims, S, T=1, full_vocab=False):
    """Algorithmic forward FLOPs of one document (SURVEY §8d formula, MAC x 2)."""
    d, di, ff = dims.d_model, dims.d_inner, dims.d_ff
    Le, Ld = dims.n_layers, dims.n_dec_layers
    R = dims.vocab if full_vocab else 2
    return 2 * (Le * (4 * d * di + 2 * d * ff) * S + Le * 2 * S * S * di + Ld * 2 * d * di * S
                + Ld * (6 * d * di + 2 * d * ff) * T + Ld * (2 * T * T + 2 * T * S) * di + R * d * T)


def random_state_dict(dims, seed=0, device="cuda"):
    """Random-init weights under HuggingFace state_dict names with HF's initialiser std
    (T5PreTrainedModel._init_weights, initializer_factor 1.0): what `T5ForConditionalGeneration(config)`
    gives the reference when no checkpoint is loaded (there is no network for real checkpoints)."""
    g = torch.Generator(device=device).manual_seed(int(seed))
    d, di, ff = dims.d_model, dims.d_inner, dims.d_ff

    def normal(shape, std):
        return torch.randn(shape, generator=g, dtype=torch.float32, device=device) * std

    w = {"shared.weight": normal((dims.vocab, d), 1.0)}
    for stack, L, dec in (("encoder", dims.n_layers, False), ("decoder", dims.n_dec_layers, True)):
        for i in range(L):
            p = f"{stack}.block.{i}.layer."
            for pre in (["0.SelfAttention."] + (["1.EncDecAttention."] if dec else [])):
                w[p + pre + "q.weight"] = normal((di, d), (d * dims.d_kv) ** -0.5)
                w[p + pre + "k.weight"] = normal((di, d), d ** -0.5)
                w[p + pre + "v.weight"] = normal((di, d), d ** -0.5)
                w[p + pre + "o.weight"] = normal((d, di), di ** -0.5)
            if i == 0:
                w[p + "0.SelfAttention.relative_attention_bias.weight"] = normal(
                    (dims.rel_buckets, dims.n_heads), d ** -0.5)
            ffi = 2 if dec else 1
            w[p + f"{ffi}.DenseReluDense.wi.weight"] = normal((ff, d), d ** -0.5)
            w[p + f"{ffi}.DenseReluDense.wo.weight"] = normal((d, ff), ff ** -0.5)
            for j in range(ffi + 1):
                w[p + f"{j}.layer_norm.weight"] = torch.ones(d, dtype=torch.float32, device=device)
        w[f"{stack}.final_layer_norm.weight"] = torch.ones(d, dtype=torch.float32, device=device)
    return w


class MonoT5Scorer:
    """dims: T5Dims; state_dict: {hf_name: tensor} (any float dtype, CPU or CUDA)."""

    def __init__(self, dims, state_dict, rel_id=REL_ID, nrel_id=NREL_ID, chunk_docs=128, cta_group=2,
                 train=False, xattn=True):
        _lib.require_gpu()
        self.dims, self.rel_id, self.nrel_id = dims, int(rel_id), int(nrel_id)
        self.chunk_docs = int(chunk_docs)
        lib = _lib.lib()
        cfg = _Cfg(dims.vocab, dims.d_model, dims.d_kv, dims.d_ff, dims.n_layers, dims.n_dec_layers,
                   dims.n_heads, dims.rel_buckets, dims.rel_max_dist, dims.pad_id, dims.eps)
        h = C.c_void_p()
        _lib.check(lib.pn_t5_create(C.byref(cfg), C.byref(h)))
        self._h = h
        _lib.check(lib.pn_t5_set_option(self._h, b"cta_group", cta_group))
        # xattn: cross-attention as X (Wk^T q) / Wv (P^T X) instead of projecting every encoder row to K and V
        _lib.check(lib.pn_t5_set_option(self._h, b"xattn", int(bool(xattn))))
        if train:   # also keep the transposed operands of the data-gradient GEMMs (must precede the weights)
            _lib.check(lib.pn_t5_set_option(self._h, b"train", 1))
        for name, t in state_dict.items():
            if not torch.is_tensor(t) or not t.is_floating_point():
                continue
            src = t.detach().to(device="cuda", dtype=torch.float32).contiguous()
            _lib.check(lib.pn_t5_load_weight(self._h, name.encode(), _lib.ptr(src), src.numel(),
                                             _lib.stream()))
        _lib.check(lib.pn_t5_finalize(self._h, _lib.stream()))
        self.vocab_pad = int(lib.pn_t5_vocab_pad(self._h))
        self._ws, self._ws_key = None, None

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _lib.lib().pn_t5_destroy(h)
            except Exception:
                pass

    def _workspace(self, docs, S, T):
        key = (docs, S, T)
        need = int(_lib.lib().pn_t5_workspace_bytes(self._h, docs, S, T, 0))
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device="cuda")
        return self._ws, need

    @staticmethod
    def _ids(input_ids):
        ids = torch.as_tensor(input_ids)
        if not ids.is_cuda:
            ids = ids.cuda(non_blocking=True)
        if ids.dtype != torch.int32:
            ids = ids.to(torch.int32)
        return ids.contiguous()

    def score(self, input_ids, return_z2=False):
        """input_ids [N, S] (int32/int64, CPU or CUDA) -> score fp32 [N] (= log_softmax over the
        'true'/'false' logits of the first decoder step), optionally z2 [N, 2]."""
        ids = self._ids(input_ids)
        N, S = ids.shape
        score = torch.empty(N, dtype=torch.float32, device="cuda")
        z2 = torch.empty(N, 2, dtype=torch.float32, device="cuda")
        lib = _lib.lib()
        for lo in range(0, N, self.chunk_docs):
            n = min(self.chunk_docs, N - lo)
            ws, need = self._workspace(min(self.chunk_docs, N), S, 1)
            _lib.check(lib.pn_t5_score(self._h, _lib.ptr(ids[lo:lo + n]), n, S, self.rel_id,
                                       self.nrel_id, _lib.ptr(score[lo:lo + n]), _lib.ptr(z2[lo:lo + n]),
                                       _lib.ptr(ws), ws.numel(), _lib.stream()))
        return (score, z2) if return_z2 else score

    def logits(self, input_ids, labels, want_logits=True):
        """model(input_ids=, labels=).logits for T = 1 or 2 label positions.
        Returns (logits fp32 [N, T, vocab] | None, z2 fp32 [N, T, 2])."""
        ids = self._ids(input_ids)
        lab = torch.as_tensor(labels).to(device="cuda", dtype=torch.int64).contiguous()
        N, S = ids.shape
        T = lab.shape[1]
        z2 = torch.empty(N * T, 2, dtype=torch.float32, device="cuda")
        logits = torch.empty(N * T, self.vocab_pad, dtype=torch.float32, device="cuda") if want_logits else None
        lib = _lib.lib()
        for lo in range(0, N, self.chunk_docs):
            n = min(self.chunk_docs, N - lo)
            ws, need = self._workspace(min(self.chunk_docs, N), S, T)
            lg = logits[lo * T:(lo + n) * T] if want_logits else None
            _lib.check(lib.pn_t5_logits(self._h, _lib.ptr(ids[lo:lo + n]), _lib.ptr(lab[lo:lo + n]), n, S,
                                        T, self.rel_id, self.nrel_id, _lib.ptr(lg),
                                        _lib.ptr(z2[lo * T:(lo + n) * T]), _lib.ptr(ws), ws.numel(),
                                        _lib.stream()))
        if want_logits:
            logits = logits.view(N, T, self.vocab_pad)[:, :, :self.dims.vocab]
        return logits, z2.view(N, T, 2)

    def rerank(self, input_ids, docids=None, rels=None, K=10):
        """Score Q x C candidate prompts (input_ids [Q, C, S]), rank each query's candidates and, when
        ``rels`` [Q, C] is given, evaluate nDCG@K / RR / RR@K / AP (eval.py:26,41-43)."""
        ids = torch.as_tensor(input_ids)
        Q, Cn, S = ids.shape
        scores = self.score(ids.reshape(Q * Cn, S)).view(Q, Cn)
        out = topk_metrics(scores, docids, rels, K)
        out["scores"] = scores
        return out

