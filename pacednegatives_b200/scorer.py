"""monoT5 relevance scorer on the B200 — the model side of the reference's hot path
(``T5ForConditionalGeneration`` as called at lceT5.py:132-133,149-150, pairwrapper.py:223-247 and the
re-rank formula of distill/wrapper.py:34-38 / pyterrier_t5.MonoT5ReRanker in eval.py:14,23).

``MonoT5Scorer`` owns a ``PnT5`` handle of libpn_b200.so (tcgen05 GEMMs + fused attention); weights are
handed over by HuggingFace state_dict name, so a ``save_pretrained`` directory's tensors load as they are.
"""
import ctypes as C
import json
import os
from dataclasses import dataclass

import torch
import torch.distributed as dist

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


_TIED = ("encoder.embed_tokens.weight", "decoder.embed_tokens.weight", "lm_head.weight")


def dims_from_hf_config(cfg):
    """T5Dims from a HuggingFace ``config.json`` dict (the checkpoints lceT5.py:106 / eval.py:14 load)."""
    proj = cfg.get("feed_forward_proj", "relu")
    if proj != "relu":
        raise ValueError(f"feed_forward_proj={proj!r}: only the ReLU feed-forward of t5-* / monot5-* is built")
    if not cfg.get("tie_word_embeddings", True):
        raise ValueError("untied lm_head is not supported (t5-* / monot5-* tie it to the embedding)")
    return T5Dims(vocab=cfg["vocab_size"], d_model=cfg["d_model"], d_kv=cfg["d_kv"], d_ff=cfg["d_ff"],
                  n_layers=cfg["num_layers"], n_dec_layers=cfg.get("num_decoder_layers") or cfg["num_layers"],
                  n_heads=cfg["num_heads"], rel_buckets=cfg.get("relative_attention_num_buckets", 32),
                  rel_max_dist=cfg.get("relative_attention_max_distance", 128),
                  eps=cfg.get("layer_norm_epsilon", 1e-6), pad_id=cfg.get("pad_token_id", 0) or 0)


def hf_config(dims, base=None, dropout_rate=None):
    """The ``config.json`` a ``T5ForConditionalGeneration.from_pretrained`` needs to rebuild these dims.  ``base``:
    the config the weights were loaded from — every field of it is preserved (dropout_rate, task_specific_params,
    n_positions ...) and only the architecture fields are (re)written from ``dims``; without one, HF's t5 defaults
    (dropout_rate 0.1, what t5-* / monot5-* ship) are written."""
    cfg = dict(base) if base else {"architectures": ["T5ForConditionalGeneration"], "model_type": "t5",
                                   "dropout_rate": 0.1, "feed_forward_proj": "relu", "dense_act_fn": "relu",
                                   "is_gated_act": False, "is_encoder_decoder": True, "initializer_factor": 1.0,
                                   "tie_word_embeddings": True, "use_cache": True, "eos_token_id": 1}
    cfg.update({"vocab_size": dims.vocab, "d_model": dims.d_model, "d_kv": dims.d_kv, "d_ff": dims.d_ff,
                "num_layers": dims.n_layers, "num_decoder_layers": dims.n_dec_layers, "num_heads": dims.n_heads,
                "relative_attention_num_buckets": dims.rel_buckets,
                "relative_attention_max_distance": dims.rel_max_dist, "layer_norm_epsilon": dims.eps,
                "pad_token_id": dims.pad_id, "decoder_start_token_id": dims.pad_id})
    if dropout_rate is not None:
        cfg["dropout_rate"] = float(dropout_rate)
    return cfg


def _tensor_shape(name, numel, dims):
    """Shape of a parameter from its HF name (the flat buffer only records sizes)."""
    d, di, ff = dims.d_model, dims.d_inner, dims.d_ff
    if name == "shared.weight":
        return (dims.vocab, d)
    if name.endswith("relative_attention_bias.weight"):
        return (dims.rel_buckets, dims.n_heads)
    if name.endswith("layer_norm.weight"):
        return (d,)
    leaf = name.rsplit(".", 2)[-2]
    shape = {"q": (di, d), "k": (di, d), "v": (di, d), "o": (d, di), "wi": (ff, d), "wo": (d, ff)}.get(leaf)
    if shape is None or shape[0] * shape[1] != numel:
        raise KeyError(f"cannot infer the shape of {name} ({numel} elements)")
    return shape


def load_checkpoint(path, return_config=False):
    """(T5Dims, state_dict[, config dict]) of a local ``save_pretrained`` directory: config.json +
    model.safetensors or pytorch_model.bin.  No network access: ``path`` must exist."""
    if not os.path.isdir(path):
        raise FileNotFoundError(f"{path!r} is not a local checkpoint directory (there is no hub access: pass the "
                                "path of a save_pretrained directory with config.json + model.safetensors)")
    with open(os.path.join(path, "config.json")) as f:
        cfg = json.load(f)
    dims = dims_from_hf_config(cfg)
    st = os.path.join(path, "model.safetensors")
    if os.path.exists(st):
        from safetensors.torch import load_file
        sd = load_file(st)
    else:
        sd = torch.load(os.path.join(path, "pytorch_model.bin"), map_location="cpu", weights_only=True)
    return (dims, sd, cfg) if return_config else (dims, sd)


def _world(group=None):
    return dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1


def gather_rows(t, group=None, equal_shards=True):
    """Rank-major concatenation along dim 0 of every rank's `t` (the re-rank score / metric gather of SURVEY 8e
    collective 3: the reference evaluates ONE table over all queries, eval.py:23-26,41-43).  One
    all_gather_into_tensor when every rank holds the same number of rows (the sharding bench.py and
    MonoT5Scorer.rerank use); otherwise the row counts are exchanged first and the shards padded to the longest
    (one host sync to trim)."""
    W = _world(group)
    t = t.contiguous()
    if W == 1:
        return t
    if equal_shards:
        out = t.new_empty((W * t.shape[0],) + tuple(t.shape[1:]))
        dist.all_gather_into_tensor(out, t, group=group)
        return out
    n = torch.tensor([t.shape[0]], dtype=torch.int64, device=t.device)
    counts = torch.empty(W, dtype=torch.int64, device=t.device)
    dist.all_gather_into_tensor(counts, n, group=group)
    counts = counts.tolist()
    mx = max(counts)
    pad = t.new_zeros((mx,) + tuple(t.shape[1:]))
    pad[:t.shape[0]] = t
    out = t.new_empty((W * mx,) + tuple(t.shape[1:]))
    dist.all_gather_into_tensor(out, pad, group=group)
    return torch.cat([out[r * mx:r * mx + c] for r, c in enumerate(counts)])


def reduce_metric_means(metrics, group=None):
    """Global means of per-query measures [Q_local, M] over every rank's queries: one (M+1)-float SUM all-reduce
    of (sum of each measure, query count) — the metrics-only form of the gather above."""
    sums = torch.cat([metrics.double().sum(0), metrics.new_full((1,), metrics.shape[0], dtype=torch.float64)])
    if _world(group) > 1:
        dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
    return (sums[:-1] / sums[-1].clamp(min=1)).float()


class _DeviceFloats:
    """Zero-copy torch view of a device buffer the library owns."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<f4", "data": (int(ptr), False),
                                         "version": 2}


class MonoT5Scorer:
    """dims: T5Dims; state_dict: {hf_name: tensor} (any float dtype, CPU or CUDA)."""

    def __init__(self, dims, state_dict, rel_id=REL_ID, nrel_id=NREL_ID, chunk_docs=128, cta_group=2,
                 train=False, xattn=True, options=None):
        _lib.require_gpu()
        self.dims, self.rel_id, self.nrel_id = dims, int(rel_id), int(nrel_id)
        self.chunk_docs = int(chunk_docs)
        # measurement knob (PN_OVERLAP_CHUNKS=1): score() alternates its chunks between two streams
        self.overlap_chunks = os.environ.get("PN_OVERLAP_CHUNKS", "0") not in ("", "0")
        self._side, self._side_ws = None, [None, None]
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
        for oname, oval in (options or {}).items():     # pn_t5_set_option: dgrad_bf16, grad_ranges ...
            _lib.check(lib.pn_t5_set_option(self._h, oname.encode(), int(oval)))
        self.hf_config_base = None      # config.json of the checkpoint the weights came from (from_pretrained)
        known = {n for n, _, _ in self.named_offsets()} | set(_TIED)
        skipped = []
        for name, t in state_dict.items():
            if not torch.is_tensor(t) or not t.is_floating_point():
                continue
            if name not in known:
                # e.g. decoder.block.0.layer.1.EncDecAttention.relative_attention_bias.weight, which stock t5-* /
                # monot5-* checkpoints still carry and HF drops (_keys_to_ignore_on_load_unexpected); a MISSING
                # weight still fails in pn_t5_finalize
                skipped.append(name)
                continue
            src = t.detach().to(device="cuda", dtype=torch.float32).contiguous()
            _lib.check(lib.pn_t5_load_weight(self._h, name.encode(), _lib.ptr(src), src.numel(),
                                             _lib.stream()))
        if skipped:
            import warnings
            warnings.warn(f"MonoT5Scorer: ignored {len(skipped)} checkpoint tensor(s) outside the model: "
                          + ", ".join(skipped[:4]) + (" ..." if len(skipped) > 4 else ""))
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

    @classmethod
    def from_pretrained(cls, path, **kw):
        """Build from a local HuggingFace checkpoint directory (``T5ForConditionalGeneration.from_pretrained``
        at lceT5.py:106, ``MonoT5ReRanker(model=...)`` at eval.py:14)."""
        dims, sd, cfg = load_checkpoint(path, return_config=True)
        m = cls(dims, sd, **kw)
        m.hf_config_base = cfg
        return m

    def parameters_flat(self):
        """fp32 view [num_params] of the library's master parameters (HF state_dict order); shares memory."""
        lib = _lib.lib()
        ptr = lib.pn_t5_params(self._h)
        if not ptr:
            raise RuntimeError(lib.pn_last_error().decode())
        return torch.as_tensor(_DeviceFloats(ptr, lib.pn_t5_num_params(self._h)), device="cuda")

    def named_offsets(self):
        """[(hf_name, offset, numel)] of the flat parameter / gradient buffers."""
        lib = _lib.lib()
        out = []
        name = C.create_string_buffer(256)
        off, numel = C.c_int64(), C.c_int64()
        for i in range(int(lib.pn_t5_num_tensors(self._h))):
            _lib.check(lib.pn_t5_tensor_info(self._h, i, name, 256, C.byref(off), C.byref(numel)))
            out.append((name.value.decode(), off.value, numel.value))
        return out

    def state_dict(self, tied=False):
        """{hf_name: fp32 CPU tensor} of the current master parameters; ``tied`` adds the three aliases of
        shared.weight that older HF checkpoints carry."""
        flat = self.parameters_flat()
        sd = {n: flat[o:o + k].view(_tensor_shape(n, k, self.dims)).cpu() for n, o, k in self.named_offsets()}
        if tied:
            for alias in _TIED:
                sd[alias] = sd["shared.weight"]
        return sd

    def save_pretrained(self, path, safe_serialization=True):
        """Write config.json + model.safetensors (or pytorch_model.bin) that ``T5ForConditionalGeneration
        .from_pretrained(path)`` loads — train/train_lce.py:103 calls this on the trained model."""
        os.makedirs(path, exist_ok=True)
        with open(os.path.join(path, "config.json"), "w") as f:
            # no source config (weights given directly): HF's t5 default rate unless this model trained with one
            rate = getattr(self, "dropout_rate", None) or None
            json.dump(hf_config(self.dims, self.hf_config_base, rate if self.hf_config_base is None else None), f,
                      indent=2)
        sd = {k: v.contiguous() for k, v in self.state_dict().items()}
        if safe_serialization:
            from safetensors.torch import save_file
            save_file(sd, os.path.join(path, "model.safetensors"), metadata={"format": "pt"})
        else:
            torch.save(sd, os.path.join(path, "pytorch_model.bin"))

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

    @staticmethod
    def _lengths(attention_mask, lengths, N, S):
        """int32 [N] device lengths from the tokenizer's attention_mask (right padding, as T5Tokenizer pads) or
        explicit lengths; None when neither is given or nothing is padded."""
        if lengths is None and attention_mask is None:
            return None
        if lengths is None:
            m = torch.as_tensor(attention_mask)
            if tuple(m.shape) != (N, S):
                raise ValueError(f"attention_mask must be [{N}, {S}], got {tuple(m.shape)}")
            m = m.to(device="cuda") != 0
            lengths = m.sum(1)
            prefix = torch.arange(S, device="cuda")[None, :] < lengths[:, None]
            if not bool((m == prefix).all()):
                raise NotImplementedError("attention_mask must be a prefix of ones per row (right padding, what "
                                          "T5Tokenizer produces); left padding / holes are not supported")
        lens = torch.as_tensor(lengths).to(device="cuda", dtype=torch.int32).contiguous()
        if lens.numel() != N:
            raise ValueError(f"lengths must have {N} entries")
        if bool((lens < 1).any()) or bool((lens > S).any()):
            raise ValueError("every prompt needs 1 <= length <= S")
        return lens

    def score(self, input_ids, return_z2=False, attention_mask=None, lengths=None):
        """input_ids [N, S] (int32/int64, CPU or CUDA) -> score fp32 [N] (= log_softmax over the
        'true'/'false' logits of the first decoder step), optionally z2 [N, 2].

        ``attention_mask`` [N, S] (or ``lengths`` [N]): the tokenizer's mask for prompts right-padded to the batch
        maximum, as pyterrier_t5.MonoT5ReRanker (eval.py:14,23) and distill/wrapper.py:36 pass it — padded keys are
        excluded from the encoder self-attention and the cross-attention.  Without one every position is attended
        (the training path of the reference drops the mask: lceT5.py:53, pairwrapper.py:79-80)."""
        ids = self._ids(input_ids)
        N, S = ids.shape
        lens = self._lengths(attention_mask, lengths, N, S)
        score = torch.empty(N, dtype=torch.float32, device="cuda")
        z2 = torch.empty(N, 2, dtype=torch.float32, device="cuda")
        lib = _lib.lib()
        if self.overlap_chunks and N > self.chunk_docs:
            self._score_two_streams(ids, lens, N, S, score, z2)
            return (score, z2) if return_z2 else score
        for lo in range(0, N, self.chunk_docs):
            n = min(self.chunk_docs, N - lo)
            ws, need = self._workspace(min(self.chunk_docs, N), S, 1)
            _lib.check(lib.pn_t5_score_lens(self._h, _lib.ptr(ids[lo:lo + n]),
                                            _lib.ptr(lens[lo:lo + n]) if lens is not None else None, n, S,
                                            self.rel_id, self.nrel_id, _lib.ptr(score[lo:lo + n]),
                                            _lib.ptr(z2[lo:lo + n]), _lib.ptr(ws), ws.numel(), _lib.stream()))
        return (score, z2) if return_z2 else score

    def _score_two_streams(self, ids, lens, N, S, score, z2):
        """Chunks alternate between two side streams with a workspace each (option ``overlap_chunks``): the decoder of
        chunk i — ~240 launches on a handful of CTAs — shares the GPU with the encoder of chunk i + 1 instead of
        running alone.  Same kernels, same results; the caller's stream waits for both side streams at the end."""
        lib = _lib.lib()
        cur = torch.cuda.current_stream()
        if self._side is None:
            self._side = [torch.cuda.Stream(), torch.cuda.Stream()]
        need = int(lib.pn_t5_workspace_bytes(self._h, self.chunk_docs, S, 1, 0))
        for i in range(2):
            if self._side_ws[i] is None or self._side_ws[i].numel() < need:
                self._side_ws[i] = torch.empty(need, dtype=torch.uint8, device="cuda")
        start = torch.cuda.Event()
        start.record(cur)
        for st in self._side:
            st.wait_event(start)
        for i, lo in enumerate(range(0, N, self.chunk_docs)):
            n = min(self.chunk_docs, N - lo)
            ws, st = self._side_ws[i & 1], self._side[i & 1]
            _lib.check(lib.pn_t5_score_lens(self._h, _lib.ptr(ids[lo:lo + n]),
                                            _lib.ptr(lens[lo:lo + n]) if lens is not None else None, n, S,
                                            self.rel_id, self.nrel_id, _lib.ptr(score[lo:lo + n]),
                                            _lib.ptr(z2[lo:lo + n]), _lib.ptr(ws), ws.numel(), st.cuda_stream))
        for st in self._side:
            done = torch.cuda.Event()
            done.record(st)
            cur.wait_event(done)

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

    def rerank(self, input_ids, docids=None, rels=None, K=10, group=None, gather=None, equal_shards=True,
               attention_mask=None, lengths=None):
        """Score Q x C candidate prompts (input_ids [Q, C, S]), rank each query's candidates and, when
        ``rels`` [Q, C] is given, evaluate nDCG@K / RR / RR@K / AP (eval.py:26,41-43).  ``attention_mask``
        [Q, C, S] / ``lengths`` [Q, C]: the tokenizer's padding mask (see ``score``).

        Data parallel (queries sharded over the ranks of ``group``; all C candidates of a query stay on one rank):
        ``gather='scores'`` adds ``scores_all`` [sum Q, C] (+ ``top_ids_all``, ``top_scores_all``, ``metrics_all``) —
        the one global result table the reference evaluates (eval.py:23-26) — and ``mean_metrics``;
        ``gather='metrics'`` only all-reduces the (M+1) floats behind ``mean_metrics``.  Default: 'scores' when a
        process group with more than one rank is initialised, else nothing."""
        ids = torch.as_tensor(input_ids)
        Q, Cn, S = ids.shape
        if attention_mask is not None:
            attention_mask = torch.as_tensor(attention_mask).reshape(Q * Cn, S)
        if lengths is not None:
            lengths = torch.as_tensor(lengths).reshape(Q * Cn)
        scores = self.score(ids.reshape(Q * Cn, S), attention_mask=attention_mask, lengths=lengths).view(Q, Cn)
        out = topk_metrics(scores, docids, rels, K)
        out["scores"] = scores
        if gather is None:
            gather = "scores" if _world(group) > 1 else "none"
        if gather == "scores":
            out["scores_all"] = gather_rows(scores, group, equal_shards)
            out["top_ids_all"] = gather_rows(out["top_ids"], group, equal_shards)
            out["top_scores_all"] = gather_rows(out["top_scores"], group, equal_shards)
            if "metrics" in out:
                out["metrics_all"] = gather_rows(out["metrics"], group, equal_shards)
        if gather in ("scores", "metrics") and "metrics" in out:
            out["mean_metrics"] = reduce_metric_means(out["metrics"], group)
        return out

