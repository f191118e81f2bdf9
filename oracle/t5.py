"""Oracle: monoT5 forward in fp32 on CPU (test infrastructure only; see oracle/__init__.py).

The model arithmetic of the reference lives in HuggingFace ``transformers`` (unpinned by the reference;
5.5.0 in the build image), called at lceT5.py:132-133,149-150, pairwrapper.py:223-247 and
distill/wrapper.py:34-38.  This restates ``T5ForConditionalGeneration.forward`` (HF = transformers/
models/t5/modeling_t5.py): RMSNorm HF:55-69, FFN HF:84-103, relative position buckets HF:188-234,
attention HF:253-345 (no 1/sqrt(d) scaling, additive learned bias, fp32 softmax), shift_right
HF:595-614, tied head scaled by d_model**-0.5 HF:1107-1110.  Dropout off (eval mode), no attention
mask (the reference passes none).  ``tests/test_oracle_golden.py`` checks this against HF
itself whenever transformers is importable, and against vectors generated from HF in
tests/golden/t5_small_golden.npz.
"""
from dataclasses import dataclass
import math

import numpy as np
import torch


@dataclass
class T5Dims:
    vocab: int = 32128
    d_model: int = 512
    d_kv: int = 64
    d_ff: int = 2048
    n_layers: int = 6       # encoder layers
    n_dec_layers: int = 6
    n_heads: int = 8
    rel_buckets: int = 32
    rel_max_dist: int = 128
    eps: float = 1e-6
    pad_id: int = 0         # also decoder_start_token_id

    @property
    def d_inner(self):
        return self.n_heads * self.d_kv


T5_SMALL = T5Dims()
T5_BASE = T5Dims(d_model=768, d_ff=3072, n_layers=12, n_dec_layers=12, n_heads=12)
T5_LARGE = T5Dims(d_model=1024, d_ff=4096, n_layers=24, n_dec_layers=24, n_heads=16)


def init_weights(dims: T5Dims, seed: int = 0, scale: float = 1.0):
    """Random weights under HF state_dict names with HF's init std (T5PreTrainedModel._init_weights,
    initializer_factor 1.0).  ``scale`` shrinks the embedding std (1.0 in HF)."""
    g = torch.Generator().manual_seed(seed)
    d, di, ff = dims.d_model, dims.d_inner, dims.d_ff
    w = {}

    def normal(shape, std):
        return torch.randn(shape, generator=g, dtype=torch.float32) * std

    w["shared.weight"] = normal((dims.vocab, d), 1.0 * scale)
    for stack, L, dec in (("encoder", dims.n_layers, False), ("decoder", dims.n_dec_layers, True)):
        for i in range(L):
            p = f"{stack}.block.{i}.layer."
            w[p + "0.SelfAttention.q.weight"] = normal((di, d), (d * dims.d_kv) ** -0.5)
            w[p + "0.SelfAttention.k.weight"] = normal((di, d), d ** -0.5)
            w[p + "0.SelfAttention.v.weight"] = normal((di, d), d ** -0.5)
            w[p + "0.SelfAttention.o.weight"] = normal((d, di), di ** -0.5)
            if i == 0:
                w[p + "0.SelfAttention.relative_attention_bias.weight"] = normal(
                    (dims.rel_buckets, dims.n_heads), d ** -0.5)
            w[p + "0.layer_norm.weight"] = 1.0 + 0.1 * normal((d,), 1.0)
            ffi = 1
            if dec:
                w[p + "1.EncDecAttention.q.weight"] = normal((di, d), (d * dims.d_kv) ** -0.5)
                w[p + "1.EncDecAttention.k.weight"] = normal((di, d), d ** -0.5)
                w[p + "1.EncDecAttention.v.weight"] = normal((di, d), d ** -0.5)
                w[p + "1.EncDecAttention.o.weight"] = normal((d, di), di ** -0.5)
                w[p + "1.layer_norm.weight"] = 1.0 + 0.1 * normal((d,), 1.0)
                ffi = 2
            w[p + f"{ffi}.DenseReluDense.wi.weight"] = normal((ff, d), d ** -0.5)
            w[p + f"{ffi}.DenseReluDense.wo.weight"] = normal((d, ff), ff ** -0.5)
            w[p + f"{ffi}.layer_norm.weight"] = 1.0 + 0.1 * normal((d,), 1.0)
        w[f"{stack}.final_layer_norm.weight"] = 1.0 + 0.1 * normal((d,), 1.0)
    return w


def relative_bucket(rel, bidirectional, num_buckets=32, max_distance=128):
    """HF:188-234 in exact integer arithmetic.  For n >= max_exact HF computes
    max_exact + trunc(log(n/max_exact)/log(max_distance/max_exact) * (num_buckets - max_exact)) in
    fp32; with the T5 constants (max_distance/max_exact = 16, num_buckets - max_exact = 8 or 16) this
    equals the integer expression below for every n (checked against HF in the tests)."""
    rel = np.asarray(rel, dtype=np.int64)
    ret = np.zeros_like(rel)
    nb = num_buckets
    if bidirectional:
        nb //= 2
        ret = ret + (rel > 0).astype(np.int64) * nb
        n = np.abs(rel)
    else:
        n = -np.minimum(rel, 0)
    max_exact = nb // 2
    with np.errstate(divide="ignore"):
        val = max_exact + (np.log(np.maximum(n, 1).astype(np.float32) / np.float32(max_exact))
                           / np.float32(math.log(max_distance / max_exact))
                           * np.float32(nb - max_exact)).astype(np.int64)
    val = np.minimum(val, nb - 1)
    return ret + np.where(n < max_exact, n, val)


def _rms(x, w, eps):
    var = x.pow(2).mean(-1, keepdim=True)
    return w * (x * torch.rsqrt(var + eps))


def _key_mask(attention_mask):
    """HF's extended attention mask (modeling_utils get_extended_attention_mask, used at HF:730-760): padded keys
    get the dtype's most negative value added to their scores, in the encoder self-attention and, through
    encoder_attention_mask, in every cross-attention.  [N, S] 0/1 -> additive [N, 1, 1, S] (None -> None)."""
    if attention_mask is None:
        return None
    m = torch.as_tensor(np.asarray(attention_mask), dtype=torch.float32)
    return ((1.0 - m) * torch.finfo(torch.float32).min)[:, None, None, :]


def _attn(xq, xkv, w, prefix, dims, bias):
    N, Tq, _ = xq.shape
    Tk = xkv.shape[1]
    H, dk = dims.n_heads, dims.d_kv
    q = (xq @ w[prefix + "q.weight"].T).view(N, Tq, H, dk).transpose(1, 2)
    k = (xkv @ w[prefix + "k.weight"].T).view(N, Tk, H, dk).transpose(1, 2)
    v = (xkv @ w[prefix + "v.weight"].T).view(N, Tk, H, dk).transpose(1, 2)
    s = q @ k.transpose(2, 3)
    if bias is not None:
        s = s + bias
    p = torch.softmax(s, dim=-1)
    o = (p @ v).transpose(1, 2).reshape(N, Tq, H * dk)
    return o @ w[prefix + "o.weight"].T


def encoder_bias(w, dims, S):
    pos = np.arange(S)
    b = relative_bucket(pos[None, :] - pos[:, None], True, dims.rel_buckets, dims.rel_max_dist)
    tab = w["encoder.block.0.layer.0.SelfAttention.relative_attention_bias.weight"]
    return tab[torch.from_numpy(b)].permute(2, 0, 1).unsqueeze(0)  # [1,H,S,S]


def decoder_bias(w, dims, T):
    pos = np.arange(T)
    rel = pos[None, :] - pos[:, None]
    b = relative_bucket(rel, False, dims.rel_buckets, dims.rel_max_dist)
    tab = w["decoder.block.0.layer.0.SelfAttention.relative_attention_bias.weight"]
    bias = tab[torch.from_numpy(b)].permute(2, 0, 1).unsqueeze(0)
    causal = torch.full((T, T), float("-inf")).triu(1)
    return bias + causal


@torch.no_grad()
def encode(w, dims, input_ids, attention_mask=None):
    ids = torch.as_tensor(np.asarray(input_ids), dtype=torch.long)
    x = w["shared.weight"][ids]
    bias = encoder_bias(w, dims, ids.shape[1])
    km = _key_mask(attention_mask)
    if km is not None:
        bias = bias + km
    for i in range(dims.n_layers):
        p = f"encoder.block.{i}.layer."
        h = _rms(x, w[p + "0.layer_norm.weight"], dims.eps)
        x = x + _attn(h, h, w, p + "0.SelfAttention.", dims, bias)
        h = _rms(x, w[p + "1.layer_norm.weight"], dims.eps)
        x = x + torch.relu(h @ w[p + "1.DenseReluDense.wi.weight"].T) @ w[p + "1.DenseReluDense.wo.weight"].T
    return _rms(x, w["encoder.final_layer_norm.weight"], dims.eps)


@torch.no_grad()
def decode(w, dims, enc, decoder_input_ids, attention_mask=None):
    """Returns the decoder output after the final norm and the d_model**-0.5 head scaling, [N,T,d]."""
    ids = torch.as_tensor(np.asarray(decoder_input_ids), dtype=torch.long)
    x = w["shared.weight"][ids]
    bias = decoder_bias(w, dims, ids.shape[1])
    km = _key_mask(attention_mask)
    for i in range(dims.n_dec_layers):
        p = f"decoder.block.{i}.layer."
        h = _rms(x, w[p + "0.layer_norm.weight"], dims.eps)
        x = x + _attn(h, h, w, p + "0.SelfAttention.", dims, bias)
        h = _rms(x, w[p + "1.layer_norm.weight"], dims.eps)
        x = x + _attn(h, enc, w, p + "1.EncDecAttention.", dims, km)
        h = _rms(x, w[p + "2.layer_norm.weight"], dims.eps)
        x = x + torch.relu(h @ w[p + "2.DenseReluDense.wi.weight"].T) @ w[p + "2.DenseReluDense.wo.weight"].T
    x = _rms(x, w["decoder.final_layer_norm.weight"], dims.eps)
    return x * (dims.d_model ** -0.5)


def shift_right(labels, start_id=0, pad_id=0):
    """HF:595-614."""
    labels = np.asarray(labels, dtype=np.int64)
    out = np.zeros_like(labels)
    out[:, 1:] = labels[:, :-1]
    out[:, 0] = start_id
    out[out == -100] = pad_id
    return out


@torch.no_grad()
def logits(w, dims, input_ids, labels):
    """model(input_ids=, labels=).logits  -> float32 [N, T, V]."""
    enc = encode(w, dims, input_ids)
    y = decode(w, dims, enc, shift_right(labels, dims.pad_id, dims.pad_id))
    return (y @ w["shared.weight"].T).numpy()


@torch.no_grad()
def score(w, dims, input_ids, rel_id, nrel_id, attention_mask=None):
    """monoT5 re-rank score (distill/wrapper.py:34-38) with one decoder step; ``attention_mask`` [N, S] as
    pyterrier_t5.MonoT5ReRanker / distill/wrapper.py:36 pass it (``self.model(**enc)`` with padding='longest').
    Returns (score float32[N], z2 float32[N,2] = (z_true, z_false))."""
    enc = encode(w, dims, input_ids, attention_mask)
    N = enc.shape[0]
    y = decode(w, dims, enc, np.full((N, 1), dims.pad_id, dtype=np.int64), attention_mask)[:, 0]
    rows = w["shared.weight"][[rel_id, nrel_id]]
    z2 = (y @ rows.T)
    s = torch.log_softmax(z2, dim=1)[:, 0]
    return s.numpy(), z2.numpy()
