"""Oracle: paced negative selection (test infrastructure only; see oracle/__init__.py).

Restates
  * TripletDataset.get_items            /root/reference/pacednegatives/dataloader.py:29-33
  * PairLoader / LevelLoader.get_batch  /root/reference/pacednegatives/dataloader.py:46-56,69-79
  * LCEDataset.sample                   /root/reference/pacednegatives/dataloader.py:134-151
  * level bump / interpolation          /root/reference/old/levels.py:74-78, old/interp.py:10-18

``LCEDataset.sample`` bottoms out in two third-party routines that are not in the reference tree:
``scipy.stats.binom.pmf`` (scipy, unpinned by the reference; 1.18.1 in the build image) and
``numpy.random.RandomState.choice(replace=False, p=...)`` (numpy, unpinned; 2.3.5 here).  ``legacy_choice``
restates numpy's published legacy algorithm (numpy/random/mtrand.pyx, ``choice``) with the uniform
draws made explicit, so a GPU kernel can consume the very same doubles.
"""
from math import ceil, floor, lgamma, log, log1p, exp

import numpy as np


def level_index(weight, C, use_ceil=False):
    """dataloader.py:19,22,32: n_neg = len(neg_idx[0]) - 1; r = round(weight * n_neg) with Python
    floats.  Returns -1 where Python would raise IndexError or wrap a negative index."""
    n_neg = C - 1
    x = float(weight) * n_neg
    r = ceil(x) if use_ceil else floor(x)
    return int(r) if 0 <= r < C else -1


def level_select(cand, weights, use_ceil=False):
    """cand int32 [Q, C]; weights scalar or [Q].  Returns (idx [Q], doc [Q])."""
    cand = np.asarray(cand)
    Q, C = cand.shape
    w = np.broadcast_to(np.asarray(weights), (Q,))
    idx = np.array([level_index(float(x), C, use_ceil) for x in w], dtype=np.int32)
    doc = np.where(idx >= 0, cand[np.arange(Q), np.maximum(idx, 0)], -1).astype(np.int32)
    return idx, doc


def level_bump(difficulty, n_neg):
    """old/levels.py:77: difficulty = min(1., difficulty + 1/n_neg)."""
    return min(1.0, difficulty + 1.0 / n_neg)


def interpolate_scalar(start_value, end_value, num_steps):
    """old/interp.py:10-18."""
    step_size = (end_value - start_value) / num_steps

    def get(step):
        if step > num_steps:
            return end_value
        return start_value + step * step_size

    return get


def binom_pmf_lgamma(C, mean):
    """The device-side pmf formula (used when scipy's table is not supplied)."""
    n = C - 1
    lp, lq, lgn = log(mean), log1p(-mean), lgamma(n + 1.0)
    return np.array([
        exp(lgn - lgamma(k + 1.0) - lgamma(n - k + 1.0) + k * lp + (n - k) * lq) for k in range(C)
    ])


def binomial_table(mean, C, var, clip_min=1e-10, clip_max=1.0 - 1e-10, pmf=None):
    """dataloader.py:135-145 followed by the first-round cdf of numpy's choice.

    Returns (p float64[C], cdf float64[C], n_nonzero).  ``pmf`` overrides scipy's table."""
    mean = np.clip(mean, clip_min, clip_max)
    if pmf is None:
        from scipy.stats import binom
        pmf = binom.pmf(np.arange(C), C - 1, mean)
    p = np.asarray(pmf, dtype=np.float64)
    adjusted = p / p.sum()
    adjusted_variance = np.var(adjusted)
    with np.errstate(divide="ignore", invalid="ignore"):
        scaling_factor = np.sqrt(var / adjusted_variance)
        adjusted = adjusted * scaling_factor
        adjusted = adjusted / adjusted.sum()
    cdf = np.cumsum(adjusted)
    cdf = cdf / cdf[-1]
    return adjusted, cdf, int(np.count_nonzero(adjusted > 0))


def legacy_choice(p, k, uniforms):
    """numpy RandomState.choice(len(p), size=(k,), replace=False, p=p) with explicit uniforms.

    Returns (indices int64[k], consumed).  Raises ValueError like numpy when fewer than k entries of p
    are non-zero, IndexError when ``uniforms`` runs out."""
    p = np.array(p, dtype=np.float64, copy=True)
    if np.count_nonzero(p > 0) < k:
        raise ValueError("Fewer non-zero entries in p than size")
    found = np.zeros(k, dtype=np.int64)
    n_uniq = 0
    pos = 0
    while n_uniq < k:
        m = k - n_uniq
        if pos + m > len(uniforms):
            raise IndexError("uniform stream exhausted")
        x = np.asarray(uniforms[pos:pos + m], dtype=np.float64)
        pos += m
        if n_uniq > 0:
            p[found[0:n_uniq]] = 0
        cdf = np.cumsum(p)
        cdf /= cdf[-1]
        new = cdf.searchsorted(x, side="right")
        _, unique_indices = np.unique(new, return_index=True)
        unique_indices.sort()
        new = new.take(unique_indices)
        found[n_uniq:n_uniq + new.size] = new
        n_uniq += new.size
    return found, pos


def binomial_select(cand, mean, var, k, uniforms, stream_mode=0, pmf=None):
    """cand [Q, C]; uniforms [Q, U] (stream_mode 0) or flat [U] consumed in query order (mode 1).
    Returns (idx [Q,k] int32, doc [Q,k] int32, consumed [Q] int32)."""
    cand = np.asarray(cand)
    Q, C = cand.shape
    p, _, _ = binomial_table(mean, C, var, pmf=pmf)
    idx = np.zeros((Q, k), dtype=np.int32)
    consumed = np.zeros(Q, dtype=np.int32)
    pos = 0
    for q in range(Q):
        if stream_mode == 0:
            f, c = legacy_choice(p, k, uniforms[q])
        else:
            f, c = legacy_choice(p, k, uniforms[pos:])
            pos += c
        idx[q] = f
        consumed[q] = c
    doc = np.take_along_axis(cand, idx.astype(np.int64), axis=1).astype(np.int32)
    return idx, doc, consumed


def np_pairwise_sum(a):
    """numpy's pairwise float64 reduction order (what the device table kernel re-implements)."""
    a = np.asarray(a, dtype=np.float64)
    n = len(a)
    if n < 8:
        res = np.float64(0.0)
        for v in a:
            res = res + v
        return res
    if n <= 128:
        r = [a[j] for j in range(8)]
        i = 8
        while i < n - (n % 8):
            for j in range(8):
                r[j] = r[j] + a[i + j]
            i += 8
        res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]))
        while i < n:
            res = res + a[i]
            i += 1
        return res
    n2 = n // 2
    n2 -= n2 % 8
    return np_pairwise_sum(a[:n2]) + np_pairwise_sum(a[n2:])


def format_prompt(q, d):
    """dataloader.py:43-44 / 66-67 / 153-155."""
    return 'Query: ' + q + ' Document: ' + d + ' Relevant:'


def pack_prompts(prefix, middle, suffix, q_tokens, d_tokens, q_idx, d_idx, S=None, max_len=0, pad_id=0):
    """Token ids of ``tokenizer([format_prompt(q, d), ...], padding=True)`` (lceT5.py:46-47, pairwrapper.py:79-80)
    assembled from pre-tokenised pieces: prefix = ids of 'Query:', middle = 'Document:', suffix = 'Relevant:' + </s>,
    q_tokens / d_tokens = per-query / per-document id lists (no </s>).  SentencePiece segments whitespace-separated
    words independently, so concatenating the pieces' ids equals tokenising the concatenated string.  max_len > 0
    drops the document's tail so the row still ends in the suffix (the reference never truncates: max_len = 0).
    Returns (ids int32 [n, S] padded with pad_id, lengths int32 [n]); S defaults to the longest row."""
    rows = []
    for qi, di in zip(q_idx, d_idx):
        q, d = list(q_tokens[qi]), list(d_tokens[di])
        fixed = len(prefix) + len(middle) + len(suffix) + len(q)
        if max_len > 0 and fixed + len(d) > max_len:
            d = d[:max(0, max_len - fixed)]
        rows.append(list(prefix) + q + list(middle) + d + list(suffix))
    lengths = np.array([len(r) for r in rows], dtype=np.int32)
    S = int(lengths.max()) if S is None else S
    out = np.full((len(rows), S), pad_id, dtype=np.int32)
    for i, r in enumerate(rows):
        if len(r) <= S:
            out[i, :len(r)] = r
    return out, lengths
