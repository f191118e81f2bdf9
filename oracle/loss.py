"""Oracle: LCE loss (test infrastructure only; see oracle/__init__.py).

Restates
  * nn.CrossEntropyLoss(ignore_index, reduction='none') as used at lceT5.py:110,117-118
  * LCEModel.pair_loss                /root/reference/pacednegatives/lceT5.py:116-123
  * init_LCEcrossentropy(...)         /root/reference/pacednegatives/utilities/loss.py:3-21
  * the re-rank score                 /root/reference/pacednegatives/distill/wrapper.py:34-38
  * check_success                     /root/reference/old/levels.py:35-39
"""
import numpy as np


def token_ce(logits, target, ignore_index=-100):
    """ce[r] = logsumexp(logits[r]) - logits[r, target[r]]; 0 where target == ignore_index.
    Returns (ce float32[rows], lse float64[rows])."""
    x = np.asarray(logits, dtype=np.float64)
    t = np.asarray(target, dtype=np.int64).reshape(-1)
    x = x.reshape(-1, x.shape[-1])
    m = x.max(axis=1)
    lse = m + np.log(np.exp(x - m[:, None]).sum(axis=1))
    live = t != ignore_index
    picked = x[np.arange(len(t)), np.where(live, t, 0)]
    ce = np.where(live, lse - picked, 0.0)
    return ce.astype(np.float32), lse


def token_ce_grad(logits, target, g, ignore_index=-100):
    """d ce[r] * g[r] / d logits = g[r] * (softmax - onehot)."""
    x = np.asarray(logits, dtype=np.float64)
    x = x.reshape(-1, x.shape[-1])
    t = np.asarray(target, dtype=np.int64).reshape(-1)
    p = np.exp(x - x.max(axis=1, keepdims=True))
    p /= p.sum(axis=1, keepdims=True)
    live = t != ignore_index
    p[np.arange(len(t))[live], t[live]] -= 1.0
    p *= (np.asarray(g, dtype=np.float64) * live)[:, None]
    return p.astype(np.float32)


def group_reduce(pce, nce, k, use_mean):
    """lceT5.py:119-121: nce.view(-1, k) -> mean|sum(dim=1); ce = pce + nce."""
    pce = np.asarray(pce, dtype=np.float32)
    nce = np.asarray(nce, dtype=np.float32).reshape(-1, k)
    red = nce.mean(axis=1, dtype=np.float32) if use_mean else nce.sum(axis=1, dtype=np.float32)
    return (pce + red).astype(np.float32)


def pair_loss(plogits, nlogits, op, on, use_mean, ignore_index=-100):
    """LCEModel.pair_loss.  ``nlogits`` is the caller's re-viewed tensor ``[-1, n, V]`` (lceT5.py:134),
    so ``n.size(-2)`` is the group size; the result has Q*T entries."""
    nlogits = np.asarray(nlogits)
    k = nlogits.shape[-2]
    pce, _ = token_ce(plogits, op, ignore_index)
    nce, _ = token_ce(nlogits, on, ignore_index)
    return group_reduce(pce, nce, k, use_mean)


def lce_crossentropy(plogits, nlogits, op, on, use_mean=True, ignore_index=-100, weights=None):
    """utilities/loss.py:5-20.  ``weights`` = callable ce -> v (or None)."""
    ce = pair_loss(plogits, nlogits, op, on, use_mean, ignore_index)
    if weights is not None:
        v = weights(ce)
        return np.float32(np.mean(ce * v, dtype=np.float32))
    return np.float32(np.mean(ce, dtype=np.float32))


def rel_score(z_true, z_false):
    """distill/wrapper.py:37-38: log_softmax(logits[:, 0, (rel, nrel)], dim=1)[:, 0]."""
    zt = np.asarray(z_true, dtype=np.float64)
    zf = np.asarray(z_false, dtype=np.float64)
    return (zt - np.logaddexp(zt, zf)).astype(np.float32)


def check_success(z2_pos, z2_neg):
    """old/levels.py:35-39: mean(softmax(pos)[:,0] > softmax(neg)[:,0])."""
    pp = np.exp(rel_score(z2_pos[:, 0], z2_pos[:, 1]).astype(np.float64))
    pn = np.exp(rel_score(z2_neg[:, 0], z2_neg[:, 1]).astype(np.float64))
    return float(np.mean(pp > pn))


def lce_restricted(z2, G):
    """Canonical LCE over the two-row head: loss[q] = logsumexp_j s[q,j] - s[q,0]."""
    z2 = np.asarray(z2, dtype=np.float64).reshape(-1, G, 2)
    s = z2[..., 0] - np.logaddexp(z2[..., 0], z2[..., 1])
    m = s.max(axis=1)
    lse = m + np.log(np.exp(s - m[:, None]).sum(axis=1))
    return (lse - s[:, 0]).astype(np.float32), s.astype(np.float32).reshape(-1)


def lce_restricted_grad(z2, G, g):
    z2 = np.asarray(z2, dtype=np.float64).reshape(-1, G, 2)
    s = z2[..., 0] - np.logaddexp(z2[..., 0], z2[..., 1])
    p = np.exp(s - s.max(axis=1, keepdims=True))
    p /= p.sum(axis=1, keepdims=True)
    p[:, 0] -= 1.0
    ds = p * np.asarray(g, dtype=np.float64)[:, None]
    pf = 1.0 - np.exp(s)  # ds/dz_true = sigmoid(zf - zt)
    out = np.stack([ds * pf, -ds * pf], axis=-1)
    return out.astype(np.float32).reshape(-1, 2)
