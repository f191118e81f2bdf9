"""Oracle: self-paced weights (test infrastructure only; see oracle/__init__.py).

Restates, in numpy float32 with the reference's operation order:
  * lceT5.LCEWeights.forward           /root/reference/pacednegatives/lceT5.py:91-99
  * weights.LCEWeights.forward/no_grad /root/reference/pacednegatives/weights.py:59-78
  * weights.EtaWeights.forward/no_grad /root/reference/pacednegatives/weights.py:26-45
  * the NewWrapper per-sample table update, pairwrapper.py:219-239 + weights.py:87-88
"""
import numpy as np

LCE, ETA = 0, 1


def spl_weights(loss, eta, mask=0.0, variant=LCE):
    """Per element: ``if loss[i] > eta: w = loss[i]*mask*eta else: w = weighting(loss[i], eta)``.

    weighting = x/y (LCE, lceT5.py:88-89) or (-x/y)+1 (Eta, weights.py:18).  float32 throughout.
    Returns (w float32[n], keep bool[n]) with keep = not (loss > eta).
    """
    loss = np.asarray(loss, dtype=np.float32)
    eta = np.float32(eta)
    mask = np.float32(mask)
    over = loss > eta
    with np.errstate(divide="ignore", invalid="ignore"):
        q = loss / eta
        keep_w = q if variant == LCE else (-q) + np.float32(1.0)
        over_w = (loss * mask) * eta
    w = np.where(over, over_w, keep_w).astype(np.float32)
    return w, ~over


def spl_stats(loss, eta, mask=0.0, variant=LCE):
    """fp64 reductions the fused kernel emits next to the weights.

    [sum w, sum w*l, d(sum w*l)/d eta, #(l < eta), sum l, #keep].  The eta-derivative is the one
    autograd produces for ``(weights*loss).sum()`` at lceT5.py:137-141 (divide by n for the mean).
    ``#(l < eta)`` is check_success_rate's numerator, old/eta_bound.py:57-58.
    """
    w, keep = spl_weights(loss, eta, mask, variant)
    l = np.asarray(loss, dtype=np.float32).astype(np.float64)
    e = float(np.float32(eta))
    m = float(np.float32(mask))
    t = l * l / (e * e)
    d_keep = -t if variant == LCE else t
    d = np.where(keep, d_keep, m * l * l)
    return np.array([
        w.astype(np.float64).sum(),
        (w.astype(np.float64) * l).sum(),
        d.sum(),
        float((np.asarray(loss, dtype=np.float32) < np.float32(eta)).sum()),
        l.sum(),
        float(keep.sum()),
    ])


def table_update(row, ce_p, ce_n, inv_len, lr):
    """``v <- sigmoid(v - lr * grad)`` with grad = d/dv [sum(ce*v)/len ... - sum(v)].

    pairwrapper.py:219-239: NewWrapper builds ``v`` twice (two different leaf tensors), so the
    gradient it takes only sees the negative term: grad = ce_n/len - 1.  Pass ce_p=None for that;
    passing both gives the intended (ce_p+ce_n)/len - 1 (StdWrapper's objective, pairwrapper.py:128-140).
    """
    row = np.asarray(row, dtype=np.float32)
    acc = np.zeros_like(row)
    if ce_p is not None:
        acc = np.asarray(ce_p, np.float32) * np.float32(inv_len)
    if ce_n is not None:
        t = np.asarray(ce_n, np.float32) * np.float32(inv_len)
        acc = acc + t if ce_p is not None else t
    g = acc + np.float32(-1.0)
    x = (row - np.float32(lr) * g).astype(np.float32)
    return (1.0 / (1.0 + np.exp(-x.astype(np.float64)))).astype(np.float32)
