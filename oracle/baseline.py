"""CPU loops in the reference's own shape, timed by bench.py's ``cpu_baseline`` / ``--impl reference`` legs of the
selection workload (test / measurement infrastructure only; see oracle/__init__.py).  Unlike the vectorised
restatements in oracle/weights.py and oracle/select.py (which exist to CHECK results), these keep the reference's
per-element Python control flow, because that is what its CPU path costs:

  * lce_weights_loop   lceT5.LCEWeights.forward            /root/reference/pacednegatives/lceT5.py:91-99
  * sample_loop        LCEDataset.sample (+ .get indexing)  /root/reference/pacednegatives/dataloader.py:128-151
  * get_items_loop     TripletDataset.get_items             /root/reference/pacednegatives/dataloader.py:29-33
"""
from math import ceil, floor

import numpy as np
import torch


def lce_weights_loop(loss, eta):
    """``for i in range(len(loss)): if loss[i] > eta: ... else: weighting(loss[i], eta)`` on torch tensors, eta a
    parameter that requires grad (every element builds autograd nodes, as in the reference)."""
    weight = torch.zeros(loss.size()).to(loss.device)
    for i in range(len(loss)):
        if loss[i] > eta:
            weight[i] = loss[i] * torch.zeros(1, requires_grad=True).to(loss.device) * eta
        else:
            weight[i] = loss[i] / eta
    return weight


def sample_loop(neg_idx, k, mean, var, calls, lo=1e-10, hi=1.0 - 1e-10):
    """`calls` times LCEDataset.sample(mean) followed by the candidate gather of LCEDataset.get: the pmf is
    recomputed per item (twice, as the reference does) although `mean` is shared."""
    from scipy.stats import binom
    C = len(neg_idx[0])
    out = []
    for q in range(calls):
        m = np.clip(mean, lo, hi)
        idx = np.arange(C)
        p = binom.pmf(idx, C - 1, m)
        p = binom.pmf(idx, C - 1, m)
        adj = p / p.sum()
        adj *= np.sqrt(var / np.var(adj))
        adj /= adj.sum()
        s = np.random.choice(idx, size=(k,), replace=False, p=adj)
        out.append(neg_idx[q % len(neg_idx)][s].tolist())
    return out


def get_items_loop(neg_idx, weights, use_ceil=False):
    """TripletDataset.get_items for item j with difficulty weights[j]: r = round(w * n_neg), one list lookup."""
    rnd = ceil if use_ceil else floor
    n_neg = len(neg_idx[0]) - 1
    out = []
    for j, w in enumerate(weights):
        out.append(neg_idx[j % len(neg_idx)][rnd(w * n_neg)])
    return out
