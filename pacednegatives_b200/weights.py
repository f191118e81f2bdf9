"""Self-paced weights on the B200 — drop-in for ``pacednegatives.weights`` (weights.py:10-94) and
``pacednegatives.lceT5.LCEWeights`` (lceT5.py:82-99).

Same class names, constructor arguments, attributes (``eta`` registered as ``eta_value``, ``mask``)
and methods (``forward``, ``no_grad``, ``clamp``, ``set_mask``, ``set_weight``, ``__getitem__``).  The
reference's per-element Python loop (one host sync per element) is one fused kernel launch here
(csrc/pn_weights.cu); the weights and the keep mask are bit-identical to the reference's fp32 values.
"""
import math

import torch
import torch.nn as nn

from . import _lib

LCE, ETA = 0, 1


def spl_weights(loss, eta, mask=0.0, variant=LCE, want_keep=False, want_stats=False):
    """One launch of pn_weights_fwd.  Returns (w, keep | None, stats | None); stats is a device
    float64[6] = [sum w, sum w*l, d(sum w*l)/d eta, #(l<eta), sum l, #keep] (no host sync)."""
    _lib.require_gpu()
    loss = loss.detach().contiguous().float()
    eta_t = eta.detach().reshape(-1)[:1].contiguous().float()
    w = torch.empty_like(loss)
    keep = torch.empty(loss.shape, dtype=torch.uint8, device=loss.device) if want_keep else None
    stats = torch.empty(13, dtype=torch.float64, device=loss.device) if want_stats else None   # [6:] = scratch
    _lib.check(_lib.lib().pn_weights_fwd(_lib.ptr(loss), loss.numel(), _lib.ptr(eta_t), float(mask),
                                         variant, _lib.ptr(w), _lib.ptr(keep), _lib.ptr(stats),
                                         _lib.stream()))
    return w, keep, (stats[:6] if stats is not None else None)


class _SplWeightFn(torch.autograd.Function):
    """weights.forward(loss) with gradients to eta (and loss) from pn_weights_bwd."""

    @staticmethod
    def forward(ctx, loss, eta, mask, variant):
        w, _, _ = spl_weights(loss, eta, mask, variant)
        ctx.save_for_backward(loss.detach(), eta.detach())
        ctx.mask, ctx.variant = mask, variant
        return w.view_as(loss)

    @staticmethod
    def backward(ctx, g):
        loss, eta = ctx.saved_tensors
        lossc = loss.contiguous().float()
        g = g.contiguous().float()
        eta_t = eta.reshape(-1)[:1].contiguous().float()
        dloss = torch.empty_like(lossc) if ctx.needs_input_grad[0] else None
        deta = torch.empty(1, dtype=torch.float64, device=loss.device)
        _lib.check(_lib.lib().pn_weights_bwd(_lib.ptr(lossc), lossc.numel(), _lib.ptr(eta_t),
                                             float(ctx.mask), ctx.variant, _lib.ptr(g),
                                             _lib.ptr(dloss), _lib.ptr(deta), _lib.stream()))
        return (dloss.view_as(loss) if dloss is not None else None,
                deta.to(eta.dtype).view_as(eta), None, None)


class _EtaBase(nn.Module):
    _variant = LCE

    def __init__(self, eta: float, device=None):
        super().__init__()
        _lib.require_gpu()
        self.device = device if device else torch.device("cuda")
        self.eta = nn.Parameter(torch.tensor([eta], dtype=torch.float32, device=self.device),
                                requires_grad=True)
        self.register_parameter("eta_value", self.eta)
        self.mask = torch.tensor([0.0], device=self.device)
        self._mask_f = 0.0

    def set_mask(self, mask: float):
        self.mask = torch.tensor([mask], device=self.device)
        self._mask_f = float(mask)

    def no_grad(self, loss, eta):
        if not torch.is_tensor(eta):
            eta = torch.tensor([float(eta)], device=loss.device)
        with torch.no_grad():
            w, _, _ = spl_weights(loss, eta.to(loss.device), self._mask_f, self._variant)
        return w.view_as(loss)

    def forward(self, loss=None):
        return _SplWeightFn.apply(loss, self.eta, self._mask_f, self._variant)

    def fused(self, loss):
        """(w, keep, stats) in one launch, no autograd — what the training step uses."""
        return spl_weights(loss, self.eta, self._mask_f, self._variant, True, True)


class EtaWeights(_EtaBase):
    """weights.py:10-45: w = 1 - l/eta below the threshold."""
    _variant = ETA

    def __init__(self, eta: float, device=None, min=-math.log(0.5) * 0.5, max=10):
        super().__init__(eta, device)
        self._min, self._max = float(min), float(max)
        self.clamp_func = lambda x: torch.clamp(x, min=min, max=max)
        self.weighting = lambda x, y: (-x / y) + 1

    def clamp(self):
        self.eta.data = self.clamp_func(self.eta.data)


class LCEWeights(_EtaBase):
    """weights.py:47-78 / lceT5.py:82-99: w = l/eta below the threshold.  Accepts either the float
    ``eta`` of weights.LCEWeights or the hparams object/dict of lceT5.LCEWeights."""
    _variant = LCE

    def __init__(self, eta, device=None):
        if isinstance(eta, dict):
            eta = eta["eta"]
        elif hasattr(eta, "eta"):
            eta = eta.eta
        super().__init__(float(eta), device)

    def weighting(self, x, y):
        return x / y


class Weights(nn.Module):
    """weights.py:80-94: per-sample table [num_batches, batch]; rows are updated in place by
    pn_weights_table_update (the closed form of NewWrapper.meta_loop, pairwrapper.py:219-239)."""

    def __init__(self, shape, device=None):
        super().__init__()
        _lib.require_gpu()
        self.shape = shape
        self.device = device if device else torch.device("cuda")
        self.weight = nn.Parameter(torch.ones(shape, device=self.device), requires_grad=False)

    def set_weight(self, idx, val):
        self.weight[idx] = torch.sigmoid(val).to(self.device)

    def __getitem__(self, idx):
        return self.weight[idx].detach().clone().requires_grad_(True)

    def forward(self, loss=None, idx=0):
        return self[idx]

    def meta_update(self, idx, ce_n, lr, ce_p=None):
        """W[idx] <- sigmoid(v - lr * (ce_n/len (+ ce_p/len) - 1)) in one launch."""
        row = self.weight.data[idx]
        assert row.is_contiguous()
        ce_n = ce_n.detach().contiguous().float()
        ce_p = ce_p.detach().contiguous().float() if ce_p is not None else None
        _lib.check(_lib.lib().pn_weights_table_update(_lib.ptr(row), _lib.ptr(ce_p), _lib.ptr(ce_n),
                                                      row.numel(), 1.0 / row.numel(), float(lr),
                                                      _lib.stream()))
        return row
