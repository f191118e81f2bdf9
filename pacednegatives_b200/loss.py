"""LCE loss on the B200 — drop-in for ``pacednegatives.utilities.loss.init_LCEcrossentropy``
(utilities/loss.py:3-21) and ``LCEModel.pair_loss`` (lceT5.py:116-123).

Logits in, per-group loss out, same shapes and ordering as the reference (including the [Q*T] layout
quirk of SURVEY §0.4).  The full-vocabulary log-softmax is one streaming pass per row
(pn_token_ce_fwd) instead of materialised log-probabilities; autograd is wired to pn_token_ce_bwd /
pn_group_loss_bwd.  ``lce_restricted`` is the two-row-head LCE used by the fast training mode.
"""
import torch

from . import _lib


def _dtype_code(t):
    if t.dtype == torch.float32:
        return 0
    if t.dtype == torch.bfloat16:
        return 1
    raise ValueError(f"logits must be float32 or bfloat16, got {t.dtype}")


class _TokenCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target, ignore_index):
        _lib.require_gpu()
        V = logits.shape[-1]
        x = logits.detach().reshape(-1, V).contiguous()
        t = target.reshape(-1).contiguous().long()
        rows = x.shape[0]
        ce = torch.empty(rows, dtype=torch.float32, device=x.device)
        lse = torch.empty(rows, dtype=torch.float32, device=x.device)
        _lib.check(_lib.lib().pn_token_ce_fwd(_lib.ptr(x), _dtype_code(x), rows, V, _lib.ptr(t),
                                              int(ignore_index), _lib.ptr(ce), _lib.ptr(lse),
                                              _lib.stream()))
        ctx.save_for_backward(x, t, lse)
        ctx.ignore_index, ctx.shape, ctx.dtype = int(ignore_index), logits.shape, logits.dtype
        return ce

    @staticmethod
    def backward(ctx, g):
        x, t, lse = ctx.saved_tensors
        rows, V = x.shape
        d = torch.empty(rows, V, dtype=torch.float32, device=x.device)
        g = g.contiguous().float()
        _lib.check(_lib.lib().pn_token_ce_bwd(_lib.ptr(x), _dtype_code(x), rows, V, _lib.ptr(t),
                                              ctx.ignore_index, _lib.ptr(lse), _lib.ptr(g),
                                              _lib.ptr(d), _lib.stream()))
        return d.view(ctx.shape).to(ctx.dtype), None, None


def token_ce(logits, target, ignore_index=-100):
    """nn.CrossEntropyLoss(ignore_index, reduction='none')(logits.view(-1,V), target.view(-1))."""
    return _TokenCE.apply(logits, target, ignore_index)


class _GroupReduce(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pce, nce, k, use_mean):
        pce = pce.contiguous().float()
        nce = nce.contiguous().float()
        m = pce.numel()
        if nce.numel() != m * k:
            raise RuntimeError(f"shape mismatch: {m} positive terms need {m * k} negative terms, "
                               f"got {nce.numel()}")
        out = torch.empty(m, dtype=torch.float32, device=pce.device)
        _lib.check(_lib.lib().pn_group_loss_fwd(_lib.ptr(pce), _lib.ptr(nce), m, k, int(use_mean),
                                                _lib.ptr(out), None, _lib.stream()))
        ctx.k, ctx.use_mean, ctx.m = k, int(use_mean), m
        return out

    @staticmethod
    def backward(ctx, g):
        g = g.contiguous().float()
        dp = torch.empty(ctx.m, dtype=torch.float32, device=g.device)
        dn = torch.empty(ctx.m * ctx.k, dtype=torch.float32, device=g.device)
        _lib.check(_lib.lib().pn_group_loss_bwd(_lib.ptr(g), ctx.m, ctx.k, ctx.use_mean,
                                                _lib.ptr(dp), _lib.ptr(dn), _lib.stream()))
        return dp, dn, None, None


def pair_loss(p, n, op, on, use_mean=True, ignore_index=-100):
    """LCEModel.pair_loss(p, n, op, on): ``n`` is the caller's ``[-1, k, V]`` view (lceT5.py:134)."""
    pce = token_ce(p, op, ignore_index)
    nce = token_ce(n, on, ignore_index)
    return _GroupReduce.apply(pce, nce, n.size(-2), use_mean)


def init_LCEcrossentropy(ignore_index=-100, use_mean=True):
    def LCEcrossentropy(plogits, nlogits, op, on, weights=None, backward=True):
        ce = pair_loss(plogits, nlogits, op, on, use_mean, ignore_index)
        if weights is not None:
            v = weights.forward(ce) if backward else weights.no_grad(ce, weights.eta)
            return torch.mean(ce * v)
        return torch.mean(ce)
    return LCEcrossentropy


class _LceRestricted(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z2, G):
        _lib.require_gpu()
        z = z2.detach().reshape(-1, 2).contiguous().float()
        Q = z.shape[0] // G
        loss = torch.empty(Q, dtype=torch.float32, device=z.device)
        score = torch.empty(Q * G, dtype=torch.float32, device=z.device)
        _lib.check(_lib.lib().pn_lce_restricted_fwd(_lib.ptr(z), Q, G, _lib.ptr(loss),
                                                    _lib.ptr(score), _lib.stream()))
        ctx.save_for_backward(z)
        ctx.G, ctx.shape = G, z2.shape
        ctx.mark_non_differentiable(score)
        return loss, score

    @staticmethod
    def backward(ctx, g, _gs):
        (z,) = ctx.saved_tensors
        Q = z.shape[0] // ctx.G
        dz = torch.empty_like(z)
        g = g.contiguous().float()
        _lib.check(_lib.lib().pn_lce_restricted_bwd(_lib.ptr(z), Q, ctx.G, _lib.ptr(g), _lib.ptr(dz),
                                                    _lib.stream()))
        return dz.view(ctx.shape), None


def lce_restricted(z2, G):
    """z2 [Q*G, 2] = (z_true, z_false) per document, positive first in each group.
    Returns (loss [Q], score [Q*G])."""
    return _LceRestricted.apply(z2, G)


def check_success(pos, neg, rel_id, nrel_id):
    """old/levels.py:35-39 / old/eta_bound.py:51-55: mean(softmax(pos[:, 0, (REL, NREL)])[:, 0] >
    softmax(neg[:, 0, (REL, NREL)])[:, 0]) over paired rows, as a device scalar (one kernel, four logits per pair
    read).  pos / neg: fp32 logits [n, T, V] (any V >= the token ids)."""
    _lib.require_gpu()
    pos, neg = pos.detach(), neg.detach()
    if pos.dim() != 3 or neg.dim() != 3 or pos.shape[0] != neg.shape[0]:
        raise ValueError(f"check_success: logits must be [n, T, V] with equal n, got {tuple(pos.shape)} / {tuple(neg.shape)}")
    pos = pos if pos.dtype == torch.float32 and pos.stride(2) == 1 else pos.float().contiguous()
    neg = neg if neg.dtype == torch.float32 and neg.stride(2) == 1 else neg.float().contiguous()
    n = pos.shape[0]
    count = torch.empty(1, dtype=torch.int32, device="cuda")
    import ctypes as C
    if not (pos.is_cuda and neg.is_cuda):
        raise ValueError("expected CUDA tensors")
    # rows may be views into vocab-padded logits: only element stride 1 within a row is required
    _lib.check(_lib.lib().pn_check_success(C.c_void_p(pos.data_ptr()), C.c_void_p(neg.data_ptr()), n,
                                           pos.stride(0), neg.stride(0),
                                           int(rel_id), int(nrel_id), _lib.ptr(count), _lib.stream()))
    return count.float() / max(n, 1)
