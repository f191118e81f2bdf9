"""Oracle: the training pass's dropout masks, and their replay inside HuggingFace T5 (test infrastructure only; see
oracle/__init__.py).

The reference trains under Lightning in train() mode, so HF T5 applies nn.Dropout(dropout_rate) in both forwards of
lceT5.py:131-133,149-150 (HF modeling_t5.py: embedding output :734, attention probabilities :332, sub-layer outputs
:149,375,406, FFN hidden :95, final norms :768).  torch's own Philox stream cannot be reproduced by another
implementation, so parity is defined on the MASKS: libpn_b200 draws them from a counter hash (csrc/pn_dropout.cuh),
this file restates that hash in numpy, and ``HFDropoutReplay`` patches ``torch.nn.functional.dropout`` so that HF itself
runs with exactly those masks.  Everything else in the golden step is the reference's unmodified code.
"""
import numpy as np
import torch

M64 = (1 << 64) - 1
GOLD = 0x9E3779B97F4A7C15

EMBED, FINAL = 900, 901
SELF_P, SELF_OUT, CROSS_P, CROSS_OUT, FF_HID, FF_OUT = 0, 1, 2, 3, 4, 5


def site(stack, layer, kind):
    """pn_drop_site: stack (0 encoder, 1 decoder) * 1000 + (kind if kind >= 900 else layer * 10 + kind)."""
    return stack * 1000 + (kind if kind >= 900 else layer * 10 + kind)


def _splitmix64(x):
    x = (x + GOLD) & M64
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & M64
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & M64
    return x ^ (x >> 31)


def _mix32(x):
    x = x & np.uint64(0xFFFFFFFF)
    x = x ^ (x >> np.uint64(16))
    x = (x * np.uint64(0x7FEB352D)) & np.uint64(0xFFFFFFFF)
    x = x ^ (x >> np.uint64(15))
    x = (x * np.uint64(0x846CA68B)) & np.uint64(0xFFFFFFFF)
    return x ^ (x >> np.uint64(16))


def multipliers(seed, site_id, p, idx):
    """float32 multipliers (0 or 1/(1-p)) of the elements with flat indices ``idx`` (any shape, integer)."""
    idx = np.asarray(idx).astype(np.uint64)
    key = _splitmix64((int(seed) & M64) ^ (((site_id + 1) * GOLD) & M64))
    key_lo, key_hi = np.uint64(key & 0xFFFFFFFF), np.uint64(key >> 32)
    pair = idx >> np.uint64(1)
    hh = _mix32((pair >> np.uint64(32)) ^ key_hi)
    h = _mix32((pair & np.uint64(0xFFFFFFFF)) ^ key_lo ^ hh)
    u = np.where((idx & np.uint64(1)) != 0, h >> np.uint64(16), h & np.uint64(0xFFFF))
    thresh = np.uint64(int((1.0 - float(p)) * 65536.0 + 0.5))
    scale = np.float32(1.0) / (np.float32(1.0) - np.float32(p))
    return np.where(u < thresh, scale, np.float32(0.0)).astype(np.float32)


def forward_sites(n_layers, n_dec_layers):
    """(site id, is attention-probability tensor) of every dropout call of one T5ForConditionalGeneration forward,
    in HF's call order."""
    out = [(site(0, 0, EMBED), False)]
    for l in range(n_layers):
        out += [(site(0, l, SELF_P), True), (site(0, l, SELF_OUT), False), (site(0, l, FF_HID), False),
                (site(0, l, FF_OUT), False)]
    out.append((site(0, 0, FINAL), False))
    out.append((site(1, 0, EMBED), False))
    for l in range(n_dec_layers):
        out += [(site(1, l, SELF_P), True), (site(1, l, SELF_OUT), False), (site(1, l, CROSS_P), True),
                (site(1, l, CROSS_OUT), False), (site(1, l, FF_HID), False), (site(1, l, FF_OUT), False)]
    out.append((site(1, 0, FINAL), False))
    return out


class HFDropoutReplay:
    """Context manager: while active, every ``torch.nn.functional.dropout`` call (nn.Dropout included) multiplies
    its input by the counter-hash mask of the matching site.  ``schedule`` is a list of (seed, doc_offset), one entry
    per model forward in the order the forwards run (e.g. the reference's training_step: no-grad p, no-grad n,
    grad p, grad n)."""

    def __init__(self, p, n_layers, n_dec_layers, schedule):
        self.p, self.sites, self.schedule = float(p), forward_sites(n_layers, n_dec_layers), list(schedule)
        self.calls = 0

    def _dropout(self, x, p=0.5, training=True, inplace=False):
        if not training or p == 0.0:
            return x
        assert abs(p - self.p) < 1e-12
        fwd, k = divmod(self.calls, len(self.sites))
        self.calls += 1
        seed, doc_offset = self.schedule[fwd]
        site_id, is_attn = self.sites[k]
        shape = tuple(x.shape)
        per_doc_shape = shape[1:]
        if is_attn:     # [N, H, Tq, Sk]: even row pitch
            skp = shape[3] + (shape[3] & 1)
            per_doc = shape[1] * shape[2] * skp
            n, h, i, j = np.meshgrid(*[np.arange(s, dtype=np.uint64) for s in shape], indexing="ij", sparse=True)
            idx = ((((n + np.uint64(doc_offset)) * np.uint64(shape[1]) + h) * np.uint64(shape[2]) + i)
                   * np.uint64(skp) + j)
            del per_doc
        else:
            per_doc = int(np.prod(per_doc_shape))
            idx = (np.arange(int(np.prod(shape)), dtype=np.uint64) + np.uint64(doc_offset * per_doc)).reshape(shape)
        m = torch.from_numpy(multipliers(seed, site_id, p, idx)).to(x.dtype)
        return x * m

    def __enter__(self):
        self._orig = torch.nn.functional.dropout
        torch.nn.functional.dropout = self._dropout
        return self

    def __exit__(self, *exc):
        torch.nn.functional.dropout = self._orig
        return False
