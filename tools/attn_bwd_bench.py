"""Timing of the attention backward kernels alone (PN_ATTN_BWD_ONLY=0 -> dK/dV/dbias pass, 1 -> dQ pass)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pacednegatives_b200 import _lib
_lib.require_gpu()
lib = _lib.lib()
torch.manual_seed(0)
docs, S, H = int(os.environ.get("DOCS", 256)), 512, 12
iters = int(os.environ.get("ITERS", 5))
di, rows = H * 64, docs * S
qkv = (torch.randn(rows, 3 * di, device="cuda") * 0.5).bfloat16()
bias = torch.randn(H, 1023, device="cuda")
DC = int(os.environ.get("DC", 91))
if DC:
    bias[:, 511 + DC:] = bias[:, 1022:1023]
    bias[:, :511 - DC + 1] = bias[:, 0:1]
dctx = (torch.randn(rows, di, device="cuda") * 0.5).bfloat16()
ctx = torch.zeros(rows, di, device="cuda", dtype=torch.bfloat16)
lse = torch.zeros(rows, H, device="cuda")
_lib.check(lib.pn_t5_self_attention_lse(_lib.ptr(qkv), docs, S, H, _lib.ptr(bias), _lib.ptr(ctx), _lib.ptr(lse),
                                        _lib.stream()))
dqkv = torch.zeros(rows, 3 * di, device="cuda", dtype=torch.bfloat16)
diag = torch.zeros(H, 1023, device="cuda")
dvec = torch.zeros(rows, H, device="cuda")


def call():
    _lib.check(lib.pn_t5_self_attention_bwd(_lib.ptr(qkv), docs, S, H, _lib.ptr(bias), _lib.ptr(ctx), _lib.ptr(dctx),
                                            _lib.ptr(lse), None, _lib.ptr(dqkv), _lib.ptr(diag), _lib.ptr(dvec),
                                            _lib.stream()))


call()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record()
for _ in range(iters):
    call()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
only = os.environ.get("PN_ATTN_BWD_ONLY", "both")
nmm = {"0": 8.0, "1": 6.0}.get(only, 14.0)
print(f"only={only} docs={docs} ms={ms:.3f} executed TFLOP/s={nmm * docs * H * S * S * 64 / ms / 1e9:.1f}", flush=True)
