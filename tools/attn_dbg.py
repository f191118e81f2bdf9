"""Dump the clock64 timeline of CTA 0 (softmax warp 4 and the stream-A MMA issuer) of the attention kernel."""
import os, sys, ctypes, torch
os.environ["PN_ATTN_DBG"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pacednegatives_b200 import _lib
_lib.require_gpu()
lib = _lib.lib()
docs, S, H = 250, 512, 12
di = H * 64
qkv = (torch.randn(docs * S, 3 * di, device="cuda") * 0.5).bfloat16()
bias = torch.randn(H, 1023, device="cuda")
out = torch.zeros(docs * S, di, device="cuda", dtype=torch.bfloat16)
for _ in range(2):
    _lib.check(lib.pn_t5_self_attention(_lib.ptr(qkv), docs, S, H, _lib.ptr(bias), _lib.ptr(out), _lib.stream()))
buf = (ctypes.c_longlong * (2 * 64 * 8))()
lib.pn_attn_debug_read.argtypes = [ctypes.c_void_p]
print("rc", lib.pn_attn_debug_read(buf))
t0 = min(v for v in buf if v > 0)
names0 = ["top", "sfull", "sfree", "pfree_w", "pfree_ok", "exp_done", "pready", "ofull"]
names1 = ["sfree_seen", "qk_issued", "pready_seen", "pv_issued"]
for n in range(24):
    sm = [buf[(0 * 64 + n) * 8 + e] for e in range(8)]
    ms = [buf[(1 * 64 + n) * 8 + e] for e in range(4)]
    print(n, "softmax:", " ".join(f"{names0[e]}={sm[e]-t0 if sm[e] else -1}" for e in range(8)))
    print(n, "  issuer:", " ".join(f"{names1[e]}={ms[e]-t0 if ms[e] else -1}" for e in range(4)))
