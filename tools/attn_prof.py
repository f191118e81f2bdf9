"""One shape of the encoder self-attention kernel, for ncu captures."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pacednegatives_b200 import _lib
_lib.require_gpu()
lib = _lib.lib()
torch.manual_seed(0)
docs, S, H = 250, 512, 12
di = H * 64
qkv = (torch.randn(docs * S, 3 * di, device="cuda") * 0.5).bfloat16()
bias = torch.randn(H, 1023, device="cuda")
out = torch.zeros(docs * S, di, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    _lib.check(lib.pn_t5_self_attention(_lib.ptr(qkv), docs, S, H, _lib.ptr(bias), _lib.ptr(out), _lib.stream()))
torch.cuda.synchronize()
print("ok")
