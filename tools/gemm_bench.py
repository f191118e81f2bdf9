"""Times the tcgen05 GEMM on the T5-base shapes (CUDA events, L2 flushed between iterations)."""
import json
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pacednegatives_b200 import _lib


def main():
    _lib.require_gpu()
    rows = 256 * 512
    shapes = [(rows, 2304, 768, 0), (rows, 768, 768, 1), (rows, 3072, 768, 0), (rows, 768, 3072, 1),
              (rows, 1536, 768, 0)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for (M, N, K, mode) in shapes:
        A = torch.randn(M, K, device="cuda").bfloat16()
        B = torch.randn(N, K, device="cuda").bfloat16()
        out = torch.zeros(M, N, device="cuda", dtype=torch.float32 if mode else torch.bfloat16)
        for cg in (1, 2):
            ts = []
            for it in range(6):
                flush.zero_()
                e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
                e0.record()
                _lib.check(_lib.lib().pn_gemm_bf16(_lib.ptr(A), K, _lib.ptr(B), K, _lib.ptr(out), N, M, N,
                                                   K, mode, 0, None, cg, _lib.stream()))
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            t = min(ts[2:])
            print(json.dumps({"M": M, "N": N, "K": K, "mode": mode, "cta_group": cg, "ms": round(t, 4),
                              "tflops": round(2 * M * N * K / t / 1e9, 1)}))
        # cuBLAS reference point
        ts = []
        for it in range(6):
            flush.zero_()
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record()
            torch.matmul(A, B.t())
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        print(json.dumps({"M": M, "N": N, "K": K, "impl": "cublas", "ms": round(min(ts[2:]), 4),
                          "tflops": round(2 * M * N * K / min(ts[2:]) / 1e9, 1)}))


if __name__ == "__main__":
    main()
