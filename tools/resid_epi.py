"""Isolated timing of the fp32 residual-epilogue GEMM (the narrow output projections) with CUDA events:
PN_GEMM_TMA_RES=0|1 python tools/resid_epi.py   (the knob is read once per process)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pacednegatives_b200 import _lib


def main():
    _lib.require_gpu()
    for (M, N, K) in [(131072, 768, 768), (524288, 768, 768), (131072, 1024, 1024), (131072, 768, 3072)]:
        A = torch.randn(M, K, device="cuda").bfloat16()
        B = (torch.randn(N, K, device="cuda") * 0.02).bfloat16()
        out = torch.zeros(M, N, device="cuda")
        call = lambda: _lib.check(_lib.lib().pn_gemm_bf16(_lib.ptr(A), K, _lib.ptr(B), K, _lib.ptr(out), N, M, N, K, 1, 0,
                                                          None, 2, _lib.stream()))
        for _ in range(3):
            call()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        torch.cuda.synchronize()
        ev[0].record()
        for _ in range(10):
            call()
        ev[1].record()
        torch.cuda.synchronize()
        us = ev[0].elapsed_time(ev[1]) * 100.0
        alg = 2 * M * K + 2 * N * K + M * N * 8
        print(f"PN_GEMM_TMA_RES={os.environ.get('PN_GEMM_TMA_RES', '(default)')} M={M} N={N} K={K}: {us:8.1f} us  "
              f"{alg / us / 1e3:7.0f} GB/s  {2.0 * M * N * K / us / 1e6:7.0f} TFLOP/s")


if __name__ == "__main__":
    main()
