"""One launch of the tcgen05 GEMM per (shape, epilogue mode, cta_group) for an ncu DRAM-traffic capture:
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --csv python tools/gemm_traffic.py"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pacednegatives_b200 import _lib


def main():
    _lib.require_gpu()
    rows = 256 * 512
    order = []
    for (M, N, K) in [(rows, 768, 3072), (rows, 768, 768), (rows, 3072, 768), (4 * rows, 768, 3072), (rows, 768, 1536),
                      (rows, 1536, 3072)]:
        A = torch.randn(M, K, device="cuda").bfloat16()
        B = torch.randn(N, K, device="cuda").bfloat16()
        for mode in (0, 1):
            out = torch.zeros(M, N, device="cuda", dtype=torch.float32 if mode else torch.bfloat16)
            for cg in (2, 1):
                torch.cuda.synchronize()
                _lib.check(_lib.lib().pn_gemm_bf16(_lib.ptr(A), K, _lib.ptr(B), K, _lib.ptr(out), N, M, N, K, mode, 0,
                                                   None, cg, _lib.stream()))
                torch.cuda.synchronize()
                alg = 2 * M * K + 2 * N * K + M * N * (8 if mode else 2)
                order.append((M, N, K, mode, cg, alg))
    for o in order:
        print("LAUNCH", *o)


if __name__ == "__main__":
    main()
