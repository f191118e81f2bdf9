"""pacednegatives_b200 — the PacedNegatives LCE / self-paced monoT5 hot path on NVIDIA B200 (sm_100a).

Drop-in for the reference's hot-path call surface (lceT5.py, pairwrapper.py, weights.py, dataloader.py,
utilities/loss.py): same class and function names, hand-written CUDA underneath (libpn_b200.so, C-ABI in
include/pn_b200.h).  No CPU fallback: importing is cheap, but every operation raises unless the library
is built and a B200 is present.
"""
__version__ = "0.1.0"
