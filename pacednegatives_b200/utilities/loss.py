"""Import-path mirror of pacednegatives/utilities/loss.py."""
from ..loss import init_LCEcrossentropy, pair_loss, token_ce, lce_restricted  # noqa: F401
