"""basebandboard_b200: B200-native noise / test-pattern datapath of adamgreig/basebandboard.

Hot path only: LUTOPT uniform RNG -> CLT Gaussian RNG, PRBS generate /
error-detect, and rnghunt's candidate evaluation, as hand-written sm_100a CUDA behind a C ABI
(include/bbb_b200.h, built into basebandboard_b200/libbbb_b200.so).  There is no CPU fallback:
importing `basebandboard_b200.lib` without the built library raises.
"""
from . import recurrences  # noqa: F401

__all__ = ["recurrences"]
