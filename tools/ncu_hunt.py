#!/usr/bin/env python3
"""One rnghunt sweep of 2^16 candidates at n = 192, for ncu."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from basebandboard_b200 import lib  # noqa: E402
c = lib.Context(0)
r = c.rnghunt_sweep(192, 1, 0, 1 << 16, k=16)
print(r["best"][:4], c.last_device_ms)
c.close()
