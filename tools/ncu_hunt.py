import os, sys
sys.path.insert(0, "/root/repo")
from basebandboard_b200 import lib
c = lib.Context(0)
r = c.rnghunt_sweep(192, 1, 0, 1 << 20, k=16)
print(c.last_kernel, c.last_device_ms, r["best"][:3])
c.close()
