#!/usr/bin/env python3
"""Quick on-box probe: LOP3 issue peak and CLT fused-histogram throughput (not the bench)."""
import json, sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from basebandboard_b200 import lib

c = lib.Context(0)
out = {"lop3_peak_thread_instr_per_s": [c.measure_lop3_peak() for _ in range(3)]}
for n, logc in ((256, 24), (256, 30), (256, 33)):
    c.lutopt_load_shipped(n)
    for rep in range(2):
        t0 = time.time()
        h, s, ss = c.clt_hist(1, 0, 1 << logc, 0)
        wall = time.time() - t0
        ms = c.last_device_ms
        out["clt_n%d_2^%d_rep%d" % (n, logc, rep)] = {
            "device_ms": ms, "wall_s": wall, "gsamples_per_s": (1 << logc) / ms / 1e6,
            "kernel": c.last_kernel, "launches": c.last_launches, "sum": s, "sumsq": ss}
print(json.dumps(out, indent=1))
