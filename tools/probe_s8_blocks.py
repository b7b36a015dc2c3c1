#!/usr/bin/env python3
"""Experiment: int8 sample stream throughput against the number of concurrently open streams
(clt_spec_max_blocks) and kernel variant.  One JSON line per setting."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from basebandboard_b200 import lib
c = lib.Context(0)
c.lutopt_load_shipped(256)
count = 1 << 32
buf = torch.empty(count, dtype=torch.uint8, device="cuda")
for variant in (2, 1):
    c.set_tuning("clt_spec_variant", variant)
    for blocks in (37, 74, 148, 296, 592):
        c.set_tuning("clt_spec_max_blocks", blocks)
        ms = []
        for _ in range(3):
            c.clt_stream_dev(1, 0, count, 0, 1, buf.data_ptr(), 1)
            c.sync()
            ms.append(c.last_device_ms)
        print(json.dumps({"variant": variant, "blocks": blocks, "ms": [round(m, 2) for m in ms], "kernel": c.last_kernel,
                          "launches": c.last_launches, "Gsamples_s": round(count / min(ms) / 1e6, 1)}), flush=True)
