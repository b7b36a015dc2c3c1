#!/usr/bin/env python3
"""One short launch of every secondary kernel, for `ncu --set full` (kept small: ncu replays each ~40x)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from basebandboard_b200 import lib  # noqa: E402

c = lib.Context(0)
buf = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")             # 1 GiB
c.prbs_stream_dev(31, 1 << 31, buf.data_ptr())                            # prbs_stream_kernel<31,28,2>
c.sync()
c.prbs_detect(31, 1 << 32, m=10, seed=1)                                  # prbs_detect_kernel<31,28> x2
c.rnghunt_sweep(192, 1, 0, 1 << 16, k=16)                                 # hunt_front_kernel<6,u8>, hunt_check_kernel<7>
c.lutopt_load_shipped(256)
c.set_tuning("clt_spec_min_steps", 256)
c.clt_stream_dev(1, 0, 296 * 4096 * 256, 0, 1, buf.data_ptr(), 1)         # clt256_s8v2_kernel (full grid, 256 clocks)
c.sync()
from basebandboard_b200 import shaper  # noqa: E402
c.tx_stream(31, 1 << 25, shaper.tx_sets()[16], dev_ptr=buf.data_ptr())    # shaper_stream_kernel, noise off (HBM-write-bound)
c.sync()
c.prbs_shaped_slicer_detect(31, 1 << 22, shaper.tx_sets()[16], noise_var=15, delay=4)   # shaped_slicer_err_kernel + detector on error words
c.sync()
c.lutopt_stream_dev(1, 0, 148 * 4096 * 32, buf.data_ptr())                # clt256_words7_kernel (2^19 streams, 36 clocks x 32 B)
c.sync()
print("ok", c.last_kernel)
c.close()
