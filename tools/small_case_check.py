#!/usr/bin/env python3
"""Small invocations of the kernels written or rewritten in round 2, at small sizes (compute-sanitizer is closed on this pool, so this is the small-case check against the oracle):
hunt_front_pow_kernel + the sync-free back end, hunt_front_warp_kernel, clt256_words2_kernel (pipelined drain),
the shaper kernels, the chunked host-buffer delivery.  Results are compared with the oracle as usual."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from basebandboard_b200 import lib, recurrences, shaper  # noqa: E402
from oracle import oracle  # noqa: E402

c = lib.Context(0)
for n, cnt in ((32, 600), (100, 300), (192, 200), (256, 100)):
    r = c.rnghunt_sweep(n, 1, 0, cnt, k=4, want_stages=True)
    ref = oracle.rnghunt_sweep(n, 1, 0, cnt, threads=8)
    assert np.array_equal(r["stages"], ref[0]), n
c.set_tuning("hunt_smem_taps", 2)
r = c.rnghunt_sweep(192, 1, 0, 100, k=4, want_stages=True)
assert np.array_equal(r["stages"], oracle.rnghunt_sweep(192, 1, 0, 100, threads=8)[0])
c.set_tuning("hunt_smem_taps", 0)
r = c.rnghunt_sweep(64, 1, 0, 2000, k=8, want_polys=True)
print("rnghunt ok", r["best"][:3], r["weights"][:3])
c.lutopt_load_shipped(256)
c.set_tuning("clt_spec_min_steps", 8)
c.set_tuning("clt_spec_max_blocks", 1)
got = c.lutopt_stream(1, 0, 4096 * 9 + 5)
ref = oracle.clt_run(recurrences.packed(256), 1, 0, 4096 * 9 + 5, model=1, want=("words",))["words"]
assert np.array_equal(got.reshape(-1), ref)
g = c.clt_stream(1, 0, 4096 * 40 + 100)
assert np.array_equal(g, oracle.clt_run(recurrences.packed(256), 1, 0, g.size, model=0, want=("out8",))["out8"])
c.set_tuning("clt_spec_min_steps", 512)
c.set_tuning("clt_spec_max_blocks", 0)
print("streams ok")
coeff = shaper.tx_sets()[16]
x = c.tx_stream(9, 64 * 20, coeff, init=1, noise_var=8)
xp = c.tx_stream(lib.TX_SRC_PULSE, 64 * 20, coeff)
ctr = c.prbs_shaped_slicer_detect(15, 64 * 64, coeff, init=1, noise_var=15, delay=4)
c.set_tuning("stream_chunk_bytes", 1 << 16)
b = c.prbs_stream(31, 64 * 4096 * 3)
assert np.array_equal(b, oracle.prbs_bits(31, 64 * 4096 * 3)[0])
print("tx ok", x[:4], xp[:4], ctr)
c.close()
