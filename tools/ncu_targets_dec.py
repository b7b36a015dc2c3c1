#!/usr/bin/env python3
"""Runs the kernels added with the decimated noise stream once each at benchmark geometry, for ncu:
clt256_dec_kernel (via the slicer loop and the shaped slicer loop), slicer_err_kernel,
shaped_slicer_err_kernel, shaper_stream_kernel.  Prints the CUDA-event time of each call."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402,F401
from basebandboard_b200 import lib, shaper  # noqa: E402

ctx = lib.Context(0)
ctx.lutopt_load_shipped(256)
log2 = int(os.environ.get("LOG2_BITS", "32"))
nbits = 1 << log2
coeff = shaper.tx_sets()[16]
for rep in range(2):
    det = lib.DetState(0, 1, (1 << 31) - 1, 0)
    c1 = ctx.prbs_slicer_detect(31, nbits, init=1, amplitude=256, noise_var=8, spb=8, delay=0, det=det)
    ms1, l1 = ctx.last_device_ms, ctx.last_launches
    det = lib.DetState(0, 1, (1 << 31) - 1, 0)
    c2 = ctx.prbs_shaped_slicer_detect(31, nbits, coeff, init=1, noise_var=8, delay=4, det=det)
    ms2, l2 = ctx.last_device_ms, ctx.last_launches
    print("rep %d: slicer loop 2^%d bits %.3f ms = %.2f Gbit/s (%d launches) %s" % (rep, log2, ms1, nbits / ms1 * 1e-6, l1, c1))
    print("rep %d: shaped slicer loop 2^%d bits %.3f ms = %.2f Gbit/s (%d launches) %s" % (rep, log2, ms2, nbits / ms2 * 1e-6, l2, c2))
buf = torch.empty((1 << 29) * 8, dtype=torch.int16, device="cuda")
ctx.tx_stream(31, 1 << 29, coeff, first_bit=0, dev_ptr=buf.data_ptr())
ctx.sync()
print("tx_stream 2^29 bits %.3f ms = %.1f GB/s" % (ctx.last_device_ms, (1 << 33) / ctx.last_device_ms * 1e-6))
ctx.close()
ctx = lib.Context(0)
ctx.lutopt_load_shipped(256)
wbuf = torch.empty((1 << 26) * 4, dtype=torch.int64, device="cuda")
ctx.lutopt_stream_dev(1, 0, 1 << 26, wbuf.data_ptr())
ctx.sync()
print("lutopt_stream 2^26 steps %.3f ms = %.1f GB/s [%s]" % (ctx.last_device_ms, (1 << 31) / ctx.last_device_ms * 1e-6, ctx.last_kernel))
ctx.close()
