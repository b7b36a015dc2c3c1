"""Times the seeding of a word-stream call (seed kernel alone via a short L) and checks both seeding kernels agree."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from basebandboard_b200 import lib
ctx = lib.Context(0)
ctx.lutopt_load_shipped(256)
n = (1 << 19) * 8                       # 2^19 streams x 8 clocks: the call is almost only seeding
a = torch.empty(n * 4, dtype=torch.int64, device="cuda")
b = torch.empty(n * 4, dtype=torch.int64, device="cuda")
for single in (0, 1, 0, 1):
    ctx.set_tuning("seed_single", single)
    buf = b if single else a
    for rep in range(3):
        ctx.lutopt_stream_dev(1, 12345, n, buf.data_ptr()); ctx.sync()
    print("seed_single %d: call %.3f ms [%s, %d launches]" % (single, ctx.last_device_ms, ctx.last_kernel, ctx.last_launches))
print("equal", bool(torch.equal(a, b)))
n = 1 << 28
buf = torch.empty(n * 4, dtype=torch.int64, device="cuda")
for single in (0, 1, 0, 1):
    ctx.set_tuning("seed_single", single)
    for rep in range(2):
        ctx.lutopt_stream_dev(1, 0, n, buf.data_ptr()); ctx.sync()
    print("seed_single %d: 2^28 steps %.3f ms = %.1f GB/s" % (single, ctx.last_device_ms, n * 32 / ctx.last_device_ms * 1e-6))
ctx.close()
