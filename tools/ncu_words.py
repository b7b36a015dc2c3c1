import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from basebandboard_b200 import lib
ctx = lib.Context(0)
ctx.lutopt_load_shipped(256)
ctx.set_tuning("words_variant", int(os.environ.get("WORDS_VARIANT", "0")))
n = 1 << 27
buf = torch.empty(n * 4, dtype=torch.int64, device="cuda")
for _ in range(2):
    ctx.lutopt_stream_dev(1, 0, n, buf.data_ptr()); ctx.sync()
    print("lutopt_stream 2^27 steps %.3f ms = %.1f GB/s [%s]" % (ctx.last_device_ms, n * 32 / ctx.last_device_ms * 1e-6, ctx.last_kernel))
ctx.close()
