import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from basebandboard_b200 import lib
ctx = lib.Context(0)
ctx.lutopt_load_shipped(256)
n = 1 << 28
buf = torch.empty(n * 4, dtype=torch.int64, device="cuda")
variant = int(os.environ.get("WORDS_VARIANT", "0"))
ctx.set_tuning("words_variant", variant)
for blocks in [int(x) for x in os.environ.get("BLOCKS", "148,111,74,37,148").split(",")]:
    ctx.set_tuning("clt_spec_max_blocks", blocks)
    for rep in range(2):
        ctx.lutopt_stream_dev(1, 0, n, buf.data_ptr()); ctx.sync()
    print("blocks %3d: %.3f ms = %.1f GB/s [%s, %d launches]" % (blocks, ctx.last_device_ms, n * 32 / ctx.last_device_ms * 1e-6, ctx.last_kernel, ctx.last_launches))
ctx.close()
