"""bbb_clt_hist_dev_static: the no-allocation variant of bbb_clt_hist_dev (static __device__ scratch, no
cudaMalloc / cudaFree in the call) against the existing entry point, through the C ABI, two calls in a row.
No counterpart in the reference; integer work, so the bar
is bit-exact.  Collected last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture()
def ctx():
    from basebandboard_b200 import lib
    c = lib.Context(0)
    yield c
    c.close()


def _acc(torch, n):
    return torch.zeros(n + 3, dtype=torch.int64, device="cuda")


@pytest.mark.parametrize("n,model", [(256, 0), (256, 1), (64, 0), (512, 0)])
def test_static_scratch_histogram_equals_the_allocating_entry_point_twice(ctx, n, model):
    import torch
    ctx.lutopt_load_shipped(n)
    # two calls in a row with different ranges: the second finds the first one's planes, operators and bins
    # in the shared array (ragged counts: several rounds with different clocks per stream)
    cases = [(1, 12345, (1 << 21) + 4099), (1, (1 << 40) + 7, (1 << 19) + 33), (1, 0, 1000)]
    want = _acc(torch, n)
    got = _acc(torch, n)
    for init, first, count in cases:
        ctx.clt_hist_dev(init, first, count, model, want.data_ptr())
        ctx.sync()
        ctx.clt_hist_dev_static(init, first, count, model, got.data_ptr())
        ctx.sync()
        assert ctx.last_kernel == "generic_static"
        assert bool(torch.equal(want, got)), (n, model, first, count)
    assert int(want[: n + 1].sum()) == sum(c[2] for c in cases)


def test_static_scratch_against_the_oracle(ctx, orc):
    import torch
    from basebandboard_b200 import recurrences
    ctx.lutopt_load_shipped(256)
    first, count = 999, (1 << 18) + 77
    ref = orc.clt_run(recurrences.packed(256), 1, first, count, 0, 1, ("hist",))
    for _ in range(2):
        got = _acc(torch, 256)
        ctx.clt_hist_dev_static(1, first, count, 0, got.data_ptr())
        ctx.sync()
        g = got.cpu().numpy()
        assert np.array_equal(g[:256].astype(np.uint64), ref["hist"][:256])
        assert int(g[257]) == ref["sum"] and int(g[258]) == ref["sumsq"]


def test_static_scratch_rejects_counts_beyond_its_maximum(ctx):
    import torch
    from basebandboard_b200 import lib
    ctx.lutopt_load_shipped(256)
    got = _acc(torch, 256)
    with pytest.raises(lib.BBBError) as e:
        ctx.clt_hist_dev_static(1, 0, (1 << 32) + 1, 0, got.data_ptr())
    assert "maximum" in str(e.value)
    assert int(got.sum()) == 0


def test_static_scratch_call_allocates_no_device_memory(ctx):
    """free device memory is the same before and after a first call in a fresh context's life"""
    import torch
    ctx.lutopt_load_shipped(256)
    got = _acc(torch, 256)
    # warm-up with the same geometry: the driver loads each kernel's code on its first launch (lazy module
    # loading), which is device memory too; the ctx caches jump matrices on the HOST
    ctx.clt_hist_dev_static(1, 0, (1 << 22) + 5, 0, got.data_ptr())
    ctx.sync()
    free0, _ = torch.cuda.mem_get_info()
    ctx.clt_hist_dev_static(1, 5, (1 << 22) + 5, 0, got.data_ptr())
    ctx.sync()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0
