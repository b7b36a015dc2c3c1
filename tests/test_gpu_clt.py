"""GPU parity: LUTOPT -> CLTGRNG through the C ABI vs the CPU oracle (bit-exact).

Reference behaviour: gateware/bbb/rng.py:21-40 (LUTOPT), :70-108 (CLTGRNG), tx.py:75-76 (noise
scaling), clt-grng-evaluate.py:18-21 (statistics).  The oracle is pinned to the reference's numpy
models by tests/test_oracle_gen.py."""
import numpy as np
import pytest

from basebandboard_b200 import lib, recurrences

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from basebandboard_b200 import lib
    c = lib.Context(0)
    yield c
    c.close()


def _oracle_hist(orc, n, init, first, count, model):
    r = orc.clt_run(recurrences.packed(n), init, first, count, model=model, threads=8 if count > 50000 else 1)
    return r["hist"], r["sum"], r["sumsq"]


@pytest.mark.parametrize("n", [16, 32, 64, 128, 256, 512])
@pytest.mark.parametrize("count", [1, 31, 33, 64, 1000, 5000])
def test_hist_small_counts_all_sizes(ctx, orc, n, count):
    ctx.lutopt_load_shipped(n)
    for model in (0, 1):
        h, s, ss = ctx.clt_hist(1, 0, count, model)
        oh, os_, oss = _oracle_hist(orc, n, 1, 0, count, model)
        assert np.array_equal(h, oh)
        assert (s, ss) == (os_, oss)
        assert int(h.sum()) == count


def test_hist_golden_first_samples(ctx, golden):
    """First samples of the reference numpy model (tests/golden/golden.json)."""
    for key in ("16", "32", "64", "128", "256"):
        n = int(key)
        g = golden["lutopt_clt"][key]
        want = np.array(g["clt_numpy"])
        ctx.lutopt_load_shipped(n)
        got = ctx.clt_stream(g["init"], 0, len(want), model=1, noise_var=1)
        assert np.array_equal(got.astype(np.int64), want)


@pytest.mark.parametrize("n,count", [(256, 1 << 20), (32, 1 << 20), (16, 300000)])
def test_hist_medium(ctx, orc, n, count):
    """n=16/32 reach the +n/2 wrap corner (SURVEY 7 'hard parts') at these counts."""
    ctx.lutopt_load_shipped(n)
    for model in (0, 1):
        h, s, ss = ctx.clt_hist(1, 0, count, model)
        oh, os_, oss = _oracle_hist(orc, n, 1, 0, count, model)
        assert np.array_equal(h, oh)
        assert (s, ss) == (os_, oss)
    if n == 16:
        assert oh[16] > 0          # the corner really occurred (numpy-model bin n)


def test_hist_offset_and_init(ctx, orc):
    ctx.lutopt_load_shipped(256)
    init = 0x123456789ABCDEF0FEDCBA9876543210 | (1 << 255)
    first = (1 << 40) + 12345
    h, s, ss = ctx.clt_hist(init, first, 70000, 0)
    oh, os_, oss = _oracle_hist(orc, 256, init, first, 70000, 0)
    assert np.array_equal(h, oh) and (s, ss) == (os_, oss)


def test_hist_shard_consistency(ctx):
    """hist(0..N) == sum of hist over shards, for shard sizes that change the stream geometry."""
    ctx.lutopt_load_shipped(256)
    N = 3_000_017
    whole, s, ss = ctx.clt_hist(1, 0, N, 0)
    for cuts in ([0, 1_000_000, 1_000_001, 2_500_000, N], [0, 17, 4096, N]):
        acc = np.zeros_like(whole)
        ts, tss = 0, 0
        for a, b in zip(cuts[:-1], cuts[1:]):
            h, s1, s2 = ctx.clt_hist(1, a, b - a, 0)
            acc += h
            ts += s1
            tss += s2
        assert np.array_equal(acc, whole) and (ts, tss) == (s, ss)


@pytest.mark.parametrize("n", [16, 64, 256])
def test_sample_stream_int8_int16(ctx, orc, n):
    ctx.lutopt_load_shipped(n)
    count = 20011
    ref = orc.clt_run(recurrences.packed(n), 1, 777, count, model=0, noise_var=8, want=("out8", "out16"))
    got8 = ctx.clt_stream(1, 777, count, model=0)
    got16 = ctx.clt_stream(1, 777, count, model=0, noise_var=8)
    assert np.array_equal(got8, ref["out8"])
    assert np.array_equal(got16, ref["out16"])          # tx.py:75-76


@pytest.mark.parametrize("n", [16, 32, 192, 256, 512])
def test_lutopt_word_stream(ctx, orc, n, golden):
    ctx.lutopt_load_shipped(n)
    count = 3000
    ref = orc.clt_run(recurrences.packed(n), 1, 0, count, model=1, want=("words",))["words"]
    got = ctx.lutopt_stream(1, 0, count)
    assert np.array_equal(got.reshape(-1), ref)
    if str(n) in golden["lutopt_clt"]:
        from basebandboard_b200.lib import state_int
        want = [int(h, 16) for h in golden["lutopt_clt"][str(n)]["states_hex"]]
        assert [state_int(r) for r in got[:len(want)]] == want


def test_custom_matrix_uses_generic_kernel(ctx, orc):
    """A recurrence that is not one of the shipped ones goes through the data-driven kernel."""
    rng = np.random.default_rng(5)
    n = 64
    packed = [sorted(rng.choice(n, size=rng.choice([3, 4]), replace=False).tolist()) for _ in range(n)]
    ctx.lutopt_load(packed)
    h, s, ss = ctx.clt_hist(0xDEADBEEFCAFEF00D, 5, 40000, 0)
    r = orc.clt_run(packed, 0xDEADBEEFCAFEF00D, 5, 40000, model=0)
    assert np.array_equal(h, r["hist"]) and (s, ss) == (r["sum"], r["sumsq"])
    assert ctx.last_kernel == "generic"


def test_jump_matches_oracle(ctx, orc):
    ctx.lutopt_load_shipped(256)
    for e in (0, 1, 2, 1000, (1 << 40) + 3, (1 << 64) - 1):
        assert ctx.lutopt_jump(0xABCDEF0123456789 << 100 | 1, e) == \
            orc.lutopt_jump(recurrences.packed(256), 0xABCDEF0123456789 << 100 | 1, e)


def test_stats_match_numpy(ctx):
    """clt-grng-evaluate.py:18-21: np.mean / np.var of the samples == stats from the histogram."""
    from basebandboard_b200 import lib
    ctx.lutopt_load_shipped(256)
    n = 200000
    samples = ctx.clt_stream(1, 0, n, model=0).astype(np.float64)
    h, _, _ = ctx.clt_hist(1, 0, n, 0)
    st = lib.clt_stats(256, h, 0)
    assert st["mean"] == pytest.approx(np.mean(samples), rel=1e-9, abs=1e-12)
    assert st["variance"] == pytest.approx(np.var(samples), rel=1e-9)
    m = samples - samples.mean()
    assert st["kurtosis"] == pytest.approx(np.mean(m ** 4) / np.var(samples) ** 2 - 3.0, rel=1e-9, abs=1e-12)
    assert st["variance_theory"] == 64.0


# ---- the specialised n256 kernel (clt_spec.cu) -------------------------------------------------
FUSED_NAME = {1: "clt256_fused", 2: "clt256_fused2"}


@pytest.mark.parametrize("variant", [1, 2])
@pytest.mark.parametrize("model", [0, 1])
@pytest.mark.parametrize("blocks,min_steps,count", [(1, 40, 4096 * 57 + 123), (3, 33, 3 * 4096 * 70 + 4099),
                                                   (0, 16, 5_000_011)])
def test_spec256_matches_oracle(ctx, orc, model, blocks, min_steps, count, variant):
    """clt256_fused / clt256_fused2 (straight-line bit-sliced kernels, two-pass and one-pass) + generic
    tail == oracle, both models; the count is large enough that popcounts outside [96,160) occur
    (the atomic slow path)."""
    ctx.lutopt_load_shipped(256)
    ctx.set_tuning("clt_spec_variant", variant)
    ctx.set_tuning("clt_spec_min_steps", min_steps)
    ctx.set_tuning("clt_spec_max_blocks", blocks)
    try:
        init = 1 if blocks != 3 else (0x5DEECE66D << 190) | 0xB
        first = 0 if blocks != 3 else (1 << 41) + 9
        h, s, ss = ctx.clt_hist(init, first, count, model)
        assert ctx.last_kernel == FUSED_NAME[variant] or count < 4096 * min_steps
        oh, os_, oss = _oracle_hist(orc, 256, init, first, count, model)
        assert np.array_equal(h, oh)
        assert (s, ss) == (os_, oss)
        if count > 2_000_000:
            assert int(oh[:96].sum() + oh[160:].sum()) > 0
    finally:
        ctx.set_tuning("clt_spec_min_steps", 512)
        ctx.set_tuning("clt_spec_max_blocks", 0)
        ctx.set_tuning("clt_spec_variant", 2)


def test_spec256_equals_generic_large(ctx):
    """2^31 samples: both specialised kernels == data-driven kernel (bit-exact histograms and moments)."""
    ctx.lutopt_load_shipped(256)
    a = ctx.clt_hist(1, 12345, 1 << 31, 0)
    assert ctx.last_kernel == "clt256_fused2"
    ctx.set_tuning("clt_spec_variant", 1)
    try:
        a1 = ctx.clt_hist(1, 12345, 1 << 31, 0)
        assert ctx.last_kernel == "clt256_fused"
    finally:
        ctx.set_tuning("clt_spec_variant", 2)
    ctx.set_tuning("clt_no_spec", 1)
    try:
        b = ctx.clt_hist(1, 12345, 1 << 31, 0)
        assert ctx.last_kernel == "generic"
    finally:
        ctx.set_tuning("clt_no_spec", 0)
    assert np.array_equal(a[0], b[0]) and a[1:] == b[1:]
    assert np.array_equal(a1[0], b[0]) and a1[1:] == b[1:]
    assert int(a[0].sum()) == 1 << 31


# ---- the specialised n256 stream kernels (clt256_s8, clt256_words) ---------------------------------
@pytest.mark.parametrize("variant", [1, 2])
@pytest.mark.parametrize("blocks,min_steps,count,first", [(1, 8, 4096 * 24 + 77, 0), (2, 16, 2 * 4096 * 40 + 13, (1 << 33) + 5),
                                                         (1, 8, 4096 * 75 + 31, 64), (3, 8, 3 * 4096 * 130 + 5, (1 << 40) + 32)])
def test_spec256_int8_stream_matches_oracle(ctx, orc, blocks, min_steps, count, first, variant):
    """Straight-line kernels (two-pass clt256_s8, one-pass clt256_s8v2) with on-chip bit->byte transpose
    and 64-bit stores + generic tail == oracle."""
    ctx.lutopt_load_shipped(256)
    ctx.set_tuning("clt_spec_variant", variant)
    ctx.set_tuning("clt_spec_min_steps", min_steps)
    ctx.set_tuning("clt_spec_max_blocks", blocks)
    try:
        got = ctx.clt_stream(1, first, count, model=0)
        assert ctx.last_launches >= 2          # a specialised launch happened (then the generic tail)
        ref = orc.clt_run(recurrences.packed(256), 1, first, count, model=0, want=("out8",))
        assert np.array_equal(got, ref["out8"])
    finally:
        ctx.set_tuning("clt_spec_min_steps", 512)
        ctx.set_tuning("clt_spec_max_blocks", 0)
        ctx.set_tuning("clt_spec_variant", 2)


# ---- decimated sample stream (clt256_dec: the CLT tree on every spb-th clock only) ------------------
@pytest.mark.parametrize("spb,delay,blocks,count,first", [(8, 0, 1, 4096 * 40 + 77, 0), (8, 5, 2, 2 * 4096 * 70 + 13, (1 << 33) + 5),
                                                          (4, 3, 1, 4096 * 33, 64), (2, 1, 3, 3 * 4096 * 100 + 5, (1 << 40) + 32),
                                                          (16, 15, 1, 4096 * 32 + 1, 7), (3, 2, 1, 4096 * 64 + 9, 1),
                                                          (8, 7, 1, 1000, 12345), (1, 0, 1, 4096 * 40 + 3, 99)])
def test_decimated_stream_matches_oracle(ctx, orc, spb, delay, blocks, count, first):
    """out[j] = g[first + j*spb + delay] (the sample of a bit period the receiver keeps, rx.py:18-19):
    clt256_dec_kernel + ragged tail == every spb-th sample of the oracle's full stream; and the
    full-stream-then-pick path (tuning knob) gives the same bytes."""
    ctx.lutopt_load_shipped(256)
    ctx.set_tuning("clt_spec_max_blocks", blocks)
    try:
        got = ctx.clt_stream_dec(1, first, count, spb, delay)
        if spb > 1 and count >= 4096 * 32:
            assert ctx.last_kernel in ("clt256_dec", "generic", "clt256_s8v2") and ctx.last_launches >= 2
        full = orc.clt_run(recurrences.packed(256), 1, first, count * spb, model=0, want=("out8",))["out8"]
        assert np.array_equal(got, full[delay::spb])
        ctx.set_tuning("dec_full_stream", 1)
        assert np.array_equal(ctx.clt_stream_dec(1, first, count, spb, delay), got)
    finally:
        ctx.set_tuning("dec_full_stream", 0)
        ctx.set_tuning("clt_spec_max_blocks", 0)
    with pytest.raises(lib.BBBError):
        ctx.clt_stream_dec(1, 0, 64, 8, 8)               # delay must be < spb
    with pytest.raises(lib.BBBError):
        ctx.clt_stream_dec(1, 0, 64, 0, 0)


def test_decimated_stream_other_matrices(ctx, orc):
    """A recurrence without a straight-line kernel takes the full stream + pick path."""
    for n in (32, 64):
        ctx.lutopt_load_shipped(n)
        got = ctx.clt_stream_dec(1, 1000, 5000, 8, 3)
        full = orc.clt_run(recurrences.packed(n), 1, 1000, 5000 * 8, model=0, want=("out8",))["out8"]
        assert np.array_equal(got, full[3::8])
    ctx.lutopt_load_shipped(256)


def test_decimated_stream_equals_full_stream_large(ctx):
    """Full geometry (296 CTAs, several launches + tail): clt256_dec == every 8th byte of clt256_s8v2."""
    import torch
    ctx.lutopt_load_shipped(256)
    nd = 296 * 4096 * 64 + 4096 * 35 + 11
    a = torch.empty(nd, dtype=torch.int8, device="cuda")
    b = torch.empty(nd * 8, dtype=torch.int8, device="cuda")
    ctx.clt_stream_dec_dev(1, 1 << 36, nd, 8, 6, a.data_ptr())
    ctx.sync()
    assert ctx.last_launches >= 4
    ctx.clt_stream_dev(1, 1 << 36, nd * 8, 0, 1, b.data_ptr(), 1)
    ctx.sync()
    assert bool(torch.equal(a, b[6::8]))


@pytest.mark.parametrize("variant", [1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16])
@pytest.mark.parametrize("blocks,min_steps,count,first", [(1, 8, 6144 * 9 + 5, 0), (2, 8, 2 * 6144 * 11 + 3, (1 << 35) + 1)])
def test_spec256_word_stream_matches_oracle(ctx, orc, blocks, min_steps, count, first, variant):
    """raw URNG words: two-pass clt256_words (1), one-pass clt256_words2 (2), the 192-thread experiment
    clt256_words3 (3), the two-CTA clt256_words4 (4), clt256_words5 (5: L2 slab) the producer / consumer clt256_words6 (6) and the default clt256_words7 (7: the stage in tensor memory; 8: without the
    line-completing evict_first stores; 9: clt256_words8, un-slicing on the drain warps;
    10: words7 with the transposes split between compute and drain warps; 11: words7 with the right shifts of the
    transposes as multiply-high on the FMA pipe; 12: the eight transposes straight-line; 13: clt256_words9, three staged clocks drained in pairs;
    14: words7 with the byte-granular transpose stages on the drain warps (the default), 15: and the last sub-byte stage;
    16: clt256_words10, two drain warps per compute warp with setmaxnreg), each + the data-driven tail, == oracle"""
    ctx.lutopt_load_shipped(256)
    ctx.set_tuning("clt_spec_variant", 1 if variant == 1 else 2)
    ctx.set_tuning("words_b192", 1 if variant == 3 else 0)
    ctx.set_tuning("words_variant", variant if variant >= 4 else 2)
    ctx.set_tuning("clt_spec_min_steps", min_steps)
    ctx.set_tuning("clt_spec_max_blocks", blocks)
    try:
        got = ctx.lutopt_stream(1, first, count)
        assert ctx.last_kernel in ("generic", ("clt256_words", "clt256_words2", "clt256_words3", "clt256_words4", "clt256_words5", "clt256_words6", "clt256_words7", "clt256_words7", "clt256_words8", "clt256_words7", "clt256_words7", "clt256_words7", "clt256_words9", "clt256_words7", "clt256_words7", "clt256_words10")[variant - 1])
        ref = orc.clt_run(recurrences.packed(256), 1, first, count, model=1, want=("words",))["words"]
        assert np.array_equal(got.reshape(-1), ref)
    finally:
        ctx.set_tuning("clt_spec_min_steps", 512)
        ctx.set_tuning("clt_spec_max_blocks", 0)
        ctx.set_tuning("clt_spec_variant", 2)
        ctx.set_tuning("words_b192", 0)
        ctx.set_tuning("words_variant", 0)


def test_spec256_streams_equal_generic_large(ctx):
    """Full-geometry launches (148 CTAs): specialised stream kernels == data-driven kernel, and the
    int8 stream's histogram == the fused kernel's histogram over the same range."""
    import torch
    ctx.lutopt_load_shipped(256)
    count = 148 * 4096 * 512 + 4096 * 3 + 7
    a = torch.empty(count, dtype=torch.int8, device="cuda")
    b = torch.empty(count, dtype=torch.int8, device="cuda")
    ctx.clt_stream_dev(1, 99, count, 0, 1, a.data_ptr(), 1)
    ctx.sync()
    assert ctx.last_kernel in ("clt256_s8v2", "generic") and ctx.last_launches >= 2
    ctx.set_tuning("clt_spec_variant", 1)
    try:
        ctx.clt_stream_dev(1, 99, count, 0, 1, b.data_ptr(), 1)
        ctx.sync()
        assert ctx.last_kernel in ("clt256_s8", "generic")
    finally:
        ctx.set_tuning("clt_spec_variant", 2)
    assert bool(torch.equal(a, b))
    b.zero_()
    ctx.set_tuning("clt_no_spec", 1)
    try:
        ctx.clt_stream_dev(1, 99, count, 0, 1, b.data_ptr(), 1)
        ctx.sync()
    finally:
        ctx.set_tuning("clt_no_spec", 0)
    assert bool(torch.equal(a, b))
    h, s, ss = ctx.clt_hist(1, 99, count, 0)
    hist = torch.bincount(a.to(torch.int64) + 128, minlength=256).cpu().numpy().astype(np.uint64)
    assert np.array_equal(hist, h) and int(a.to(torch.int64).sum()) == s
    del a, b
    wc = 148 * 4096 * 512 // 8 + 333
    wa = torch.empty(wc * 4, dtype=torch.int64, device="cuda")
    wb = torch.empty(wc * 4, dtype=torch.int64, device="cuda")
    ctx.set_tuning("clt_spec_min_steps", 64)
    try:
        ctx.lutopt_stream_dev(1, 5, wc, wa.data_ptr())
        ctx.sync()
        # the product path: 112 CTAs of clt256_words7 (tensor-memory stage), 84 clocks, a ragged tail of 262 477 steps on
        # the third stream
        assert ctx.last_kernel == "clt256_words7" and ctx.last_launches == 4
        ctx.set_tuning("clt_no_spec", 1)
        ctx.lutopt_stream_dev(1, 5, wc, wb.data_ptr())
        ctx.sync()
    finally:
        ctx.set_tuning("clt_no_spec", 0)
        ctx.set_tuning("clt_spec_min_steps", 512)
    assert bool(torch.equal(wa, wb))


@pytest.mark.parametrize("n", [128, 256])
def test_four_group_seeding_equals_single_group_seeding(ctx, n):
    """seed4_kernel (four stream groups per warp, used from 8 groups per SM up) == seed_kernel, through the word
    stream of launches big enough to take it (data-driven kernel for n = 128, both kernels for n = 256)."""
    import torch
    ctx.lutopt_load_shipped(n)
    count = (1 << 23) + 4099
    words = n // 64
    a = torch.empty(count * words, dtype=torch.int64, device="cuda")
    b = torch.empty(count * words, dtype=torch.int64, device="cuda")
    try:
        for no_spec in ((0, 1) if n == 256 else (0,)):
            ctx.set_tuning("clt_no_spec", no_spec)
            ctx.set_tuning("seed_single", 0)
            ctx.lutopt_stream_dev(1, (1 << 37) + 3, count, a.data_ptr())
            ctx.sync()
            ctx.set_tuning("seed_single", 1)
            ctx.lutopt_stream_dev(1, (1 << 37) + 3, count, b.data_ptr())
            ctx.sync()
            assert bool(torch.equal(a, b)), (n, no_spec)
    finally:
        ctx.set_tuning("seed_single", 0)
        ctx.set_tuning("clt_no_spec", 0)
        ctx.lutopt_load_shipped(256)


@pytest.mark.parametrize("variant", [1, 2])
@pytest.mark.parametrize("noise_var", [0, 1, 8, 15])
def test_spec256_int16_noise_stream_matches_oracle(ctx, orc, noise_var, variant):
    """tx.py:75-76: noise = g * noise_var as int16, straight-line kernels + generic tail == oracle."""
    ctx.lutopt_load_shipped(256)
    ctx.set_tuning("clt_spec_variant", variant)
    ctx.set_tuning("clt_spec_min_steps", 16)
    ctx.set_tuning("clt_spec_max_blocks", 2)
    try:
        count, first = 2 * 4096 * 72 + 21, 1234567
        got = ctx.clt_stream(1, first, count, model=0, noise_var=noise_var)
        assert ctx.last_kernel in ("clt256_s16" if variant == 1 else "clt256_s16v2", "generic")
        ref = orc.clt_run(recurrences.packed(256), 1, first, count, model=0, noise_var=noise_var, want=("out16",))
        assert np.array_equal(got, ref["out16"])
    finally:
        ctx.set_tuning("clt_spec_min_steps", 512)
        ctx.set_tuning("clt_spec_max_blocks", 0)
        ctx.set_tuning("clt_spec_variant", 2)
