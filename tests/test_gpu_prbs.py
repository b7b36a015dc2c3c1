"""GPU parity: PRBS generator and PRBSErrorDetector through the C ABI vs the CPU oracle (bit-exact).

Reference behaviour: gateware/bbb/prbs.py:14 (TAPS), :23-35 (PRBS; model :112-113), :43-99
(PRBSErrorDetector; test protocol :124-165).  The oracle is pinned to those model lines and to the
reference's detector property by tests/test_oracle_gen.py."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ALL_K = [7, 9, 11, 15, 20, 23, 31]


@pytest.fixture(scope="module")
def ctx():
    from basebandboard_b200 import lib
    c = lib.Context(0)
    yield c
    c.close()


def _bits_to_str(words, nbits):
    return "".join(str((int(words[i >> 6]) >> (i & 63)) & 1) for i in range(nbits))


@pytest.mark.parametrize("k", ALL_K)
def test_stream_golden_prefix(ctx, golden, k):
    """The reference model's own first bits (tests/golden/golden.json, prbs.py:112-113)."""
    want = golden["prbs"][str(k)]
    got = ctx.prbs_stream(k, 64 * ((len(want) + 63) // 64))
    assert _bits_to_str(got, len(want)) == want


@pytest.mark.parametrize("k", ALL_K)
@pytest.mark.parametrize("nbits", [0, 64, 128, 64 * 63, 64 * 64 * 31 + 64, 1 << 20])
def test_stream_matches_oracle(ctx, orc, k, nbits):
    got = ctx.prbs_stream(k, nbits)
    want, _ = orc.prbs_bits(k, nbits)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("k", [7, 15, 23, 31])
def test_stream_offsets_and_states(ctx, orc, k):
    """Arbitrary (not word-aligned) first_bit, non-reset state0, forced multi-warp geometry."""
    base, _ = orc.prbs_bits(k, 1 << 18, state=5)
    for first in (1, 63, 64, 1000, 65537):
        ctx.set_tuning("prbs_stream_reps", 1)
        got = ctx.prbs_stream(k, 1 << 16, state0=5, first_bit=first)
        ctx.set_tuning("prbs_stream_reps", 0)
        want = np.zeros(1 << 10, dtype=np.uint64)
        for i in range(1 << 10):   # re-pack the oracle's bits at the shifted offset
            a = first + 64 * i
            lo = int(base[a >> 6]) >> (a & 63)
            hi = (int(base[(a >> 6) + 1]) << (64 - (a & 63))) if (a & 63) else 0
            want[i] = (lo | hi) & 0xFFFFFFFFFFFFFFFF
        assert np.array_equal(got, want)


def test_stream_unaligned_device_pointer(ctx, orc):
    """An 8-byte aligned output pointer takes the 64-bit-store kernel; same bits."""
    import torch
    buf = torch.zeros(4096 + 1, dtype=torch.int64, device="cuda")
    ctx.prbs_stream_dev(31, 4096 * 64, buf.data_ptr() + 8)
    ctx.sync()
    want, _ = orc.prbs_bits(31, 4096 * 64)
    assert np.array_equal(buf[1:].cpu().numpy().view(np.uint64), want)
    assert int(buf[0]) == 0


def test_stream_into_pinned_and_pageable_host_buffers(ctx, orc):
    """bbb_prbs_stream delivers the same words into a page-locked caller buffer (chunks copied straight into it) and
    into a pageable one (pinned staging + host copy); several chunks, ragged last chunk."""
    import torch
    nbits = 64 * (3 * 8192 + 777)
    want, _ = orc.prbs_bits(23, nbits)
    ctx.set_tuning("stream_chunk_bytes", 1 << 16)
    try:
        pinned = torch.empty(nbits // 8, dtype=torch.uint8, pin_memory=True).numpy().view(np.uint64)
        pinned[:] = 0
        got_p = ctx.prbs_stream(23, nbits, out=pinned)
        got_h = ctx.prbs_stream(23, nbits, out=np.ones(nbits // 64, dtype=np.uint64))
    finally:
        ctx.set_tuning("stream_chunk_bytes", 0)
    assert np.array_equal(got_p, want) and np.array_equal(got_h, want)


def test_stream_large_window_properties(ctx):
    """2^33 bits (1 GiB) of PRBS-31 on the device: every word obeys W[j] = W[j-31] ^ W[j-28]
    and the stream continues exactly where a second call starts."""
    import torch
    nw = 1 << 27
    buf = torch.zeros(nw, dtype=torch.int64, device="cuda")
    ctx.prbs_stream_dev(31, nw * 64, buf.data_ptr())
    ctx.sync()
    assert bool(torch.equal(buf[31:], buf[:-31] ^ buf[3:-28]))
    tail = torch.zeros(1024, dtype=torch.int64, device="cuda")
    ctx.prbs_stream_dev(31, 1024 * 64, tail.data_ptr(), first_bit=(nw - 1024) * 64)
    ctx.sync()
    assert bool(torch.equal(tail, buf[-1024:]))


# ------------------------------------------------------------------------------------------------
# detector
# ------------------------------------------------------------------------------------------------
def _same_det(a, b):
    return (a.bit_in, a.prbs, a.err_sr, a.reload_ctr) == (b.bit_in, b.prbs, b.err_sr, b.reload_ctr)


@pytest.mark.parametrize("k", ALL_K)
@pytest.mark.parametrize("m", [0, 10, 6, 3, 1])
def test_detect_counters_match_oracle(ctx, orc, k, m):
    from basebandboard_b200 import lib
    nbits = 1 << 18
    want, _, wdet = orc.prbs_detect(k, nbits, m=m, seed=1)
    for chunk in (0, 32, 1000):
        ctx.set_tuning("prbs_chunk_words", chunk)
        det = lib.DetState(0, 1, (1 << k) - 1, 0)
        got = ctx.prbs_detect(k, nbits, m=m, seed=1, det=det)
        assert got == want, (k, m, chunk)
        assert _same_det(det, wdet)
    ctx.set_tuning("prbs_chunk_words", 0)


def test_detect_no_errors_is_startup_transient_only(ctx, orc):
    for k in ALL_K:
        got = ctx.prbs_detect(k, 1 << 20, m=0)
        want, _, _ = orc.prbs_detect(k, 1 << 20, m=0)
        assert got == want
        assert got["reload_events"] == 1 and got["injected"] == 0   # prbs.py:80,91-94


@pytest.mark.parametrize("nbits", [1, 63, 65, 1000, 64 * 32 + 1, 100003])
def test_detect_ragged_lengths(ctx, orc, nbits):
    for k in (7, 31):
        got = ctx.prbs_detect(k, nbits, m=4, seed=9)
        want, _, _ = orc.prbs_detect(k, nbits, m=4, seed=9)
        assert got == want


def test_detect_resume_and_offsets(ctx, orc):
    """Splitting a run at arbitrary bit positions and carrying `det` gives the one-shot result."""
    from basebandboard_b200 import lib
    k, m, seed = 15, 5, 3
    want, _, wdet = orc.prbs_detect(k, 300001, m=m, seed=seed)
    det = lib.DetState(0, 1, (1 << k) - 1, 0)
    total = {}
    first = 0
    for n in (1, 777, 64, 100000, 199159):
        got = ctx.prbs_detect(k, n, m=m, seed=seed, first_bit=first, det=det)
        for key, v in got.items():
            total[key] = total.get(key, 0) + v
        first += n
    assert first == 300001 and total == want and _same_det(det, wdet)


def test_detect_speculation_rounds(ctx, orc):
    """Low error rate: the locked-state guess is right everywhere (2 rounds = run + check).  Heavy
    error rate: chunks are re-run until the fixed point; the answer is still the oracle's."""
    ctx.set_tuning("prbs_chunk_words", 32)
    got = ctx.prbs_detect(31, 1 << 20, m=10, seed=1)
    assert ctx.last_rounds == 2
    assert got == orc.prbs_detect(31, 1 << 20, m=10, seed=1)[0]
    got = ctx.prbs_detect(7, 1 << 20, m=2, seed=1)
    assert ctx.last_rounds >= 2
    assert got == orc.prbs_detect(7, 1 << 20, m=2, seed=1)[0]
    ctx.set_tuning("prbs_chunk_words", 0)


def test_spec_state_is_the_locked_state(ctx, orc):
    """bbb_prbs_spec_state == the oracle's detector state after running from reset (no reload pending)."""
    for k in (7, 23, 31):
        for first in (4096, 100000, 123457):
            _, _, det = orc.prbs_detect(k, first, m=10, seed=1)
            if det.reload_ctr == 0:
                assert _same_det(ctx.prbs_spec_state(k, first, m=10, seed=1), det)


def test_detect_reference_property(ctx):
    """prbs.py:155-163: wherever the detector is not reloading, detected errors == injected errors.
    With p = 2^-10 and k = 31 no reload fires after start-up, so err_valid counts every injected
    error that arrived after the start-up reload window (k + k//2 cycles, + the 1-cycle input register)."""
    nbits = 1 << 24
    got = ctx.prbs_detect(31, nbits, m=10, seed=1)
    assert got["reload_events"] == 1
    assert got["reload_cycles"] == 31 + 15
    early = ctx.prbs_detect(31, 31 + 15 + 1, m=10, seed=1)
    assert got["err_valid"] == got["injected"] - early["injected"] + early["err_valid"]


def test_bad_arguments(ctx):
    from basebandboard_b200 import lib
    with pytest.raises(lib.BBBError):
        ctx.prbs_stream(8, 64)
    with pytest.raises(lib.BBBError):
        ctx.prbs_stream(7, 100)
    with pytest.raises(lib.BBBError):
        ctx.prbs_stream(7, 64, state0=1 << 7)
    with pytest.raises(lib.BBBError):
        ctx.prbs_detect(7, 64, m=17)


# ---- caller-supplied error words, and the TX noise -> slicer -> detector loop --------
@pytest.mark.parametrize("k", [7, 15, 31])
def test_detect_with_supplied_error_words(ctx, orc, k):
    import torch
    from basebandboard_b200 import lib
    rng = np.random.default_rng(k)
    nbits = (1 << 18) + 64 * 3
    nw = nbits // 64
    ew = np.zeros(nw, dtype=np.uint64)
    hits = rng.integers(0, nbits, size=400)
    for h in hits:
        ew[h >> 6] |= np.uint64(1 << int(h & 63))
    ew[nw // 2: nw // 2 + 3] = np.uint64(0xFFFFFFFFFFFFFFFF)      # a burst: forces reloads
    first = 12345
    tx, _ = orc.prbs_bits(k, nbits, state=orc.lib().orc_prbs_advance(k, 1, first))
    want, wdet = orc.det_run_bits(k, tx ^ ew, nbits, tx_words=tx)
    d_ew = torch.from_numpy(ew.view(np.int64)).cuda()
    for chunk in (0, 32, 1000):
        ctx.set_tuning("prbs_chunk_words", chunk)
        det = lib.DetState(0, 1, (1 << k) - 1, 0)
        got = ctx.prbs_detect_errs_dev(k, nbits, d_ew.data_ptr(), first_bit=first, det=det)
        assert got == want, (k, chunk)
        assert _same_det(det, wdet)
    ctx.set_tuning("prbs_chunk_words", 0)
    assert want["reload_events"] >= 2 and want["err_valid"] > 100


def test_slicer_loop_sampled_noise_equals_full_stream_large(ctx):
    """Full geometry: the loop over the decimated noise stream (clt256_dec, 1 byte per bit) gives the counters
    and detector state of the loop over the full 8-samples-per-bit stream (tuning knob dec_full_stream)."""
    from basebandboard_b200 import lib
    ctx.lutopt_load_shipped(256)
    k, nbits, first = 31, (1 << 26) + 64 * 777, 64 * 5 + 3
    res = []
    for full in (0, 1):
        ctx.set_tuning("dec_full_stream", full)
        try:
            det = lib.DetState(0, 1, (1 << k) - 1, 0)
            got = ctx.prbs_slicer_detect(k, nbits, init=1, amplitude=256, noise_var=12, spb=8, delay=3, first_bit=first, det=det)
            res.append((got, (det.bit_in, det.prbs, det.err_sr, det.reload_ctr)))
        finally:
            ctx.set_tuning("dec_full_stream", 0)
    assert res[0] == res[1]
    assert res[0][0]["injected"] > 100000


@pytest.mark.parametrize("k,amp,nv,spb,delay", [(7, 300, 8, 8, 0), (15, 200, 15, 8, 5), (31, 900, 15, 4, 3),
                                                (23, 64, 1, 1, 0), (9, 2047, 15, 2, 1)])
def test_slicer_loop_matches_oracle(ctx, orc, k, amp, nv, spb, delay):
    """PRBS -> (+-A + CLT-GRNG*noise_var, 12-bit) -> slicer -> PRBSErrorDetector (tx.py:70-81, rx.py:28-46)
    == oracle composition, including chunked passes with detector hand-off."""
    from basebandboard_b200 import lib, recurrences
    ctx.lutopt_load_shipped(256)
    first, nbits = 64 * 1000 + 17, 64 * 3000
    g = orc.clt_run(recurrences.packed(256), 1, first * spb, nbits * spb, model=0, want=("out8",))["out8"]
    tx, _ = orc.prbs_bits(k, nbits, state=orc.lib().orc_prbs_advance(k, 1, first))
    rx = orc.slicer_rx_bits(tx, g, nbits, amp, nv, spb, delay)
    want, wdet = orc.det_run_bits(k, rx, nbits, tx_words=tx)
    for chunk_bits in (0, 64 * 700):
        ctx.set_tuning("slicer_chunk_bits", chunk_bits)
        det = lib.DetState(0, 1, (1 << k) - 1, 0)
        got = ctx.prbs_slicer_detect(k, nbits, init=1, amplitude=amp, noise_var=nv, spb=spb, delay=delay,
                                     first_bit=first, det=det)
        assert got == want, (k, chunk_bits)
        assert _same_det(det, wdet)
    ctx.set_tuning("slicer_chunk_bits", 0)
    if (k, amp) == (15, 200):
        assert want["injected"] > 1000   # sigma = 8 * 15 = 120: a 1.7 sigma slicer margin
    with pytest.raises(lib.BBBError):
        ctx.prbs_slicer_detect(k, 100, spb=8)            # not a multiple of 64
    with pytest.raises(lib.BBBError):
        ctx.prbs_slicer_detect(k, 128, spb=3)
