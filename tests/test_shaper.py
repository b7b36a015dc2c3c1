"""pulse shaper and the full TX word.

Reference: gateware/bbb/bitshaper.py:12-91 (PRBSShaper), :93-109 (from_rcf), :112-165 (its test: the
lfilter relation), gateware/bbb/tx.py:52-81 (set list, noise scaling, 12-bit output register).
tests/golden/shaper_rcf.json is produced by executing the reference's own lines
(tests/golden/make_golden_shaper.py)."""
import json
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def gold():
    return json.load(open(os.path.join(ROOT, "tests", "golden", "shaper_rcf.json")))


def _unpack(words, nbits):
    return np.unpackbits(np.ascontiguousarray(words, dtype="<u8").view(np.uint8), bitorder="little")[:nbits].astype(np.int64)


def test_coefficient_sets_equal_the_references(gold):
    from basebandboard_b200 import shaper
    assert shaper.tx_sets() == gold["tx_sets"]                       # tx.py:53-57, 32 raised-cosine sets
    assert shaper.from_rcf([0.5]) == gold["half_plus_rect"]          # + the rectangular set, bitshaper.py:107-108
    assert all(-256 < v <= 255 for s in gold["tx_sets"] for v in s)  # 9-bit signed ROM words, bitshaper.py:19-20
    assert all(s[32] == 254 for s in gold["tx_sets"])


def test_oracle_sequence_is_the_reference_tests_relation(gold, orc):
    """bitshaper.py:150-155: shaped == lfilter(c, 1, +-1 pulses at the middle of each bit) once the
    pipeline is full (the test compares f[60:]); before that the HDL's zeroed shift register reads as
    -1 pulses, which lfilter's zero history does not model."""
    bits = gold["lfilter_prbs9"]["bits"]
    f = np.array(gold["lfilter_prbs9"]["f"])
    S = orc.shaper_stream([0] * 8 + bits, gold["half_plus_rect"][0])
    assert np.array_equal(S[60:], f[60:])
    c = np.array(gold["half_plus_rect"][0])
    assert S[0] == -int(c[4::8].sum())            # t = 0: all eight register bits are 0 -> eight -c terms


def test_tx_word_wraps_at_12_bits(orc):
    x = orc.tx_words([2000, -2000, 100], [127, -128, 0], 15)
    assert x.tolist() == [2000 + 1905 - 4096, -2000 - 1920 + 4096, 100]


@pytest.mark.gpu
@pytest.mark.parametrize("k,sel,first,nv", [(9, 16, 0, 0), (7, 0, 5, 8), (31, 31, 64 * 1000 + 3, 15), (15, 32, 70, 3)])
def test_tx_stream_matches_oracle(gold, orc, k, sel, first, nv):
    from basebandboard_b200 import lib, recurrences
    sets = gold["tx_sets"] + [gold["half_plus_rect"][1]]
    coeff = sets[sel]
    nbits = 64 * 40 + 17
    with lib.Context(0) as c:
        c.lutopt_load_shipped(256)
        c.set_tuning("slicer_chunk_bits", 64 * 16)                  # several passes
        got = c.tx_stream(k, nbits, coeff, first_bit=first, init=1 if nv else None, noise_var=nv)
        got_off = c.tx_stream(k, nbits, coeff, first_bit=first, bit_en=0, init=1, noise_var=max(nv, 1))
    start = max(first - 8, 0)
    words, _ = orc.prbs_bits(k, first - start + nbits, state=orc.lib().orc_prbs_advance(k, 1, start))
    hist = [0] * (8 - (first - start)) + _unpack(words, first - start + nbits).tolist()
    S = orc.shaper_stream(hist, coeff)
    g = orc.clt_run(recurrences.packed(256), 1, first * 8, nbits * 8, model=0, want=("out8",))["out8"]
    assert np.array_equal(got, orc.tx_words(S, g, nv))
    assert np.array_equal(got_off, orc.tx_words(S, g, max(nv, 1), bit_en=0))


@pytest.mark.gpu
@pytest.mark.parametrize("sel,first,nv", [(16, 0, 0), (0, 250, 8), (31, 64 * 1000 + 3, 15)])
def test_tx_stream_pulser_source_matches_oracle(gold, orc, sel, first, nv):
    """tx.py:59-65 with src_sel = 1: the Pulser (tx.py:20-31; one 1 bit every 256 bit periods) through the same
    shaper -- bbb_tx_stream with k = BBB_TX_SRC_PULSE."""
    from basebandboard_b200 import lib, recurrences
    coeff = gold["tx_sets"][sel]
    nbits = 64 * 40 + 17
    with lib.Context(0) as c:
        c.lutopt_load_shipped(256)
        c.set_tuning("slicer_chunk_bits", 64 * 16)
        got = c.tx_stream(lib.TX_SRC_PULSE, nbits, coeff, first_bit=first, init=1 if nv else None, noise_var=nv)
    start = max(first - 8, 0)
    hist = [0] * (8 - (first - start)) + orc.pulser_bits(start, first - start + nbits).tolist()
    S = orc.shaper_stream(hist, coeff)
    g = orc.clt_run(recurrences.packed(256), 1, first * 8, nbits * 8, model=0, want=("out8",))["out8"]
    assert np.array_equal(got, orc.tx_words(S, g, nv))
    if first == 0 and nv == 0:
        # between pulses every register bit is 0: the steady word is minus the sum of the taps at phase t % 8
        cc = np.array(coeff)
        assert got[8 * 20] == -int(cc[4::8].sum())


@pytest.mark.gpu
@pytest.mark.parametrize("k,sel,delay,nv,first", [(15, 10, 4, 15, 0), (7, 31, 2, 15, 64 * 50), (31, 0, 7, 8, 12345), (23, 32, 0, 15, 1)])
def test_shaped_slicer_loop_matches_oracle(gold, orc, k, sel, delay, nv, first):
    """decisions on x[8(J+4)+delay] vs b[J], then PRBSErrorDetector: == oracle composition"""
    from basebandboard_b200 import lib, recurrences
    sets = gold["tx_sets"] + [gold["half_plus_rect"][1]]
    coeff = sets[sel]
    nbits = 64 * 600
    start = max(first - 8, 0)
    total = first - start + nbits + 8
    words, _ = orc.prbs_bits(k, total, state=orc.lib().orc_prbs_advance(k, 1, start))
    b = _unpack(words, total)
    S = orc.shaper_stream([0] * (8 - (first - start)) + b.tolist(), coeff)        # S[8*first ..]
    g = orc.clt_run(recurrences.packed(256), 1, first * 8, (nbits + 8) * 8, model=0, want=("out8",))["out8"]
    x = orc.tx_words(S[: (nbits + 8) * 8], g, nv).astype(np.int64)
    rx = (x[8 * 4 + delay::8][:nbits] >= 0).astype(np.uint8)
    tx = b[first - start: first - start + nbits].astype(np.uint8)
    pack = lambda v: np.packbits(v, bitorder="little").view("<u8").astype(np.uint64)
    want, wdet = orc.det_run_bits(k, pack(rx), nbits, tx_words=pack(tx))
    with lib.Context(0) as c:
        c.lutopt_load_shipped(256)
        for chunk in (0, 64 * 128):
            c.set_tuning("slicer_chunk_bits", chunk)
            det = lib.DetState(0, 1, (1 << k) - 1, 0)
            got = c.prbs_shaped_slicer_detect(k, nbits, coeff, init=1, noise_var=nv, delay=delay, first_bit=first, det=det)
            assert got == want, (k, chunk)
            assert (det.bit_in, det.prbs, det.err_sr, det.reload_ctr) == (wdet.bit_in, wdet.prbs, wdet.err_sr, wdet.reload_ctr)
    if delay == 0:
        assert want["injected"] > 100      # sampling between two bits: the eye is closed
