"""Pins the generator oracle (oracle/gen.c) against golden vectors produced by executing the
reference's own numpy/Python model lines (tests/golden/make_golden.py):
LUTOPT rng.py:134-135, CLT tree rng.py:173-181 and clt-grng-evaluate.py:10-15, PRBS prbs.py:112-113;
and checks the PRBSErrorDetector restatement (prbs.py:61-99) with the reference's own test
protocol (prbs.py:124-165)."""
import numpy as np
import pytest

from basebandboard_b200 import recurrences
from oracle import oracle as O

SURVEY_GOLD = {   # SURVEY.md section 7.1c, transcribed independently of golden.json
    16: ([0x2124, 0xbf2a, 0x6474, 0x9ddf, 0x15d1, 0xcc16], [-2, 2, -1, 0, 1, -3]),
    32: ([0x24080, 0x40105234, 0xe64686f3, 0x88264e3a, 0xe39ec875, 0x1ae774dd], [-1, 2, 3, 1, 4, -1]),
}


@pytest.mark.parametrize("key", ["16", "32", "64", "128", "256", "16_init_ace1"])
def test_lutopt_and_clt_match_reference_model(golden, key):
    g = golden["lutopt_clt"][key]
    n = int(key.split("_")[0])
    packed = recurrences.packed(n)
    x = g["init"]
    for want_hex, want_clt in zip(g["states_hex"], g["clt_numpy"]):
        x = O.lutopt_step(packed, x)
        assert x == int(want_hex, 16)
        assert O.clt(n, x, O.MODEL_NUMPY) == want_clt
        gw = O.clt(n, x, O.MODEL_GATEWARE)
        assert gw == (want_clt if want_clt != n // 2 else -n // 2)


def test_survey_transcribed_values(golden):
    for n, (states, clts) in SURVEY_GOLD.items():
        g = golden["lutopt_clt"][str(n)]
        assert [int(h, 16) for h in g["states_hex"][:6]] == states
        assert g["clt_numpy"][:6] == clts
    assert golden["lutopt_clt"]["256"]["clt_numpy"][:6] == [-4, -6, 0, 1, 2, -4]
    assert int(golden["lutopt_clt"]["256"]["states_hex"][0], 16) == \
        0x20000000000040000000000000000000800000000000000000000000010


def test_numpy_restatement_agrees_with_c():
    rng = np.random.default_rng(7)
    for n in (16, 32, 64, 256, 512):
        packed = recurrences.packed(n)
        init = int.from_bytes(rng.bytes(n // 8), "little") | 1
        states = O.np_lutopt_states(packed, 12, init)
        x = init
        for s in states:
            x = O.lutopt_step(packed, x)
            assert x == s
            assert O.clt(n, x, O.MODEL_NUMPY) == O.np_clt(n, x)


def test_clt_closed_form_and_wrap_corner():
    """g = popcount(x XOR M) - n/2 with M = odd-parity indices (SURVEY 7 'hard parts'); the one
    value +n/2 wraps to -n/2 at gateware widths (rng.py:78,108) and stays +n/2 in the numpy model."""
    rng = np.random.default_rng(11)
    for n in (16, 32, 256):
        M = sum(1 << i for i in range(n) if bin(i).count("1") & 1)
        for _ in range(50):
            x = int.from_bytes(rng.bytes(n // 8), "little")
            assert O.clt(n, x, O.MODEL_NUMPY) == bin(x ^ M).count("1") - n // 2
        top = ((1 << n) - 1) ^ M                 # every +1 input set, every -1 input clear
        assert O.clt(n, top, O.MODEL_NUMPY) == n // 2 == O.np_clt(n, top)
        assert O.clt(n, top, O.MODEL_GATEWARE) == -n // 2
        assert O.clt(n, M, O.MODEL_GATEWARE) == -n // 2 == O.clt(n, M, O.MODEL_NUMPY)


def test_clt_run_consistency():
    packed = recurrences.packed(256)
    a = O.clt_run(packed, 1, 0, 5000, want=("hist", "out8", "out16", "words"), noise_var=8)
    assert a["hist"].sum() == 5000
    vals = a["out8"].astype(np.int64)
    assert a["sum"] == vals.sum() and a["sumsq"] == (vals * vals).sum()
    assert np.array_equal(np.bincount(vals + 128, minlength=256).astype(np.uint64), a["hist"])
    assert np.array_equal(a["out16"], (vals * 8).astype(np.int16))          # tx.py:75-76
    assert O.state_int(a["words"][:4]) == 0x20000000000040000000000000000000800000000000000000000000010
    # a window reached by jump-ahead equals the same window of the sequential run
    b = O.clt_run(packed, 1, 3000, 2000, want=("out8",))
    assert np.array_equal(b["out8"], a["out8"][3000:])
    # multi-threaded split gives the same histogram
    c = O.clt_run(packed, 1, 0, 5000, threads=3)
    assert np.array_equal(c["hist"], a["hist"]) and c["sum"] == a["sum"] and c["sumsq"] == a["sumsq"]


def test_jump_matches_stepping():
    for n in (16, 64, 192, 512):
        packed = recurrences.packed(n)
        x = 1
        for t in range(1, 70):
            x = O.lutopt_step(packed, x)
            if t in (1, 2, 31, 64, 69):
                assert O.lutopt_jump(packed, 1, t) == x


@pytest.mark.parametrize("k", [7, 9, 11, 15, 20, 23, 31])
def test_prbs_matches_reference_model(golden, k):
    want = golden["prbs"][str(k)]
    words, _ = O.prbs_bits(k, len(want))
    got = "".join(str((int(words[i // 64]) >> (i % 64)) & 1) for i in range(len(want)))
    assert got == want
    assert got == "".join(str(b) for b in O.py_prbs_bits(k, len(want)))


def test_prbs_survey_prefixes(golden):
    assert golden["prbs"]["7"][:40] == "0000011000010100011110010001011001110101"
    assert golden["prbs"]["15"][:40] == "0000000000000110000000000001010000000000"
    assert golden["prbs"]["23"][:40] == "0000000000000000010000100000000000010000"
    assert golden["prbs"]["31"][:40] == "0000000000000000000000000001001000000000"


def _detector_protocol(k, tx_errors):
    """prbs.py:137-163 with the oracle's cycle model in place of the migen simulation."""
    L = O.lib()
    import ctypes as C
    det = O.DetState()
    L.orc_det_reset(k, C.byref(det))
    bits = O.py_prbs_bits(k, len(tx_errors))
    rx, valid = [], []
    # migen testbench writes (`yield source.eq(..)`, prbs.py:147) commit AT the next clock edge,
    # like a register: the detector samples value i at edge i+1 (hence the reference's [1:] lag).
    source = 0
    for b, e in zip(bits, tx_errors):
        o = L.orc_det_clock(k, C.byref(det), source)
        source = b ^ e
        rx.append(o & 1)
        valid.append(1 - ((o >> 1) & 1))
    valid = np.array(valid, dtype=bool)[:-1]
    return np.array(tx_errors)[:-1][valid].tolist(), np.array(rx)[1:][valid].tolist(), valid


@pytest.mark.parametrize("k", [7, 9, 11, 15, 20, 23, 31])
def test_prbs_error_detector_reference_protocol(k):
    """The reference's property: outside reload windows detected == injected (one-cycle lag).
    Error positions are seeded and kept >= 2k apart so that no error falls inside a reload window
    (the reference test's own unseeded draw is flaky in exactly that case)."""
    nbits = min((1 << k) - 1, 512)
    rng = np.random.default_rng(k)
    tx = np.zeros(nbits, dtype=int)
    pos = 4 * k + 3                                             # k=7 re-fires once at start-up
    while pos < nbits:
        tx[pos] = 1
        pos += int(rng.integers(2 * k + 2, 4 * k))
    tx[:2 * k] = 0                                              # prbs.py:133
    tx[nbits // 2:nbits // 2 + 3 * k] = 1                       # prbs.py:136 burst
    tx[nbits // 2 + 3 * k:nbits // 2 + 8 * k] = 0               # (reference: 5k) re-sync guard
    want, got, valid = _detector_protocol(k, tx.tolist())
    assert want == got
    assert valid[4 * k:nbits // 2].all()                        # locked before the burst
    assert not valid[nbits // 2 + k:nbits // 2 + 2 * k].all()   # the burst forces a reload


def test_detector_startup_transient_and_counters():
    for k in (7, 15, 23, 31):
        ctr, gs, det = O.prbs_detect(k, 4096, m=0)
        # first edge always fires a reload (err_sr resets full, prbs.py:80,91-94)
        assert ctr["reload_events"] >= 1 and ctr["injected"] == 0
        assert ctr["err_valid"] == 0                  # error-free channel: nothing once locked
        assert det.reload_ctr == 0
        # chunked run == single run (state hand-off)
        c1, g1, d1 = O.prbs_detect(k, 1000, m=0)
        c2, g2, d2 = O.prbs_detect(k, 3096, m=0, first_bit=1000, gen_state=g1, det=d1)
        assert {x: c1[x] + c2[x] for x in c1} == ctr and g2 == gs
        assert (d2.prbs, d2.err_sr, d2.reload_ctr, d2.bit_in) == (det.prbs, det.err_sr, det.reload_ctr, det.bit_in)


def test_error_word_rate():
    L = O.lib()
    ones = sum(bin(L.orc_err_word(1, w, 4)).count("1") for w in range(4096))
    assert abs(ones / (4096 * 64) - 2 ** -4) < 0.003
    assert L.orc_err_word(1, 5, 0) == 0xFFFFFFFFFFFFFFFF
    ctr, _, _ = O.prbs_detect(15, 1 << 16, m=6, seed=1)
    assert ctr["injected"] > 500 and ctr["err"] >= ctr["err_valid"] > 0


def test_detector_on_supplied_bits_equals_keyed_channel():
    """orc_det_run_bits (the detector alone) fed with generator ^ keyed error words reproduces
    orc_prbs_detect exactly; and the slicer model reduces to it: zero noise -> zero errors."""
    L = O.lib()
    for k, m in ((7, 3), (15, 6), (31, 4)):
        nbits = 20000
        tx, _ = O.prbs_bits(k, nbits)
        ew = np.array([L.orc_err_word(1, w, m) for w in range((nbits + 63) // 64)], dtype=np.uint64)
        want, _, wdet = O.prbs_detect(k, nbits, m=m, seed=1)
        # the keyed channel injects only inside [0, nbits): mask the ragged tail the same way
        tail = nbits % 64
        if tail:
            ew[-1] &= np.uint64((1 << tail) - 1)
        got, gdet = O.det_run_bits(k, tx ^ ew, nbits, tx_words=tx)
        assert got == want
        assert (gdet.prbs, gdet.err_sr, gdet.reload_ctr, gdet.bit_in) == (wdet.prbs, wdet.err_sr, wdet.reload_ctr, wdet.bit_in)
    tx, _ = O.prbs_bits(9, 640)
    g = np.zeros(640 * 8, dtype=np.int8)
    assert np.array_equal(O.slicer_rx_bits(tx, g, 640, 100, 8, 8, 3), tx)
    # 12-bit wrap (tx.py:80-81): +2047 + 127*15 overflows to a negative word -> sliced 0
    ones = np.full(10, 0xFFFFFFFFFFFFFFFF, dtype=np.uint64)
    g = np.full(640, 127, dtype=np.int8)
    assert int(O.slicer_rx_bits(ones, g, 640, 2047, 15, 1, 0).sum()) == 0
