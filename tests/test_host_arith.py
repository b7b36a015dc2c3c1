"""Host-side arithmetic of the C ABI that needs no GPU: bbb_clt_value (CLTGRNG of one URNG word,
gateware/bbb/rng.py:96-108) against the oracle and against the reference's own model lines (rng.py:173-181),
and the hardware start-up sequence (rng.py:70-72,189-190) the --hw-latency switch of cli/clt_grng_evaluate.py
prepends."""
import random

import numpy as np
import pytest

from basebandboard_b200 import lib


@pytest.mark.parametrize("n", [2, 16, 32, 64, 256, 512])
def test_clt_value_matches_oracle_and_numpy_model(orc, n):
    rnd = random.Random(n)
    for _ in range(100):
        x = rnd.getrandbits(n)
        assert lib.clt_value(n, x, lib.MODEL_NUMPY) == orc.clt(n, x, 1) == orc.np_clt(n, x)
        assert lib.clt_value(n, x, lib.MODEL_GATEWARE) == orc.clt(n, x, 0)
    # the +n/2 corner: every even-parity index set, no odd one (rng.py:78 truncates it to -n/2 in gateware)
    top = sum(1 << i for i in range(n) if bin(i).count("1") % 2 == 0)
    assert lib.clt_value(n, top, lib.MODEL_NUMPY) == n // 2
    assert lib.clt_value(n, top, lib.MODEL_GATEWARE) == (-(n // 2) if n > 2 else lib.clt_value(n, top, 0))
    assert lib.clt_value(n, 0, 0) == 0


def test_clt_value_rejects_bad_arguments():
    with pytest.raises(lib.BBBError):
        lib.clt_value(24, 1)            # not a power of two (rng.py:73-74)


def test_hw_latency_sequence_split():
    """hw[h]: log2(n) zeros, clt(x0), g[0], g[1], ...  (the HDL's pipeline registers reset to 0, rng.py:88-91;
    test_cltgrng compares grng.x with the model delayed by logn, rng.py:189-190)."""
    n, logn = 32, 5
    for first in range(0, 12):
        for count in range(0, 12):
            z, has, g0, gc = lib.hw_latency_split(n, first, count)
            seq = []
            for h in range(first, first + count):
                seq.append("z" if h < logn else "x0" if h == logn else h - logn - 1)
            assert z == seq.count("z") and has == ("x0" in seq)
            gs = [v for v in seq if isinstance(v, int)]
            assert gc == len(gs) and (not gs or (g0 == gs[0] and gs == list(range(g0, g0 + gc))))


def test_hw_latency_against_a_cycle_model_of_the_pipeline(orc):
    """Clock the HDL's register pipeline (rng.py:96-108) level by level from reset on the shipped n32 recurrence and
    compare what it shows after every edge with zeros | clt(x0) | g[i]."""
    from basebandboard_b200 import recurrences
    n, logn = 32, 5
    packed = recurrences.packed(n)
    x = 1
    levels = [[0] * (n >> (lv + 1)) for lv in range(logn)]
    shown = []
    for _ in range(40):
        shown.append(levels[-1][0])                       # grng.x before this edge
        new = [[0] * len(lv) for lv in levels]
        bits = [(x >> i) & 1 for i in range(n)]
        for j in range(n // 2):
            new[0][j] = bits[2 * j] - bits[2 * j + 1]     # rng.py:96-98
        for lv in range(1, logn):
            for j in range(len(levels[lv])):
                new[lv][j] = levels[lv - 1][2 * j] - levels[lv - 1][2 * j + 1]   # rng.py:101-105
        levels = new
        x = orc.lutopt_step(packed, x)
    ref = orc.clt_run(packed, 1, 0, 40, model=1, want=("out16",))["out16"]
    want = [0] * logn + [lib.clt_value(n, 1, lib.MODEL_NUMPY)] + [int(v) for v in ref]
    assert shown == want[:40]
