"""CPU: the C-ABI library loads, exports every symbol include/bbb_b200.h declares, and has no CPU
fallback (no compute call is made here: there is no GPU in this container)."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "bbb_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bbb_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound():
    from basebandboard_b200 import lib
    L = lib.load()
    names = _declared()
    assert len(names) >= 30
    for name in names:
        assert hasattr(L, name), "libbbb_b200.so does not export %s" % name
        assert name in lib.SIGNATURES, "basebandboard_b200.lib has no prototype for %s" % name
    assert set(lib.SIGNATURES) <= set(names)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from basebandboard_b200 import lib
    with pytest.raises(lib.BBBError):
        lib.Context(0)
    h = ctypes.c_void_p()
    assert lib.load().bbb_create(0, ctypes.byref(h)) != 0 and not h.value


def test_host_only_entry_points(orc):
    """bbb_prbs_tap (prbs.py:14), bbb_clt_stats (clt-grng-evaluate.py:18-21) and
    bbb_prbs_spec_state are plain host arithmetic."""
    from basebandboard_b200 import lib
    L = lib.load()
    assert [L.bbb_prbs_tap(k) for k in (7, 9, 11, 15, 20, 23, 31, 8)] == [6, 5, 9, 14, 3, 18, 28, 0]
    rng = np.random.default_rng(1)
    g = np.clip(np.rint(rng.normal(0, 8, 200000)), -128, 127).astype(np.int64)
    hist = np.bincount(g + 128, minlength=256).astype(np.uint64)
    st = lib.clt_stats(256, hist)
    assert abs(st["mean"] - g.mean()) <= 1e-9 * max(1.0, abs(g.mean()))
    assert abs(st["variance"] - g.var()) <= 1e-9 * g.var()
    m4 = ((g - g.mean()) ** 4).mean()
    assert abs(st["kurtosis"] - (m4 / g.var() ** 2 - 3.0)) <= 1e-9
    assert st["variance_theory"] == 64.0
    # the locked-state helper agrees with the oracle's detector after a clean start
    for k, first in ((7, 640), (31, 100000)):
        _, _, det = orc.prbs_detect(k, first, m=8, seed=2)
        got = lib.prbs_spec_state(k, first, 8, 2)
        if det.reload_ctr == 0:
            assert (got.bit_in, got.prbs, got.err_sr) == (det.bit_in, det.prbs, det.err_sr)
