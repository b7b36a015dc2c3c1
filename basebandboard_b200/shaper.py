"""Pulse-shape coefficient sets of the baseband transmitter (host side of bbb_tx_stream).

Restates PRBSShaper.from_rcf (reference gateware/bbb/bitshaper.py:93-109) and the set list TX builds
(tx.py:53-57: betas = np.linspace(0, 1, 32)): raised-cosine pulses sampled at t = -32..31 with T = 8,
scaled by T*254 and truncated toward zero to integers.  tests/golden/shaper_rcf.json holds the tables the
reference's own lines produce under this image's numpy; tests/test_shaper.py requires equality.
"""
import numpy as np

T = 8
NTAPS = 64


def rcf_pulse(beta):
    """One raised-cosine pulse (bitshaper.py:101-108) -> list of 64 ints."""
    t = np.arange(-NTAPS // 2, NTAPS // 2)
    hole = None
    if beta != 0.0:
        hole = np.where(np.abs(t) == T / (2 * beta))      # the 0/0 points of the closed form
        t[hole] = 0
    with np.errstate(divide="ignore", invalid="ignore"):
        c = 1 / T * np.sinc(t / T) * np.cos(np.pi * beta * t / T) / (1 - (2 * beta * t / T) ** 2)
    if hole is not None:
        c[hole] = np.pi / (4 * T) * np.sinc(1 / (2 * beta))
    return (c * T * 254).astype(np.int64).tolist()


def rect_pulse():
    """The rectangular set from_rcf appends when fewer than 32 betas are given (bitshaper.py:109-110)."""
    return [0] * 30 + [254] * 4 + [0] * 30


def tx_sets():
    """The 32 sets TX instantiates (tx.py:53-57); index = shape_sel."""
    return [rcf_pulse(float(b)) for b in np.linspace(0, 1, 32).tolist()]


def from_rcf(betas):
    cc = [rcf_pulse(float(b)) for b in betas]
    if len(cc) < 32:
        cc.append(rect_pulse())
    return cc
