#!/usr/bin/env python3
"""Generate tests/golden/golden.json by EXECUTING the reference's own model lines.

Run in the build container (needs /root/reference; it does not exist on the GPU box, so the
resulting JSON is committed).  Nothing is copied: the model statements are read from the
reference checkout at run time, by line number, dedented and exec'd:

  gateware/bbb/rng.py:134-135     LUTOPT numpy model (x = A.x mod 2, integer with bit0 = LSB)
  gateware/bbb/rng.py:173-181     CLT pairwise-difference tree on one URNG value
  gateware/bbb/prbs.py:112-113    PRBS generator model
  software/rnghunt/util/lfsr.py   generator of the Berlekamp-Massey test strings
  software/clt-grng/clt-grng-evaluate.py:10-15  the script's tree (on LUTOPT states, since its
                                  own input is an unseeded np.random.randint)

The matrices are `gateware/bbb/rng_recurrences.py` (imported as-is: pure data).
"""
import contextlib
import importlib.util
import io
import json
import os
import sys
import textwrap

import numpy as np

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def ref_lines(rel, first, last):
    lines = open(os.path.join(REF, rel)).read().split("\n")
    return textwrap.dedent("\n".join(lines[first - 1:last]))


def load_recurrences():
    spec = importlib.util.spec_from_file_location(
        "rng_recurrences", os.path.join(REF, "gateware/bbb/rng_recurrences.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def dense(packed):
    # restates LUTOPT.from_packed (rng.py:50-54) to build `a`
    k = len(packed)
    a = np.zeros((k, k), dtype=np.uint8)
    for row in range(k):
        for idx in packed[row]:
            a[row, idx] = 1
    return a


LUTOPT_MODEL = ref_lines("gateware/bbb/rng.py", 134, 135)
CLT_MODEL = ref_lines("gateware/bbb/rng.py", 175, 180)          # the level loop of :173-181
CLT_SCRIPT = ref_lines("software/clt-grng/clt-grng-evaluate.py", 10, 15)
PRBS_MODEL = ref_lines("gateware/bbb/prbs.py", 112, 113)
assert "np.mod(np.dot(a, x), 2)" in LUTOPT_MODEL and "x_int" in LUTOPT_MODEL
assert "y[pair//2] = x[pair] - x[pair+1]" in CLT_MODEL and "y[pair//2] = x[pair] - x[pair+1]" in CLT_SCRIPT
assert "lfsr" in PRBS_MODEL


def lutopt_and_clt(packed, nsteps, init=1):
    a = dense(packed)
    k = len(packed)
    x = np.zeros((k, 1), dtype=np.uint8)
    for i in range(k):
        x[i] = (init >> i) & 1                                   # rng.py:126 generalised
    states, clts, clts_script = [], [], []
    for _ in range(nsteps):
        ns = {"np": np, "a": a, "x": x}
        exec(LUTOPT_MODEL, ns)
        x, x_int = ns["x"], ns["x_int"]
        states.append("%x" % x_int)
        n = k
        logn = int(np.log2(n))
        # rng.py:173-174 (bit list of the URNG integer, LSB first), then the level loop
        bits = np.array([int(c) for c in bin(int(x_int))[2:].rjust(n, "0")[::-1]])
        ns2 = {"np": np, "x": bits, "logn": logn}
        exec(CLT_MODEL, ns2)
        clts.append(int(ns2["x"][0]))                            # rng.py:181
        ns3 = {"np": np, "x": bits.astype(np.int16), "logn": logn}
        exec(CLT_SCRIPT, ns3)
        clts_script.append(int(ns3["x"][0]))                     # clt-grng-evaluate.py:16
    assert clts == clts_script
    return states, clts


def prbs(k, taps, nbits):
    ns = {"lfsr": 1, "k": k, "TAPS": taps}
    out = []
    for _ in range(nbits):
        exec(PRBS_MODEL, ns)
        out.append(str(ns["bit"]))
    return "".join(out)


def main():
    rr = load_recurrences()
    gold = {"_made_by": "tests/golden/make_golden.py (executes reference model lines)"}
    gold["lutopt_clt"] = {}
    for n, steps in ((16, 100), (32, 100), (64, 40), (128, 30), (192, 0), (256, 64)):
        if steps == 0:
            continue
        states, clts = lutopt_and_clt(getattr(rr, "n%d" % n), steps)
        gold["lutopt_clt"][str(n)] = {"init": 1, "states_hex": states, "clt_numpy": clts}
    # a non-trivial init as well (state = all ones) for n = 16 so the +n/2 corner is exercised:
    # clt(all ones XOR-mask) -- see tests for the wrap case built directly.
    states, clts = lutopt_and_clt(rr.n16, 50, init=0xACE1)
    gold["lutopt_clt"]["16_init_ace1"] = {"init": 0xACE1, "states_hex": states, "clt_numpy": clts}

    taps = {7: 6, 9: 5, 11: 9, 15: 14, 20: 3, 23: 18, 31: 28}
    prbs_src = open(os.path.join(REF, "gateware/bbb/prbs.py")).read()
    assert "TAPS = {7: 6, 9: 5, 11: 9, 15: 14, 20: 3, 23: 18, 31: 28}" in prbs_src
    gold["prbs"] = {str(k): prbs(k, taps, min((1 << k) - 1, 512)) for k in taps}

    # util/lfsr.py prints the two long BM test strings (berlekamp_massey.rs:50,63)
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        exec(open(os.path.join(REF, "software/rnghunt/util/lfsr.py")).read(), {})
    gold["bm_strings_from_lfsr_py"] = buf.getvalue().split()

    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(gold, f, indent=0, sort_keys=True)
    print("wrote golden.json:", {k: (len(v) if hasattr(v, "__len__") else v) for k, v in gold.items()})


if __name__ == "__main__":
    main()
