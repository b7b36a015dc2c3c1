#!/usr/bin/env python3
"""Generate tests/golden/shaper_rcf.json by EXECUTING the reference's own lines (run in the build
container; /root/reference does not exist on the GPU box, so the JSON is committed):

  gateware/bbb/bitshaper.py:96-108    body of PRBSShaper.from_rcf (raised-cosine sets, float -> int), with
                                      `np.int` (removed from numpy 1.24+) bound to int
  gateware/bbb/bitshaper.py:150-154   the relation the reference test asserts between the PRBS and the
                                      shaped output (scipy.signal.lfilter), evaluated on PRBS-9 with
                                      beta = 0.5 -> "lfilter_prbs9" (the f[60:-13] slice it compares)
"""
import json
import os
import sys
import textwrap

import numpy as np
import scipy.signal

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
src = open(os.path.join(REF, "gateware/bbb/bitshaper.py")).read().split("\n")
body = textwrap.dedent("\n".join(src[95:108]))           # bitshaper.py:96-108
assert body.startswith("T = 8") and "cc.append((c * T * 254).astype(np.int).tolist())" in body
np.int = int                                             # the reference predates its removal


def from_rcf(betas):
    ns = {"np": np, "betas": betas}
    exec(body, ns)
    return ns["cc"]


out = {"tx_sets": from_rcf(np.linspace(0, 1, 32).tolist()),          # tx.py:53-57
       "half_plus_rect": from_rcf([0.5])}
# the PRBS-9 bit sequence by the reference's PRBS model lines (prbs.py:112-113 semantics restated in
# tests/golden/golden.json already); here only the filter relation is evaluated
gold = json.load(open(os.path.join(HERE, "golden.json")))
bits = [int(ch) for ch in gold["prbs"]["9"][:40]]        # produced by prbs.py:112-113 (make_golden.py)
prbs = np.repeat(np.array(bits), 8)                      # prbs.x held for 8 sys clocks per bit
c = out["half_plus_rect"][0]
rel = textwrap.dedent("\n".join(src[149:154]))           # bitshaper.py:150-154
assert rel.startswith("b = np.array(c)") and "scipy.signal.lfilter(b, [1], y)" in rel
ns = {"np": np, "scipy": scipy, "c": c, "prbs": prbs.tolist()}
exec(rel, ns)
out["lfilter_prbs9"] = {"bits": bits, "f": [int(v) for v in ns["f"]]}
json.dump(out, open(os.path.join(HERE, "shaper_rcf.json"), "w"))
print("wrote shaper_rcf.json", len(out["tx_sets"]), len(out["lfilter_prbs9"]["f"]))
