#!/usr/bin/env python3
"""Generate the fixtures for cli/pack.py and cli/verify.py by RUNNING the reference's own scripts.

Run in the build container (needs /root/reference; it does not exist on the GPU box, so the
resulting files are committed).  Nothing is copied:

  software/rnghunt/util/pack.py             executed as a subprocess on matrices/16 and matrices/64
                                            -> pack_16.txt, pack_64.txt (its whole stdout)
  software/rnghunt/util/verify.py:37-52     the `outnums` writer, read from the checkout by line
                                            number and exec'd on matrices/32 with a FIXED start vector
                                            (the script's own is an unseeded np.random.randint) and
                                            nsamples = 64 -> outnums_32.txt
"""
import os
import subprocess
import sys
import textwrap

import numpy as np

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
UTIL = os.path.join(REF, "software/rnghunt/util")
MATS = os.path.join(REF, "software/rnghunt/matrices")

for n in (16, 64):
    out = subprocess.run([sys.executable, os.path.join(UTIL, "pack.py"), os.path.join(MATS, str(n))],
                         stdout=subprocess.PIPE, check=True, cwd=HERE).stdout.decode()
    open(os.path.join(HERE, "pack_%d.txt" % n), "w").write(out)

src = open(os.path.join(UTIL, "verify.py")).read().split("\n")
writer = textwrap.dedent("\n".join(src[36:52]))          # verify.py:37-52
assert writer.startswith("nsamples = 200000") and "numbit: 32" in writer and "int(bitstring[i:i+32], 2)" in writer
writer = writer.replace("nsamples = 200000", "nsamples = 64").replace('open("outnums", "w")', "open(OUT, 'w')")
n = 32
rows = [ln.strip() for ln in open(os.path.join(MATS, str(n)))]
a = np.array([[int(ch) for ch in r] for r in rows], dtype=np.uint8)
b = np.zeros((n, 1), dtype=np.uint8)
b[0, 0] = 1                                              # LUTOPT reset value init = 1: state bit 0 set
exec(writer, {"np": np, "a": a, "b": b, "n": n, "OUT": os.path.join(HERE, "outnums_32.txt")})
print("wrote pack_16.txt pack_64.txt outnums_32.txt")
