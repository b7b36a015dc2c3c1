#!/usr/bin/env python3
"""GPU stand-in for the reference's software/rnghunt/util/verify.py.

The reference script loads an `out` matrix, prints its weights, runs 2n steps from a random start
and hands the sequence to an external `./ppsearch` binary, saves a picture, and writes `outnums`: a
dieharder ASCII input file of the URNG words of 200000 steps (verify.py:38-52).  Here:

  * the same four header lines on stdout (verify.py:10-17);
  * the verdict comes from the device (bbb_rnghunt_eval: Berlekamp-Massey degree == n and primitive,
    rnghunt.rs:36-44) instead of ppsearch, which is not part of the reference tree;
  * `outnums` has the reference's exact header and word format -- per step n/32 decimal words, word j =
    state bits 32j .. 32j+31 with bit 32j as the MOST significant one (verify.py:48-52) -- and its
    words come from the device LUTOPT stream (bbb_lutopt_stream), `--nsamples` steps from `--init`
    (the reference draws an unseeded random start vector, so its numbers are not reproducible;
    the default here is LUTOPT's reset value 1, rng.py:21,30);
  * the PNG is left out (no imaging library on this box).

    python cli/verify.py out [--nsamples 200000] [--init 1] [--outnums outnums]
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pack import read_matrix, tap_lists  # noqa: E402
from basebandboard_b200 import recurrences  # noqa: E402

HEADER = """#==================================================================
# generator rnghunt  seed = 0
#==================================================================
type: d
count: {}
numbit: 32
"""


def dieharder_words(state_words, n):
    """state_words: uint64 [steps, ceil(n/64)], state bit i = weight 2^i (rng.py:135).
    -> uint32 [steps, n//32]: word j = bits 32j..32j+31, bit 32j most significant."""
    w32 = np.ascontiguousarray(state_words).view("<u4")[:, : n // 32].copy()
    # reverse the 32 bits of every word
    w = w32.astype(np.uint32)
    w = ((w >> 1) & 0x55555555) | ((w & 0x55555555) << 1)
    w = ((w >> 2) & 0x33333333) | ((w & 0x33333333) << 2)
    w = ((w >> 4) & 0x0F0F0F0F) | ((w & 0x0F0F0F0F) << 4)
    w = ((w >> 8) & 0x00FF00FF) | ((w & 0x00FF00FF) << 8)
    return ((w >> 16) | (w << 16)).astype(np.uint32)


def write_outnums(path, words):
    with open(path, "w") as f:
        f.write(HEADER.format(words.size))
        flat = words.reshape(-1)
        for i in range(0, flat.size, 1 << 16):
            f.write("\n".join(map(str, flat[i:i + (1 << 16)].tolist())) + "\n")


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("matrix")
    ap.add_argument("--nsamples", type=int, default=200000)
    ap.add_argument("--init", type=lambda s: int(s, 0), default=1)
    ap.add_argument("--outnums", default="outnums")
    ap.add_argument("--gpu", type=int, default=0)
    args = ap.parse_args(argv)
    n, a = read_matrix(args.matrix)
    print("Detected n={}".format(n))
    print("Loaded matrix")
    print("Row weights:", a.sum(axis=1))
    print("Col weights:", a.sum(axis=0))
    taps = tap_lists(a)
    from basebandboard_b200 import lib
    with lib.Context(args.gpu) as c:
        stage, poly = c.rnghunt_eval(n, recurrences.to_binary_matrix(taps), want_poly=True)
        ok = int(stage[0]) == lib.STAGE_PRIMITIVE
        print("Verdict: %s" % ("maximum period (degree %d, primitive)" % n if ok
                               else "rejected at stage '%s'" % lib.STAGE_NAMES[int(stage[0])]))
        if ok:
            # the one quality figure the shipped matrices are compared by (103 terms at n = 192)
            print("Characteristic polynomial weight: %d" % sum(bin(int(w)).count("1") for w in poly[0]))
        c.lutopt_load(taps)
        words = dieharder_words(c.lutopt_stream(args.init, 0, args.nsamples), n)
    write_outnums(args.outnums, words)
    print("Wrote %d words to %s" % (words.size, args.outnums))
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
