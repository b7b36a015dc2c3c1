#!/usr/bin/env python3
"""Drop-in for the reference's software/rnghunt/util/pack.py: turns an `out` /
`matrices/N` text matrix (n lines of n '0'/'1' characters, rnghunt.rs:46-53) into the packed tap
lists that gateware/bbb/rng_recurrences.py holds.

    python cli/pack.py out                 same stdout as the reference script (pack.py:9-23)
    python cli/pack.py out --module        additionally a ready-to-paste `nN = [...]` block in the
                                           layout of rng_recurrences.py:8-16
    python cli/pack.py out --check         also evaluates the matrix on the GPU (BM degree + primitivity,
                                           rnghunt.rs:36-44) and prints the verdict

Pure host work except --check; numpy is used for the weight lines so that they print exactly as the
reference's `print("Row weights:", a.sum(axis=1))` does.
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from basebandboard_b200 import recurrences  # noqa: E402


def read_matrix(path):
    """-> (n, uint8 array [n, n]); row r = line r, column c = character c."""
    with open(path) as f:
        rows = [ln.strip() for ln in f if ln.strip()]
    n = len(rows[0])
    if len(rows) != n or any(len(r) != n or set(r) - {"0", "1"} for r in rows):
        raise SystemExit("%s: not an n x n 0/1 matrix" % path)
    a = (np.frombuffer("".join(rows).encode(), dtype=np.uint8) - ord("0")).reshape(n, n)
    return n, a


def tap_lists(a):
    return [np.flatnonzero(row).tolist() for row in a]


def packed_block(taps, per_line=3):
    """the "Packed:" block of pack.py:20-23: rows three to a line, each line ending in a comma"""
    out = ["["]
    for i in range(0, len(taps), per_line):
        out.append(", ".join(str(t) for t in taps[i:i + per_line]) + ",")
    out.append("]")
    return "\n".join(out)


def module_block(n, taps, per_line=3):
    """rng_recurrences.py:8-16 layout: `nN = [` / 4-space indented rows / `]`"""
    out = ["n%d = [" % n]
    for i in range(0, len(taps), per_line):
        out.append("    " + ", ".join(str(t) for t in taps[i:i + per_line]) + ",")
    out.append("]")
    return "\n".join(out)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("matrix")
    ap.add_argument("--module", action="store_true")
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--gpu", type=int, default=0)
    args = ap.parse_args(argv)
    n, a = read_matrix(args.matrix)
    print("Detected n={}".format(n))
    print("Loaded matrix")
    print("Row weights:", a.sum(axis=1))
    print("Col weights:", a.sum(axis=0))
    taps = tap_lists(a)
    print("Packed:\n" + packed_block(taps))
    if args.module:
        print()
        print(module_block(n, taps))
    if args.check:
        from basebandboard_b200 import lib
        with lib.Context(args.gpu) as c:
            stage = int(c.rnghunt_eval(n, recurrences.to_binary_matrix(taps))[0])
        print("Verdict: %s" % ("maximum period (degree %d, primitive)" % n if stage == lib.STAGE_PRIMITIVE
                               else "rejected at stage '%s'" % lib.STAGE_NAMES[stage]))
    return 0


if __name__ == "__main__":
    sys.exit(main())
