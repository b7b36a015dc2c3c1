#!/usr/bin/env python3
"""prbs -- PRBS generator / error-detector runs on the B200 library (the reference has no PRBS tool;
gateware/bbb/prbs.py is the specification).

  prbs_tool.py gen    --k 31 --nbits N [--first-bit F] [--state0 S] [--out FILE]   packed bits, LSB first
  prbs_tool.py detect --k 31 --nbits N [--err-m M] [--seed S] [--first-bit F]       counters as JSON
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("cmd", choices=["gen", "detect"])
    ap.add_argument("--k", type=int, default=31)
    ap.add_argument("--nbits", type=int, default=1 << 20)
    ap.add_argument("--first-bit", type=int, default=0)
    ap.add_argument("--state0", type=lambda s: int(s, 0), default=1)
    ap.add_argument("--err-m", type=int, default=0, help="channel error probability 2^-M (0 = none)")
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--out", default=None)
    ap.add_argument("--gpu", type=int, default=0)
    a = ap.parse_args()
    from basebandboard_b200 import lib
    with lib.Context(a.gpu) as c:
        if a.cmd == "gen":
            words = c.prbs_stream(a.k, (a.nbits + 63) // 64 * 64, state0=a.state0, first_bit=a.first_bit)
            if a.out:
                words.tofile(a.out)
            else:
                print("".join(str((int(words[i >> 6]) >> (i & 63)) & 1) for i in range(min(a.nbits, 4096))))
        else:
            ctr = c.prbs_detect(a.k, a.nbits, m=a.err_m, seed=a.seed, first_bit=a.first_bit)
            ctr.update(k=a.k, nbits=a.nbits, err_m=a.err_m, seed=a.seed, device_ms=c.last_device_ms,
                       rounds=c.last_rounds)
            print(json.dumps(ctr))


if __name__ == "__main__":
    main()
