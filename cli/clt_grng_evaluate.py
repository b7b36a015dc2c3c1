#!/usr/bin/env python3
"""clt-grng-evaluate -- software/clt-grng/clt-grng-evaluate.py on the B200 library.

With no arguments: n = 256, 100 000 samples (clt-grng-evaluate.py:4-6) and the reference's two
stdout lines (:30-31).  The reference feeds the CLT tree with unseeded numpy random bits and
plots; here the bits come from the gateware's own uniform generator, LUTOPT(n256, init=1)
(gateware/bbb/tx.py:70-71), the samples are produced and histogrammed on the GPU, and the plot is
replaced by numbers (--hist prints the unit-width histogram the reference would draw, :34).

  --n N --nsamp COUNT --init X --first I --model gateware|numpy --hist --stats --gpu D --hw-latency
  --hw-latency  the sequence the hardware emits after reset instead of g[first..]: log2(n) zeros (the reset
                value of the pipeline registers), clt(x0), then g[0], g[1], ... (gateware/bbb/rng.py:70-72,189-190)
"""
import argparse
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=256)
    ap.add_argument("--nsamp", type=int, default=100000)
    ap.add_argument("--init", type=lambda s: int(s, 0), default=1)
    ap.add_argument("--first", type=int, default=0)
    ap.add_argument("--model", choices=["gateware", "numpy"], default="numpy")
    ap.add_argument("--hist", action="store_true")
    ap.add_argument("--stats", action="store_true")
    ap.add_argument("--gpu", type=int, default=0)
    ap.add_argument("--hw-latency", action="store_true")
    a = ap.parse_args()
    from basebandboard_b200 import lib
    logn = int(math.log2(a.n))
    model = lib.MODEL_NUMPY if a.model == "numpy" else lib.MODEL_GATEWARE
    zeros, has_x0, g_first, g_count = (0, False, a.first, a.nsamp)
    if a.hw_latency:
        zeros, has_x0, g_first, g_count = lib.hw_latency_split(a.n, a.first, a.nsamp)
    with lib.Context(a.gpu) as c:
        c.lutopt_load_shipped(a.n)
        hist, s1, s2 = c.clt_hist(a.init, g_first, g_count, model)
    extra = [0] * zeros + ([lib.clt_value(a.n, a.init, model)] if has_x0 else [])
    for g in extra:                                        # the start-up samples the device stream does not hold
        hist[g + a.n // 2] += 1
        s1 += g
        s2 += g * g
    mu_s = s1 / a.nsamp                                    # np.mean  (:19)
    var_s = s2 / a.nsamp - mu_s * mu_s                     # np.var, ddof = 0  (:21)
    mu_t, var_t = 0, float(2 ** (logn - 2))                # :18, :20
    print("Theoretical mean μ={:.4e}, variance σ²={:.4e}.".format(mu_t, var_t))
    print("Sample mean μ={:.4e}, variance σ²={:.4e}.".format(mu_s, var_s))
    if a.stats:
        st = lib.clt_stats(a.n, hist, model)
        print("excess kurtosis={kurtosis:.6e} chi2(normal)={chi2_normal:.6e} chi2(binomial)={chi2_binomial:.6e}"
              .format(**st), file=sys.stderr)
    if a.hist:
        for b, h in enumerate(hist):
            if h:
                print("%d %d" % (b - a.n // 2, h))


if __name__ == "__main__":
    main()
