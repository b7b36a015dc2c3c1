#!/usr/bin/env python3
"""Load-order optimiser for the one-pass clt256 kernel (tools/gen_clt256_v2.py).

One LUTOPT(n256) clock (reference gateware/bbb/rng.py:38-40, rng_recurrences.py:172-259) is done in a
single pass over the old state: state plane `bit` is loaded at time pos[bit] (0..255); a row's first
three taps are folded by one LOP3 when the third arrives, a fourth tap by a second one.  The new
plane of a row is stored in place, i.e. at slot pos[row], as soon as that slot has been read.

What the order decides is how many values are live in registers at a time:
  * a loaded plane, from its load until the last row that still needs it in raw form has acted;
  * the partial sum of a 4-tap row, between its third and fourth tap;
  * a finished row whose slot has not been read yet (it waits for the load of that slot).
live_profile() computes that count per time step, anneal() minimises its peak (then its mean).
"""
import math
import random

import numpy as np

N = 256


def row_times(rows, pos):
    """per row: sorted tap times"""
    return [sorted(pos[t] for t in rows[r]) for r in range(N)]


def live_profile(rows, pos, lead=0):
    """live[t] = values held in registers after step t's work (t = 0..255).
    lead: slots are read `lead` steps ahead of use (prefetch), so slot s is free from time s-lead."""
    diff = np.zeros(N + 2, dtype=np.int64)
    death = list(range(N))                      # by position: last raw use
    inv = [0] * N
    for b, p in enumerate(pos):
        inv[p] = b
    for r in range(N):
        ts = sorted(pos[t] for t in rows[r])
        fold = ts[2] if len(ts) >= 3 else ts[-1]
        for t in ts[:3]:
            if death[t] < fold:
                death[t] = fold
        done = ts[-1]
        if len(ts) == 4 and ts[3] > ts[2]:
            diff[ts[2]] += 1                    # partial sum live on [t3, t4)
            diff[ts[3]] -= 1
        free = pos[r] - lead
        if free > done:                         # finished row waits for its slot
            diff[done] += 1
            diff[free] -= 1
    for p in range(N):
        if death[p] > p:
            diff[p] += 1
            diff[death[p]] -= 1
    return np.cumsum(diff)[:N]


def cost_of(rows, pos, lead=0):
    lv = live_profile(rows, pos, lead)
    return float(lv.max()) + float(lv.mean()) / 64.0, int(lv.max()), float(lv.mean())


def anneal(rows, iters, seed=2024, lead=0, verbose=False, start=None):
    rnd = random.Random(seed)
    pos = list(start) if start is not None else list(range(N))
    cur, _, _ = cost_of(rows, pos, lead)
    best, best_pos = cur, pos[:]
    T0, T1 = 2.0, 0.02
    for it in range(iters):
        T = T0 * (T1 / T0) ** (it / iters)
        a, b = rnd.randrange(N), rnd.randrange(N)
        if a == b:
            continue
        pos[a], pos[b] = pos[b], pos[a]
        c, _, _ = cost_of(rows, pos, lead)
        d = c - cur
        if d <= 0 or rnd.random() < math.exp(-d / T):
            cur = c
            if cur < best:
                best, best_pos = cur, pos[:]
        else:
            pos[a], pos[b] = pos[b], pos[a]
        if verbose and it % 20000 == 0:
            print("it", it, "cur %.3f best %.3f" % (cur, best), flush=True)
    return best_pos, best


if __name__ == "__main__":
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from basebandboard_b200 import recurrences
    rows = recurrences.packed(N)
    print("identity", cost_of(rows, list(range(N))))
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
    p, c = anneal(rows, iters, verbose=True)
    print("annealed", cost_of(rows, p))
