#!/usr/bin/env python3
"""Generates basebandboard_b200/csrc/clt256_v2_gen.inc: ONE-PASS straight-line body of a LUTOPT(n256)
clock + CLTGRNG popcount network + histogram indices on 32 bit-sliced streams per thread
(reference gateware/bbb/rng.py:38-40, :96-105; matrix rng_recurrences.py:172-259).

Differences from tools/gen_clt256.py (two passes, 128 fixed accumulators, 768 shared-memory words
per clock): the old state is read ONCE (256 words) in the load order tools/clt256_anneal.c chose,
rows are folded as their taps arrive (3 taps -> one LOP3, the 4th -> a second one: 492 LOP3, the
minimum), new planes are stored IN PLACE as soon as their slot has been read (256 words), and
planes that finish late and are needed early in the next clock may stay in registers across the
loop edge ("resident" planes: no store, no load).

Run:  python tools/gen_clt256_v2.py      (deterministic: fixed annealing seed; rewrites the .inc)
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from basebandboard_b200 import recurrences  # noqa: E402
from gen_clt256 import Emitter, BitHeap, gen_transpose, parity  # noqa: E402
import clt256_sched  # noqa: E402

N = 256
LEAD = int(os.environ.get("CLT256V2_LEAD", "6"))          # loads are issued this many steps ahead
ITERS = int(os.environ.get("CLT256V2_ITERS", "12000000"))
SEED = int(os.environ.get("CLT256V2_SEED", "1"))
NRES = int(os.environ.get("CLT256V2_RESIDENT", "0"))      # planes kept in registers across clocks


def run_annealer(rows):
    exe = os.path.join(ROOT, "build", "clt256_anneal")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    subprocess.check_call(["gcc", "-O3", "-o", exe, os.path.join(ROOT, "tools", "clt256_anneal.c"), "-lm"])
    inp = "".join("%d %s\n" % (len(r), " ".join(map(str, r))) for r in rows)
    out = subprocess.run([exe, str(ITERS), str(SEED), str(LEAD)], input=inp.encode(), stdout=subprocess.PIPE,
                         check=True).stdout.decode().split()
    pos = [int(v) for v in out]
    assert sorted(pos) == list(range(N))
    return pos


def pick_residents(rows, pos, nres):
    """Planes whose register life across the loop edge is shortest: finished at `done`, last raw use
    in the next clock at `death` -> wrap length (N - done) + death."""
    if nres <= 0:
        return set()
    times = clt256_sched.row_times(rows, pos)
    death = {b: pos[b] for b in range(N)}
    for r in range(N):
        ts = times[r]
        fold = ts[2] if len(ts) >= 3 else ts[-1]
        for b in rows[r]:
            if pos[b] <= fold and pos[b] in ts[:3]:
                death[b] = max(death[b], fold)
    wrap = sorted(range(N), key=lambda b: (N - times[b][-1]) + death[b])
    return set(wrap[:nres])


def rmw_hooks_v2():
    """32 read-modify-writes of the private histogram (previous clock's samples, words z0..z7), one
    every 8 steps: the load at step 8k+1, the store at step 8k+5."""
    hooks = {}
    k = 0
    for zi in range(8):
        for m in range(4):
            nm = "z%d_%d" % (zi, m)
            hooks.setdefault(8 * k + 1, []).append("CLT256V2_RMW_LD(%s, z%d, %d)" % (nm, zi, m))
            hooks.setdefault(8 * k + 5, []).append("CLT256V2_RMW_ST(%s)" % nm)
            k += 1
    return hooks


def body_hist_v2(rows, pos, resident):
    em = Emitter()
    perm = [0] * N
    for b, p in enumerate(pos):
        perm[p] = b
    taps_at = [[] for _ in range(N)]          # time -> rows tapping the plane loaded then
    for r in range(N):
        for b in rows[r]:
            taps_at[pos[b]].append(r)
    hooks = rmw_hooks_v2()
    heap = BitHeap(em)
    pending = {r: [] for r in range(N)}       # raw planes not yet folded
    acc = {}                                  # row -> partial sum expr
    left = {r: len(rows[r]) for r in range(N)}
    store_at = {}                             # time -> [(slot, expr)]
    old_uses_left = {b: sum(1 for r in range(N) if b in rows[r]) for b in resident}
    assigned = set()

    def wexpr(t):
        b = perm[t]
        return ("r%d" % b) if b in resident else ("w%d" % t)

    def load(t):
        if t < N and perm[t] not in resident:
            em.emit("const uint32_t w%d = st[%d * CLT256V2_B];" % (t, t))
        for slot, ex in store_at.pop(t, []):
            em.emit("st[%d * CLT256V2_B] = %s;" % (slot, ex))

    def finish(r, ex, t):
        if r in resident:
            assert old_uses_left[r] == 0, "resident plane %d overwritten while its old value is needed" % r
            em.emit("r%d = %s;" % (r, ex))
            assigned.add(r)
            return
        slot = pos[r]
        if slot - LEAD <= t:
            em.emit("st[%d * CLT256V2_B] = %s;" % (slot, ex))
        else:
            store_at.setdefault(slot, []).append((slot, ex))

    for t in range(min(LEAD, N)):
        load(t)
    for t in range(N):
        load(t + LEAD)
        for ln in hooks.get(t, []):
            em.emit(ln)
        w = wexpr(t)
        heap.add(w, parity(perm[t]) == 1)
        heap.reduce()
        for r in taps_at[t]:
            pending[r].append(w)
            left[r] -= 1
            if perm[t] in resident:
                old_uses_left[perm[t]] -= 1
            if r not in acc and len(pending[r]) == 3:
                nm = em.tmp()
                em.emit("const uint32_t %s = %s ^ %s ^ %s;" % (nm, *pending[r]))
                em.ops["rec"] += 1
                pending[r] = []
                if left[r] == 0:
                    finish(r, nm, t)
                else:
                    acc[r] = nm
            elif left[r] == 0:
                terms = ([acc[r]] if r in acc else []) + pending[r]
                assert 2 <= len(terms) <= 3
                nm = em.tmp()
                em.emit("const uint32_t %s = %s;" % (nm, " ^ ".join(terms)))
                em.ops["rec"] += 1
                finish(r, nm, t)
    assert not store_at, store_at
    assert assigned == set(resident)
    planes = heap.finish()
    for wgt in range(9):
        em.emit("const uint32_t pc%d = %s;" % (wgt, planes[wgt]))
    em.emit("const uint32_t inr = ~pc8 & ((~pc7 & pc6 & pc5) | (pc7 & ~pc6 & ~pc5));   // 96 <= popcount < 160")
    em.emit("CLT256V2_AFTER_COUNT(pc0, pc1, pc2, pc3, pc4, pc5, pc6, pc7, pc8, inr)")
    q = []
    for b, src in enumerate(["pc0", "pc1", "pc2", "pc3", "pc4", "pc7"]):
        nm = "q%d" % b
        em.emit("const uint32_t %s = %s & inr;" % (nm, src))
        q.append(nm)
    gen_transpose(em, q)
    return em, perm


def main():
    rows = recurrences.packed(N)
    pos = run_annealer(rows)
    peak = int(clt256_sched.live_profile(rows, pos, LEAD).max())
    resident = pick_residents(rows, pos, NRES)
    em, perm = body_hist_v2(rows, pos, resident)
    out = os.path.join(ROOT, "basebandboard_b200", "csrc", "clt256_v2_gen.inc")
    with open(out, "w") as f:
        f.write("// GENERATED by tools/gen_clt256_v2.py -- do not edit.  One-pass LUTOPT(n256) clock + CLTGRNG popcount\n")
        f.write("// network + histogram indices on 32 bit-sliced streams (rng.py:38-40, :96-105; rng_recurrences.py:172-259).\n")
        f.write("// recurrence LOP3: %d, carry-save adders: %d, half adders: %d, modelled peak live values: %d, lead %d, resident %d\n"
                % (em.ops["rec"], em.ops["csa"], em.ops["ha"], peak, LEAD, len(resident)))
        f.write("#ifdef CLT256V2_SECTION_PERM\n// slot p (= load time p) holds state bit CLT256V2_PERM[p]\n")
        f.write("static __constant__ uint8_t CLT256V2_PERM[256] = {%s};\n" % ", ".join(str(v) for v in perm))
        f.write("#endif\n#ifdef CLT256V2_SECTION_RESIDENT_DECL\n")
        for b in sorted(resident):
            f.write("    uint32_t r%d = st[%d * CLT256V2_B];\n" % (b, pos[b]))
        f.write("#endif\n#ifdef CLT256V2_SECTION_BODY\n")
        f.write("\n".join(em.lines))
        f.write("\n#endif\n")
    print("wrote", out, "ops", em.ops, "peak live", peak, "resident", len(resident), file=sys.stderr)


if __name__ == "__main__":
    main()
