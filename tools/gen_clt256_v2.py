#!/usr/bin/env python3
"""Generates basebandboard_b200/csrc/clt256_v2_gen.inc: ONE-PASS straight-line body of a LUTOPT(n256)
clock + CLTGRNG popcount network + histogram indices on 32 bit-sliced streams per thread
(reference gateware/bbb/rng.py:38-40, :96-105; matrix rng_recurrences.py:172-259).

Differences from tools/gen_clt256.py (two passes, 128 fixed accumulators, 768 shared-memory words
per clock): the old state is read ONCE (256 words) in the load order tools/clt256_anneal.c chose,
rows are folded as their taps arrive (3 taps -> one LOP3, the 4th -> a second one: 492 LOP3, the
minimum without sub-sums shared between rows), new planes are stored IN PLACE as soon as their slot has been read (256 words), and
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
LEAD = int(os.environ.get("CLT256V2_LEAD", "10"))         # loads are issued this many steps ahead
ITERS = int(os.environ.get("CLT256V2_ITERS", "12000000"))
SEED = int(os.environ.get("CLT256V2_SEED", "1"))
NRES = int(os.environ.get("CLT256V2_RESIDENT", "104"))    # most planes kept in registers across clocks
RES_BUDGET = int(os.environ.get("CLT256V2_RES_BUDGET", "200"))   # modelled live values allowed with them
NRES_PLAIN = int(os.environ.get("CLT256V2_PLAIN_RESIDENT", "256"))   # resident planes of the plain-clock body P2
PLAIN_BUDGET = int(os.environ.get("CLT256V2_PLAIN_BUDGET", "205"))
# the committed load order (annealer output, 60M iterations, seed 3, lead 6; the settings here were picked
# by a measured sweep on the B200, gpurun_out/variants.log); CLT256V2_REANNEAL=1 recomputes an order
POS_FILE = None if os.environ.get("CLT256V2_REANNEAL") else \
    os.environ.get("CLT256V2_POS", os.path.join(ROOT, "tools", "clt256_v2_order.txt"))


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


def pick_residents(rows, pos, nres, budget, overlap=False, start=()):
    """Planes that stay in registers across the loop edge instead of going through shared memory.
    Plane b is finished at `done` and last needed in raw form at `death` of the next clock; it is
    eligible when done > death (the old value is dead before the new one exists) and costs one
    register over [done, N) and [0, pos[b]).  Greedy by shortest wrap, while the modelled live count
    stays within `budget`."""
    if nres <= 0:
        return set()
    import numpy as np
    times = clt256_sched.row_times(rows, pos)
    perm = [0] * N
    for b, p in enumerate(pos):
        perm[p] = b
    death = {b: pos[b] for b in range(N)}
    for r in range(N):
        ts = times[r]
        fold = ts[2] if len(ts) >= 3 else ts[-1]
        for t in ts[:3]:
            death[perm[t]] = max(death[perm[t]], fold)
    # overlap: also planes whose new value exists before the old one is dead (two registers for a while; the body
    # assigns those at the end of the clock).  start: planes that are resident in any case (their cost comes first).
    cand = sorted(((N - times[b][-1]) + death[b], b) for b in range(N) if overlap or times[b][-1] > death[b])
    cand = [(k, b) for k, b in cand if b in start] + [(k, b) for k, b in cand if b not in start]
    cur = clt256_sched.live_profile(rows, pos, LEAD).copy()
    chosen = set()
    for _, b in cand:
        if len(chosen) >= nres:
            break
        budget_b = (1 << 30) if b in start else budget
        done = times[b][-1]
        add = np.zeros(N, dtype=np.int64)
        add[done:] += 1
        add[:pos[b]] += 1
        fr = pos[b] - LEAD
        if fr > done:
            add[done:fr] -= 1          # the wait for its slot is already in the profile
        if (cur + add).max() <= budget_b:
            cur += add
            chosen.add(b)
    assert set(start) <= chosen
    return chosen


def slot_map(pos, resident):
    """load time -> compact shared-memory slot, in load order, skipping resident planes"""
    perm = [0] * N
    for b, p in enumerate(pos):
        perm[p] = b
    m, k = {}, 0
    for t in range(N):
        if perm[t] not in resident:
            m[t] = k
            k += 1
    return m


RMW_FIRST = int(os.environ.get("CLT256V2_RMW_FIRST", "60"))   # first step with a histogram update
RMW_EVERY = 6
RMW_GAP = int(os.environ.get("CLT256V2_RMW_GAP", "5"))         # steps between its load and its store
TAIL_STEPS = 40                                                # steps the previous clock's tail is spread over


def rmw_hooks_v2():
    """32 read-modify-writes of the private histogram (words z0..z7 of the PREVIOUS clock, produced by
    the tail code in the first TAIL_STEPS steps), one every RMW_EVERY steps."""
    hooks = {}
    k = 0
    for zi in range(8):
        for m in range(4):
            nm = "z%d_%d" % (zi, m)
            t0 = RMW_FIRST + RMW_EVERY * k
            assert t0 + RMW_GAP < N and RMW_GAP < RMW_EVERY
            hooks.setdefault(t0, []).append("CLT256V2_RMW_LD(%s, z%d, %d)" % (nm, zi, m))
            hooks.setdefault(t0 + RMW_GAP, []).append("CLT256V2_RMW_ST(%s)" % nm)
            k += 1
    return hooks


def tail_lines(shape, kind="hist"):
    """Finishes the popcount network of the previous clock from the carried planes k<w>_<i>
    (shape[w] = [negated?, ...]) and un-slices into z0..z7: the histogram indices (kind "hist"), or the
    int8 samples g = popcount - 128 in two's complement, gateware width rng.py:78 (kind "s8")."""
    em = Emitter()
    em.ntmp = 100000
    heap = BitHeap(em)
    for w, flags in enumerate(shape):
        heap.heap[w] = [("k%d_%d" % (w, i), neg) for i, neg in enumerate(flags)]
    planes = heap.finish()
    for wgt in range(9):
        em.emit("const uint32_t pc%d = %s;" % (wgt, planes[wgt]))
    if kind == "s8":
        em.emit("const uint32_t pc7n = ~pc7;")
        gen_transpose(em, ["pc0", "pc1", "pc2", "pc3", "pc4", "pc5", "pc6", "pc7n"], decl="const uint32_t ",
                      bsel="CLT256V2_BSEL(%s, %s, %s)")
        em.emit("CLT256V2_S8_STAGE(z0, z1, z2, z3, z4, z5, z6, z7)")
        return em.lines, em.ops
    em.emit("const uint32_t inr = ~pc8 & ((~pc7 & pc6 & pc5) | (pc7 & ~pc6 & ~pc5));   // 96 <= popcount < 160")
    em.emit("CLT256V2_AFTER_COUNT(pc0, pc1, pc2, pc3, pc4, pc5, pc6, pc7, pc8, inr)")
    q = []
    for b, src in enumerate(["pc0", "pc1", "pc2", "pc3", "pc4", "pc7"]):
        nm = "q%d" % b
        em.emit("const uint32_t %s = %s & inr;" % (nm, src))
        q.append(nm)
    # right shifts as multiply-high by a power of two held in a register: FMA pipe instead of the ALU pipe
    gen_transpose(em, q, decl="const uint32_t ", shr="CLT256V2_SHR(%s, %d)", bsel="CLT256V2_BSEL(%s, %s, %s)")
    return em.lines, em.ops


def carried_init(shape):
    """Values of the carried planes before the first clock: a popcount of exactly 128 in every stream
    (in range, private bin 32; the kernel takes those 32 counts back at the end)."""
    need = 128
    init = {}
    for w in range(9, -1, -1):
        for i, neg in enumerate(shape[w] if w < len(shape) else []):
            one = need >= (1 << w)
            if one:
                need -= 1 << w
            init["k%d_%d" % (w, i)] = 0xFFFFFFFF if (one != neg) else 0
    assert need == 0, "carried shape cannot represent 128"
    return init


def body_hist_v2(rows, pos, resident, shape=None, kind="hist", slot_resident=None):
    """shape None: dry run (returns the carried-plane shape); else the real body."""
    em = Emitter()
    perm = [0] * N
    for b, p in enumerate(pos):
        perm[p] = b
    taps_at = [[] for _ in range(N)]          # time -> rows tapping the plane loaded then
    for r in range(N):
        for b in rows[r]:
            taps_at[pos[b]].append(r)
    hooks = rmw_hooks_v2() if kind == "hist" else {}
    if kind == "words":
        # the 32 streams' words of the PREVIOUS clock leave the shared-memory stage while this clock computes:
        # one drain macro (one stream index j) every 8 steps
        hooks = {8 * j + 4: ["CLT256V2_W_DRAIN(%d)" % j] for j in range(32)}
    tail, tail_ops = tail_lines(shape, kind) if (shape is not None and kind not in ("words", "plain")) else ([], {})
    per_step = -(-len(tail) // TAIL_STEPS) if tail else 0
    heap = BitHeap(em)
    pending = {r: [] for r in range(N)}       # raw planes not yet folded
    acc = {}                                  # row -> partial sum expr
    left = {r: len(rows[r]) for r in range(N)}
    store_at = {}                             # time -> [(slot, expr)]
    old_uses_left = {b: sum(1 for r in range(N) if b in rows[r]) for b in resident}
    assigned = set()

    # load time -> shared-memory slot (resident planes have none; a body with MORE resident planes than the kernel's
    # slot layout -- the plain clocks of the decimated stream -- passes the layout's set as slot_resident)
    slot_of = slot_map(pos, resident if slot_resident is None else slot_resident)
    deferred = []

    def wexpr(t):
        b = perm[t]
        return ("r%d" % b) if b in resident else ("w%d" % t)

    def load(t):
        if t < N and perm[t] not in resident:
            em.emit("const uint32_t w%d = st[%d * CLT256V2_B];" % (t, slot_of[t]))
        for tt, ex in store_at.pop(t, []):
            em.emit("st[%d * CLT256V2_B] = %s;" % (slot_of[tt], ex))

    def used(exprs):
        # the old value of a resident plane is read where the fold is emitted, not when its tap "arrives"
        for ex in exprs:
            if ex[0] == "r":
                old_uses_left[int(ex[1:])] -= 1

    def finish(r, ex, t):
        if r in resident:
            if old_uses_left[r] == 0:
                em.emit("r%d = %s;" % (r, ex))
            else:                             # the old value is still needed: the register changes hands at the end
                em.emit("const uint32_t n%d = %s;" % (r, ex))
                deferred.append(r)
            assigned.add(r)
            return
        slot = pos[r]
        if slot - LEAD <= t:
            em.emit("st[%d * CLT256V2_B] = %s;" % (slot_of[slot], ex))
        else:
            store_at.setdefault(slot, []).append((slot, ex))

    for t in range(min(LEAD, N)):
        load(t)
    for t in range(N):
        load(t + LEAD)
        for ln in tail[:per_step]:
            em.lines.append(ln)
        del tail[:per_step]
        for ln in hooks.get(t, []):
            em.emit(ln)
        w = wexpr(t)
        if kind not in ("words", "plain"):
            heap.add(w, parity(perm[t]) == 1)
            heap.reduce()
        for r in taps_at[t]:
            pending[r].append(w)
            left[r] -= 1
            if r not in acc and len(pending[r]) == 3:
                nm = em.tmp()
                em.emit("const uint32_t %s = %s ^ %s ^ %s;" % (nm, *pending[r]))
                em.ops["rec"] += 1
                used(pending[r])
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
                used(pending[r])
                finish(r, nm, t)
    assert not store_at, store_at
    assert not tail
    assert assigned == set(resident)
    for r in deferred:
        em.emit("r%d = n%d;" % (r, r))
    if kind == "plain":
        # the clock alone (no popcount network, nothing emitted): the clocks between two sampled ones
        return em, perm
    if kind == "words":
        # raw URNG words: the state AFTER the clock, un-sliced 32 planes at a time (resident planes from
        # their registers, the others from their slots, where the clock has just stored them)
        # BEGIN / END let a kernel wrap the eight emits in ONE loop (a switch picks the 32 source planes of word wd,
        # the transpose and the staging code exist once): the straight-line version was 8 x 300 instructions
        em.emit("CLT256V2_W_BEGIN")
        for wd in range(N // 32):
            src = []
            for b in range(32 * wd, 32 * wd + 32):
                src.append(("r%d" % b) if b in resident else "st[%d * CLT256V2_B]" % slot_of[pos[b]])
            em.emit("CLT256V2_W_EMIT(%d, %s)" % (wd, ", ".join(src)))
        em.emit("CLT256V2_W_END")
        return em, perm
    new_shape = [[neg for _, neg in heap.heap[w]] for w in range(10)]
    if shape is None:
        return new_shape
    assert new_shape == shape
    for w in range(10):
        for i, (ex, _) in enumerate(heap.heap[w]):
            em.emit("k%d_%d = %s;" % (w, i, ex))
    for k in ("csa", "ha"):
        em.ops[k] += tail_ops.get(k, 0)
    return em, perm


def main():
    rows = recurrences.packed(N)
    pos = [int(v) for v in open(POS_FILE).read().split()] if POS_FILE else run_annealer(rows)
    peak = int(clt256_sched.live_profile(rows, pos, LEAD).max())
    resident = pick_residents(rows, pos, NRES, RES_BUDGET)
    shape = body_hist_v2(rows, pos, resident)
    em, perm = body_hist_v2(rows, pos, resident, shape)
    init = carried_init(shape)
    tail, _ = tail_lines(shape)
    em8, _ = body_hist_v2(rows, pos, resident, shape, "s8")
    emw, _ = body_hist_v2(rows, pos, resident, shape, "words")
    tail8, _ = tail_lines(shape, "s8")
    emp, _ = body_hist_v2(rows, pos, resident, shape, "plain")
    # the plain clocks of the decimated stream with MORE resident planes (the popcount network's registers are free
    # there): same slot layout; the extra planes are spilled to / filled from their slots around a sampled clock
    res_p = pick_residents(rows, pos, NRES_PLAIN, PLAIN_BUDGET, overlap=True, start=resident)
    emp2, _ = body_hist_v2(rows, pos, res_p, shape, "plain", slot_resident=resident)
    # CLT256V2_OUT / CLT256V2_PREFIX: a second instance with its own settings and macro prefix (the decimated-stream
    # kernel's bodies, clt256_dec_gen.inc, are generated with more resident planes: tools/gen_clt256_dec.sh)
    out = os.environ.get("CLT256V2_OUT") or os.path.join(ROOT, "basebandboard_b200", "csrc", "clt256_v2_gen.inc")
    prefix = os.environ.get("CLT256V2_PREFIX", "CLT256V2")
    import io
    real_out, f = out, io.StringIO()
    if True:
        f.write("// GENERATED by tools/gen_clt256_v2.py -- do not edit.  One-pass LUTOPT(n256) clock + CLTGRNG popcount\n")
        f.write("// network + histogram indices on 32 bit-sliced streams (rng.py:38-40, :96-105; rng_recurrences.py:172-259).\n")
        f.write("// recurrence LOP3: %d, carry-save adders: %d, half adders: %d, modelled peak live values: %d, lead %d, resident %d\n"
                % (em.ops["rec"], em.ops["csa"], em.ops["ha"], peak, LEAD, len(resident)))
        sm = slot_map(pos, resident)
        slot_bit = [perm[t] for t in range(N) if t in sm]
        f.write("#ifdef CLT256V2_SECTION_PERM\n// shared-memory slot s holds state bit CLT256V2_SLOT_BIT[s]; the other planes stay in registers\n")
        f.write("#define CLT256V2_NSLOTS %d\n" % len(slot_bit))
        f.write("static __constant__ uint8_t CLT256V2_SLOT_BIT[%d] = {%s};\n" % (len(slot_bit), ", ".join(str(v) for v in slot_bit)))
        f.write("#endif\n#ifdef CLT256V2_SECTION_RESIDENT_DECL\n")
        for b in sorted(resident):
            f.write("    uint32_t r%d = planes[(size_t)%d * T + t];\n" % (b, b))
        for nm, v in init.items():
            f.write("    uint32_t %s = 0x%08xu;\n" % (nm, v))
        f.write("#endif\n#ifdef CLT256V2_SECTION_BODY\n")
        f.write("\n".join(em.lines))
        f.write("\n#endif\n#ifdef CLT256V2_SECTION_TAIL\n// the last clock's popcount planes -> z0..z7\n")
        f.write("\n".join(tail))
        f.write("\n#endif\n#ifdef CLT256V2_SECTION_S8_BODY\n// same clock, int8 samples instead of histogram indices\n")
        f.write("\n".join(em8.lines))
        f.write("\n#endif\n#ifdef CLT256V2_SECTION_S8_TAIL\n")
        f.write("\n".join(tail8))
        f.write("\n#endif\n#ifdef CLT256V2_SECTION_W_BODY\n// same clock without the popcount network, then the new state as raw URNG words (rng.py:27)\n")
        f.write("\n".join(emw.lines))
        f.write("\n#endif\n#ifdef CLT256V2_SECTION_P_BODY\n// the clock alone: no popcount network, nothing emitted (decimated sample streams)\n")
        f.write("\n".join(emp.lines))
        f.write("\n#endif\n#ifdef CLT256V2_SECTION_P2_DECL\n// %d more resident planes for the plain clocks (P2_BODY)\n" % (len(res_p) - len(resident)))
        extra = sorted(res_p - resident)
        for b in extra:
            f.write("    uint32_t r%d = planes[(size_t)%d * T + t];\n" % (b, b))
        f.write("#endif\n#ifdef CLT256V2_SECTION_P2_SPILL\n")
        for b in extra:
            f.write("    st[%d * CLT256V2_B] = r%d;\n" % (sm[pos[b]], b))
        f.write("#endif\n#ifdef CLT256V2_SECTION_P2_FILL\n")
        for b in extra:
            f.write("    r%d = st[%d * CLT256V2_B];\n" % (b, sm[pos[b]]))
        f.write("#endif\n#ifdef CLT256V2_SECTION_P2_BODY\n// the clock alone with %d resident planes\n" % len(res_p))
        f.write("\n".join(emp2.lines))
        f.write("\n#endif\n")
    with open(real_out, "w") as g:
        g.write(f.getvalue().replace("CLT256V2_", prefix + "_"))
    print("wrote", out, "ops", em.ops, "peak live", peak, "resident", len(resident), file=sys.stderr)


if __name__ == "__main__":
    main()
