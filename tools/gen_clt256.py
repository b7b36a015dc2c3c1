#!/usr/bin/env python3
"""Generates basebandboard_b200/csrc/clt256_gen.inc: the straight-line body of one LUTOPT(n256)
clock + CLTGRNG popcount tree in bit-sliced form (32 streams per thread), specialised to the
shipped n256 recurrence (reference gateware/bbb/rng_recurrences.py:172-259).

Scheme (DESIGN.md "clt256 kernel"):
  * a thread keeps its 256 state planes in shared memory as 64 uint4 groups, in a storage order
    `perm` chosen here (storage position p holds state bit perm[p]);
  * a clock is two passes over the old state in storage order; pass 1 accumulates the new planes
    stored at positions 0..127 in registers and feeds every loaded plane to the carry-save
    popcount network, pass 2 accumulates the new planes of positions 128..255;
  * loaded planes stay live for WINDOW groups so that two or three taps of one row that are close
    in storage order are folded into one LOP3; `perm` is optimised (annealing, fixed seed) to make
    that happen often.
Run:  python tools/gen_clt256.py            (rewrites the .inc; deterministic)
"""
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from basebandboard_b200 import recurrences  # noqa: E402

N = 256
HALF = 128
WINDOW = 3          # groups of 4 planes kept live


def parity(i):
    return bin(i).count("1") & 1


def row_cost(ts, W=WINDOW):
    """LOP3/MOV count to accumulate a row whose taps arrive at sorted storage positions ts.
    A tap loaded in group g is live until group g+W-1 has been processed.  A row acts when its
    oldest pending tap is about to die: one instruction folds the oldest pending taps, up to 3
    for the first instruction of the row and 2 afterwards (accumulator + 2 = one LOP3)."""
    groups = [t // 4 for t in ts]
    ops = 0
    i = 0
    n = len(ts)
    while i < n:
        dead = groups[i] + W - 1          # last group during which tap i is live
        j = i
        while j < n and groups[j] <= dead:
            j += 1
        take = min(j - i, 2 if ops else 3)
        ops += 1
        i += take
    return ops


def total_cost(rows, pos, W=WINDOW):
    """rows: list of tap lists (state-bit indices); pos[bit] = storage position."""
    c = 0
    for r in range(N):
        ts = sorted(pos[t] for t in rows[r])
        c += row_cost(ts, W)
    return c


def optimise_perm(rows, iters, seed=12345, W=WINDOW, verbose=False):
    """Anneal pos[] (bit -> storage position).  The half a row is stored in does not change its
    cost, only the storage order of its taps does."""
    rnd = random.Random(seed)
    pos = list(range(N))
    uses = [[] for _ in range(N)]           # bit -> rows that tap it
    for r in range(N):
        for t in rows[r]:
            uses[t].append(r)
    rc = [row_cost(sorted(pos[t] for t in rows[r]), W) for r in range(N)]
    cur = sum(rc)
    best, best_pos = cur, pos[:]
    T0, T1 = 1.5, 0.05
    for it in range(iters):
        T = T0 * (T1 / T0) ** (it / iters)
        a, b = rnd.randrange(N), rnd.randrange(N)
        if a == b:
            continue
        aff = set(uses[a]) | set(uses[b])
        pos[a], pos[b] = pos[b], pos[a]
        new = {r: row_cost(sorted(pos[t] for t in rows[r]), W) for r in aff}
        d = sum(new.values()) - sum(rc[r] for r in aff)
        if d <= 0 or rnd.random() < pow(2.718281828, -d / T):
            for r, v in new.items():
                rc[r] = v
            cur += d
            if cur < best:
                best, best_pos = cur, pos[:]
        else:
            pos[a], pos[b] = pos[b], pos[a]
        if verbose and it % 50000 == 0:
            print("it", it, "cur", cur, "best", best, file=sys.stderr)
    return best_pos, best


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "analyse":
    rows = recurrences.packed(N)
    ident = list(range(N))
    for W in (1, 2, 3, 4, 6):
        print("W", W, "identity cost", total_cost(rows, ident, W))
    for W in (2, 3, 4):
        p, c = optimise_perm(rows, 300000, W=W, verbose=False)
        print("W", W, "optimised cost", c)
    sys.exit(0)


# ------------------------------------------------------------------------------------------------
# code emission
# ------------------------------------------------------------------------------------------------
XYZW = "xyzw"


class Emitter:
    def __init__(self):
        self.lines = []
        self.ntmp = 0
        self.ops = {"rec": 0, "csa": 0, "ha": 0}

    def emit(self, s):
        self.lines.append("    " + s)

    def tmp(self):
        self.ntmp += 1
        return "t%d" % self.ntmp


class BitHeap:
    """Carry-save popcount network: heap[w] = planes of weight 2^w, each (expr, negated)."""

    def __init__(self, em):
        self.em = em
        self.heap = [[] for _ in range(10)]

    @staticmethod
    def _x(p):
        return ("~" + p[0]) if p[1] else p[0]

    def add(self, expr, neg):
        self.heap[0].append((expr, neg))

    def csa(self, w):
        a, b, c = self.heap[w][:3]
        del self.heap[w][:3]
        s, k = self.em.tmp(), self.em.tmp()
        xa, xb, xc = self._x(a), self._x(b), self._x(c)
        self.em.emit("const uint32_t %s = %s ^ %s ^ %s;" % (s, xa, xb, xc))
        self.em.emit("const uint32_t %s = (%s & %s) | (%s & %s) | (%s & %s);" % (k, xa, xb, xa, xc, xb, xc))
        self.em.ops["csa"] += 1
        self.heap[w].append((s, False))
        self.heap[w + 1].append((k, False))

    def ha(self, w):
        a, b = self.heap[w][:2]
        del self.heap[w][:2]
        s, k = self.em.tmp(), self.em.tmp()
        xa, xb = self._x(a), self._x(b)
        self.em.emit("const uint32_t %s = %s ^ %s;" % (s, xa, xb))
        self.em.emit("const uint32_t %s = %s & %s;" % (k, xa, xb))
        self.em.ops["ha"] += 1
        self.heap[w].append((s, False))
        self.heap[w + 1].append((k, False))

    def reduce(self):
        again = True
        while again:
            again = False
            for w in range(9):
                while len(self.heap[w]) >= 3:
                    self.csa(w)
                    again = True

    def finish(self):
        for w in range(9):
            while len(self.heap[w]) >= 2:
                if len(self.heap[w]) >= 3:
                    self.csa(w)
                else:
                    self.ha(w)
        out = []
        for w in range(9):
            if self.heap[w]:
                out.append(self._x(self.heap[w][0]))
            else:
                out.append("0u")
        assert not self.heap[9]
        return out


def gen_pass(em, rows, pos, perm, which, with_clt, hooks, W=WINDOW, prefetch=2):
    """One pass over the old state.  which = 0: rows stored at positions 0..127, 1: 128..255.
    hooks: dict slot -> list of C lines emitted at the start of that group slot (histogram
    read-modify-writes interleaved with the pass)."""
    tag = "p%d" % which
    mine = [r for r in range(N) if (pos[r] >= HALF) == bool(which)]
    taps_at = [[] for _ in range(N)]      # storage position -> rows of this pass tapping it
    for r in mine:
        for t in rows[r]:
            taps_at[pos[t]].append(r)
    pending = {r: [] for r in mine}      # row -> [(group, expr)]
    started = set()
    heap = BitHeap(em) if with_clt else None

    def vname(g):
        return "v%s_%d" % (tag, g)

    def load(g):
        src = "lo_old[%d * CLT256_B]" % g if g < 32 else "hi[%d * CLT256_B]" % (g - 32)
        em.emit("const uint4 %s = %s;" % (vname(g), src))

    def act(r):
        p = pending[r]
        take = min(len(p), 2 if r in started else 3)
        ex = " ^ ".join(e for _, e in p[:take])
        a = "a%d" % (pos[r] - (HALF if which else 0))
        if r in started:
            em.emit("%s ^= %s;" % (a, ex))
        else:
            em.emit("%s = %s;" % (a, ex))
            started.add(r)
        em.ops["rec"] += 1
        del p[:take]

    for g in range(min(prefetch, 64)):
        load(g)
    for g in range(64):
        for ln in hooks.get(g, []):
            em.emit(ln)
        if g + prefetch < 64:
            load(g + prefetch)
        for k in range(4):
            c = 4 * g + k
            ex = "%s.%s" % (vname(g), XYZW[k])
            for r in taps_at[c]:
                pending[r].append((g, ex))
            if with_clt:
                heap.add(ex, parity(perm[c]) == 1)
        if with_clt:
            heap.reduce()
        dying = g - W + 1
        if dying >= 0:
            for r in mine:
                while pending[r] and pending[r][0][0] <= dying:
                    act(r)
    for r in mine:
        while pending[r]:
            act(r)
    assert started == set(mine)
    planes = heap.finish() if with_clt else None
    return planes


def rmw_hooks(words, first_slot=0):
    """16 read-modify-writes of the private histogram, 4 per z word, spread over the 64 slots."""
    hooks = {}
    k = 0
    for zi in words:
        for m in range(4):
            sh = 8 * m - 7
            src = "(z%d << 7)" % zi if sh < 0 else "(z%d >> %d)" % (zi, sh)
            nm = "z%d_%d" % (zi, m)
            hooks.setdefault(4 * k, []).append(
                "const uint32_t o%s = %s & 0x1f80u; const uint32_t h%s = *(const uint32_t *)(hpb + o%s);" % (nm, src, nm, nm))
            hooks.setdefault(4 * k + 3, []).append("*(uint32_t *)(hpb + o%s) = h%s + 1u;" % (nm, nm))
            k += 1
    assert k == 16
    return hooks


def gen_transpose(em, q, out="z%d", decl="", shr=None, bsel=None):
    """8x8 bit-block transpose of up to 8 planes q[0..] (missing ones are 0): on return byte m of
    word s holds the value (bit b = plane b) of stream 8m+s.  Writes out % s for s = 0..7."""
    x = list(q) + ["0u"] * (8 - len(q))
    for sh, mask in ((4, "0x0f0f0f0fu"), (2, "0x33333333u"), (1, "0x55555555u")):
        y = [None] * 8
        for i in range(8):
            if i & sh:
                continue
            lo, hi = x[i], x[i + sh]
            nl, nh = em.tmp(), em.tmp()
            sh_lo = (shr % (lo, sh)) if shr else "(%s >> %d)" % (lo, sh)
            if bsel and hi != "0u":
                # (a & m) | (b & ~m) as one LOP3 with the mask as its immediate (nvcc emits two for the C form)
                em.emit("const uint32_t %s = %s;" % (nl, bsel % (mask, lo, "(%s << %d)" % (hi, sh))))
                em.emit("const uint32_t %s = %s;" % (nh, bsel % (mask, sh_lo, hi)))
            elif bsel:
                em.emit("const uint32_t %s = %s & %s;" % (nl, lo, mask))
                em.emit("const uint32_t %s = %s & %s;" % (nh, sh_lo, mask))
            else:
                em.emit("const uint32_t %s = (%s & %s) | ((%s & %s) << %d);" % (nl, lo, mask, hi, mask, sh))
                em.emit("const uint32_t %s = (%s & %s) | (%s & ~%s);" % (nh, sh_lo, mask, hi, mask))
            y[i], y[i + sh] = nl, nh
        x = y
    for s in range(8):
        em.emit("%s%s = %s;" % (decl, out % s, x[s]))


def optimise_perm_words(rows, iters, seed=777, W=WINDOW):
    """Like optimise_perm, with the constraint the raw-word stream kernel needs: state bits 0..127
    (output words 0..3) are stored at positions 0..127, bits 128..255 at 128..255, so that all 32
    planes of an output word are accumulators of the same pass.  Swaps stay inside a half."""
    rnd = random.Random(seed)
    pos = list(range(N))
    uses = [[] for _ in range(N)]
    for r in range(N):
        for t in rows[r]:
            uses[t].append(r)
    rc = [row_cost(sorted(pos[t] for t in rows[r]), W) for r in range(N)]
    cur = sum(rc)
    best, best_pos = cur, pos[:]
    T0, T1 = 1.5, 0.05
    for it in range(iters):
        T = T0 * (T1 / T0) ** (it / iters)
        a = rnd.randrange(N)
        b = rnd.randrange(HALF) + (HALF if a >= HALF else 0)
        if a == b:
            continue
        aff = set(uses[a]) | set(uses[b])
        pos[a], pos[b] = pos[b], pos[a]
        new = {r: row_cost(sorted(pos[t] for t in rows[r]), W) for r in aff}
        d = sum(new.values()) - sum(rc[r] for r in aff)
        if d <= 0 or rnd.random() < pow(2.718281828, -d / T):
            for r, v in new.items():
                rc[r] = v
            cur += d
            if cur < best:
                best, best_pos = cur, pos[:]
        else:
            pos[a], pos[b] = pos[b], pos[a]
    return best_pos, best


def emit_store(em, dst):
    for g in range(32):
        em.emit("%s[%d * CLT256_B] = make_uint4(a%d, a%d, a%d, a%d);" % (dst, g, 4 * g, 4 * g + 1, 4 * g + 2, 4 * g + 3))


def body_hist(rows, pos, perm):
    em = Emitter()
    em.emit("uint32_t " + ", ".join("a%d" % i for i in range(HALF)) + ";")
    em.emit("// ---- pass 1: new planes of storage positions 0..127 + popcount network ----")
    planes = gen_pass(em, rows, pos, perm, 0, True, rmw_hooks([4, 5, 6, 7]))
    emit_store(em, "lo_new")
    em.emit("// ---- popcount planes -> histogram index planes ----")
    for w in range(9):
        em.emit("const uint32_t pc%d = %s;" % (w, planes[w]))
    em.emit("const uint32_t inr = ~pc8 & ((~pc7 & pc6 & pc5) | (pc7 & ~pc6 & ~pc5));   // 96 <= popcount < 160")
    em.emit("CLT256_AFTER_COUNT(pc0, pc1, pc2, pc3, pc4, pc5, pc6, pc7, pc8, inr)")
    q = []
    for b, src in enumerate(["pc0", "pc1", "pc2", "pc3", "pc4", "pc7"]):
        nm = "q%d" % b
        em.emit("const uint32_t %s = %s & inr;" % (nm, src))
        q.append(nm)
    gen_transpose(em, q)
    em.emit("// ---- pass 2: new planes of storage positions 128..255 ----")
    gen_pass(em, rows, pos, perm, 1, False, rmw_hooks([0, 1, 2, 3]))
    emit_store(em, "hi")
    return em


def body_s8(rows, pos, perm):
    """int8 sample stream: the low 8 popcount planes with bit 7 flipped are g = popcount - 128 in
    two's complement (gateware width, rng.py:78); un-sliced to one byte per stream."""
    em = Emitter()
    em.emit("uint32_t " + ", ".join("a%d" % i for i in range(HALF)) + ";")
    planes = gen_pass(em, rows, pos, perm, 0, True, {})
    emit_store(em, "lo_new")
    for w in range(9):
        em.emit("const uint32_t pc%d = %s;" % (w, planes[w]))
    em.emit("const uint32_t pc7n = ~pc7;")
    gen_transpose(em, ["pc0", "pc1", "pc2", "pc3", "pc4", "pc5", "pc6", "pc7n"], out="zs%d", decl="const uint32_t ")
    em.emit("CLT256_S8_STAGE(zs0, zs1, zs2, zs3, zs4, zs5, zs6, zs7, pc8)")
    gen_pass(em, rows, pos, perm, 1, False, {})
    emit_store(em, "hi")
    return em


def body_words(rows, pos, perm):
    """raw URNG words: after each pass the 128 new planes (state bits 0..127, then 128..255) are stored
    to shared memory and handed, in state-bit order, to CLT256_W_EMIT for the 32x32 transposes."""
    em = Emitter()
    em.emit("uint32_t " + ", ".join("a%d" % i for i in range(HALF)) + ";")
    for which, dst in ((0, "lo_new"), (1, "hi")):
        gen_pass(em, rows, pos, perm, which, False, {})
        emit_store(em, dst)
        for w in range(4):
            bits = range(128 * which + 32 * w, 128 * which + 32 * w + 32)
            regs = ", ".join("a%d" % (pos[b] - 128 * which) for b in bits)
            em.emit("CLT256_W_EMIT(%d, %s)" % (4 * which + w, regs))
        em.emit("CLT256_W_STORE(%d)" % which)
    return em


def write_inc(name, title, perm, em, cost, prefix):
    out = os.path.join(ROOT, "basebandboard_b200", "csrc", name)
    with open(out, "w") as f:
        f.write("// GENERATED by tools/gen_clt256.py -- do not edit.  %s\n" % title)
        f.write("// on 32 bit-sliced streams (rng.py:38-40, :96-105; rng_recurrences.py:172-259).\n")
        f.write("// recurrence instructions: %d (cost model %d), carry-save adders: %d, half adders: %d, window %d\n"
                % (em.ops["rec"], cost, em.ops["csa"], em.ops["ha"], WINDOW))
        if perm is not None:
            f.write("#ifdef %s_SECTION_PERM\n// storage position p holds state bit %s_PERM[p]\n" % (prefix, prefix))
            f.write("static __constant__ uint8_t %s_PERM[256] = {%s};\n" % (prefix, ", ".join(str(v) for v in perm)))
            f.write("#endif\n")
        f.write("#ifdef %s_SECTION_BODY\n" % prefix)
        f.write("\n".join(em.lines))
        f.write("\n#endif\n")
    print("wrote", out, "ops", em.ops, "cost model", cost, file=sys.stderr)


def main():
    rows = recurrences.packed(N)
    iters = int(os.environ.get("CLT256_ANNEAL_ITERS", "1500000"))
    pos, cost = optimise_perm(rows, iters, W=WINDOW, verbose=True)
    perm = [0] * N
    for b, p in enumerate(pos):
        perm[p] = b
    write_inc("clt256_gen.inc", "One LUTOPT(n256) clock + CLTGRNG popcount network + histogram indices",
              perm, body_hist(rows, pos, perm), cost, "CLT256")
    write_inc("clt256_s8_gen.inc", "One LUTOPT(n256) clock + CLTGRNG popcount network -> int8 samples (same storage order as clt256_gen.inc)",
              None, body_s8(rows, pos, perm), cost, "CLT256S8")
    wpos, wcost = optimise_perm_words(rows, iters, W=WINDOW)
    wperm = [0] * N
    for b, p in enumerate(wpos):
        wperm[p] = b
    write_inc("clt256_words_gen.inc", "One LUTOPT(n256) clock -> raw URNG words",
              wperm, body_words(rows, wpos, wperm), wcost, "CLT256W")


if __name__ == "__main__":
    main()
