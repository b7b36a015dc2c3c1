#!/usr/bin/env python3
"""Host-side check of the generated clock bodies (no GPU): evaluates the P_BODY (plain clock, 104 resident planes) and
the P2_BODY (151 resident planes, with its DECL / SPILL / FILL sections) of clt256_v2_gen.inc as Python on random
32-stream words and requires identical state after several clocks, in slots and registers, and against the
recurrence itself (rng.py:38-40)."""
import os
import random
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from basebandboard_b200 import recurrences  # noqa: E402

src = open(os.path.join(ROOT, "basebandboard_b200", "csrc", "clt256_v2_gen.inc")).read()


def section(name):
    return re.search(r"#ifdef CLT256V2_SECTION_%s\n(.*?)\n#endif" % name, src, re.S).group(1)


def tr(code):
    out = []
    for ln in code.split("\n"):
        ln = ln.strip()
        if not ln or ln.startswith("//"):
            continue
        ln = ln.replace("const uint32_t ", "").replace("uint32_t ", "")
        ln = re.sub(r"st\[(\d+) \* CLT256V2_B\]", r"st[\1]", ln)
        ln = re.sub(r"planes\[\(size_t\)(\d+) \* T \+ t\]", r"planes[\1]", ln)
        ln = re.sub(r"0x([0-9a-f]+)u", r"0x\1", ln)
        out.append(ln.rstrip(";"))
    return "\n".join(out)


def main():
    slot_bit = [int(x) for x in re.search(r"CLT256V2_SLOT_BIT\[\d+\] = \{(.*?)\}", src).group(1).split(",")]
    rows = recurrences.packed(256)
    random.seed(1)
    planes = [random.getrandbits(32) for _ in range(256)]

    def fresh():
        env = {"planes": planes, "st": [planes[b] for b in slot_bit]}
        exec(tr(section("RESIDENT_DECL")), env)
        return env

    def state(env):
        x = [None] * 256
        for s, b in enumerate(slot_bit):
            x[b] = env["st"][s]
        for k, v in env.items():
            if re.fullmatch(r"r\d+", k) and x[int(k[1:])] is None:
                x[int(k[1:])] = v
        return x

    clocks = 5
    want = planes[:]
    for _ in range(clocks):
        nxt = []
        for r in range(256):
            v = 0
            for t in rows[r]:
                v ^= want[t]
            nxt.append(v)
        want = nxt
    a = fresh()
    for _ in range(clocks):
        exec(tr(section("P_BODY")), a)
    assert state(a) == want, "P_BODY differs from the recurrence"
    b = fresh()
    exec(tr(section("P2_DECL")), b)
    for i in range(clocks):
        exec(tr(section("P2_BODY")), b)
        if i == 2:                       # a sampled clock in between: spill, (any body on the slots), fill
            exec(tr(section("P2_SPILL")), b)
            exec(tr(section("P2_FILL")), b)
    exec(tr(section("P2_SPILL")), b)
    assert state(b) == want, "P2_BODY differs from the recurrence"
    print("ok: P_BODY and P2_BODY == %d clocks of the recurrence" % clocks)


if __name__ == "__main__":
    main()
