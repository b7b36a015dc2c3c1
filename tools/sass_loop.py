#!/usr/bin/env python3
"""Prints the hot loop of a kernel (cuobjdump -sass text) as: index, stall count, instruction.
usage: sass_loop.py file.sass [start [count]]"""
import re, sys
lines = open(sys.argv[1]).read().split("\n")
pat = re.compile(r"^\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);\s+/\* (0x[0-9a-f]{16}) \*/")
pat2 = re.compile(r"^\s+/\* (0x[0-9a-f]{16}) \*/")
ins = []
i = 0
while i < len(lines) - 1:
    m = pat.match(lines[i]); m2 = pat2.match(lines[i + 1]) if m else None
    if m and m2:
        hi = int(m2.group(1), 16)
        ins.append((int(m.group(1), 16), m.group(2).strip(), (hi >> 41) & 0xf))
        i += 2
    else:
        i += 1
best = None
for a, t, _ in ins:
    m = re.search(r"BRA.*0x([0-9a-f]+)", t)
    if m:
        tg = int(m.group(1), 16)
        if tg < a and (best is None or a - tg > best[1] - best[0]):
            best = (tg, a)
loop = [x for x in ins if best[0] <= x[0] <= best[1]]
s = int(sys.argv[2]) if len(sys.argv) > 2 else 0
n = int(sys.argv[3]) if len(sys.argv) > 3 else 80
for k, (a, t, st) in enumerate(loop[s:s + n]):
    print("%4d %d %s" % (s + k, st, t))
