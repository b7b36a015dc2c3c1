#!/usr/bin/env python3
"""Static issue-cycle estimate of a kernel's hot loop from `cuobjdump -sass` (sm_100 control bits:
stall count = bits 41..44 of the second encoding word).  Prints the sum of stall counts between the
loop head and the backward branch, per opcode, and the most common (opcode -> next opcode, stall) pairs.
usage: sass_stalls.py file.sass"""
import re
import sys
from collections import Counter, defaultdict

ins = []   # (addr, text, stall)
lines = open(sys.argv[1]).read().split("\n")
i = 0
pat = re.compile(r"^\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);\s+/\* (0x[0-9a-f]{16}) \*/")
pat2 = re.compile(r"^\s+/\* (0x[0-9a-f]{16}) \*/")
while i < len(lines):
    m = pat.match(lines[i])
    if m and i + 1 < len(lines):
        m2 = pat2.match(lines[i + 1])
        if m2:
            hi = int(m2.group(1), 16)
            ins.append((int(m.group(1), 16), m.group(2).strip(), (hi >> 41) & 0xf, (hi >> 45) & 1))
            i += 2
            continue
    i += 1
# the hot loop = the longest backward branch span
best = None
for k, (a, t, s, y) in enumerate(ins):
    m = re.search(r"BRA\S*\s+(?:\S+\s+)?`?\(?\.?L?_?x?_?(\w+)\)?", t)
    m = re.search(r"BRA.*0x([0-9a-f]+)", t)
    if m:
        tgt = int(m.group(1), 16)
        if tgt < a and (best is None or a - tgt > best[1] - best[0]):
            best = (tgt, a)
lo, hi = best
loop = [x for x in ins if lo <= x[0] <= hi]
print("loop: %d instructions, sum of stall counts = %d cycles" % (len(loop), sum(x[2] for x in loop)))
byop = defaultdict(lambda: [0, 0])
for a, t, s, y in loop:
    op = t.split()[1].split(".")[0] if t.startswith("@") else t.split()[0].split(".")[0]
    byop[op][0] += 1
    byop[op][1] += s
for op, (n, s) in sorted(byop.items(), key=lambda kv: -kv[1][1]):
    print("  %-8s n=%4d stall sum=%5d avg %.2f" % (op, n, s, s / n))
ALU = {"LOP3", "SHF", "PRMT", "VIADD", "IADD3", "LEA", "ISETP", "SEL", "MOV"}
alu_run = Counter()
prev_alu = False
for a, t, s, y in loop:
    op = t.split()[1].split(".")[0] if t.startswith("@") else t.split()[0].split(".")[0]
    alu_run[(op in ALU, s)] += 1
print("  (is_alu, stall) histogram:", sorted(alu_run.items()))
