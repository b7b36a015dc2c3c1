#!/usr/bin/env python3
"""Aggregates `ncu --page source --csv --print-source sass` by opcode for the hot loop:
instructions executed, samples, and the not-issued stall reasons.  usage: ncu_stalls.py file.csv [min_exec]"""
import csv
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
body = rows[2:]
execs = [int(r[ix["Instructions Executed"]] or 0) for r in body]
thr = max(execs) // 2 if len(sys.argv) < 3 else int(sys.argv[2])
stall_cols = [h for h in hdr if h.startswith("stall_") and h.endswith("(Not Issued)")]
agg = defaultdict(lambda: defaultdict(int))
tot = defaultdict(int)
n_hot = 0
for r, e in zip(body, execs):
    if e < thr:
        continue
    n_hot += 1
    op = r[ix["Source"]].split()[0]
    if op.startswith("@"):
        op = r[ix["Source"]].split()[1]
    op = op.split(".")[0]
    agg[op]["n"] += 1
    agg[op]["samples"] += int(r[ix["# Samples"]] or 0)
    tot["samples"] += int(r[ix["# Samples"]] or 0)
    for c in stall_cols:
        v = int(r[ix[c]] or 0)
        agg[op][c] += v
        tot[c] += v
print("hot instructions:", n_hot, "total samples:", tot["samples"])
print("not-issued stall totals:", {c[6:-13]: v for c, v in sorted(tot.items(), key=lambda kv: -kv[1]) if c != "samples" and v})
for op, d in sorted(agg.items(), key=lambda kv: -kv[1]["samples"]):
    top = sorted(((c[6:-13], v) for c, v in d.items() if c.startswith("stall_") and v), key=lambda kv: -kv[1])[:4]
    print("%-10s n=%4d samples=%7d (%.2f/inst) %s" % (op, d["n"], d["samples"], d["samples"] / d["n"], top))
