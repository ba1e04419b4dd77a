#!/usr/bin/env python
"""SASS-level view of an .ncu-rep (no GPU needed): splits the kernel's instruction stream into phases at
barriers and reports, per phase, the stall samples by reason and the instruction mix.
  python tools/ncu_sass.py rep.ncu-rep [--dump]"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
col = {c: i for i, c in enumerate(h)}
stalls = [c for c in h if c.startswith("stall_") and "Not Issued" not in c]
body = [r for r in rows[hi + 1:] if len(r) == len(h)]
tot = sum(float(r[col["# Samples"]]) for r in body) or 1
phases, cur = [], {"n": 0, "samples": 0.0, "ops": collections.Counter(), "st": collections.Counter(), "exec": 0.0, "first": 0}
for k, r in enumerate(body):
    op = r[col["Source"]].split()
    op = op[1] if op[0].startswith("@") else op[0]
    base = op.split(".")[0]
    s = float(r[col["# Samples"]])
    cur["n"] += 1; cur["samples"] += s; cur["ops"][base] += 1; cur["exec"] += float(r[col["Instructions Executed"]])
    for c in stalls:
        cur["st"][c] += float(r[col[c]])
    if "--dump" in sys.argv:
        top = max(stalls, key=lambda c: float(r[col[c]]))
        print("%5d %6.2f%% %-70s %s" % (k, 100 * s / tot, r[col["Source"]].strip()[:70], top if s > 0.001 * tot else ""))
    if base in ("BAR", "WARPSYNC", "BRA", "EXIT"):
        cur["last"] = k; phases.append(cur)
        cur = {"n": 0, "samples": 0.0, "ops": collections.Counter(), "st": collections.Counter(), "exec": 0.0, "first": k + 1}
phases.append(cur)
print("total samples %d, %d SASS instructions" % (tot, len(body)))
for p in phases:
    if p["samples"] < 0.01 * tot:
        continue
    st = ", ".join("%s %.0f%%" % (c[6:], 100 * v / p["samples"]) for c, v in p["st"].most_common(5) if v > 0.03 * p["samples"])
    ops = ", ".join("%s %d" % kv for kv in p["ops"].most_common(7))
    print("[%4d..%4d] %3d instr %5.1f%% of samples | %s | %s" % (p["first"], p.get("last", 0), p["n"], 100 * p["samples"] / tot, st, ops))
