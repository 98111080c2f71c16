"""Condensed SASS listing with stall samples for one kernel instance of `ncu --page source --csv` output.
Shows sync / TMA / MMA / TMEM / branch instructions plus anything above a sample threshold, so the warp
roles (producer, MMA issuer, epilogue) and where each one waits can be read off.
usage: ncu_sass_roles.py src.csv <instance-index> [min_pct]"""
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
inst = int(sys.argv[2])
min_pct = float(sys.argv[3]) if len(sys.argv) > 3 else 0.4
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
starts = starts[::2] if len(starts) > 1 and rows[starts[0]][1] == rows[starts[1]][1] else starts
s = starts[inst]
e = starts[inst + 1] if inst + 1 < len(starts) else len(rows)
hdr = rows[s + 1]
body = [r for r in rows[s + 2:e] if len(r) == len(hdr) and r[0] != "Address"]
# the page lists each kernel twice: keep the first copy
half = len(body) // 2
if half and [r[1] for r in body[:half]] == [r[1] for r in body[half:]]:
    body = body[:half]
SI, SRC, EX = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
reasons = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
RI = [hdr.index(h) for h in reasons]
tot = sum(int(r[SI]) for r in body) or 1
print(rows[s][1], "samples", tot, "instructions", len(body))
key = re.compile(r"SYNCS|UTMA|UTC|LDTM|STTM|BAR|EXIT|BRA|ELECT|UBLKCP|WARPSYNC|REDG|ATOM|RED\.|CALL|RET")
for i, r in enumerate(body):
    n = int(r[SI])
    pct = 100.0 * n / tot
    if pct >= min_pct or key.search(r[SRC]):
        top = max(zip([int(r[j] or 0) for j in RI], reasons))
        print(f"{i:5d} {n:6d} {pct:5.1f}% ex={r[EX]:>8s} {r[SRC].strip()[:86]:86s} {top[1] if n else ''}")
