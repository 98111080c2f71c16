"""Where do the warp roles of conv_tc_kernel wait?  Aggregates ncu stall samples of the mbarrier try_wait
loops by barrier class (decoded from the shared-memory offset of the SYNCS.PHASECHK operand).
usage: ncu_barrier_waits.py src.csv [n_warps=12]"""
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
nw = int(sys.argv[2]) if len(sys.argv) > 2 else 12
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
BASE = int(sys.argv[3], 16) if len(sys.argv) > 3 else 0x1210
CLASSES = [("a_full", 0, 8), ("a_empty", 8, 16), ("b_full", 16, 24), ("b_empty", 24, 32), ("t_full", 32, 36),
           ("t_empty", 36, 40), ("w_full", 40, 41)]  # ub_conv_tc.cuh: bars[] order (kMaxSA = kMaxSB = 8, kMaxAcc = 4)


def classify(off):
    i = (off - BASE) // 8
    for n, lo, hi in CLASSES:
        if lo <= i < hi:
            return n
    return f"?{off:#x}"


seen = 0
for k, s in enumerate(starts):
    if k % 2:  # each kernel is listed twice
        continue
    e = starts[k + 1] if k + 1 < len(starts) else len(rows)
    hdr = rows[s + 1]
    body = [r for r in rows[s + 2:e] if len(r) == len(hdr) and r[0] != "Address"]
    SI, SRC = hdr.index("# Samples"), hdr.index("Source")
    tot = sum(int(r[SI]) for r in body) or 1
    agg = {}
    cur = None
    for i, r in enumerate(body):
        m = re.search(r"SYNCS\.PHASECHK\.TRANS64\.TRYWAIT\s+\w+, \[(\w+)(?:\+(\w+))?(?:\+(0x[0-9a-f]+))?\]", r[SRC])
        if m:
            offs = [g for g in m.groups() if g and g.startswith("0x")]
            cur = classify(int(offs[-1], 16)) if offs else "dyn"
            agg[cur] = agg.get(cur, 0) + int(r[SI])
            # the wait loop = this instruction and the following few (branch back)
            for j in range(1, 12):
                if i + j < len(body):
                    s2 = body[i + j][SRC]
                    if "SYNCS.PHASECHK" in s2 or "UTMALDG" in s2 or "UTCHMMA" in s2 or "LDTM" in s2:
                        break
                    if "BRA" in s2:
                        agg[cur] += int(body[i + j][SI])
                        break
                    agg[cur] += int(body[i + j][SI])
    name = rows[s][1].split("conv_tc_kernel")[-1].split("(ub")[0]
    per_warp = tot / nw
    txt = "  ".join(f"{n}={v / per_warp:4.2f}" for n, v in sorted(agg.items()))
    ex = sum(int(r[SI]) for r in body if "EXIT" in r[SRC] or "BAR.SYNC" in r[SRC])
    print(f"[{k // 2:2d}] {name:40s} samples {tot:6d}  (in warp-equivalents) {txt}  exit/bar={ex / per_warp:4.2f}")
