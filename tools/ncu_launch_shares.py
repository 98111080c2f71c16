"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel: launches, total time, share.
usage: ncu_launch_shares.py launches.csv "<description of the profiled command>" """
import csv
import sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
sc = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}
agg = defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    name = r[kn].split("(")[0].replace("ub::", "")
    a = agg[name]
    a[0] += 1
    a[1] += float(r[mv].replace(",", "")) * sc.get(r[mu], 1.0)
tot = sum(a[1] for a in agg.values())
print(f"kernel launches total_us share   ({sys.argv[2] if len(sys.argv) > 2 else ''})")
for name, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{name} {n} {t:.1f} {t / tot:.3f}")
