"""Summarises an ncu launch list (--metrics gpu__time_duration.sum --csv) of tools/time_clip_vit.py <n> ours: per kernel
class, launches and time of the LAST pass.  Usage: python tools/vit_launch_summary.py gpurun_out/xx_vit_launches.csv"""
import collections
import csv
import sys


def main(path):
    with open(path) as f:
        rows = list(csv.DictReader(l for l in f if not l.startswith("==")))
    names = [r["Kernel Name"] for r in rows]
    first = max(i for i, n in enumerate(names) if "clip_patch" in n)
    agg, total = collections.OrderedDict(), 0.0
    for r in rows[first:]:
        t = float(r["Metric Value"]) / 1e3
        a = agg.setdefault(r["Kernel Name"][:64], [0, 0.0])
        a[0] += 1
        a[1] += t
        total += t
    for n, (c, t) in agg.items():
        print(f"{n:66s} x{c:3d} {t:9.1f} us  ({t / c:6.1f} each)")
    print(f"total {total:.1f} us")


if __name__ == "__main__":
    main(sys.argv[1])
