"""Print the SASS instructions with the most warp-stall samples from `ncu --page source --csv` output."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 14
print(rows[0][1])
hdr = rows[1]
data = [r for r in rows[2:] if len(r) == len(hdr) and r[0] != "Address"]
si, src = hdr.index("# Samples"), hdr.index("Source")
tot = sum(int(r[si]) for r in data)
print("total samples", tot, "instructions", len(data))
top = sorted(enumerate(data), key=lambda t: -int(t[1][si]))[:n]
for i, r in sorted(top):
    print(f"{i:5d} {int(r[si]):7d} {100*int(r[si])/max(tot,1):5.1f}%  {r[src].strip()[:64]:64s} <- {data[i-1][src].strip()[:44]} <- {data[i-2][src].strip()[:36]}")
