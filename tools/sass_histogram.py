"""SASS opcode histogram of the built library: the mnemonics that prove the Blackwell-native path (the PTX names never
appear in SASS): tcgen05.mma -> UTC*MMA, tcgen05.ld/st -> LDTM/STTM, cp.async.bulk.tensor -> UTMALDG, tcgen05.commit ->
UTCBAR, cp.async -> LDGSTS, mbarrier try_wait -> SYNCS.  Runs anywhere cuobjdump is (no GPU needed).
usage: python tools/sass_histogram.py [path/to/lib.so] > profiles/rNN_sass_histogram.txt"""
import collections
import os
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                         "image-segmentation_b200", "csrc", "libunetb200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
pat = re.compile(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)")
interesting = re.compile(r"^(UTC\w*MMA|UTCBAR|UTCATOM|LDTM|STTM|UTMALDG|UTMAPF|UTMASTG|LDGSTS|SYNCS|UCGABAR|ELECT|FFMA2|FADD2|REDUX|ATOMS|ATOMG|RED)")
total = collections.Counter()
per_kernel = collections.defaultdict(collections.Counter)
kernel = None
n_kernels = 0
for line in sass.splitlines():
    m = re.match(r"\s+Function : (\S+)", line)
    if m:
        kernel = m.group(1)
        n_kernels += 1
        continue
    m = pat.match(line)
    if m and kernel:
        op = m.group(1)
        if interesting.match(op):
            total[op] += 1
            per_kernel[kernel][op.split(".")[0]] += 1
print(f"# cuobjdump -sass {os.path.basename(lib)}: {n_kernels} kernels (sm_100a); counts of the Blackwell-specific and epilogue-relevant opcodes")
for op, n in sorted(total.items(), key=lambda kv: (-kv[1], kv[0])):
    print(f"{n:7d}  {op}")
print("\n# kernels that issue tensor-core MMAs (UTC*MMA), with their TMEM loads, TMA loads and commits")
for k, c in sorted(per_kernel.items()):
    mma = sum(v for o, v in c.items() if o.startswith("UTC") and o.endswith("MMA"))
    if mma:
        name = subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip().split("(")[0]
        print(f"{name:70s} MMA {mma:4d}  LDTM {c.get('LDTM', 0):4d}  UTMALDG {c.get('UTMALDG', 0):3d}  UTCBAR {c.get('UTCBAR', 0):3d}  LDGSTS {c.get('LDGSTS', 0):3d}")
