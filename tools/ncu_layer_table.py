"""Per-launch table from `ncu -i X.ncu-rep --page raw --csv` output: duration, achieved TFLOP/s, tensor-pipe
activity, DRAM bytes, L2 / L1TEX throughput.  usage: ncu_layer_table.py raw.csv layer,layer,..."""
import csv
import sys

GMAC = {'enc1.0': 1.208, 'enc1.3': 2.416, 'enc2.0': 1.208, 'enc2.3': 2.416, 'enc3.0': 1.208, 'enc3.3': 2.416,
        'bott.0': 1.208, 'bott.3': 2.416, 'dec1.up': 0.537, 'dec1.0': 1.208, 'dec1.3': 0.604, 'dec2.up': 0.134,
        'dec2.0': 1.208, 'dec2.3': 0.604, 'dec3.up': 0.134, 'dec3.0': 1.208, 'dec3.3': 0.604, 'dec4.up': 0.134,
        'dec4.0': 1.208, 'dec4.3': 0.604}
ORDER = ['enc1.0', 'enc1.3', 'enc2.0', 'enc2.3', 'enc3.0', 'enc3.3', 'bott.0', 'bott.3', 'dec1.up', 'dec1.0', 'dec1.3',
         'dec2.up', 'dec2.0', 'dec2.3', 'dec3.up', 'dec3.0', 'dec3.3', 'dec4.up', 'dec4.0', 'dec4.3']

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
first = sys.argv[2] if len(sys.argv) > 2 else 'enc1.0'
mb = int(sys.argv[3]) if len(sys.argv) > 3 else 32
layers = [ORDER[(ORDER.index(first) + i) % len(ORDER)] for i in range(len(data))]


def ci(s):
    if s in hdr:
        return hdr.index(s)
    return [i for i, h in enumerate(hdr) if h.endswith('.' + s)][0]


T, HM, CY = ci('gpu__time_duration.sum'), ci('sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg'), ci('sm__cycles_elapsed.max')
RD, WR = ci('dram__bytes_read.sum'), ci('dram__bytes_write.sum')
LT, L1 = ci('lts__throughput.avg.pct_of_peak_sustained_elapsed'), ci('l1tex__throughput.avg.pct_of_peak_sustained_elapsed')
GH = ci('sm__cycles_elapsed.max.per_second')
KN = ci('Kernel Name')
scale = {'Mbyte': 1.0, 'Kbyte': 1e-3, 'Gbyte': 1e3, 'byte': 1e-6}
print(f"{'layer':8s} {'kernel':22s} {'t_us':>7s} {'TF/s':>6s} {'tensor%':>7s} {'rdMB':>7s} {'wrMB':>7s} {'GB/s':>6s} {'lts%':>5s} {'l1%':>5s} GHz")
for r, l in zip(data, layers):
    t = float(r[T]) * {'us': 1.0, 'ns': 1e-3, 'ms': 1e3}[units[T]]
    tf = GMAC[l] * 2 * mb / t * 1e3
    try:
        tp = float(r[HM]) / 4 / float(r[CY]) * 100
    except ValueError:
        tp = float('nan')
    try:
        rd, wr = float(r[RD]) * scale[units[RD]], float(r[WR]) * scale[units[WR]]
    except ValueError:
        rd = wr = float('nan')
    k = r[KN].split('conv_tc_kernel')[-1].split('(')[0]
    print(f"{l:8s} {k:22s} {t:7.1f} {tf:6.0f} {tp:7.1f} {rd:7.1f} {wr:7.1f} {(rd + wr) / t * 1e3:6.0f} {r[LT][:5]:>5s} {r[L1][:5]:>5s} {r[GH][:4]}")
