"""Per-launch table from an `ncu --page raw --csv` export of one micro-batch (all kernels of the plan).
usage: ncu_table2.py raw.csv first_layer mb > table.txt"""
import csv, sys
ORDER = ['stem', 'enc1.0', 'enc1.3', 'enc2.0', 'enc2.3', 'enc3.0', 'enc3.3', 'bott.0', 'bott.3', 'dec1.up', 'resample', 'dec1.0',
         'dec1.3', 'dec2.up', 'dec2.0', 'dec2.3', 'dec3.up', 'dec3.0', 'dec3.3', 'dec4.up', 'dec4.0', 'dec4.3']
GM = {'enc1.0': 1.208, 'enc1.3': 2.416, 'enc2.0': 1.208, 'enc2.3': 2.416, 'enc3.0': 1.208, 'enc3.3': 2.416, 'bott.0': 1.208,
      'bott.3': 2.416, 'dec1.up': 0.537, 'dec1.0': 1.208, 'dec1.3': 0.604, 'dec2.up': 0.134, 'dec2.0': 1.208, 'dec2.3': 0.604,
      'dec3.up': 0.134, 'dec3.0': 1.208, 'dec3.3': 0.604, 'dec4.up': 0.134, 'dec4.0': 1.208, 'dec4.3': 0.604}
rows = list(csv.reader(open(sys.argv[1])))
hdr, u = rows[0], rows[1]
first, mb = sys.argv[2], int(sys.argv[3])
names = [ORDER[(ORDER.index(first) + i) % len(ORDER)] for i in range(len(rows) - 2)]
ci = hdr.index
T, CY, RD, WR = ci('gpu__time_duration.sum'), ci('sm__cycles_elapsed.max'), ci('dram__bytes_read.sum'), ci('dram__bytes_write.sum')
HM = [i for i, h in enumerate(hdr) if h.endswith('sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg')][0]
LT, GH = ci('lts__throughput.avg.pct_of_peak_sustained_elapsed'), ci('sm__cycles_elapsed.max.per_second')
sc = {'Mbyte': 1, 'Kbyte': 1e-3, 'Gbyte': 1e3, 'byte': 1e-6}
print(f"# ncu --set full, one micro-batch of {mb} images (UB_NO_PDL=1, cold cache, serialised); tensor% = hmma subpipe cycles / elapsed")
print(f"{'layer':9s} {'kernel <N,BK,KS,WS,TW,CG2>':30s} {'t_us':>7s} {'TF/s':>6s} {'tensor%':>7s} {'rdMB':>7s} {'wrMB':>7s} {'GB/s':>6s} {'lts%':>5s} GHz")
tot = 0.0
for i, (r, n) in enumerate(zip(rows[2:], names)):
    t = float(r[T]); tot += t if i < len(ORDER) else 0.0   # one pass over the plan; later rows repeat layers
    k = r[ci('Kernel Name')]
    k = k.split('conv_tc_kernel')[-1].split('(')[0] if 'conv_tc' in k else k.split('(')[0]
    try: tp = float(r[HM]) / 4 / float(r[CY]) * 100
    except ValueError: tp = float('nan')
    try: rd, wr = float(r[RD]) * sc[u[RD]], float(r[WR]) * sc[u[WR]]
    except ValueError: rd = wr = float('nan')
    print(f"{n:9s} {k:30s} {t:7.1f} {GM.get(n, 0) * 2 * mb / t * 1e3:6.0f} {tp:7.1f} {rd:7.1f} {wr:7.1f} {(rd + wr) / t * 1e3:6.0f} {r[LT][:5]:>5s} {r[GH][:4]}")
print(f"total {tot:.1f} us per {mb} images -> {mb / tot * 1e6:.0f} img/s serialised")
