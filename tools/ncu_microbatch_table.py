"""Per-launch table of one micro-batch of the default UNet path from `ncu -i X.ncu-rep --page raw --csv` (captured with
tools/run_forward_once.py): duration, TFLOP/s, tensor-pipe activity, DRAM bytes; writes the per-image DRAM traffic JSON that
bench.py quotes as roofline.traffic.  usage: ncu_microbatch_table.py raw.csv n_images [traffic.json]"""
import csv
import json
import sys

LAYERS = [("pack", 0.0), ("enc1 (conv.0+conv.3+pool, one kernel)", 1.208 + 2.416), ("enc2.0", 1.208), ("enc2.3", 2.416), ("enc3.0", 1.208),
          ("enc3.3", 2.416), ("bott.0", 1.208), ("bott.3", 2.416), ("dec1.up", 0.537), ("resample", 0.0), ("dec1.0", 1.208),
          ("dec1.3", 0.604), ("dec2.up", 0.134), ("dec2.0", 1.208), ("dec2.3", 0.604), ("dec3.up", 0.134), ("dec3.0", 1.208),
          ("dec3.3", 0.604), ("dec4.up", 0.134), ("dec4 (conv.0+conv.3+head, one kernel)", 1.208 + 0.604)]   # GMAC per image
# every layer's input once + output once, minus the tensors inside the two fused conv pairs (written + read: 2 x 8.39 MB in
# enc1, 2 x 4.19 MB in dec4)
ALGORITHMIC = 89.2e6 - 2 * 8388608.0 - 2 * 4194304.0
# builds from r02 call 69 on: ConvTranspose + resize of dec1 in one epilogue (19 launches: no resample; dec1.up writes the
# 32 x 32 x 256 tensor instead of 64 x 64 x 256: -1.57 MB per image) and the stem's readers fetch one plane of the packed image
# instead of two (enc1 and dec4: -1.05 MB per image each)
LAYERS_19 = [l for l in LAYERS if l[0] != "resample"]
LAYERS_19[8] = ("dec1.up (+resize in the epilogue)", 0.537)
ALGORITHMIC_19 = ALGORITHMIC - (2097152.0 - 524288.0) - 2 * 1048576.0


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units, data = rows[0], rows[1], rows[2:]
    n = int(sys.argv[2])
    col = lambda name: hdr.index(name) if name in hdr else [i for i, h in enumerate(hdr) if h.endswith("." + name)][0]
    T, RD, WR = col("gpu__time_duration.sum"), col("dram__bytes_read.sum"), col("dram__bytes_write.sum")
    TP = col("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active")
    HM, CY = col("sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg"), col("sm__cycles_elapsed.max")   # as in the r01/r02_call11 tables
    GH, KN = col("sm__cycles_elapsed.avg.per_second"), col("Kernel Name")
    scale = {"Mbyte": 1.0, "Kbyte": 1e-3, "Gbyte": 1e3, "byte": 1e-6}
    tus = {"us": 1.0, "ns": 1e-3, "ms": 1e3}[units[T]]
    layers, algorithmic = (LAYERS_19, ALGORITHMIC_19) if len(data) == 19 else (LAYERS, ALGORITHMIC)
    assert len(data) == len(layers), (len(data), len(layers))
    print(f"# ncu --set full, one micro-batch of {n} images (cold cache, serialised)")
    print(f"{'layer':40s} {'kernel':34s} {'t_us':>7s} {'TF/s':>6s} {'tensor%':>7s} {'hmma%':>6s} {'rdMB':>7s} {'wrMB':>7s} {'GB/s':>6s} GHz")
    tot_t = tot_b = gemm_b = gemm_t = w_tp = w_hm = 0.0
    gemms = 0
    for r, (name, gmac) in zip(data, layers):
        t = float(r[T]) * tus
        rd, wr = float(r[RD]) * scale[units[RD]], float(r[WR]) * scale[units[WR]]
        try:
            tpv = float(r[TP])
        except ValueError:
            tpv = 0.0
        k = r[KN].replace("void ", "").replace("ub::", "").split("(")[0][:34]
        try:
            hm = float(r[HM]) / 4 / float(r[CY]) * 100   # hmma sub-pipe cycles / 4 sub-pipes / elapsed cycles
        except ValueError:
            hm = 0.0
        print(f"{name:40s} {k:34s} {t:7.1f} {gmac * 2 * n / t * 1e3:6.0f} {tpv:7.1f} {hm:6.1f} {rd:7.1f} {wr:7.1f} "
              f"{(rd + wr) / t * 1e3:6.0f} {r[GH][:4]}")
        tot_t += t
        tot_b += rd + wr
        if gmac:
            gemm_b += rd + wr
            gemms += 1
            gemm_t += t
            w_tp += t * tpv
            w_hm += t * hm
    print(f"total {tot_t:.1f} us per {n} images -> {n / tot_t * 1e6:.0f} img/s serialised; DRAM {tot_b / n:.2f} MB per image "
          f"({gemm_b / n:.2f} in the {gemms} tensor-core launches); time-weighted tensor-pipe activity of those launches: "
          f"{w_tp / gemm_t:.1f} % (pipe active / SM active), {w_hm / gemm_t:.1f} % (hmma sub-pipe cycles / 4 / elapsed, the r01 tables' metric)")
    if len(sys.argv) > 3:
        json.dump({"source": f"{sys.argv[1]} (ncu --set full, one micro-batch of {n} images, final default build of round 2)",
                   "microbatch": n, "gemm_launches_per_microbatch": gemms, "dram_bytes_per_image": gemm_b / n * 1e6,
                   "algorithmic_bytes_per_image": algorithmic}, open(sys.argv[3], "w"), indent=1)


if __name__ == "__main__":
    main()
