"""ncu csv with gpu__time_duration / dram__bytes_read / dram__bytes_write / lts__t_bytes per launch (tools/gpu_evidence.sh)
-> per-kernel table: launches, time, DRAM and L2 bytes, DRAM / L2 throughput and the fraction of the measured HBM peak.

    python tools/summarize_memory.py gpurun_out/step_mem.csv profiles/r02_memory.csv"""
import collections
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3,
        "usecond": 1.0, "msecond": 1e3}


def main(src, dst):
    rows = list(csv.reader(l for l in open(src) if l.startswith('"')))
    hdr = rows[0]
    ci = {h: i for i, h in enumerate(hdr)}
    per = collections.OrderedDict()
    for r in rows[1:]:
        if len(r) < len(hdr):
            continue
        key = (r[ci["ID"]], r[ci["Kernel Name"]].split("(")[0])
        v = float(r[ci["Metric Value"]].replace(",", "")) * UNIT.get(r[ci["Metric Unit"]], 1.0)
        per.setdefault(key, {})[r[ci["Metric Name"]]] = v
    agg = collections.OrderedDict()
    for (_, name), m in per.items():
        a = agg.setdefault(name, [0, 0.0, 0.0, 0.0, 0.0])
        a[0] += 1
        a[1] += m.get("gpu__time_duration.sum", 0.0)
        a[2] += m.get("dram__bytes_read.sum", 0.0)
        a[3] += m.get("dram__bytes_write.sum", 0.0)
        a[4] += m.get("lts__t_bytes.sum", 0.0)
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6650.0
    tot = sum(a[1] for a in agg.values())
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel", "launches", "time_us", "share", "dram_read_MB", "dram_write_MB", "l2_MB", "dram_GBs", "l2_GBs",
                    "dram_frac_of_measured_hbm_peak"])
        for name, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            t = max(a[1], 1e-9)
            w.writerow([name[:70], a[0], round(a[1], 1), round(a[1] / tot, 4), round(a[2] / 1e6, 2), round(a[3] / 1e6, 2),
                        round(a[4] / 1e6, 2), round((a[2] + a[3]) / t / 1e3, 1), round(a[4] / t / 1e3, 1),
                        round((a[2] + a[3]) / t / 1e3 / peak, 3)])
    print("launches %d, summed kernel time %.1f us, HBM peak used %.1f GB/s -> %s" % (sum(a[0] for a in agg.values()), tot, peak, dst))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
