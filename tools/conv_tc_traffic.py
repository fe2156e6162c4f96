"""ncu `--set full` capture of the k_conv_tc launches of one step -> profiles/conv_tc_traffic.json (+ a per-launch CSV).

    ncu -i gpurun_out/conv_tc_full.ncu-rep --page raw --csv > /tmp/raw.csv
    python tools/conv_tc_traffic.py /tmp/raw.csv profiles/r01_conv_tc_step5.csv "<state the capture was taken at>"

bench.py reads `dram_bytes_per_launch` from the JSON for `roofline.traffic`."""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0,
         "nsecond": 1e-3, "msecond": 1e3}


def main(raw, out_csv, note):
    rows = list(csv.reader(open(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}

    def val(r, name):
        i = col[name]
        return float(r[i].replace(",", "")) * SCALE.get(units[i], 1.0)

    tensor = "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"
    out, tot = [], dict(rd=0.0, wr=0.0, lts=0.0, t=0.0, tw=0.0)
    for n, r in enumerate(data):
        t, rd, wr = val(r, "gpu__time_duration.sum"), val(r, "dram__bytes_read.sum"), val(r, "dram__bytes_write.sum")
        lts = (val(r, "lts__t_bytes.sum") if "lts__t_bytes.sum" in col else
               32.0 * val(r, "lts__t_sectors.sum") if "lts__t_sectors.sum" in col else 0.0)
        tp = val(r, tensor) if tensor in col else 0.0
        out.append([n, r[col["Kernel Name"]].split("(")[0], r[col["Grid Size"]].replace(" ", ""), round(t, 2), int(rd),
                    int(wr), int(lts), round(tp, 2), r[col["launch__registers_per_thread"]]])
        tot["rd"] += rd
        tot["wr"] += wr
        tot["lts"] += lts
        tot["t"] += t
        tot["tw"] += tp * t
    with open(out_csv, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["launch", "kernel", "grid", "time_us", "dram_read_bytes", "dram_write_bytes", "l2_bytes",
                    "tensor_pipe_active_pct", "registers"])
        w.writerows(out)
    n = len(out)
    from bench import kernel_digest       # bench.py refuses a capture taken from another version of the kernel source
    js = {"kernel": "k_conv_tc + k_conv_tcp (all instantiations)", "kernel_digest": kernel_digest(), "launches": n, "dram_bytes_read": tot["rd"], "dram_bytes_write": tot["wr"],
          "dram_bytes_per_launch": (tot["rd"] + tot["wr"]) / n, "lts_bytes_per_launch": tot["lts"] / n,
          "time_us_cold": round(tot["t"], 2), "tensor_pipe_active_pct_time_weighted": tot["tw"] / tot["t"],
          "source": "ncu --set full --clock-control none -k regex:k_conv_tc, one inference step (BENCH_PROFILE=1 python bench.py); "
                    + note + "; per launch: " + os.path.relpath(out_csv, ROOT)}
    json.dump(js, open(os.path.join(ROOT, "profiles", "conv_tc_traffic.json"), "w"), indent=1)
    print(json.dumps(js, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
