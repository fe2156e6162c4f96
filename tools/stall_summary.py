"""`ncu --page source --csv` of one kernel launch (tools/gpu_capture_layers.sh) -> the SASS instructions with the most warp-stall
samples and the stall reasons of each, plus totals per reason.

    python tools/stall_summary.py gpurun_out/cap_ts8_source.csv "<title>" [top] > profiles/r02_conv_tcp_ts8_stalls.txt"""
import collections
import csv
import sys

path, title = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
rows = list(csv.reader(open(path, errors="ignore")))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
h = rows[hi]
ix = {n: i for i, n in enumerate(h)}
stall_cols = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
data = []
tot = collections.Counter()
for r in rows[hi + 1:]:
    if len(r) < len(h):
        continue
    n = int(float(r[ix["Warp Stall Sampling (All Samples)"]] or 0))
    if n == 0:
        continue
    reasons = {c[6:]: int(float(r[ix[c]] or 0)) for c in stall_cols}
    reasons = {k: v for k, v in reasons.items() if v}
    tot.update(reasons)
    data.append((n, r[ix["Source"]].strip(), reasons))
data.sort(key=lambda t: -t[0])
allsamp = sum(t[0] for t in data)
print("# %s" % title)
print("# kernel: %s" % (rows[0][1][:110] if rows and len(rows[0]) > 1 else "?"))
print("# warp-stall samples per SASS instruction (top %d of %d samples); totals per reason: %s" %
      (top, allsamp, ", ".join("%s %d" % kv for kv in tot.most_common(8))))
for n, src, reasons in data[:top]:
    rs = ", ".join("%s %d" % kv for kv in sorted(reasons.items(), key=lambda kv: -kv[1])[:3])
    print("%6d  %-88s {%s}" % (n, src[:88], rs))
