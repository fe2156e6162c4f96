#!/bin/bash
# ncu --set full --import-source on of single convolution launches on real layer shapes (tools/bench_conv.py): the persistent
# kernel on ts8 and k_conv_tc on ts32.  The reports stay on the box; their source pages (csv) and the head of their details
# pages come back in gpurun_out/ (tools/stall_summary.py turns the csv into profiles/r02_*_stalls.txt).
mkdir -p gpurun_out
for lv in ts8 ts32; do
  ITERS=1 timeout 300 ncu --set full --import-source on --clock-control none -k regex:k_conv_tc --launch-skip 3 -c 1 \
      -o gpurun_out/cap_$lv -f python tools/bench_conv.py $lv > gpurun_out/cap_$lv.log 2>&1; echo "$lv capture exit $?"
  ncu -i gpurun_out/cap_$lv.ncu-rep --page source --csv > gpurun_out/cap_${lv}_source.csv 2> /dev/null
  ncu -i gpurun_out/cap_$lv.ncu-rep --page details 2> /dev/null | head -n 140 > gpurun_out/cap_${lv}_details.txt
  rm -f gpurun_out/cap_$lv.ncu-rep
done
du -sh gpurun_out
