#!/bin/bash
# ncu launch list of ONE inference step + one full capture of the top kernel (each after a plain run exited 0).
mkdir -p gpurun_out
export BENCH_PROFILE=1
python bench.py > gpurun_out/plain.log 2>&1 &&
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches.csv python bench.py > gpurun_out/ncu_list.log 2>&1
echo "launch list exit $?"
python bench.py > gpurun_out/plain2.log 2>&1 &&
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:${NCU_KERNEL:-k_conv_tc} -c ${NCU_COUNT:-6} \
    -o gpurun_out/prof_top python bench.py > gpurun_out/ncu_full.log 2>&1
echo "full capture exit $?"
ls -la gpurun_out | tail -8
