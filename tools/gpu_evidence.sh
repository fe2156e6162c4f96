#!/bin/bash
# Evidence pass of a round: the driver's own test command, smoke(), the full bench line, then ncu (each only after its plain
# command exited 0): launch list + DRAM / L2 bytes of one inference step, full capture of the tcgen05 convolution launches,
# launch list of one training step.  Logs land in gpurun_out/ (copy what is to be judged into profiles/).
mkdir -p gpurun_out
if [ -z "$SKIP_TESTS" ]; then
  timeout 1500 python -m pytest tests/ -x -q -m gpu --no-header -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
  echo "pytest -m gpu exit $?: $(tail -n 1 gpurun_out/pytest_gpu.log)"
fi
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?: $(tail -n 1 gpurun_out/smoke.log)"
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_full.log 2> gpurun_out/bench_full.err; echo "bench exit $?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "bench reference exit $?"
export BENCH_PROFILE=1
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum
timeout 600 ncu --profile-from-start off --metrics $M --clock-control none --csv \
    --log-file gpurun_out/step_mem.csv python bench.py --configs none > gpurun_out/ncu_mem.log 2>&1; echo "memory list exit $?"
timeout 900 ncu --profile-from-start off --set full --clock-control none -k regex:k_conv_tc -c 60 \
    -o gpurun_out/conv_tc_full -f python bench.py --configs none > gpurun_out/ncu_full.log 2>&1; echo "conv_tc full capture exit $?"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches_train.csv python bench.py --mode train > gpurun_out/ncu_train.log 2>&1; echo "train launch list exit $?"
# the report itself is larger than what gpurun brings back: keep its raw page as csv
ncu -i gpurun_out/conv_tc_full.ncu-rep --page raw --csv > gpurun_out/conv_tc_raw.csv 2> /dev/null; rm -f gpurun_out/conv_tc_full.ncu-rep
du -sh gpurun_out
