#!/bin/bash
# Full GPU pass: every GPU test file under its own timeout, smoke(), the bench lines of all workloads, then the ncu launch
# lists of one inference step and one training step (each after its plain run exited 0).  Logs land in gpurun_out/.
mkdir -p gpurun_out
rc=0
for t in test_coord_gpu test_head_gpu test_nms_gpu test_conv_gpu test_detector_gpu test_train_gpu test_fullsize_gpu test_surface_gpu test_data_gpu test_zz_no_alloc; do
  timeout 420 python -m pytest tests/$t.py -m gpu -q -x --no-header -p no:cacheprovider > gpurun_out/$t.log 2>&1
  r=$?; echo "$t exit $r: $(tail -n 1 gpurun_out/$t.log)"; [ $r -ne 0 ] && rc=1
done
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"
timeout 900 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"
timeout 600 python bench.py --workload sunrgbd > gpurun_out/bench_sun.log 2>> gpurun_out/bench.err; echo "bench sunrgbd exit $?"
timeout 600 python bench.py --workload s3dis1m > gpurun_out/bench_s3dis.log 2>> gpurun_out/bench.err; echo "bench s3dis exit $?"
timeout 600 python bench.py --mode train > gpurun_out/bench_train.log 2>> gpurun_out/bench.err; echo "bench train exit $?"
timeout 600 python bench.py --mode train --workload sunrgbd --steps 5 > gpurun_out/bench_train_sun.log 2>> gpurun_out/bench.err; echo "bench train sunrgbd exit $?"
timeout 600 python bench.py --mode train --dtype fp32 --steps 5 > gpurun_out/bench_train_fp32.log 2>> gpurun_out/bench.err; echo "bench train fp32 exit $?"
export BENCH_PROFILE=1
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches.csv python bench.py > gpurun_out/ncu_list.log 2>&1; echo "launch list exit $?"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches_train.csv python bench.py --mode train > gpurun_out/ncu_train.log 2>&1; echo "train launch list exit $?"
exit $rc
