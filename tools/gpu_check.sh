#!/bin/bash
# First-contact GPU run: every GPU test file under its own hard timeout (a hung kernel must not eat the box),
# then smoke() and a short bench.  Logs land in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
rc=0
for t in test_coord_gpu test_head_gpu test_nms_gpu test_conv_gpu test_detector_gpu; do
  timeout 420 python -m pytest tests/$t.py -m gpu -q -x --no-header -p no:cacheprovider > gpurun_out/$t.log 2>&1
  r=$?; echo "$t exit $r"; tail -n 25 gpurun_out/$t.log; [ $r -ne 0 ] && rc=1
done
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -n 5 gpurun_out/smoke.log
timeout 600 python bench.py --steps 5 --warmup 3 ${BENCH_ARGS} > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"
tail -n 3 gpurun_out/bench.log; tail -n 8 gpurun_out/bench.err
exit $rc
