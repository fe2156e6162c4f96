#!/bin/bash
# Iteration run: the GPU test files given in $TESTS (default: the ones that touch the neck / maps), then A/B bench lines
# (device-resident value only, no CPU baseline) for the switches named in the log.  Logs land in gpurun_out/.
mkdir -p gpurun_out
rc=0
for t in ${TESTS:-test_coord_gpu test_conv_gpu test_head_gpu test_surface_gpu test_detector_gpu}; do
  timeout 420 python -m pytest tests/$t.py -m gpu -q -x --no-header -p no:cacheprovider > gpurun_out/$t.log 2>&1
  r=$?; echo "$t exit $r: $(tail -n 1 gpurun_out/$t.log)"; [ $r -ne 0 ] && { rc=1; tail -n 30 gpurun_out/$t.log; }
done
b() { # name, env..., -- args
  name=$1; shift
  env "$@" timeout 300 python bench.py --no-cpu-baseline ${ARGS_EXTRA} > gpurun_out/bench_$name.log 2> gpurun_out/bench_$name.err
  echo "bench $name exit $?: $(python - <<P
import json
try:
    d = json.loads(open('gpurun_out/bench_$name.log').read().strip().splitlines()[-1])
    print(d['value'], d['unit'], 'ms', d['ms_per_step'], 'e2e', d.get('e2e', {}).get('value'), 'launches', d.get('gpu_launches'))
except Exception as e:
    print('no line', e)
P
)"
}
b new B2_X=1
b old_prune B2_PRUNE_FIRST=0
b old_kmap B2_KMAP_STRIDED=0
ARGS_EXTRA="--workload s3dis1m" b s3dis B2_X=1
ARGS_EXTRA="--workload s3dis1m" b s3dis_oldprune B2_PRUNE_FIRST=0
ARGS_EXTRA="--workload sunrgbd" b sun B2_X=1
exit $rc
