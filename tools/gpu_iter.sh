#!/bin/bash
# Iteration run: the GPU test files given in $TESTS (default: the ones that touch the neck / maps), then A/B bench lines
# (no CPU baseline) given as "name ENV=V[,ENV=V] [bench args]" lines in $BENCHES, then (NCU=1) the launch list of one
# inference step.  Logs land in gpurun_out/.
mkdir -p gpurun_out
rc=0
for t in ${TESTS:-test_coord_gpu test_conv_gpu test_head_gpu test_surface_gpu test_detector_gpu}; do
  timeout 420 python -m pytest tests/$t.py -m gpu -q -x --no-header -p no:cacheprovider > gpurun_out/$t.log 2>&1
  r=$?; echo "$t exit $r: $(tail -n 1 gpurun_out/$t.log)"; [ $r -ne 0 ] && { rc=1; tail -n 30 gpurun_out/$t.log; }
done
while read -r name envs args; do
  [ -z "$name" ] && continue
  env $(echo "$envs" | tr ',' ' ') timeout 300 python bench.py --no-cpu-baseline --configs none $args > gpurun_out/bench_$name.log 2> gpurun_out/bench_$name.err
  echo "bench $name exit $?: $(python - <<P
import json
try:
    d = json.loads(open('gpurun_out/bench_$name.log').read().strip().splitlines()[-1])
    r = d['roofline']
    print(round(d['value'], 1), d['unit'], 'ms', round(d['ms_per_step'], 3), 'e2e', round(d.get('e2e', {}).get('value') or 0, 1),
          'frac', round(r['frac'], 3), {k: round(v['ms'], 3) for k, v in r['other_kernels'].items()})
except Exception as e:
    print('no line', e)
P
)"
done <<< "${BENCHES:-new B2_X=1}"
if [ -n "$NCU" ]; then
  BENCH_PROFILE=1 timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches.csv python bench.py --no-cpu-baseline $NCU_ARGS > gpurun_out/ncu_list.log 2>&1; echo "launch list exit $?"
fi
exit $rc
