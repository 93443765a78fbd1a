#!/bin/bash
# Runs on the GPU box: GPU tests, then the bench with 1-4 waves x pool sizes.
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -q -x -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -6 gpurun_out/pytest_gpu.log
for o in ${OPTSETS:-waves=1 waves=2 waves=3 waves=4 waves=2,pool_paths=1048576 waves=2,pool_paths=4194304 waves=3,pool_paths=2097152 waves=2,shade_ctas=2}; do
  args=""; for kv in $(echo $o | tr ',' ' '); do args="$args --opt $kv"; done
  timeout 600 python bench.py --steps 2 --warmup 2 --no-cpu-baseline ${BENCH_ARGS} $args > gpurun_out/tmp.json 2> gpurun_out/tmp.err
  echo "$o: exit $? $(python -c "import json; d=json.load(open('gpurun_out/tmp.json')); print(round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1))" 2>&1 | tail -1)"
done
